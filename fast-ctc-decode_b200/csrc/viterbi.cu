// viterbi.cu — batched Viterbi (greedy) CTC decode for sm_100a.
//
// Replaces search::viterbi_search (src/search.rs:320-383), find_max (:303-318) and phred
// (:31-36) of the reference.  Per posterior row: argmax over labels (strict '>' from label 0, so
// ties go to the lowest index and a NaN never wins unless it is label 0), blank drop, optional
// repeat collapse, emission row ("path"), and one quality character per emitted label from the
// *sequential* f32 mean of the winning probabilities over the label's run (src/search.rs:362-365
// accumulates in time order; a tree reduction would round differently, so each run is summed by
// one thread in row order).
//
// Memory-bound streaming kernel (20 B/row at C = 5):
//   * persistent CTAs (SM count x resident CTAs), each walking reads blockIdx.x, +gridDim.x, ...
//   * posterior rows are staged through shared memory by the TMA engine: 1-D bulk async copies
//     (cp.async.bulk -> UBLKCP) of 1024-row tiles into an N-stage ring, completion on mbarriers;
//     the ring runs ahead across read boundaries, so HBM requests are always in flight
//   * each thread owns 4 consecutive rows: 5 x LDS.128 (80 B, bank-conflict free), 4 argmaxes
//   * emission flags -> warp-shuffle scan + cross-warp sums give output positions
//   * a read's labels / qualities / path are collected in shared memory and written out once,
//     coalesced, into a pool slot obtained with a single atomicAdd per read.
#include "device_utils.cuh"
#include "kernels.h"
#include "phred.cuh"

#include "../../include/fctc.h"

namespace fctc {

constexpr int VT_THREADS = 256;
constexpr int VT_RPT = 4;  // rows per thread
constexpr int VT_TILE = VT_THREADS * VT_RPT;
constexpr int VT_WARPS = VT_THREADS / 32;
constexpr int VT_MAX_STAGES = 4;
constexpr int VT_SMEM_LIMIT = 227 * 1024;

__constant__ double c_logf_invc[16] = FCTC_LOGF_INVC;
__constant__ double c_logf_logc[16] = FCTC_LOGF_LOGC;

struct RunCarry {  // a label run that reaches the end of a tile and may continue in the next one
    int active;
    int label;
    float sum;
    int count;
    int idx;  // output slot whose quality is still pending
};

struct TileShared {  // double-buffered by tile parity so that one tile needs only two barriers
    uint32_t lab32[2][VT_THREADS];  // 4 labels per word: byte r = row 4*tid + r
    float4 prob4[2][VT_THREADS];    // winning probability of the same 4 rows
    int wsum[2][VT_WARPS];
    RunCarry carry[2];
    double invc[16];
    double logc[16];
    long long off;
    int flag;
};

__device__ __forceinline__ void argmax_row5(const float* f, int& arg, float& best) {
    best = f[0];
    arg = 0;
#pragma unroll
    for (int c = 1; c < 5; ++c)
        if (f[c] > best) { best = f[c]; arg = c; }
}

template <typename QT>
__device__ __forceinline__ QT finish_quality(TileShared& S, float sum, int n, float qscale, float qbias) {
    float mean = sum / float(n);  // IEEE division (no fast-math), as src/search.rs:350
    uint32_t code = phred_code(mean, qscale, qbias, S.invc, S.logc);
    if (sizeof(QT) == 4 && !is_unicode_scalar(code)) S.flag = 1;
    return QT(code);
}

// One 1024-row tile whose labels / winning probabilities are already in S.lab32[buf] /
// S.prob4[buf] (and a __syncthreads() has been passed).  Returns the number of labels emitted by
// the tile; their records go to out_*[base_count ...].
template <typename QT>
__device__ __forceinline__ int process_tile(TileShared& S, int buf, int tid, bool first_tile, bool last_tile,
                                            int row0, uint32_t my_lab4, const float (&my_prob)[VT_RPT],
                                            int base_count, bool collapse, bool qstring, float qscale, float qbias,
                                            uint8_t* out_lab, QT* out_qual, int* out_path) {
    const int lane = tid & 31, warp = tid >> 5;
    const uint8_t* labs = reinterpret_cast<const uint8_t*>(S.lab32[buf]);
    const float* probs = reinterpret_cast<const float*>(S.prob4[buf]);

    int prev = 0;
    bool prev_valid = true;
    if (tid > 0) prev = labs[VT_RPT * tid - 1];
    else if (!first_tile) prev = reinterpret_cast<const uint8_t*>(S.lab32[buf ^ 1])[VT_TILE - 1];
    else prev_valid = false;  // last_label = None, src/search.rs:337

    int lab[VT_RPT];
    bool em[VT_RPT];
    int cnt = 0;
#pragma unroll
    for (int r = 0; r < VT_RPT; ++r) {
        lab[r] = int((my_lab4 >> (8 * r)) & 0xffu);
        em[r] = lab[r] != 0 && (!collapse || !prev_valid || lab[r] != prev);  // src/search.rs:347
        prev = lab[r];
        prev_valid = true;
        cnt += em[r] ? 1 : 0;
    }
    const int incl = warp_inclusive_scan(cnt, lane);
    if (lane == 31) S.wsum[buf][warp] = incl;
    __syncthreads();
    int wbase = 0, total = 0;
#pragma unroll
    for (int w = 0; w < VT_WARPS; ++w) {
        int s = S.wsum[buf][w];
        if (w < warp) wbase += s;
        total += s;
    }
    int pos = base_count + wbase + incl - cnt;

#pragma unroll
    for (int r = 0; r < VT_RPT; ++r) {
        if (!em[r]) continue;
        const int idx = pos++;
        out_lab[idx] = uint8_t(lab[r]);
        out_path[idx] = row0 + VT_RPT * tid + r;
        if (qstring) {
            float sum = fadd(0.0f, my_prob[r]);  // label_prob_total starts at 0.0, src/search.rs:339,354
            int n = 1;
            bool open = false;
            if (collapse) {  // without collapse every non-blank row is its own run
                int r2 = VT_RPT * tid + r + 1;
                while (r2 < VT_TILE && labs[r2] == lab[r]) {
                    sum = fadd(sum, probs[r2]);
                    ++n;
                    ++r2;
                }
                open = (r2 == VT_TILE) && !last_tile;
            }
            if (open) S.carry[buf] = RunCarry{1, lab[r], sum, n, idx};
            else out_qual[idx] = finish_quality<QT>(S, sum, n, qscale, qbias);
        }
    }
    // a run carried in from the previous tile of this read: continue it in row order
    if (qstring && tid == 0 && !first_tile) {
        const RunCarry c = S.carry[buf ^ 1];
        if (c.active) {
            float sum = c.sum;
            int n = c.count;
            int r2 = 0;
            while (r2 < VT_TILE && labs[r2] == c.label) {
                sum = fadd(sum, probs[r2]);
                ++n;
                ++r2;
            }
            if (r2 == VT_TILE && !last_tile) S.carry[buf] = RunCarry{1, c.label, sum, n, c.idx};
            else out_qual[c.idx] = finish_quality<QT>(S, sum, n, qscale, qbias);
        }
    }
    return total;
}

// Pool slot for one read: one atomicAdd, rounded up to 32 entries so that every array of the
// slot starts on a 32-byte sector (128 bytes for path) and 16-byte vector stores may overrun
// `len` inside the slot.
__device__ __forceinline__ long long reserve_slot(const ViterbiParams& p, int len) {
    unsigned long long want = (unsigned long long)((len + 31) & ~31);
    unsigned long long off = atomicAdd(p.cursor, want);
    return (off + want <= (unsigned long long)p.pool_cap) ? (long long)off : -1;
}

// ------------------------------------------------------------------------------------------
// Streaming kernel.
// ------------------------------------------------------------------------------------------
struct StreamLayout {
    int nstage;
    int stage_bytes;
    int out_lab_bytes, out_qual_bytes, out_path_bytes;
    int total_bytes;
};
constexpr int VT_FIXED_BYTES = 64 + ((int(sizeof(TileShared)) + 127) & ~127);

template <typename QT>
__global__ void __launch_bounds__(VT_THREADS, 2) viterbi_stream_kernel(const ViterbiParams p, const StreamLayout L) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    TileShared& S = *reinterpret_cast<TileShared*>(smem + 64);
    unsigned char* stage0 = smem + VT_FIXED_BYTES;
    unsigned char* outbuf = stage0 + L.nstage * L.stage_bytes;
    uint8_t* out_lab = outbuf;
    QT* out_qual = reinterpret_cast<QT*>(outbuf + L.out_lab_bytes);
    int* out_path = reinterpret_cast<int*>(outbuf + L.out_lab_bytes + L.out_qual_bytes);

    const int tid = threadIdx.x;
    const int C = p.C;
    const long long T = p.T;
    const int ntiles = int((T + VT_TILE - 1) / VT_TILE);
    const long long my_reads = (p.N > blockIdx.x) ? (p.N - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long total_q = my_reads * ntiles;
    const long long total_elems = p.N * T * C;
    const long long safe_bytes = (total_elems * 4) & ~15LL;  // bulk copies never read past this
    const long long safe_elems = safe_bytes >> 2;
    const bool collapse = p.collapse != 0, qstring = p.qstring != 0;

    auto issue = [&](long long q) {  // thread 0 only: start the bulk copy of tile q of this CTA's sequence
        const long long k = q / ntiles;
        const int ti = int(q - k * ntiles);
        const long long n = blockIdx.x + k * (long long)gridDim.x;
        const long long row0 = (long long)ti * VT_TILE;
        const long long rows = (T - row0 < VT_TILE) ? (T - row0) : VT_TILE;
        const long long e0 = (n * T + row0) * C, e1 = e0 + rows * C;
        const long long b0 = (e0 * 4) & ~15LL;
        long long b1 = (e1 * 4 + 15) & ~15LL;
        if (b1 > safe_bytes) b1 = safe_bytes;
        const uint32_t bytes = b1 > b0 ? uint32_t(b1 - b0) : 0u;
        const int s = int(q % L.nstage);
        mbar_expect_tx(&full[s], bytes);
        if (bytes) bulk_g2s(stage0 + s * L.stage_bytes, reinterpret_cast<const char*>(p.post) + b0, bytes, &full[s]);
    };

    if (tid == 0) {
        for (int s = 0; s < L.nstage; ++s) mbar_init(&full[s], 1);
        mbar_init_fence();
    }
    if (tid < 16) {
        S.invc[tid] = c_logf_invc[tid];
        S.logc[tid] = c_logf_logc[tid];
    }
    __syncthreads();
    if (tid == 0) {
        const long long pre = total_q < L.nstage ? total_q : L.nstage;
        for (long long q = 0; q < pre; ++q) issue(q);
    }

    long long q = 0;
    for (long long k = 0; k < my_reads; ++k) {
        const long long n = blockIdx.x + k * (long long)gridDim.x;
        int base = 0;
        if (tid == 0) S.flag = 0;
        for (int ti = 0; ti < ntiles; ++ti, ++q) {
            const int s = int(q % L.nstage);
            const uint32_t phase = uint32_t(q / L.nstage) & 1u;
            const int buf = int(q & 1);
            const long long row0 = (long long)ti * VT_TILE;
            const int rows = int((T - row0 < VT_TILE) ? (T - row0) : VT_TILE);
            const long long e0 = (n * T + row0) * C;
            const float* stagef = reinterpret_cast<const float*>(stage0 + s * L.stage_bytes) + int(e0 & 3);
            mbar_wait(&full[s], phase);

            float pr[VT_RPT];
            uint32_t lab4 = 0;
            const int r_base = VT_RPT * tid;
            const bool vec_ok = (C == 5) && ((e0 & 3) == 0) && ((e0 + (long long)rows * C) <= safe_elems);
            if (vec_ok && r_base + VT_RPT <= rows) {
                const float4* s4 = reinterpret_cast<const float4*>(stagef) + 5 * tid;
                const float4 a = s4[0], b = s4[1], c = s4[2], d = s4[3], e = s4[4];
                const float f[20] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y,
                                     c.z, c.w, d.x, d.y, d.z, d.w, e.x, e.y, e.z, e.w};
#pragma unroll
                for (int r = 0; r < VT_RPT; ++r) {
                    int arg;
                    argmax_row5(f + 5 * r, arg, pr[r]);
                    lab4 |= uint32_t(arg) << (8 * r);
                }
            } else {
#pragma unroll
                for (int r = 0; r < VT_RPT; ++r) {
                    const int row = r_base + r;
                    int arg = 0;
                    float best = 0.0f;
                    if (row < rows) {
                        const long long le = (long long)row * C;
                        auto ld = [&](long long i) { return (e0 + i < safe_elems) ? stagef[i] : p.post[e0 + i]; };
                        best = ld(le);
                        for (int c = 1; c < C; ++c) {
                            const float v = ld(le + c);
                            if (v > best) { best = v; arg = c; }
                        }
                    }
                    pr[r] = best;
                    lab4 |= uint32_t(arg) << (8 * r);
                }
            }
            S.lab32[buf][tid] = lab4;
            S.prob4[buf][tid] = make_float4(pr[0], pr[1], pr[2], pr[3]);
            if (tid == 0) S.carry[buf].active = 0;
            __syncthreads();  // tile's labels visible; everyone is done reading stage s
            if (tid == 0 && q + L.nstage < total_q) issue(q + L.nstage);
            base += process_tile<QT>(S, buf, tid, ti == 0, ti == ntiles - 1, int(row0), lab4, pr, base, collapse,
                                     qstring, p.qscale, p.qbias, out_lab, out_qual, out_path);
        }
        // write the read's records out, coalesced, into its pool slot
        if (tid == 0) S.off = reserve_slot(p, base);
        __syncthreads();
        const long long off = S.off;
        if (off >= 0) {
            const int n16 = (base + 15) >> 4, n4 = (base + 3) >> 2;
            uint4* gl = reinterpret_cast<uint4*>(p.labels + off);
            const uint4* sl = reinterpret_cast<const uint4*>(out_lab);
            for (int i = tid; i < n16; i += VT_THREADS) gl[i] = sl[i];
            uint4* gp = reinterpret_cast<uint4*>(p.path + off);
            const uint4* sp = reinterpret_cast<const uint4*>(out_path);
            for (int i = tid; i < n4; i += VT_THREADS) gp[i] = sp[i];
            if (qstring) {
                uint4* gq = reinterpret_cast<uint4*>(reinterpret_cast<QT*>(p.quals) + off);
                const uint4* sq = reinterpret_cast<const uint4*>(out_qual);
                const int nq = sizeof(QT) == 1 ? n16 : n4;
                for (int i = tid; i < nq; i += VT_THREADS) gq[i] = sq[i];
            }
        }
        if (tid == 0) {
            p.offsets[n] = off >= 0 ? off : 0;
            p.lens[n] = off >= 0 ? base : 0;
            p.status[n] = off < 0 ? FCTC_OUTPUT_OVERFLOW : (S.flag ? FCTC_INVALID_QUALITY_CHAR : FCTC_OK);
        }
    }
}

// ------------------------------------------------------------------------------------------
// Generic kernel: any alphabet size, any row count, no alignment requirement.  One CTA per read,
// two passes (count, then decode straight into the pool slot).  Used when the streaming kernel's
// shared-memory plan does not fit (very long reads / very wide alphabets) or the input pointer
// is not 16-byte aligned.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void argmax_global(const float* row, int C, int& arg, float& best) {
    best = row[0];
    arg = 0;
    for (int c = 1; c < C; ++c) {
        const float v = row[c];
        if (v > best) { best = v; arg = c; }
    }
}

template <typename QT>
__global__ void __launch_bounds__(VT_THREADS) viterbi_generic_kernel(const ViterbiParams p) {
    __shared__ TileShared S;
    __shared__ int s_red[VT_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = p.C;
    const long long T = p.T;
    const int ntiles = int((T + VT_TILE - 1) / VT_TILE);
    const bool collapse = p.collapse != 0, qstring = p.qstring != 0;
    if (tid < 16) {
        S.invc[tid] = c_logf_invc[tid];
        S.logc[tid] = c_logf_logc[tid];
    }
    __syncthreads();
    int parity = 0;
    for (long long n = blockIdx.x; n < p.N; n += gridDim.x) {
        const float* rd = p.post + n * T * C;
        // pass A: number of emitted labels
        int cnt = 0;
        for (long long t = tid; t < T; t += VT_THREADS) {
            int a;
            float v;
            argmax_global(rd + t * C, C, a, v);
            if (a != 0) {
                bool emit = true;
                if (collapse && t > 0) {
                    int b;
                    argmax_global(rd + (t - 1) * C, C, b, v);
                    emit = (b != a);
                }
                cnt += emit ? 1 : 0;
            }
        }
        cnt = warp_inclusive_scan(cnt, lane);
        if (lane == 31) s_red[warp] = cnt;
        if (tid == 0) S.flag = 0;
        __syncthreads();
        if (tid == 0) {
            int total = 0;
            for (int w = 0; w < VT_WARPS; ++w) total += s_red[w];
            S.off = reserve_slot(p, total);
        }
        __syncthreads();
        const long long off = S.off;
        int base = 0;
        if (off >= 0) {
            uint8_t* out_lab = p.labels + off;
            QT* out_qual = reinterpret_cast<QT*>(p.quals) + off;
            int* out_path = p.path + off;
            for (int ti = 0; ti < ntiles; ++ti, parity ^= 1) {
                const long long row0 = (long long)ti * VT_TILE;
                float pr[VT_RPT];
                uint32_t lab4 = 0;
#pragma unroll
                for (int r = 0; r < VT_RPT; ++r) {
                    const long long row = row0 + VT_RPT * tid + r;
                    int arg = 0;
                    float best = 0.0f;
                    if (row < T) argmax_global(rd + row * C, C, arg, best);
                    pr[r] = best;
                    lab4 |= uint32_t(arg) << (8 * r);
                }
                S.lab32[parity][tid] = lab4;
                S.prob4[parity][tid] = make_float4(pr[0], pr[1], pr[2], pr[3]);
                if (tid == 0) S.carry[parity].active = 0;
                __syncthreads();
                base += process_tile<QT>(S, parity, tid, ti == 0, ti == ntiles - 1, int(row0), lab4, pr, base,
                                         collapse, qstring, p.qscale, p.qbias, out_lab, out_qual, out_path);
            }
        }
        __syncthreads();
        if (tid == 0) {
            p.offsets[n] = off >= 0 ? off : 0;
            p.lens[n] = off >= 0 ? base : 0;
            p.status[n] = off < 0 ? FCTC_OUTPUT_OVERFLOW : (S.flag ? FCTC_INVALID_QUALITY_CHAR : FCTC_OK);
        }
    }
}

static bool plan_stream(const ViterbiParams& p, StreamLayout* L) {
    if ((reinterpret_cast<uintptr_t>(p.post) & 15u) != 0) return false;
    if (p.T > (1LL << 30) || p.T <= 0) return false;
    const long long qb = p.qstring ? p.qual_bytes : 0;
    const long long out_lab = (p.T + 15) & ~15LL;
    const long long out_qual = (p.T * qb + 15) & ~15LL;
    const long long out_path = (p.T * 4 + 15) & ~15LL;
    const long long rows = p.T < VT_TILE ? p.T : VT_TILE;
    const long long stage = ((rows * p.C * 4 + 15) & ~15LL) + 32;
    if (stage > (1 << 19)) return false;  // mbarrier tx-count range
    for (int ns = VT_MAX_STAGES; ns >= 2; --ns) {
        const long long total = VT_FIXED_BYTES + ns * stage + out_lab + out_qual + out_path;
        // prefer a plan that leaves room for two CTAs per SM; fall back to one
        if (total * 2 <= VT_SMEM_LIMIT - 2048 || (ns == 2 && total <= VT_SMEM_LIMIT)) {
            L->nstage = ns;
            L->stage_bytes = int(stage);
            L->out_lab_bytes = int(out_lab);
            L->out_qual_bytes = int(out_qual);
            L->out_path_bytes = int(out_path);
            L->total_bytes = int(total);
            return true;
        }
    }
    return false;
}

template <typename QT>
static cudaError_t launch_viterbi_t(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which) {
    StreamLayout L;
    if (plan_stream(p, &L)) {
        cudaError_t e = cudaFuncSetAttribute(viterbi_stream_kernel<QT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             L.total_bytes);
        if (e != cudaSuccess) return e;
        const int per_sm = (2LL * L.total_bytes <= VT_SMEM_LIMIT - 2048) ? 2 : 1;
        long long grid = (long long)sm_count * per_sm;
        if (grid > p.N) grid = p.N;
        if (which) *which = 0;
        viterbi_stream_kernel<QT><<<dim3((unsigned)grid), VT_THREADS, L.total_bytes, stream>>>(p, L);
        return cudaGetLastError();
    }
    long long grid = (long long)sm_count * 8;
    if (grid > p.N) grid = p.N;
    if (which) *which = 1;
    viterbi_generic_kernel<QT><<<dim3((unsigned)grid), VT_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_viterbi(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which) {
    if (p.N <= 0) return cudaSuccess;
    return p.qual_bytes == 4 ? launch_viterbi_t<uint32_t>(p, sm_count, stream, which)
                             : launch_viterbi_t<uint8_t>(p, sm_count, stream, which);
}

}  // namespace fctc
