// viterbi.cu — batched Viterbi (greedy) CTC decode for sm_100a.
//
// Replaces search::viterbi_search (src/search.rs:320-383), find_max (:303-318) and phred
// (:31-36) of the reference.  Per posterior row: argmax over labels (strict '>' from label 0, so
// ties go to the lowest index and a NaN never wins unless it is label 0), blank drop, optional
// repeat collapse, emission row ("path"), and one quality character per emitted label from the
// *sequential* f32 mean of the winning probabilities over the label's run (src/search.rs:362-365
// accumulates in time order; a tree reduction would round differently, so each run is summed by
// one lane in row order).
//
// Memory-bound streaming kernel (20 B/row at C = 5).  Every warp is autonomous:
//   * a warp owns whole reads (read = gw, gw + #warps, ...), so no block-level barrier exists
//     anywhere in the steady state;
//   * its posterior rows are staged through shared memory by the TMA engine: 1-D bulk async
//     copies (cp.async.bulk -> UBLKCP) of 32*RPL-row tiles into a per-warp ring of stages,
//     completion counted on mbarriers; the ring runs ahead across read boundaries, so every warp
//     always has several HBM requests in flight;
//   * each lane owns RPL consecutive rows of a tile: 5 x LDS.128 (80 B, bank-conflict free) for
//     C = 5, then RPL argmaxes in registers;
//   * emission flags -> warp-shuffle scan -> output positions; records go straight to the read's
//     slot in the result pool (slot n starts at n * round_up(T, 32): deterministic placement, no
//     atomics, no second pass);
//   * quality: the run mean is summed sequentially in shared memory; the phred character uses a
//     MUFU.LG2 fast path whose result is only trusted when it is provably on the same side of
//     every rounding boundary as glibc's log10f, else the exact binary64 re-evaluation of glibc's
//     algorithm (phred.cuh) decides.
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels.h"
#include "phred.cuh"

#include "../../include/fctc.h"

namespace fctc {

constexpr unsigned FULLMASK = 0xffffffffu;
constexpr int VW_WARPS = 4;  // warps per CTA of the streaming kernel
constexpr int VW_THREADS = VW_WARPS * 32;
constexpr int VW_MAX_STAGES = 8;
#ifndef VW_MINBLOCKS
#define VW_MINBLOCKS 6
#endif
constexpr int VT_SMEM_LIMIT = 227 * 1024;

__constant__ double c_logf_invc[16] = FCTC_LOGF_INVC;
__constant__ double c_logf_logc[16] = FCTC_LOGF_LOGC;

// ---------------------------------------------------------------------------------------------
// Quality character.  Fast path: q from MUFU.LG2; it differs from the reference's float
// evaluation (glibc log10f, three rounded f32 ops) by at most
//     D <= 6e-6*|qscale| + 4e-7*(|q - qbias| + |q|)
// (2 ulp of lg2 at |lg2 p| <= 128, times 3.0103, plus the roundings on both sides).  If q is
// farther than eps >> D from every k + 0.5 boundary, both round to the same integer and the
// fast result is exact; otherwise (about 0.2 % of labels) the binary64 path decides.
// ---------------------------------------------------------------------------------------------
struct QualityParams {  // per-launch constants of the fast path
    float qscale, qbias;
    float k1;    // -10*log10(2) * qscale
    float eps0;  // 1e-3*|qscale| + 2e-5*|qbias| + 1e-6
};
__device__ __forceinline__ QualityParams make_quality_params(float qscale, float qbias) {
    QualityParams q;
    q.qscale = qscale;
    q.qbias = qbias;
    q.k1 = -3.0102999566f * qscale;
    q.eps0 = 1e-3f * fabsf(qscale) + 2e-5f * fabsf(qbias) + 1e-6f;
    return q;
}

__device__ __forceinline__ bool phred_fast(float mean, const QualityParams& qp, uint32_t* code) {
    const float miss = __fadd_rn(1.0f, -mean);
    const float pm = (miss < 1e-4f) ? 1e-4f : miss;  // >= 1e-4 (or NaN): never denormal
    float l2;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(pm));
    const float q = fmaf(l2, qp.k1, qp.qbias);
    const float eps = fmaf(fabsf(q), 2e-5f, qp.eps0);  // >= the bound D above, with a wide margin
    const float fr = q - floorf(q);
    *code = __float2uint_rn(q) + 33u;  // cvt.rni.u32.f32 clamps to [0, 2^32-1] and maps NaN to 0
    return fabsf(fr - 0.5f) > eps;     // false for NaN / inf / huge |q|: the exact path decides
}

template <typename QT>
__device__ __forceinline__ QT quality_char(float sum, int n, const QualityParams& qp, const double* invc,
                                           const double* logc, bool& bad_char) {
    const float mean = __fdiv_rn(sum, float(n));  // IEEE division (src/search.rs:350); x / 1 == x exactly
    uint32_t code;
    if (!phred_fast(mean, qp, &code)) code = phred_code(mean, qp.qscale, qp.qbias, invc, logc);
    if (sizeof(QT) == 4 && !is_unicode_scalar(code)) bad_char = true;
    return QT(code);
}

__device__ __forceinline__ void argmax_row5(const float* f, int& arg, float& best) {
    best = f[0];
    arg = 0;
#pragma unroll
    for (int c = 1; c < 5; ++c)
        if (f[c] > best) { best = f[c]; arg = c; }
}

struct WarpLayout {  // shared memory plan of the streaming kernel (bytes)
    int nstage;
    int stage_bytes;
    int off_stage;   // from the warp's base
    int off_rec;     // (TILE + 1) x {float winning probability, int label} per row + a sentinel
    int off_orow;    // TILE bytes: tile-local row of the tile's j-th emission
    int off_open;    // 3 words: the run left open at the end of a tile {sum, count, output index}
    int warp_bytes;
    int cta_fixed;   // log tables at the start of the CTA's shared memory
};

// Explicit shared-space accesses on 32-bit addresses (keeps address arithmetic 32-bit and lets
// the hot loops skip generic->shared window conversions).
__device__ __forceinline__ void sts_rec2(uint32_t addr, float p0, int l0, float p1, int l1) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(p0), "r"(l0), "f"(p1), "r"(l1) : "memory");
}
__device__ __forceinline__ void sts_rec1(uint32_t addr, float p0, int l0) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "f"(p0), "r"(l0) : "memory");
}
__device__ __forceinline__ void lds_rec(uint32_t addr, float& prob, int& label) {
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(prob), "=r"(label) : "r"(addr) : "memory");
}
__device__ __forceinline__ int lds_u8(uint32_t addr) {
    int v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u8(uint32_t addr, int v) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

template <typename P>
__device__ __forceinline__ P* keep_in_register(P* ptr) {  // stops the compiler from re-deriving the pointer
    asm volatile("" : "+l"(ptr));
    return ptr;
}

// ------------------------------------------------------------------------------------------
// Streaming kernel: one warp per read, RPL rows per lane and tile.
// ------------------------------------------------------------------------------------------
template <typename QT, int RPL, bool COLLAPSE, bool QSTRING>
__global__ void __launch_bounds__(VW_THREADS, VW_MINBLOCKS) viterbi_stream_kernel(const ViterbiParams p, const WarpLayout L) {
    constexpr int TILE = 32 * RPL;
    extern __shared__ __align__(128) unsigned char smem[];
    double* s_invc = reinterpret_cast<double*>(smem);
    double* s_logc = s_invc + 16;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* ws = smem + L.cta_fixed + (size_t)warp * L.warp_bytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(ws);
    unsigned char* stage0 = ws + L.off_stage;
    const uint32_t rec_base = smem_u32(ws + L.off_rec);
    const uint32_t orow_base = smem_u32(ws + L.off_orow);
    float* s_open = reinterpret_cast<float*>(ws + L.off_open);
    const QualityParams qp = make_quality_params(p.qscale, p.qbias);

    const int C = p.C;
    const int T = int(p.T);
    const int ntiles = (T + TILE - 1) / TILE;
    const long long gw = (long long)blockIdx.x * VW_WARPS + warp;
    const long long nwarps = (long long)gridDim.x * VW_WARPS;
    const int my_reads = int((p.N > gw) ? (p.N - gw + nwarps - 1) / nwarps : 0);
    const long long read_elems = (long long)T * C;
    const long long total_elems = p.N * read_elems;
    const long long safe_bytes = (total_elems * 4) & ~15LL;  // bulk copies never read past this
    const long long safe_elems = safe_bytes >> 2;
    const long long stride = p.slot_stride;
    // every tile of every read starts 16-byte aligned and lies inside the bulk-copy-safe range
    const bool aligned = ((read_elems & 3) == 0) && (((TILE * C) & 3) == 0);
    const bool aligned5 = (RPL == 4) && (C == 5) && aligned;
    const int nstage = L.nstage;
    const int stage_bytes = L.stage_bytes;
    const int tile_elems = TILE * C;

    // producer state (advanced by every lane so that it stays warp-uniform; lane 0 issues)
    long long p_left = (long long)my_reads * ntiles;  // tiles not yet issued
    long long p_n = gw;                               // read of the next tile to issue
    int p_ti = 0;
    long long p_e0 = gw * read_elems;  // its first element
    int p_s = 0;
    auto issue_next = [&]() {
        if (p_left <= 0) return;
        const int rows = (T - p_ti * TILE < TILE) ? (T - p_ti * TILE) : TILE;
        if (lane == 0) {
            long long b0;
            uint32_t bytes;
            if (aligned) {
                b0 = p_e0 * 4;
                bytes = uint32_t(rows * C * 4);
            } else {
                const long long e1 = p_e0 + (long long)rows * C;
                b0 = (p_e0 * 4) & ~15LL;
                long long b1 = (e1 * 4 + 15) & ~15LL;
                if (b1 > safe_bytes) b1 = safe_bytes;
                bytes = b1 > b0 ? uint32_t(b1 - b0) : 0u;
            }
            mbar_expect_tx(&full[p_s], bytes);
            if (bytes) bulk_g2s(stage0 + p_s * stage_bytes, reinterpret_cast<const char*>(p.post) + b0, bytes, &full[p_s]);
        }
        --p_left;
        if (++p_s == nstage) p_s = 0;
        if (++p_ti == ntiles) { p_ti = 0; p_n += nwarps; p_e0 = p_n * read_elems; }
        else p_e0 += tile_elems;
    };

    if (threadIdx.x < 16) {
        s_invc[threadIdx.x] = c_logf_invc[threadIdx.x];
        s_logc[threadIdx.x] = c_logf_logc[threadIdx.x];
    }
    if (lane == 0) {
        for (int s = 0; s < nstage; ++s) mbar_init(&full[s], 1);
        mbar_init_fence();
        sts_rec1(rec_base + 8 * TILE, 0.0f, -3);  // sentinel row: ends every run walk at the tile edge
    }
    __syncthreads();  // the only block-level barrier: tables + barrier inits visible
    for (int s = 0; s < nstage; ++s) issue_next();

    int c_s = 0;  // consumer ring position
    uint32_t c_phase = 0;
    const int r_base = RPL * lane;
    for (int k = 0; k < my_reads; ++k) {
        const long long n = gw + (long long)k * nwarps;
        uint8_t* const out_lab = keep_in_register(p.labels + n * stride);
        QT* const out_qual = keep_in_register(reinterpret_cast<QT*>(p.quals) + n * stride);
        int* const out_path = keep_in_register(p.path + n * stride);
        long long e0 = n * read_elems;
        int out_count = 0;
        int last_label = -1;  // None, src/search.rs:337
        bool open_active = false;  // a run that reached the end of the previous tile (warp-uniform)
        float open_sum = 0.0f;
        int open_n = 0, open_idx = 0;
        bool bad_char = false;

        for (int ti = 0; ti < ntiles; ++ti, e0 += tile_elems) {
            const int row0 = ti * TILE;
            const int rows = (T - row0 < TILE) ? (T - row0) : TILE;
            const bool last_tile = (ti == ntiles - 1);
            const unsigned char* stage = stage0 + c_s * stage_bytes;
            mbar_wait(&full[c_s], c_phase);

            float pr[RPL];
            int lab[RPL];
            if (aligned5 && rows == TILE) {
                const float4* s4 = reinterpret_cast<const float4*>(stage) + 5 * lane;
                const float4 a = s4[0], b = s4[1], c = s4[2], d = s4[3], e = s4[4];
                const float f[20] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y,
                                     c.z, c.w, d.x, d.y, d.z, d.w, e.x, e.y, e.z, e.w};
#pragma unroll
                for (int r = 0; r < RPL; ++r) argmax_row5(f + 5 * (r & 3), lab[r], pr[r]);
            } else {
                const float* stagef = reinterpret_cast<const float*>(stage) + int(e0 & 3);
#pragma unroll
                for (int r = 0; r < RPL; ++r) {
                    const int row = r_base + r;
                    int arg = 0;
                    float best = 0.0f;
                    if (row < rows) {
                        const int le = row * C;
                        auto ld = [&](int i) { return (e0 + i < safe_elems) ? stagef[i] : p.post[e0 + i]; };
                        best = ld(le);
                        for (int c = 1; c < C; ++c) {
                            const float v = ld(le + c);
                            if (v > best) { best = v; arg = c; }
                        }
                    }
                    pr[r] = best;
                    lab[r] = arg;  // rows past the end of the read count as blank
                }
            }
            if (QSTRING) {
                if (RPL % 2 == 0) {
#pragma unroll
                    for (int r = 0; r < RPL; r += 2)
                        sts_rec2(rec_base + 8 * (r_base + r), pr[r], lab[r], pr[(r + 1) % RPL], lab[(r + 1) % RPL]);
                } else {
#pragma unroll
                    for (int r = 0; r < RPL; ++r) sts_rec1(rec_base + 8 * (r_base + r), pr[r], lab[r]);
                }
            }
            __syncwarp();  // stage c_s fully consumed; per-row arrays visible
            issue_next();  // refill the stage just freed (it is the producer's next slot by construction)
            if (++c_s == nstage) { c_s = 0; c_phase ^= 1u; }

            // emission flags (src/search.rs:347) and output positions
            int prev = __shfl_up_sync(FULLMASK, lab[RPL - 1], 1);
            if (lane == 0) prev = last_label;
            bool em[RPL];
            int cnt = 0;
#pragma unroll
            for (int r = 0; r < RPL; ++r) {
                em[r] = lab[r] != 0 && (!COLLAPSE || lab[r] != prev);
                prev = lab[r];
                cnt += em[r] ? 1 : 0;
            }
            const int incl = warp_inclusive_scan(cnt, lane);
            const int tile_total = __shfl_sync(FULLMASK, incl, 31);
            const int li0 = incl - cnt;  // tile-local index of this lane's first emission

            // labels and emission rows go straight to the read's slot
            {
                int li = li0;
#pragma unroll
                for (int r = 0; r < RPL; ++r) {
                    if (em[r]) {
                        out_lab[out_count + li] = uint8_t(lab[r]);
                        out_path[out_count + li] = row0 + r_base + r;
                        if (QSTRING) sts_u8(orow_base + li, r_base + r);
                    }
                    li += em[r] ? 1 : 0;
                }
            }

            if (QSTRING) {
                __syncwarp();
                // Dense phase: lane j owns the tile's j-th emission (plus, as item `tile_total`, the run
                // carried in from the previous tile): sum its run in row order, then either close it
                // (quality character, coalesced byte stores) or hand it on as the open run.
                const int n_items = tile_total + (open_active ? 1 : 0);
                bool my_open = false;
                for (int j = lane; j < n_items; j += 32) {
                    int r2, nn, gidx, label;
                    float sum;
                    if (COLLAPSE && j == tile_total) {
                        r2 = 0; nn = open_n; sum = open_sum; label = last_label; gidx = open_idx;
                    } else {
                        const int row = lds_u8(orow_base + j);
                        float pv;
                        lds_rec(rec_base + 8 * row, pv, label);
                        sum = __fadd_rn(0.0f, pv);  // label_prob_total starts at 0.0 (src/search.rs:339,354)
                        nn = 1;
                        r2 = row + 1;
                        gidx = out_count + j;
                    }
                    bool open = false;
                    if (COLLAPSE) {  // without collapse every non-blank row is its own run
                        for (;;) {   // the sentinel at row TILE (and blank padding rows) stop the walk
                            float p2;
                            int l2;
                            lds_rec(rec_base + 8 * r2, p2, l2);
                            if (l2 != label) break;
                            sum = __fadd_rn(sum, p2);
                            ++nn;
                            ++r2;
                        }
                        open = (r2 == TILE) && !last_tile;
                    }
                    if (open) {
                        my_open = true;
                        s_open[0] = sum;
                        s_open[1] = __int_as_float(nn);
                        s_open[2] = __int_as_float(gidx);
                    } else {
                        out_qual[gidx] = quality_char<QT>(sum, nn, qp, s_invc, s_logc, bad_char);
                    }
                }
                if (COLLAPSE) {
                    open_active = __any_sync(FULLMASK, my_open);  // at most one item touches the tile end
                    if (open_active) {
                        __syncwarp();  // orders the owner lane's s_open stores before these loads
                        open_sum = s_open[0];
                        open_n = __float_as_int(s_open[1]);
                        open_idx = __float_as_int(s_open[2]);
                    }
                }
                __syncwarp();  // dense reads done before the next tile rewrites the per-row arrays
            }
            if (COLLAPSE) last_label = __shfl_sync(FULLMASK, lab[RPL - 1], 31);
            out_count += tile_total;
        }
        if (sizeof(QT) == 4) bad_char = __any_sync(FULLMASK, bad_char);
        if (lane == 0) {
            p.offsets[n] = p.offset_base + n * stride;
            p.lens[n] = out_count;
            p.status[n] = bad_char ? FCTC_INVALID_QUALITY_CHAR : FCTC_OK;
        }
    }
}

// ------------------------------------------------------------------------------------------
// Generic kernel: any alphabet size, no alignment requirement.  One CTA per read, rows loaded
// straight from global memory.  Used when the streaming kernel's shared-memory plan does not fit
// (alphabets above 64 labels) or the input pointer is not 16-byte aligned.
// ------------------------------------------------------------------------------------------
constexpr int VG_THREADS = 256;
constexpr int VG_RPT = 4;
constexpr int VG_TILE = VG_THREADS * VG_RPT;
constexpr int VG_WARPS = VG_THREADS / 32;

struct RunCarry {  // a label run that reaches the end of a tile and may continue in the next one
    int active;
    int label;
    float sum;
    int count;
    int idx;
};

struct TileShared {  // double-buffered by tile parity so that one tile needs only two barriers
    uint32_t lab32[2][VG_THREADS];  // 4 labels per word: byte r = row 4*tid + r
    float4 prob4[2][VG_THREADS];    // winning probability of the same 4 rows
    int wsum[2][VG_WARPS];
    RunCarry carry[2];
    double invc[16];
    double logc[16];
    int flag;
};

__device__ __forceinline__ void argmax_global(const float* row, int C, int& arg, float& best) {
    best = row[0];
    arg = 0;
    for (int c = 1; c < C; ++c) {
        const float v = row[c];
        if (v > best) { best = v; arg = c; }
    }
}

template <typename QT>
__global__ void __launch_bounds__(VG_THREADS) viterbi_generic_kernel(const ViterbiParams p) {
    __shared__ TileShared S;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = p.C;
    const long long T = p.T;
    const int ntiles = int((T + VG_TILE - 1) / VG_TILE);
    const bool collapse = p.collapse != 0, qstring = p.qstring != 0;
    const QualityParams qp = make_quality_params(p.qscale, p.qbias);
    if (tid < 16) {
        S.invc[tid] = c_logf_invc[tid];
        S.logc[tid] = c_logf_logc[tid];
    }
    __syncthreads();
    int buf = 0;
    for (long long n = blockIdx.x; n < p.N; n += gridDim.x) {
        const float* rd = p.post + n * T * C;
        uint8_t* out_lab = p.labels + n * p.slot_stride;
        QT* out_qual = reinterpret_cast<QT*>(p.quals) + n * p.slot_stride;
        int* out_path = p.path + n * p.slot_stride;
        int base = 0;
        bool bad_char = false;
        if (tid == 0) S.flag = 0;
        for (int ti = 0; ti < ntiles; ++ti, buf ^= 1) {
            const long long row0 = (long long)ti * VG_TILE;
            const bool first_tile = ti == 0, last_tile = ti == ntiles - 1;
            float pr[VG_RPT];
            uint32_t lab4 = 0;
#pragma unroll
            for (int r = 0; r < VG_RPT; ++r) {
                const long long row = row0 + VG_RPT * tid + r;
                int arg = 0;
                float best = 0.0f;
                if (row < T) argmax_global(rd + row * C, C, arg, best);
                pr[r] = best;
                lab4 |= uint32_t(arg) << (8 * r);
            }
            S.lab32[buf][tid] = lab4;
            S.prob4[buf][tid] = make_float4(pr[0], pr[1], pr[2], pr[3]);
            if (tid == 0) S.carry[buf].active = 0;
            __syncthreads();
            const uint8_t* labs = reinterpret_cast<const uint8_t*>(S.lab32[buf]);
            const float* probs = reinterpret_cast<const float*>(S.prob4[buf]);
            int prev = 0;
            bool prev_valid = true;
            if (tid > 0) prev = labs[VG_RPT * tid - 1];
            else if (!first_tile) prev = reinterpret_cast<const uint8_t*>(S.lab32[buf ^ 1])[VG_TILE - 1];
            else prev_valid = false;
            int lab[VG_RPT];
            bool em[VG_RPT];
            int cnt = 0;
#pragma unroll
            for (int r = 0; r < VG_RPT; ++r) {
                lab[r] = int((lab4 >> (8 * r)) & 0xffu);
                em[r] = lab[r] != 0 && (!collapse || !prev_valid || lab[r] != prev);
                prev = lab[r];
                prev_valid = true;
                cnt += em[r] ? 1 : 0;
            }
            const int incl = warp_inclusive_scan(cnt, lane);
            if (lane == 31) S.wsum[buf][warp] = incl;
            __syncthreads();
            int wbase = 0, total = 0;
#pragma unroll
            for (int w = 0; w < VG_WARPS; ++w) {
                const int s = S.wsum[buf][w];
                if (w < warp) wbase += s;
                total += s;
            }
            int pos = base + wbase + incl - cnt;
#pragma unroll
            for (int r = 0; r < VG_RPT; ++r) {
                if (!em[r]) continue;
                const int idx = pos++;
                out_lab[idx] = uint8_t(lab[r]);
                out_path[idx] = int(row0) + VG_RPT * tid + r;
                if (qstring) {
                    float sum = __fadd_rn(0.0f, pr[r]);
                    int nn = 1;
                    bool open = false;
                    if (collapse) {
                        int r2 = VG_RPT * tid + r + 1;
                        while (r2 < VG_TILE && labs[r2] == lab[r]) {
                            sum = __fadd_rn(sum, probs[r2]);
                            ++nn;
                            ++r2;
                        }
                        open = (r2 == VG_TILE) && !last_tile;
                    }
                    if (open) S.carry[buf] = RunCarry{1, lab[r], sum, nn, idx};
                    else out_qual[idx] = quality_char<QT>(sum, nn, qp, S.invc, S.logc, bad_char);
                }
            }
            if (qstring && tid == 0 && !first_tile) {
                const RunCarry c = S.carry[buf ^ 1];
                if (c.active) {
                    float sum = c.sum;
                    int nn = c.count;
                    int r2 = 0;
                    while (r2 < VG_TILE && labs[r2] == c.label) {
                        sum = __fadd_rn(sum, probs[r2]);
                        ++nn;
                        ++r2;
                    }
                    if (r2 == VG_TILE && !last_tile) S.carry[buf] = RunCarry{1, c.label, sum, nn, c.idx};
                    else out_qual[c.idx] = quality_char<QT>(sum, nn, qp, S.invc, S.logc, bad_char);
                }
            }
            base += total;
        }
        if (bad_char) S.flag = 1;
        __syncthreads();
        if (tid == 0) {
            p.offsets[n] = p.offset_base + n * p.slot_stride;
            p.lens[n] = base;
            p.status[n] = S.flag ? FCTC_INVALID_QUALITY_CHAR : FCTC_OK;
        }
        __syncthreads();
    }
}

template <int RPL>
static bool plan_stream(const ViterbiParams& p, WarpLayout* L) {
    if ((reinterpret_cast<uintptr_t>(p.post) & 15u) != 0) return false;
    if (p.T >= (1LL << 30) / (p.C > 0 ? p.C : 1)) return false;  // 32-bit row arithmetic in the kernel
    constexpr int TILE = 32 * RPL;
    const long long rows = p.T < TILE ? p.T : TILE;
    const long long stage = ((rows * p.C * 4 + 15) & ~15LL) + 32;
    if (stage > 16 * 1024) return false;
    // enough bytes in flight per warp to cover HBM latency at ~16-20 resident warps per SM
    int ns = int((8 * 1024) / stage);
    if (const char* e = getenv("FCTC_VIT_STAGES")) ns = atoi(e) > 0 ? atoi(e) : ns;  // tuning knob
    if (ns > VW_MAX_STAGES) ns = VW_MAX_STAGES;
    if (ns < 2) ns = 2;
    L->nstage = ns;
    L->stage_bytes = int(stage);
    L->off_stage = 64;
    L->off_rec = int(L->off_stage + ns * stage);
    L->off_orow = L->off_rec + (TILE + 2) * 8;
    L->off_open = L->off_orow + ((TILE + 15) & ~15);
    L->warp_bytes = (L->off_open + 16 + 127) & ~127;
    L->cta_fixed = 256;
    return L->cta_fixed + VW_WARPS * L->warp_bytes <= VT_SMEM_LIMIT;
}

template <typename QT, int RPL, bool COLLAPSE, bool QSTRING>
static cudaError_t launch_stream(const ViterbiParams& p, const WarpLayout& L, int sm_count, cudaStream_t stream) {
    auto kernel = viterbi_stream_kernel<QT, RPL, COLLAPSE, QSTRING>;
    const int smem = L.cta_fixed + VW_WARPS * L.warp_bytes;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, VW_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sm_count * per_sm;
    const long long need = (p.N + VW_WARPS - 1) / VW_WARPS;
    if (grid > need) grid = need;
    kernel<<<dim3((unsigned)grid), VW_THREADS, smem, stream>>>(p, L);
    return cudaGetLastError();
}

template <typename QT, int RPL>
static cudaError_t launch_stream_flags(const ViterbiParams& p, const WarpLayout& L, int sm_count, cudaStream_t stream) {
    if (p.collapse) {
        return p.qstring ? launch_stream<QT, RPL, true, true>(p, L, sm_count, stream)
                         : launch_stream<QT, RPL, true, false>(p, L, sm_count, stream);
    }
    return p.qstring ? launch_stream<QT, RPL, false, true>(p, L, sm_count, stream)
                     : launch_stream<QT, RPL, false, false>(p, L, sm_count, stream);
}

template <typename QT>
static cudaError_t launch_viterbi_t(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which) {
    WarpLayout L;
    if (p.C <= 16 && plan_stream<4>(p, &L)) {
        if (which) *which = 0;
        return launch_stream_flags<QT, 4>(p, L, sm_count, stream);
    }
    if (p.C <= 64 && plan_stream<1>(p, &L)) {
        if (which) *which = 0;
        return launch_stream_flags<QT, 1>(p, L, sm_count, stream);
    }
    long long grid = (long long)sm_count * 8;
    if (grid > p.N) grid = p.N;
    if (which) *which = 1;
    viterbi_generic_kernel<QT><<<dim3((unsigned)grid), VG_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_viterbi(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which) {
    if (p.N <= 0) return cudaSuccess;
    return p.qual_bytes == 4 ? launch_viterbi_t<uint32_t>(p, sm_count, stream, which)
                             : launch_viterbi_t<uint8_t>(p, sm_count, stream, which);
}

}  // namespace fctc
