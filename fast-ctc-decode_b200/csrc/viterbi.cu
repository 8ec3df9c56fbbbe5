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
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdlib>

#include "device_utils.cuh"
#include "input.cuh"
#include "kernels.h"
#include "phred.cuh"

#include "../../include/fctc.h"

namespace fctc {

constexpr unsigned FULLMASK = 0xffffffffu;
constexpr int VW_WARPS = 4;  // warps per CTA of the streaming kernel
constexpr int VW_THREADS = VW_WARPS * 32;
constexpr int VW_MAX_STAGES = 16;
// Resident CTAs per SM the compiler must allow (register budget) and bytes of posteriors each warp
// keeps in flight.  Measured on B200 (65,536 reads x 4000 x 5): with quality strings the kernel is
// issue-bound and wants more warps (6 CTAs, 3 stages); without, it is HBM-bound and runs best with
// fewer, fatter warps (4 CTAs, 4 stages: 0.95 of the measured copy bandwidth).
#ifndef VW_MINBLOCKS_Q
#define VW_MINBLOCKS_Q 6
#endif
#ifndef VW_MINBLOCKS_NOQ
#define VW_MINBLOCKS_NOQ 4
#endif
constexpr int VT_SMEM_LIMIT = 227 * 1024;

__constant__ double c_logf_invc[16] = FCTC_LOGF_INVC;
__constant__ double c_logf_logc[16] = FCTC_LOGF_LOGC;
__constant__ uint64_t c_vit_exp2f[32] = FCTC_EXP2F_TAB;  // glibc expf's table, for the exact softmax of logits (input.cuh)

// The two halves of quality_char: the straight-line shortcut (false = its result must not be used) and the exact path.
__device__ __forceinline__ bool quality_fast(float sum, int n, const QualityParams& qp, uint32_t* code) {
    const bool moderate = (__float_as_uint(sum) - 0x21800000u) < 0x3c000000u;
    return phred_fast(fdiv_moderate(sum, float(n)), qp, code) && moderate;
}
__device__ __forceinline__ uint32_t quality_exact(float sum, int n, const QualityParams& qp, const double* invc,
                                                  const double* logc) {
    return phred_code(__fdiv_rn(sum, float(n)), qp.qscale, qp.qbias, invc, logc);
}

template <typename QT>
__device__ __forceinline__ QT quality_char(float sum, int n, const QualityParams& qp, const double* invc,
                                           const double* logc, bool& bad_char) {
    // mean = sum / n, an IEEE division (src/search.rs:350).  For a sum of moderate size (2^-60 <= sum < 2^60; n < 2^30)
    // the straight-line core of the division is exact (phred.cuh), so the common case has no range check of its own:
    // one test sends everything else — odd sums and quality values too close to a rounding boundary — to the exact path.
    uint32_t code;
    if (!quality_fast(sum, n, qp, &code)) code = quality_exact(sum, n, qp, invc, logc);
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
    int off_queue;   // (TILE + 1) x {float run sum, int run length}: slot 0 = run carried in, slot s = s-th emission
    int off_open;    // {sum, length} of the run left open at the end of the previous tile
    int warp_bytes;
    int cta_fixed;   // log tables at the start of the CTA's shared memory
    int max_ctas;    // resident CTAs per SM the plan wants at most (0: as many as fit)
};

// Explicit shared-space accesses on 32-bit addresses (keeps address arithmetic 32-bit and lets
// the hot loops skip generic->shared window conversions).
__device__ __forceinline__ void sts_run(uint32_t addr, float sum, int n) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "f"(sum), "r"(n) : "memory");
}
__device__ __forceinline__ void lds_run(uint32_t addr, float& sum, int& n) {
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(sum), "=r"(n) : "r"(addr) : "memory");
}

__device__ __forceinline__ void lds_f4(uint32_t addr, float* f) {
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[0]), "=f"(f[1]), "=f"(f[2]), "=f"(f[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ float lds_f1(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}

template <typename P>
__device__ __forceinline__ P* keep_in_register(P* ptr) {  // stops the compiler from re-deriving the pointer
    asm volatile("" : "+l"(ptr));
    return ptr;
}

// ------------------------------------------------------------------------------------------
// Streaming kernel: one warp per read, RPL rows per lane and tile.
// ------------------------------------------------------------------------------------------
// FAST5: five labels, RPL == 4 and every tile 16-byte aligned (the nanopore case), so that all the
// tile geometry is a compile-time constant.
// HALF (with FAST5 only): the values are 16 bits wide (binary16, or bfloat16 when p.in_bf16), 10 bytes per row; a
// lane's four rows are five 8-byte shared loads, widened to f32 in registers.  Everything after the load is the f32 code.
// LOGITS (with FAST5 only): the rows hold logits; each is turned into its softmax (input.cuh, p.softmax = 1 exact /
// 2 fast) right after the load, in registers, and the rest of the kernel sees posteriors.
template <typename QT, int RPL, bool COLLAPSE, bool QSTRING, bool FAST5, bool HALF, bool LOGITS>
__global__ void __launch_bounds__(VW_THREADS, QSTRING ? VW_MINBLOCKS_Q : VW_MINBLOCKS_NOQ) viterbi_stream_kernel(const ViterbiParams p, const WarpLayout L) {
    static_assert(!FAST5 || RPL == 4, "the five-label fast path is written for four rows per lane");
    static_assert(!HALF || FAST5, "16-bit input is only staged by the five-label fast path");
    static_assert(!LOGITS || FAST5, "the fused softmax is only built into the five-label fast path");
    constexpr int EB = HALF ? 2 : 4;  // bytes per posterior
    constexpr int TILE = 32 * RPL;
    extern __shared__ __align__(128) unsigned char smem[];
    double* s_invc = reinterpret_cast<double*>(smem);
    double* s_logc = s_invc + 16;
    uint64_t* s_xt = reinterpret_cast<uint64_t*>(s_logc + 16);
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(FULLMASK, threadIdx.x >> 5, 0);  // tells the compiler it is warp-uniform
    unsigned char* ws = smem + L.cta_fixed + (size_t)warp * L.warp_bytes;
    const uint32_t bar0 = smem_u32(ws);  // nstage mbarriers, then the stages (all addressed in the shared window)
    const uint32_t stage0 = smem_u32(ws + L.off_stage);
    const uint32_t q_base = smem_u32(ws + L.off_queue);
    const uint32_t open_addr = smem_u32(ws + L.off_open);
    const QualityParams qp = make_quality_params(p.qscale, p.qbias);

    const int C = FAST5 ? 5 : p.C;
    const int T = int(p.T);
    const int ntiles = (T + TILE - 1) / TILE;
    const int last_rows = T - (ntiles - 1) * TILE;
    const long long gw = (long long)blockIdx.x * VW_WARPS + warp;
    const long long nwarps = (long long)gridDim.x * VW_WARPS;
    const int my_reads = int((p.N > gw) ? (p.N - gw + nwarps - 1) / nwarps : 0);
    const long long read_elems = (long long)T * C;
    const long long total_elems = p.N * read_elems;
    const long long safe_bytes = (total_elems * EB) & ~15LL;  // bulk copies never read past this
    const long long safe_elems = safe_bytes >> 2;
    const long long stride = p.slot_stride;
    // every tile of every read starts 16-byte aligned and lies inside the bulk-copy-safe range
    const bool aligned = FAST5 || (((read_elems & 3) == 0) && (((TILE * C) & 3) == 0));
    const int nstage = L.nstage;
    const int stage_bytes = L.stage_bytes;
    const int tile_elems = TILE * C;
    const int last_elems = last_rows * C;
    const long long skip_elems = (nwarps - 1) * read_elems;  // from the end of one of my reads to my next one

    // Producer state, advanced by every lane so that it stays warp-uniform; lane 0 issues.  The
    // stage it fills is always the one the consumer has just released, so both share one ring position.
    int p_reads = my_reads;            // reads not yet completely issued
    int p_tl = ntiles;                 // tiles left in the read being issued
    long long p_e0 = gw * read_elems;  // first element of the next tile to issue
    auto issue_next = [&](uint32_t bar, uint32_t stage) {
        if (p_reads <= 0) return;
        const int elems = (p_tl == 1) ? last_elems : tile_elems;
        if (lane == 0) {
            long long b0 = p_e0 * EB;
            uint32_t bytes = uint32_t(elems) * uint32_t(EB);
            if (!aligned) {
                const long long e1 = p_e0 + elems;
                b0 &= ~15LL;
                long long b1 = (e1 * 4 + 15) & ~15LL;
                if (b1 > safe_bytes) b1 = safe_bytes;
                bytes = b1 > b0 ? uint32_t(b1 - b0) : 0u;
            }
            mbar_expect_tx_a(bar, bytes);
            if (FAST5 || bytes) bulk_g2s_a(stage, reinterpret_cast<const char*>(p.post) + b0, bytes, bar);
        }
        p_e0 += elems;
        if (--p_tl == 0) { p_tl = ntiles; p_e0 += skip_elems; --p_reads; }
    };

    if (threadIdx.x < 16) {
        s_invc[threadIdx.x] = c_logf_invc[threadIdx.x];
        s_logc[threadIdx.x] = c_logf_logc[threadIdx.x];
    }
    if (LOGITS && threadIdx.x < 32) s_xt[threadIdx.x] = c_vit_exp2f[threadIdx.x];
    if (lane == 0) {
        for (int s = 0; s < nstage; ++s) mbar_init_a(bar0 + 8 * s, 1);
        mbar_init_fence();
    }
    __syncthreads();  // the only block-level barrier: tables + barrier inits visible
    for (int s = 0; s < nstage; ++s) issue_next(bar0 + 8 * s, stage0 + s * stage_bytes);

    int r_s = 0;  // ring position
    uint32_t r_bar = bar0, r_stage = stage0, r_phase = 0;
    const int r_base = RPL * lane;
    for (int k = 0; k < my_reads; ++k) {
        const long long n = gw + (long long)k * nwarps;
        uint8_t* const out_lab = keep_in_register(p.labels + n * stride);
        QT* const out_qual = keep_in_register(reinterpret_cast<QT*>(p.quals) + n * stride);
        int* const out_path = keep_in_register(p.path + n * stride);
        int out_count = 0;
        bool open_in = false;  // the previous tile ended inside a non-blank run (warp-uniform)
        int last_label = -1;  // None, src/search.rs:337
        bool bad_char = false;

        for (int ti = 0, row0 = 0; ti < ntiles; ++ti, row0 += TILE) {
            const bool last_tile = (ti == ntiles - 1);
            const int rows = last_tile ? last_rows : TILE;
            mbar_wait_a(r_bar, r_phase);

            float pr[RPL];
            int lab[RPL];
            if (FAST5) {
                float f[20];
                if (HALF) {
                    uint32_t hw[10];
#pragma unroll
                    for (int v = 0; v < 5; ++v) lds_v2(r_stage + 40 * lane + 8 * v, hw[2 * v], hw[2 * v + 1]);
                    const bool bf16 = p.in_bf16 != 0;
#pragma unroll
                    for (int j = 0; j < 10; ++j) {
                        const float2 w = widen16x2(hw[j], bf16);
                        f[2 * j] = w.x;
                        f[2 * j + 1] = w.y;
                    }
                } else {
#pragma unroll
                    for (int v = 0; v < 5; ++v) lds_f4(r_stage + 80 * lane + 16 * v, f + 4 * v);
                }
                if (LOGITS) {
#pragma unroll
                    for (int r = 0; r < RPL; ++r) softmax_row(f + 5 * (r & 3), 5, p.softmax, s_xt);
                }
#pragma unroll
                for (int r = 0; r < RPL; ++r) argmax_row5(f + 5 * (r & 3), lab[r], pr[r]);
                if (rows != TILE) {  // rows past the end of the read (stale stage contents) count as blank
#pragma unroll
                    for (int r = 0; r < RPL; ++r)
                        if (r_base + r >= rows) { lab[r] = 0; pr[r] = 0.0f; }
                }
            } else {
                const long long e0 = n * read_elems + (long long)ti * tile_elems;
                const uint32_t stagef = r_stage + 4u * uint32_t(e0 & 3);
                const bool tile_safe = e0 + tile_elems <= safe_elems;  // warp-uniform: the whole tile came through the bulk copy
                auto ld = [&](int i) {
                    return (tile_safe || e0 + i < safe_elems) ? lds_f1(stagef + 4u * uint32_t(i)) : p.post[e0 + i];
                };
                const bool vec4 = tile_safe && (C & 3) == 0 && (e0 & 3) == 0;  // warp-uniform
                // argmax of one staged row (strict '>' from label 0: first maximum, NaN never wins unless it is label 0)
                auto row_argmax = [&](int row, int& arg, float& best) {
                    arg = 0;
                    best = 0.0f;  // rows past the end of the read count as blank
                    if (row >= rows) return;
                    const int le = row * C;
                    if (vec4) {  // rows of whole 16-byte groups: a quarter of the loads and of the bank conflicts
                        float v[4];
                        lds_f4(stagef + 4u * uint32_t(le), v);
                        best = v[0];
#pragma unroll
                        for (int k = 1; k < 4; ++k)
                            if (v[k] > best) { best = v[k]; arg = k; }
                        for (int c = 4; c < C; c += 4) {
                            lds_f4(stagef + 4u * uint32_t(le + c), v);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (v[k] > best) { best = v[k]; arg = c + k; }
                        }
                    } else {
                        best = ld(le);
                        for (int c = 1; c < C; ++c) {
                            const float v = ld(le + c);
                            if (v > best) { best = v; arg = c; }
                        }
                    }
                };
                if (RPL > 1) {
                    // Rows are read interleaved (lane, lane + 32, ...): the lanes of one load are then C words
                    // apart instead of RPL*C, which for alphabets of 4, 8, 16 labels is the difference between an
                    // 8-way and a 32-way shared-memory bank conflict.  The {probability, label} pairs go through
                    // the (idle) quality queue to reach the lanes that own them as RPL consecutive rows.
                    __syncwarp();  // the previous tile's queue reads are over
#pragma unroll
                    for (int r = 0; r < RPL; ++r) {
                        const int row = lane + 32 * r;
                        int arg;
                        float best;
                        row_argmax(row, arg, best);
                        sts_run(q_base + 8 * row, best, arg);
                    }
                    __syncwarp();
#pragma unroll
                    for (int r = 0; r < RPL; ++r) lds_run(q_base + 8 * (r_base + r), pr[r], lab[r]);
                } else {
                    row_argmax(lane, lab[0], pr[0]);
                }
            }
            __syncwarp();  // stage fully consumed (and the previous tile's queue reads are over)
            issue_next(r_bar, r_stage);  // refill the stage just freed
            r_bar += 8;
            r_stage += stage_bytes;
            if (++r_s == nstage) { r_s = 0; r_bar = bar0; r_stage = stage0; r_phase ^= 1u; }

            // emission flags (src/search.rs:347) and output positions
            int prev0 = __shfl_up_sync(FULLMASK, lab[RPL - 1], 1);
            if (lane == 0) prev0 = last_label;
            // The emissions before this lane's first, from one ballot per row position: the four votes are independent,
            // where a shuffle scan is five dependent steps (16 % of the kernel's stall samples sat on it).
            bool em[RPL];
            int li0 = 0;         // tile-local index of this lane's first emission
            int tile_total = 0;  // warp-uniform
            {
                int prev = prev0;
                const unsigned below = (1u << lane) - 1u;
#pragma unroll
                for (int r = 0; r < RPL; ++r) {
                    em[r] = lab[r] != 0 && (!COLLAPSE || lab[r] != prev);
                    prev = lab[r];
                    const unsigned votes = __ballot_sync(FULLMASK, em[r]);
                    li0 += __popc(votes & below);
                    tile_total += __popc(votes);
                }
            }

            // labels and emission rows go straight to the read's slot
            {
                int li = li0;
#pragma unroll
                for (int r = 0; r < RPL; ++r) {
                    if (em[r]) {
                        out_lab[out_count + li] = uint8_t(lab[r]);
                        out_path[out_count + li] = row0 + r_base + r;
                    }
                    li += em[r] ? 1 : 0;
                }
            }
            const int new_last = COLLAPSE ? __shfl_sync(FULLMASK, lab[RPL - 1], 31) : -1;

            if (QSTRING) {
                // Run sums, always in row order (src/search.rs:362-365), as ONE chain over the lane's rows that starts
                // anew wherever the label changes: acc[k] / len[k] = sum and length, up to row k, of the run that holds
                // row k, counted from the run's first row in this lane.  A run that ends inside the lane and began
                // here (at an emission) is complete: it goes to the queue slot of that emission — slot s (1-based)
                // holds the tile's s-th emission, slot 0 the run left open by the previous tile.  The run that holds
                // the lane's last row is finished by the NEXT lane (which adds its leading rows that continue it,
                // possibly none); a run that came in from the previous lane is finished further down.
                // (label_prob_total starts at 0.0, src/search.rs:339,354; 0.0 + p differs from p only in the sign of a
                // zero, which the quality value, a function of 1 - mean, does not see.)
                bool cont[RPL];  // row k carries on row k-1's run (k >= 1)
                float acc[RPL];
                int len[RPL];
                acc[0] = pr[0];
                len[0] = 1;
                bool all_cont = COLLAPSE;
#pragma unroll
                for (int k = 1; k < RPL; ++k) {
                    cont[k] = COLLAPSE && lab[k] == lab[k - 1];
                    acc[k] = cont[k] ? __fadd_rn(acc[k - 1], pr[k]) : pr[k];
                    len[k] = cont[k] ? len[k - 1] + 1 : 1;
                    all_cont = all_cont && cont[k];
                }
                {
                    int e = li0;  // tile-local number of the last emission at or before row k
#pragma unroll
                    for (int k = 0; k < RPL; ++k) {
                        e += em[k] ? 1 : 0;
                        const bool ends = (k + 1 < RPL) ? !cont[k + 1] : !COLLAPSE;  // without collapse every row is a run
                        if (ends && lab[k] != 0 && e != li0) sts_run(q_base + 8 * e, acc[k], len[k]);
                    }
                }
                bool open_out = false;
                if (COLLAPSE) {
                    const bool c0 = lab[0] == prev0;
                    float tS = acc[RPL - 1];  // the run holding the lane's last row, from its first row in this lane
                    int tn = len[RPL - 1];
                    const bool whole = c0 && all_cont;  // every row continues the run carried in
                    if (whole && prev0 != 0) tn = -1;   // its tail is unknown until the carry arrives
                    // The run left open by the previous tile.  The cell is only meaningful when one is open
                    // (open_in): otherwise shared memory may hold anything there — including the bit pattern
                    // of the "tail unknown" marker, which would keep the carry loop below spinning.
                    float o_S = 0.0f;
                    int o_n = 0;
                    if (lane == 0 && open_in) lds_run(open_addr, o_S, o_n);
                    float in_S;
                    int in_n;
                    for (;;) {
                        in_S = __shfl_up_sync(FULLMASK, tS, 1);
                        in_n = __shfl_up_sync(FULLMASK, tn, 1);
                        if (lane == 0) { in_S = o_S; in_n = o_n; }
                        if (tn == -1 && in_n != -1) {
                            tS = in_S;
#pragma unroll
                            for (int r = 0; r < RPL; ++r) tS = __fadd_rn(tS, pr[r]);
                            tn = in_n + RPL;
                        }
                        if (!__any_sync(FULLMASK, in_n == -1)) break;
                    }
                    if (!whole && prev0 > 0) {  // finish the run carried in: it is the emission before this lane's first
                        float hs = in_S;
                        int hn = in_n;
                        bool alive = c0;
#pragma unroll
                        for (int r = 0; r + 1 < RPL; ++r) {
                            if (alive) { hs = __fadd_rn(hs, pr[r]); ++hn; }
                            alive = alive && cont[r + 1];
                        }
                        sts_run(q_base + 8 * li0, hs, hn);
                    }
                    open_out = new_last != 0 && !last_tile;
                    if (lane == 31 && new_last != 0)  // the tile's last run: closed at the end of the read
                        sts_run(last_tile ? q_base + 8 * tile_total : open_addr, tS, tn);  // (src/search.rs:367-373)
                } else {
                    (void)prev0;
                }
                __syncwarp();  // queue complete
                const int hi = tile_total + (open_out ? 0 : 1);
                for (int j = (open_in ? 0 : 1) + lane; j < hi; j += 32) {
                    float sum;
                    int nn;
                    lds_run(q_base + 8 * j, sum, nn);
                    out_qual[out_count + j - 1] = quality_char<QT>(sum, nn, qp, s_invc, s_logc, bad_char);
                }
                open_in = open_out;
            }
            if (COLLAPSE) last_label = new_last;
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
// (alphabets of 128 labels and more) or the input pointer is not 16-byte aligned.
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
    uint64_t xt[32];
    int flag;
};

__device__ __forceinline__ float widen(float v) { return v; }
__device__ __forceinline__ float widen(__half v) { return __half2float(v); }
__device__ __forceinline__ float widen(__nv_bfloat16 v) { return __bfloat162float(v); }

// sm = 0: the row holds posteriors.  sm = 1 / 2: logits; the argmax runs over the row's softmax (input.cuh), formed
// on the fly in three passes over the row (maximum, sum of exponentials, quotients) since the row is not kept.
template <typename IN>
__device__ __forceinline__ void argmax_global(const IN* row, int C, int sm, const uint64_t* xt, int& arg, float& best) {
    if (sm == 0) {
        best = widen(row[0]);
        arg = 0;
        for (int c = 1; c < C; ++c) {
            const float v = widen(row[c]);
            if (v > best) { best = v; arg = c; }
        }
        return;
    }
    float m = widen(row[0]);
    for (int c = 1; c < C; ++c) {
        const float v = widen(row[c]);
        if (v > m) m = v;
    }
    float S = 0.0f;
    for (int c = 0; c < C; ++c) {
        const float e = softmax_exp(__fadd_rn(widen(row[c]), -m), sm, xt);
        S = (c == 0) ? e : __fadd_rn(S, e);
    }
    const float inv = __frcp_rn(S);
    for (int c = 0; c < C; ++c) {
        const float e = softmax_exp(__fadd_rn(widen(row[c]), -m), sm, xt);
        const float v = (sm == 1) ? __fdiv_rn(e, S) : __fmul_rn(e, inv);
        if (c == 0) { best = v; arg = 0; }
        else if (v > best) { best = v; arg = c; }
    }
}

template <typename QT, typename IN>
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
    if (tid < 32) S.xt[tid] = c_vit_exp2f[tid];
    __syncthreads();
    int buf = 0;
    for (long long n = blockIdx.x; n < p.N; n += gridDim.x) {
        const IN* rd = reinterpret_cast<const IN*>(p.post) + n * T * C;
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
                if (row < T) argmax_global(rd + row * C, C, p.softmax, S.xt, arg, best);
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
    const long long stage = ((rows * p.C * (p.in_f16 ? 2 : 4) + 15) & ~15LL) + 32;
    if (stage > 16 * 1024) return false;
    // Ring depth and resident CTAs.  Every warp is one stream of 2.5 KB requests into HBM, and beyond ~2,400-3,000
    // concurrent streams the memory system loses efficiency faster than the extra warps gain (1M reads x 4000 x 5 f32,
    // ms per launch as CTAs per SM x stages; with / without quality strings): 3x3 17.8 / 14.2, 4x2 16.6 / 14.2,
    // 4x3 16.7 / 14.4, 5x3 16.6 / 16.1, 6x3 17.4 / 20.6 (round 1's choice with quality strings), 4x4 18.2 / 14.7 (round
    // 1's choice without), 8x2 17.4 / 20.5, 2x4 22.7 / 13.85.  So the quality variant, which needs its warps for the
    // arithmetic, runs 5 CTAs of 3 stages and the other 3 CTAs of 3 stages; other alphabets keep what shared memory allows.
    int ns = int(((p.qstring ? 8 : 11) * 1024) / stage);
    L->max_ctas = 0;
    if (!p.in_f16 && p.C == 5 && p.softmax == 0) {  // (four labels, whose rows go through the interleaved path, still gain up to
        // 7-8 CTAs; logits, with the softmax in the kernel, keep every warp they can get)
        ns = 3;
        L->max_ctas = p.qstring ? 5 : 3;
    }
    if (p.in_f16) {  // half-size tiles: deeper rings.  With quality strings the kernel is bound by its arithmetic (6 CTAs x 6
        // stages 15.0 ms, 4 x 6 16.3, 3 x 6 18.4); without, by the streams again (3 x 6 8.31 ms, 4 x 4 8.36, 4 x 10 8.65, 8 x 4 19.5)
        ns = p.qstring ? 6 : (p.C == 5 ? 6 : 10);
        if (!p.qstring && p.C == 5 && p.softmax == 0) L->max_ctas = 3;
    }
    static const int stages_knob = env_knob("FCTC_VIT_STAGES");  // tuning knob
    if (stages_knob > 0) ns = stages_knob;
    if (ns > VW_MAX_STAGES) ns = VW_MAX_STAGES;
    if (ns < 2) ns = 2;
    L->nstage = ns;
    L->stage_bytes = int(stage);
    L->off_stage = (ns * 8 + 63) & ~63;  // ns mbarriers come first
    L->off_queue = int(L->off_stage + ns * stage);
    L->off_open = L->off_queue + (TILE + 1) * 8;
    L->warp_bytes = (L->off_open + 16 + 127) & ~127;
    L->cta_fixed = 512;  // logf tables (2 x 16 doubles) + expf table (32 x 8 bytes)
    return L->cta_fixed + VW_WARPS * L->warp_bytes <= VT_SMEM_LIMIT;
}

template <typename QT, int RPL, bool COLLAPSE, bool QSTRING, bool FAST5, bool HALF, bool LOGITS>
static cudaError_t launch_stream(const ViterbiParams& p, const WarpLayout& L, int sm_count, cudaStream_t stream) {
    auto kernel = viterbi_stream_kernel<QT, RPL, COLLAPSE, QSTRING, FAST5, HALF, LOGITS>;
    const int smem = L.cta_fixed + VW_WARPS * L.warp_bytes;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, VW_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    static const int ctas_knob = env_knob("FCTC_VIT_CTAS");  // tuning knob: resident CTAs per SM, up to what fits
    const int want = ctas_knob > 0 ? ctas_knob : L.max_ctas;
    if (want > 0 && want < per_sm) per_sm = want;
    long long grid = (long long)sm_count * per_sm;
    const long long need = (p.N + VW_WARPS - 1) / VW_WARPS;
    if (grid > need) grid = need;
    kernel<<<dim3((unsigned)grid), VW_THREADS, smem, stream>>>(p, L);
    return cudaGetLastError();
}

template <typename QT, int RPL, bool FAST5, bool HALF, bool LOGITS = false>
static cudaError_t launch_stream_flags(const ViterbiParams& p, const WarpLayout& L, int sm_count, cudaStream_t stream) {
    if (p.collapse) {
        return p.qstring ? launch_stream<QT, RPL, true, true, FAST5, HALF, LOGITS>(p, L, sm_count, stream)
                         : launch_stream<QT, RPL, true, false, FAST5, HALF, LOGITS>(p, L, sm_count, stream);
    }
    return p.qstring ? launch_stream<QT, RPL, false, true, FAST5, HALF, LOGITS>(p, L, sm_count, stream)
                     : launch_stream<QT, RPL, false, false, FAST5, HALF, LOGITS>(p, L, sm_count, stream);
}

template <typename QT>
static cudaError_t launch_viterbi_t(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which) {
    WarpLayout L;
    long long grid = (long long)sm_count * 8;
    if (grid > p.N) grid = p.N;
    // Logits: the fast softmax keeps the order of the values, so without quality strings (no probability is used) the
    // rows are decoded as they are; otherwise the five-label streaming kernel has the softmax built in and every other
    // shape takes the generic kernel.
    const bool logits = p.softmax == 1 || (p.softmax == 2 && p.qstring);
    if (p.in_f16) {
        // 16-bit rows are 10 bytes at five labels: reads (and so every 128-row tile) start 16-byte aligned when T % 8 == 0
        if (p.C == 5 && (p.T & 7) == 0 && plan_stream<4>(p, &L)) {
            if (which) *which = 0;
            return logits ? launch_stream_flags<QT, 4, true, true, true>(p, L, sm_count, stream)
                          : launch_stream_flags<QT, 4, true, true>(p, L, sm_count, stream);
        }
        if (which) *which = 1;
        ViterbiParams q = p;
        q.softmax = logits ? p.softmax : 0;
        if (p.in_bf16) viterbi_generic_kernel<QT, __nv_bfloat16><<<dim3((unsigned)grid), VG_THREADS, 0, stream>>>(q);
        else viterbi_generic_kernel<QT, __half><<<dim3((unsigned)grid), VG_THREADS, 0, stream>>>(q);
        return cudaGetLastError();
    }
    const bool fast5 = p.C == 5 && ((p.T * 5) & 3) == 0;  // every read (hence every tile) starts 16-byte aligned
    if (logits && !(fast5 && plan_stream<4>(p, &L))) {
        if (which) *which = 1;
        viterbi_generic_kernel<QT, float><<<dim3((unsigned)grid), VG_THREADS, 0, stream>>>(p);
        return cudaGetLastError();
    }
    if (p.C <= 16 && plan_stream<4>(p, &L)) {
        if (which) *which = 0;
        if (logits) return launch_stream_flags<QT, 4, true, false, true>(p, L, sm_count, stream);
        return fast5 ? launch_stream_flags<QT, 4, true, false>(p, L, sm_count, stream)
                     : launch_stream_flags<QT, 4, false, false>(p, L, sm_count, stream);
    }
    if (p.C <= 127 && plan_stream<1>(p, &L)) {
        if (which) *which = 0;
        return launch_stream_flags<QT, 1, false, false>(p, L, sm_count, stream);
    }
    if (which) *which = 1;
    ViterbiParams q = p;
    q.softmax = 0;  // (fast softmax without quality strings: order-preserving, rows decoded as they are)
    viterbi_generic_kernel<QT, float><<<dim3((unsigned)grid), VG_THREADS, 0, stream>>>(q);
    return cudaGetLastError();
}

cudaError_t launch_viterbi(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which) {
    if (p.N <= 0) return cudaSuccess;
    return p.qual_bytes == 4 ? launch_viterbi_t<uint32_t>(p, sm_count, stream, which)
                             : launch_viterbi_t<uint8_t>(p, sm_count, stream, which);
}

}  // namespace fctc
