// beam.cu — batched 1-D CTC prefix beam search for sm_100a: one warp per read.
//
// Replaces search::beam_search (src/search.rs:159-301) with its SearchPoint (:7-28), the prefix
// trie tree::SuffixTree (src/tree.rs:17-194) and its child table vec2d::Vec2D (src/vec2d.rs).
// The arithmetic is the reference's, operation for operation, in linear f32 probability space
// (compiled with -fmad=false; '/' is IEEE division): expansion with the beam_cut_threshold prune
// (blank uses '>', labels use "not <", src/search.rs:191,201), merge of hypotheses that reach the
// same trie node, ranking by label_prob + gap_prob, cut to beam_size, renormalisation by the top
// score, and a parent-pointer traceback that reports each label's first-proposal row.
//
// What is different is the organisation:
//   * a read is owned by one warp for its whole life (the search is sequential in time, reads are
//     independent); warps take reads from an atomic ticket counter;
//   * the beam lives in shared memory (two buffers of B entries per warp), posterior rows are
//     prefetched into a shared-memory ring 32 rows ahead;
//   * the trie is a per-warp node pool in HBM scratch: {parent, first-proposal row} records, a
//     label byte and a child table row per node, ids in creation order exactly as the reference
//     allocates them (beam rank major, label minor) so that tie-breaks by node id agree;
//   * merging needs no sort: a node can only receive its own blank/repeat continuation and its
//     parent's extension (<= 2 non-zero terms per field, so f32 addition order is immaterial);
//     each hypothesis looks up its parent in the beam and folds that term in itself;
//   * ranking is a warp-level selection, no comparison sort: every lane orders its own hypothesis'
//     C candidates by pairwise counting, then each output rank is one REDUX.MAX over the list
//     heads; exact score ties (two lanes holding the maximum) are detected afterwards and only
//     then ranked again with the node-id key (lowest node id first);
//   * narrow beams (beam_size * labels <= 32) take beam_small.cu instead: one lane per candidate.
#include <cuda_fp16.h>

#include "device_utils.cuh"
#include "input.cuh"
#include "kernels.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>

#include "../../include/fctc.h"

namespace fctc {

constexpr int BM_WARPS = 4;  // warps (= reads in flight) per CTA
constexpr int BM_THREADS = BM_WARPS * 32;
constexpr int BM_ROWBUF = 160;  // floats per prefetch buffer (32 rows of five labels; 7 CTAs of 4 warps fit one SM at beam 32)
constexpr unsigned FULL = 0xffffffffu;
constexpr int ROOT = -1;

// Trie pool of one warp slot in scratch: records, child table, label bytes; each 256-byte aligned.
__host__ __device__ inline size_t up256(size_t v) { return (v + 255) & ~size_t(255); }
__host__ __device__ inline size_t beam_slot_child_offset(long long cap) { return up256((size_t)cap * 8); }
__host__ __device__ inline size_t beam_slot_label_offset(long long cap, int NL) {
    return beam_slot_child_offset(cap) + up256((size_t)cap * 4 * (size_t)(NL > 0 ? NL : 1));
}
__host__ __device__ inline size_t beam_slot_bytes(long long cap, int NL) {
    return beam_slot_label_offset(cap, NL) + up256((size_t)cap);
}

struct BeamLayout {  // per-warp shared memory plan, in bytes from the warp's base
    int beam_stride;  // one beam buffer: 6 arrays of B words (the crf search's state arrays live at off_st)
    int off_newcnt, off_fold, off_key, off_node, off_lp, off_gp, off_src, off_sel, off_srec, off_rows, off_root, off_st;
    int warp_bytes;
    int rows_per_chunk;
};

struct BeamView {  // one beam buffer
    int* nd;
    int* par;
    int* dep;
    int* tip;
    float* lp;
    float* gp;
    int* st;  // crf transition state (src/search.rs:11-12); unused by the plain search
};
__device__ __forceinline__ BeamView beam_view(unsigned char* ws, const BeamLayout& L, int buf, int B) {
    BeamView v;
    v.nd = reinterpret_cast<int*>(ws + buf * L.beam_stride);
    v.par = v.nd + B;
    v.dep = v.par + B;
    v.tip = v.dep + B;
    v.lp = reinterpret_cast<float*>(v.tip + B);
    v.gp = v.lp + B;
    v.st = reinterpret_cast<int*>(ws + L.off_st) + buf * B;  // only planned (and only touched) by the crf search
    return v;
}

// CRF = true gives search::crf_beam_search (src/search.rs:38-157): posteriors are [T, S, C], every
// hypothesis carries a transition state and reads row probs[t][state][:], a label moves it to
// (state * n_base) % n_state + label, and there is no repeat-collapse case.
// CT = compile-time label count (0 = read it from the parameters): with CT = 5, the nanopore
// alphabet, every per-label loop unrolls and the slot arithmetic becomes constant.
// HALF (plain search only): the posteriors are binary16, widened exactly to f32 as they are fetched.
static __constant__ uint64_t c_beam_exp2f[32] = FCTC_EXP2F_TAB;

// Out of line, so that the posterior-input path (the common one) keeps its register allocation.
static __device__ __noinline__ void softmax_staged_rows(float* rows, int n_rows, int C, int mode, const uint64_t* xt, int lane) {
    for (int r = lane; r < n_rows; r += 32) softmax_row(rows + r * C, C, mode, xt);
}

// Launches sharing one scratch (kernels.h): the warp takes a free trie pool from the ring (slot number + 1) and puts it
// back when it has no read left.  Every lane takes every turn of the waiting loop: a loop that only lane 0 runs leaves
// the warp diverged in the compiler's eyes for the rest of the kernel, and the per-warp state moves from uniform to
// per-thread registers (48 -> 80 registers, 6 resident CTAs instead of 8).
static __device__ __noinline__ unsigned beam_take_slot(unsigned* ctr, unsigned* ring, unsigned n) {
    const bool first = (threadIdx.x & 31) == 0;
    unsigned i = 0;
    if (first) i = atomicAdd(ctr, 1u) % n;
    unsigned v;
    do {
        unsigned x = 0;
        if (first) x = atomicExch(ring + i, 0u);
        v = __shfl_sync(0xffffffffu, x, 0);
        if (v == 0u) __nanosleep(256);
    } while (v == 0u);
    __threadfence();
    return v;
}

static __device__ __noinline__ void beam_return_slot(unsigned* ctr, unsigned* ring, unsigned n, int slot) {
    __threadfence();  // the pool's stores are ordered before the slot number shows up in the ring
    __syncwarp();
    if ((threadIdx.x & 31) == 0) {
        const unsigned j = atomicAdd(ctr + 1, 1u) % n;
        while (atomicCAS(ring + j, 0u, unsigned(slot) + 1u) != 0u) __nanosleep(256);
    }
}

template <bool CRF, int CT, bool HALF>
__global__ void __launch_bounds__(BM_THREADS) beam_search_kernel(const BeamParams p, const BeamLayout L) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint64_t s_xt[32];  // glibc expf's table, for the exact softmax of logits (input.cuh)
    if (threadIdx.x < 32) s_xt[threadIdx.x] = c_beam_exp2f[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // tells the compiler it is warp-uniform
    int slot = blockIdx.x * BM_WARPS + warp;
    if (slot >= p.n_slots) return;
    if (p.slot_ring) slot = int(beam_take_slot(p.ring_ctr, p.slot_ring, unsigned(p.ring_size))) - 1;
    unsigned char* ws = smem + (size_t)warp * L.warp_bytes;
    const int B = p.beam_size, C = CT ? CT : p.C, NL = C - 1;
    const long long T = p.T;
    const float thr = p.threshold;
    const bool collapse = !CRF && p.collapse != 0;
    const int S = CRF ? p.S : 1;

    int* newcnt = reinterpret_cast<int*>(ws + L.off_newcnt);
    unsigned char* fold = ws + L.off_fold;
    uint32_t* ckey = reinterpret_cast<uint32_t*>(ws + L.off_key);
    int* cnode = reinterpret_cast<int*>(ws + L.off_node);
    float* clp = reinterpret_cast<float*>(ws + L.off_lp);
    float* cgp = reinterpret_cast<float*>(ws + L.off_gp);
    int* sel = reinterpret_cast<int*>(ws + L.off_sel);
    // per-lane ranked candidate lists: (C + 1) x {score key, candidate slot}
    const uint32_t srec_base = smem_u32(ws + L.off_srec);
    float* rowbuf = reinterpret_cast<float*>(ws + L.off_rows);
    int* root_child = reinterpret_cast<int*>(ws + L.off_root);

    // this warp's trie pool
    const long long cap = p.nodes_per_slot;
    unsigned char* pool = static_cast<unsigned char*>(p.scratch) + 256 + (size_t)slot * beam_slot_bytes(cap, NL);
    int2* rec = reinterpret_cast<int2*>(pool);                                  // {parent, first-proposal row}
    int* child = reinterpret_cast<int*>(pool + beam_slot_child_offset(cap));    // [node][label] -> node | -1
    uint8_t* nlab = reinterpret_cast<uint8_t*>(pool + beam_slot_label_offset(cap, NL));

    const int R = L.rows_per_chunk;  // rows per prefetch chunk, R*C <= BM_ROWBUF
    const int chunk_elems = R * C;

    for (int i = lane; i < B * NL; i += 32) fold[i] = 0;

    for (;;) {
        unsigned long long ticket = 0;
        if (lane == 0) ticket = atomicAdd(p.ticket, 1ull);
        ticket = __shfl_sync(FULL, ticket, 0);
        if (ticket >= (unsigned long long)p.N) break;
        const long long n = (long long)ticket;
#ifdef FCTC_BEAM_TIMING
        unsigned long long t_start;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
#endif
        const float* rd = p.post + n * T * C * S;
        const uint16_t* rdh = reinterpret_cast<const uint16_t*>(p.post) + n * T * C * S;  // the same rows when HALF
        const bool bf16 = p.in_bf16 != 0;
        auto fetch = [&](long long e) -> float { return HALF ? widen16(__ldg(rdh + e), bf16) : __ldg(rd + e); };
        // element i of a posterior row: staged rows are float32 in shared memory; the crf search reads its
        // state-dependent rows where they lie, so 16-bit values are widened here
        auto at = [&](const float* row, int i) -> float {
            if (CRF && HALF) return widen16(__ldg(rdh + (row - rd) + i), bf16);
            return row[i];
        };

        int cur = 0;
        BeamView bv = beam_view(ws, L, 0, B);
        int status = FCTC_OK;
        if (lane == 0) {  // src/search.rs:170-175
            bv.nd[0] = ROOT; bv.par[0] = -2; bv.dep[0] = 0; bv.tip[0] = -1; bv.lp[0] = 0.0f; bv.gp[0] = 1.0f;
            if (CRF) bv.st[0] = 0;
            if (CRF) {  // src/search.rs:54-59: label_prob = max(init), gap_prob = init[0], state = argmax(init)
                const float* init = p.init + n * p.n_init;
                float best = init[0];
                int arg = 0;
                bool bad = (best != best) && p.n_init > 1;
                for (int i = 1; i < p.n_init; ++i) {
                    const float v = init[i];
                    if (v != v) bad = true;
                    if (v > best) { best = v; arg = i; }
                }
                bv.lp[0] = best; bv.gp[0] = init[0]; bv.st[0] = arg;
                if (bad || arg >= S) status = FCTC_REFERENCE_PANIC;  // unwrap() on a NaN / slice out of range
            }
        }
        status = __shfl_sync(FULL, status, 0);
        for (int l = lane; l < NL; l += 32) root_child[l] = -1;
        int nbeam = 1;
        long long n_nodes = 0;

        // prefetch chunk 0 of the posterior rows (the crf search reads its state-dependent rows directly)
        float pre[BM_ROWBUF / 32];
        if (!CRF) {
            const long long total = T * C;
#pragma unroll
            for (int j = 0; j < BM_ROWBUF / 32; ++j) {
                const int e = lane + 32 * j;
                pre[j] = (e < chunk_elems && e < total) ? fetch(e) : 0.0f;
            }
        }
        __syncwarp();

        for (long long t0 = 0; t0 < T && status == FCTC_OK; t0 += R) {
            float* rows = rowbuf + ((t0 / R) & 1) * BM_ROWBUF;
            if (!CRF) {
#pragma unroll
                for (int j = 0; j < BM_ROWBUF / 32; ++j) {
                    const int e = lane + 32 * j;
                    if (e < chunk_elems) rows[e] = pre[j];
                }
            }
            if (!CRF) {   // start loading the next chunk; it lands while this one is processed
                const long long base = (t0 + R) * C, total = T * C;
#pragma unroll
                for (int j = 0; j < BM_ROWBUF / 32; ++j) {
                    const long long e = base + lane + 32 * j;
                    pre[j] = (lane + 32 * j < chunk_elems && e < total) ? fetch(e) : 0.0f;
                }
            }
            __syncwarp();
            const int rows_here = int((T - t0 < R) ? (T - t0) : R);
            if (!CRF && p.softmax != 0) {  // logits: every staged row becomes its softmax (input.cuh), one row per lane
                softmax_staged_rows(rows, rows_here, C, p.softmax, s_xt, lane);
                __syncwarp();
            }

            for (int rr = 0; rr < rows_here && status == FCTC_OK; ++rr) {
                const long long t = t0 + rr;
                const float* pr_shared = rows + rr * C;
                const float* pr_time = rd + t * S * C;  // crf: [S][C] block of this time step
                BeamView nb = beam_view(ws, L, cur ^ 1, B);

                // The child rows of the beam's nodes are needed in phase B1; with four labels a row is one
                // 16-byte load, issued here so that its latency hides behind phase A (beams of <= 32).
                int4 kids = make_int4(-1, -1, -1, -1);
                if (CT == 5 && B <= 32 && lane < nbeam) {
                    const int node = bv.nd[lane];
                    if (node != ROOT) kids = *reinterpret_cast<const int4*>(child + (size_t)node * 4);
                }
                // ---- phase A: each hypothesis' own continuation, plus its parent's extension -------
                for (int b = lane; b < nbeam; b += 32) {
                    const int node = bv.nd[b];
                    const float lp = bv.lp[b], gp = bv.gp[b];
                    const int tip = bv.tip[b];
                    const float* pr = CRF ? pr_time + (size_t)bv.st[b] * C : pr_shared;
                    const float p_blank = at(pr, 0);
                    bool has_self = false;
                    float s_lp = 0.0f, s_gp = 0.0f;
                    if (p_blank > thr) {  // src/search.rs:191-198
                        s_gp = __fmul_rn(__fadd_rn(lp, gp), p_blank);
                        has_self = true;
                    }
                    if (collapse && tip >= 0) {  // src/search.rs:205-211
                        const float pt = at(pr, tip + 1);
                        if (!(pt < thr)) {
                            s_lp = __fmul_rn(lp, pt);
                            has_self = true;
                        }
                    }
                    if (has_self && node != ROOT) {
                        // is my trie parent in the beam?  then its extension by my label lands on me
                        const int mypar = bv.par[b];
                        int r = -1;
                        for (int k = 0; k < nbeam; ++k)
                            if (bv.nd[k] == mypar) r = k;
                        if (r >= 0) {
                            const float* prp = CRF ? pr_time + (size_t)bv.st[r] * C : pr_shared;  // the parent's row
                            const float pt = at(prp, tip + 1);
                            if (!(pt < thr)) {
                                const float plp = bv.lp[r], pgp = bv.gp[r];
                                float c;
                                if (collapse && bv.tip[r] == tip) c = __fmul_rn(pgp, pt);      // :220-226
                                else c = __fmul_rn(__fadd_rn(plp, pgp), pt);                   // :233-238
                                s_lp = __fadd_rn(s_lp, c);                                       // :251
                                fold[r * NL + tip] = 1;
                            }
                        }
                    }
                    const int s0 = b * C;
                    cnode[s0] = node;
                    clp[s0] = s_lp;
                    cgp[b] = s_gp;  // one per hypothesis: a child candidate's gap probability is always 0 (src/search.rs:213,230)
                    ckey[s0] = has_self ? 1u : 0u;  // real key is filled in after the merge
                }
                __syncwarp();

                // ---- phase B1: child proposals; new nodes get provisional ids -2-k ----------------
                for (int b = lane; b < nbeam; b += 32) {
                    const int node = bv.nd[b];
                    const float lp = bv.lp[b], gp = bv.gp[b];
                    const int tip = bv.tip[b];
                    const float* pr = CRF ? pr_time + (size_t)bv.st[b] * C : pr_shared;
                    int n_new = 0;
                    for (int l = 0; l < NL; ++l) {
                        const int s = b * C + 1 + l;
                        const float pl = at(pr, l + 1);
                        uint32_t key = 0;
                        if (!(pl < thr)) {  // src/search.rs:201-203
                            const bool repeat = collapse && l == tip;
                            int x;
                            if (node == ROOT) x = root_child[l];
                            else if (CT == 5 && B <= 32) x = (l == 0) ? kids.x : (l == 1) ? kids.y : (l == 2) ? kids.z : kids.w;
                            else x = child[(size_t)node * NL + l];
                            bool take = true;
                            if (x < 0) {
                                if (repeat && !(gp > 0.0f)) take = false;  // src/search.rs:212-218
                                else x = -2 - (n_new++);
                            } else if (fold[b * NL + l]) {
                                fold[b * NL + l] = 0;  // merged into the child's own entry in phase A
                                take = false;
                            }
                            if (take) {
                                const float v = repeat ? __fmul_rn(gp, pl) : __fmul_rn(__fadd_rn(lp, gp), pl);
                                cnode[s] = x;
                                clp[s] = v;
                                key = 1u;
                            }
                        }
                        ckey[s] = key;
                    }
                    newcnt[b] = n_new;
                }
                __syncwarp();

                // ---- exclusive scan of the new-node counts in beam order (node ids = creation order) --
                int total_new = 0;
                for (int b0 = 0; b0 < nbeam; b0 += 32) {
                    const int b = b0 + lane;
                    const int v = (b < nbeam) ? newcnt[b] : 0;
                    const int inc = warp_inclusive_scan(v, lane);
                    if (b < nbeam) newcnt[b] = total_new + inc - v;
                    total_new += __shfl_sync(FULL, inc, 31);
                }
                if (n_nodes + total_new > cap) {
                    status = FCTC_NODE_POOL_OVERFLOW;
                    break;
                }
                __syncwarp();

                // ---- phase B2: materialise the new nodes, compute ranking keys ---------------------
                bool any_nan = false;
                int n_cand = 0;
                for (int b = lane; b < nbeam; b += 32) {
                    const int node = bv.nd[b];
                    const int base = int(n_nodes) + newcnt[b];
                    for (int sl = 0; sl < C; ++sl) {
                        const int s = b * C + sl;
                        if (ckey[s] == 0u) continue;
                        if (sl > 0 && cnode[s] <= -2) {
                            const int l = sl - 1;
                            const int id = base + (-2 - cnode[s]);
                            cnode[s] = id;
                            rec[id] = make_int2(node, int(t));  // src/tree.rs:138-142, data = idx
                            nlab[id] = uint8_t(l);
                            if (CT == 5) {  // four children: one 16-byte store (the child table is 16-byte aligned)
                                *reinterpret_cast<int4*>(child + (size_t)id * 4) = make_int4(-1, -1, -1, -1);
                            } else {
                                for (int j = 0; j < NL; ++j) child[(size_t)id * NL + j] = -1;  // src/tree.rs:143
                            }
                            if (node == ROOT) root_child[l] = id;
                            else child[(size_t)node * NL + l] = id;
                        }
                        const float tot = __fadd_rn(__fadd_rn(clp[s], sl == 0 ? cgp[b] : 0.0f), 0.0f);  // probability(), -0 -> +0
                        if (tot != tot) any_nan = true;
                        ckey[s] = (tot != tot) ? 1u : order_key(tot);
                        ++n_cand;
                    }
                }
                n_nodes += total_new;
                n_cand = __reduce_add_sync(FULL, n_cand);
                any_nan = __any_sync(FULL, any_nan);
                __syncwarp();
                if (n_cand >= 2 && any_nan) {  // src/search.rs:261-272 (see oracle note on n >= 2)
                    status = FCTC_INCOMPARABLE_VALUES;
                    break;
                }
                if (n_cand == 0) {  // src/search.rs:274-277
                    status = FCTC_RAN_OUT_OF_BEAM;
                    break;
                }

                // ---- ranking: select the best min(B, n_cand) candidates in order --------------------
                // order: score descending, equal scores by ascending node id (the oracle's tie rule)
                const int n_slots_c = nbeam * C;
                const int n_keep = n_cand < B ? n_cand : B;
                if (nbeam <= 32 && n_slots_c > 64) {
                    // Lane b ranks its own hypothesis' C candidates once (pairwise counting, no data-dependent
                    // moves) and lays them out in rank order as {score key, slot} records closed by a zero
                    // sentinel; then every output rank costs one REDUX.MAX (two on a score tie, which reads
                    // the node-id keys) and a head advance on the winning lane.
                    const int base = lane * C;
                    const uint32_t list0 = srec_base + uint32_t(lane) * uint32_t(C + 1) * 8u;
                    if (lane < nbeam) {
                        if (C == 5) {
                            uint32_t hi[5], lo[5];
                            int rank[5] = {0, 0, 0, 0, 0};
#pragma unroll
                            for (int j = 0; j < 5; ++j) {
                                hi[j] = ckey[base + j];
                                lo[j] = hi[j] ? 0x7FFFFFFFu - uint32_t(cnode[base + j]) : 0u;
                            }
#pragma unroll
                            for (int x = 0; x < 5; ++x)
#pragma unroll
                                for (int y = x + 1; y < 5; ++y) {
                                    const bool x_first = hi[x] > hi[y] || (hi[x] == hi[y] && lo[x] >= lo[y]);
                                    rank[y] += x_first ? 1 : 0;
                                    rank[x] += x_first ? 0 : 1;
                                }
#pragma unroll
                            for (int j = 0; j < 5; ++j) {
                                sts_v2(list0 + 8u * uint32_t(rank[j]), hi[j], uint32_t(base + j));
                            }
                        } else {
                            for (int j = 0; j < C; ++j) {
                                const uint32_t hj = ckey[base + j];
                                const uint32_t lj = hj ? 0x7FFFFFFFu - uint32_t(cnode[base + j]) : 0u;
                                int rk = 0;
                                for (int y = 0; y < C; ++y) {
                                    if (y == j) continue;
                                    const uint32_t hy = ckey[base + y];
                                    const uint32_t ly = hy ? 0x7FFFFFFFu - uint32_t(cnode[base + y]) : 0u;
                                    const bool y_first = hy > hj || (hy == hj && (ly > lj || (ly == lj && y < j)));
                                    rk += y_first ? 1 : 0;
                                }
                                sts_v2(list0 + 8u * uint32_t(rk), hj, uint32_t(base + j));
                            }
                        }
                        sts_v2(list0 + 8u * uint32_t(C), 0u, 0u);
                    }
                    uint32_t head = list0;
                    uint32_t cur_hi = 0u, cur_idx = 0u;
                    if (lane < nbeam) lds_v2(head, cur_hi, cur_idx);
                    // Fast rounds: the lane whose list head is the maximum takes the next rank (one REDUX and a
                    // compare).  Equal scores on two lanes would both take it; the winner count then exceeds
                    // n_keep and the exact rounds below (score, then lowest node id) rank the row again.
                    // (A double win can also exhaust the candidates early: then the last maximum is 0.)
                    uint32_t last_max = 0u;  // maxima never increase from round to round
                    uint32_t sel_addr = smem_u32(sel);
                    for (int r = 0; r < n_keep; ++r, sel_addr += 4) {
                        last_max = __reduce_max_sync(FULL, cur_hi);
                        asm volatile(  // if (cur_hi == max) { sel[r] = cur_idx; head += 8; load the next record } without a branch
                            "{\n\t.reg .pred p;\n\t"
                            "setp.eq.u32 p, %0, %3;\n\t"
                            "@p st.shared.b32 [%4], %1;\n\t"
                            "@p add.u32 %2, %2, 8;\n\t"
                            "@p ld.shared.v2.b32 {%0, %1}, [%2];\n\t}"
                            : "+r"(cur_hi), "+r"(cur_idx), "+r"(head)
                            : "r"(last_max), "r"(sel_addr)
                            : "memory");
                    }
                    // every win moved a list head by one record (lists end in a zero sentinel; a lane that
                    // "wins" with a zero only happens when the maximum itself is 0, which forces the re-rank)
                    const int won = __reduce_add_sync(FULL, int(head - list0) >> 3);
                    if (last_max == 0u || won != n_keep) {
                        __syncwarp();
                        head = list0;
                        cur_hi = 0u;
                        if (lane < nbeam) lds_v2(head, cur_hi, cur_idx);
                        for (int r = 0; r < n_keep; ++r) {
                            const uint32_t m1 = __reduce_max_sync(FULL, cur_hi);
                            unsigned who = __ballot_sync(FULL, cur_hi == m1);
                            if (who & (who - 1)) {  // several lanes share the top score: lowest node id wins
                                // the tie key: lower node id ranks first (list heads carry the candidate's slot)
                                const uint32_t my_lo = (cur_hi == m1 && cur_hi != 0u) ? 0x7FFFFFFFu - uint32_t(cnode[cur_idx]) : 0u;
                                const uint32_t m2 = __reduce_max_sync(FULL, my_lo);
                                who = __ballot_sync(FULL, cur_hi == m1 && my_lo == m2);
                            }
                            const int wl = __ffs(who) - 1;
                            if (lane == wl) {
                                sel[r] = int(cur_idx);
                                head += 8;
                                lds_v2(head, cur_hi, cur_idx);
                            }
                        }
                    }
                } else {
                    uint32_t best_hi = 0, best_lo = 0;
                    int best_i = -1;
                    auto rescan = [&]() {
                        best_hi = 0; best_lo = 0; best_i = -1;
                        for (int i = lane; i < n_slots_c; i += 32) {
                            const uint32_t k = ckey[i];
                            if (k == 0u) continue;
                            const uint32_t lo = 0x7FFFFFFFu - uint32_t(cnode[i]);  // lower node id ranks first
                            if (k > best_hi || (k == best_hi && lo > best_lo)) { best_hi = k; best_lo = lo; best_i = i; }
                        }
                    };
                    rescan();
                    for (int r = 0; r < n_keep; ++r) {
                        const uint32_t m1 = __reduce_max_sync(FULL, best_hi);
                        const uint32_t m2 = __reduce_max_sync(FULL, best_hi == m1 ? best_lo : 0u);
                        if (best_hi == m1 && best_lo == m2) {
                            sel[r] = best_i;
                            ckey[best_i] = 0u;
                            rescan();
                        }
                    }
                }
                __syncwarp();

                // ---- new beam, renormalised by the top score (src/search.rs:278-282) -----------------
                const int i0 = sel[0];
                const float top = __fadd_rn(clp[i0], (i0 % C == 0) ? cgp[i0 / C] : 0.0f);
                for (int r = lane; r < n_keep; r += 32) {
                    const int i = sel[r];
                    const int b = i / C, sl = i - b * C;
                    nb.nd[r] = cnode[i];
                    nb.lp[r] = div_keep_zero(clp[i], top);
                    nb.gp[r] = div_keep_zero(sl == 0 ? cgp[b] : 0.0f, top);
                    if (sl == 0) {
                        nb.par[r] = bv.par[b]; nb.dep[r] = bv.dep[b]; nb.tip[r] = bv.tip[b];
                        if (CRF) nb.st[r] = bv.st[b];
                    } else {
                        nb.par[r] = bv.nd[b]; nb.dep[r] = bv.dep[b] + 1; nb.tip[r] = sl - 1;
                        if (CRF) nb.st[r] = int(((long long)bv.st[b] * NL) % S) + (sl - 1);  // src/search.rs:97
                    }
                }
                bool bad_state = false;
                if (CRF)
                    for (int r = lane; r < n_keep; r += 32) bad_state |= (nb.st[r] >= S);
                if (CRF && __any_sync(FULL, bad_state)) status = FCTC_REFERENCE_PANIC;  // slice out of bounds in the reference
                __syncwarp();
                cur ^= 1;
                bv = nb;
                nbeam = n_keep;
            }
        }
        if (status != FCTC_OK) {
            // leave the fold flags clean for the next read
            for (int i = lane; i < B * NL; i += 32) fold[i] = 0;
        }
        __syncwarp();

        // ---- traceback (src/search.rs:285-300) -----------------------------------------------------
        if (lane == 0) {
            int len = 0;
            long long off = 0;
            if (status == FCTC_OK) {
                len = bv.dep[0];
                const unsigned long long want = (unsigned long long)((len + 31) & ~31);
                const unsigned long long o = atomicAdd(p.cursor, want);
                if (o + want > (unsigned long long)p.pool_cap) {
                    status = FCTC_OUTPUT_OVERFLOW;
                    len = 0;
                } else {
                    off = (long long)o;
                    int node = bv.nd[0];
                    for (int k = len - 1; k >= 0; --k) {
                        const int2 rc = rec[node];
                        p.labels[off + k] = uint8_t(nlab[node] + 1);
                        p.path[off + k] = rc.y;
                        node = rc.x;
                    }
                }
            }
            p.offsets[n] = off;
            p.lens[n] = len;
            p.status[n] = status;
#ifdef FCTC_BEAM_TIMING
            {   // development probe (tools/beam_timing_probe.py), in place of the read's offset: start in microseconds
                // (bits 0-31), duration in units of 8 microseconds (bits 32-47), SM (bits 48-63)
                unsigned long long t_end;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
                unsigned smid;
                asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
                unsigned long long dur = (t_end - t_start) / 8000ull;
                if (dur > 0xffffull) dur = 0xffffull;
                p.offsets[n] = (long long)(((t_start / 1000ull) & 0xffffffffull) | (dur << 32) | ((unsigned long long)smid << 48));
            }
#endif
        }
        __syncwarp();
    }
    if (p.slot_ring) beam_return_slot(p.ring_ctr, p.slot_ring, unsigned(p.ring_size), slot);
}

__global__ void beam_ring_init_kernel(unsigned int* ring, unsigned int* ctr, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ring[i] = unsigned(i) + 1u;
    if (i < 2) ctr[i] = 0u;
}

static bool plan_beam(int B, int C, bool crf, BeamLayout* L) {
    const int NL = C - 1;
    auto up = [](long long v, int a) { return (v + a - 1) / a * a; };
    long long o = 0;
    L->beam_stride = int(up(6LL * 4 * B, 16));
    o = 2LL * L->beam_stride;
    L->off_newcnt = int(o); o = up(o + 4LL * B, 16);
    L->off_fold = int(o);   o = up(o + (long long)B * (NL > 0 ? NL : 1), 16);
    const long long nc = (long long)B * C;
    L->off_key = int(o);    o = up(o + 4 * nc, 16);
    L->off_node = int(o);   o = up(o + 4 * nc, 16);
    L->off_lp = int(o);     o = up(o + 4 * nc, 16);
    L->off_gp = int(o);     o = up(o + 4LL * B, 16);  // own continuations only
    L->off_src = int(o);
    L->off_sel = int(o);    o = up(o + 4LL * B, 16);
    L->off_srec = int(o);   o = up(o + 8LL * 32 * (C + 1), 16);
    L->off_rows = int(o);   o = up(o + 2LL * BM_ROWBUF * 4, 16);
    L->off_root = int(o);   o = up(o + 4LL * (NL > 0 ? NL : 1), 16);
    L->off_st = int(o);     if (crf) o = up(o + 2LL * 4 * B, 16);  // transition states of the two beam buffers
    L->warp_bytes = int(up(o, 128));
    int R = BM_ROWBUF / C;
    if (R > 32) R = 32;
    if (R < 1) R = 1;
    L->rows_per_chunk = R;
    return o * BM_WARPS <= 200 * 1024 && C <= BM_ROWBUF;
}

static int default_nodes_per_row(int C, int B) {
    const long long bound = (long long)(C - 1) * B;  // at most (C-1) proposals per hypothesis per row
    // measured on synthetic posteriors: ~7 proposals per row at beam 5 / threshold 0.1, ~40-68 at
    // beam 32 / threshold 0 (of a bound of 128); reads that need more are flagged and retried
    long long d = 3LL * B;
    if (d < 32) d = 32;
    if (d > bound) d = bound;
    if (d < 1) d = 1;
    return int(d);
}

int beam_wide_warps_per_sm(int beam_size, int C, int64_t T) {
    if (beam_small_eligible(beam_size, C, T)) return 0;
    int per_sm = 24;
    BeamLayout L;
    if (plan_beam(beam_size, C, false, &L)) {
        const long long cta = (long long)L.warp_bytes * BM_WARPS + 256 + 1024;  // + static table + the per-CTA reserve
        long long ctas = (227LL * 1024) / cta;
        if (ctas > 8) ctas = 8;
        if (ctas < 1) ctas = 1;
        per_sm = int(ctas) * BM_WARPS;
    }
    static const int per_sm_knob = env_knob("FCTC_BEAM_WARPS_PER_SM");  // tuning knob
    if (per_sm_knob > 0) per_sm = per_sm_knob;
    return per_sm;
}

int64_t beam_scratch_bytes(int sm_count, int64_t N, int64_t T, int C, int beam_size, int nodes_per_row,
                           int64_t budget_bytes, int64_t* nodes_per_slot, int* n_slots) {
    if (nodes_per_row <= 0) nodes_per_row = default_nodes_per_row(C, beam_size);
    const long long bound = (long long)(C - 1) * beam_size;
    if (nodes_per_row > bound) nodes_per_row = int(bound > 0 ? bound : 1);
    // The kernel is latency-bound (sequential per read), so throughput follows the number of resident
    // warps: 24 per SM is what registers and shared memory allow.  One trie pool per resident warp.
    // as many 4-warp CTAs as one SM's shared memory holds (7 at beam 32 with five labels: 28 warps), 32 warps at most
    int per_sm = beam_wide_warps_per_sm(beam_size, C, T);
    if (per_sm == 0) per_sm = beam_small_warps_per_sm();  // narrow beams: lighter warps
    const int wpc = beam_warps_per_cta(beam_size, C, T);
    long long slots = (long long)sm_count * per_sm;
    if (slots > N) slots = N;
    if (slots < 1) slots = 1;
    slots = (slots + wpc - 1) / wpc * wpc;
    const long long cap = (long long)nodes_per_row * (T > 0 ? T : 1);
    const long long per_slot = (long long)beam_pool_bytes(cap, beam_size, C, T);
    if (budget_bytes > 0) {  // fewer reads in flight rather than more memory
        long long fit = (budget_bytes - 256) / per_slot;
        fit = fit / wpc * wpc;
        if (fit < wpc) fit = wpc;
        if (slots > fit) slots = fit;
    }
    if (nodes_per_slot) *nodes_per_slot = cap;
    if (n_slots) *n_slots = int(slots);
    return 256 + slots * per_slot;
}

int64_t beam_pool_bytes(int64_t cap, int beam_size, int C, int64_t T) {
    return beam_small_eligible(beam_size, C, T) ? beam_small_slot_bytes(cap) : int64_t(beam_slot_bytes(cap, C - 1));
}

int beam_warps_per_cta(int beam_size, int C, int64_t T) {
    return beam_small_eligible(beam_size, C, T) ? beam_small_warps_per_cta() : BM_WARPS;
}

cudaError_t launch_beam(const BeamParams& p, int sm_count, cudaStream_t stream) {
    (void)sm_count;
    if (p.N <= 0) return cudaSuccess;
    if (beam_small_eligible(p.beam_size, p.C, p.T)) return launch_beam_small(p, stream);
    BeamLayout L;
    if (!plan_beam(p.beam_size, p.C, p.S > 0, &L)) return cudaErrorInvalidConfiguration;
    const int smem = L.warp_bytes * BM_WARPS;
    const bool crf = p.S > 0;
    auto kernel = crf ? (p.in_f16 ? (p.C == 5 ? beam_search_kernel<true, 5, true> : beam_search_kernel<true, 0, true>)
                                  : (p.C == 5 ? beam_search_kernel<true, 5, false> : beam_search_kernel<true, 0, false>))
                      : p.in_f16 ? (p.C == 5 ? beam_search_kernel<false, 5, true> : beam_search_kernel<false, 0, true>)
                                 : (p.C == 5 ? beam_search_kernel<false, 5, false> : beam_search_kernel<false, 0, false>);
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.ticket, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.cursor, 0, sizeof(unsigned long long), stream);  // the pool fills from entry 0 on every call
    if (e != cudaSuccess) return e;
    const int grid = p.n_slots / BM_WARPS;
    kernel<<<grid, BM_THREADS, smem, stream>>>(p, L);
    return cudaGetLastError();
}

cudaError_t launch_beam_ring_init(unsigned int* slot_ring, unsigned int* ring_ctr, int ring_size, cudaStream_t stream) {
    if (ring_size <= 0) return cudaErrorInvalidValue;
    beam_ring_init_kernel<<<(ring_size + 255) / 256, 256, 0, stream>>>(slot_ring, ring_ctr, ring_size);
    return cudaGetLastError();
}

bool beam_shape_supported(int beam_size, int C) {
    BeamLayout L;
    return plan_beam(beam_size, C, true, &L);
}

}  // namespace fctc

// ---------------------------------------------------------------------------------------------
// fctc_beam_search_batch_device_static
//
// Some programs that want to decode on the GPU are not in a position to obtain device memory at
// the moment they call us: a real-time basecalling loop that has handed all HBM to its own arena
// at start-up, a plugin running inside a host application that forbids allocator calls on its
// worker threads, or code that runs while another component holds the allocator lock.  The
// regular device entry point makes those callers bring `fctc_beam_scratch_bytes()` bytes of trie
// scratch; this variant brings its own instead: a fixed array that is part of the library's device
// image.  Nothing here calls cudaMalloc/cudaFree (or any routine that allocates) — the price is a
// hard size limit and FCTC_BEAM_STATIC_SCRATCH_BYTES of HBM held, from the moment the library is
// loaded until the process exits, on every GPU that loads it.
//
// How the array is reached: in host code `g_beam_static_scratch` is only the host-side handle of
// the symbol, not an address a kernel may dereference.  The first call on each device therefore
// asks the runtime for the symbol's device address (cudaGetSymbolAddress, with that device
// current) and caches it; later calls reuse the cached pointer.  Make that first call outside any
// CUDA stream capture.
//
// Sharing: every call on a device uses the same array and finds in it whatever the previous call
// left behind.  Only the read ticket counter at its start has to be zero for a launch, and the
// launch clears it on the caller's stream; trie nodes initialise their own child rows when they
// are created, so stale pools are harmless.  Two calls that may run on the GPU at the same time
// (different streams, or different host threads) would decode into the same pools and corrupt
// each other: serialise them.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_beam_static_scratch[FCTC_BEAM_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
// device address of the array, per device (null = not looked up yet).  Atomic: two handles on one device, each on
// its own host thread, may both make the first call; they store the same address.
std::atomic<void*> g_static_scratch_ptr[kMaxDevices] = {};
}  // namespace

extern "C" int fctc_beam_search_batch_device_static(fctc_handle* h, const float* post_dev, int64_t n_reads,
                                                    int64_t n_rows, int32_t n_labels, int32_t beam_size,
                                                    float beam_cut_threshold, int32_t collapse_repeats,
                                                    uint8_t* labels_dev, int32_t* path_dev, int64_t pool_capacity,
                                                    int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                                    unsigned long long* pool_used_dev, int32_t nodes_per_row,
                                                    void* stream) {
    if (!h) return FCTC_E_ARGUMENT;
    if (int rc = fctc::check_beam_call(h, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (!post_dev || !labels_dev || !path_dev || !offsets_dev || !lens_dev || !status_dev || !pool_used_dev)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null device pointer");
    const int dev = fctc::handle_device(h);
    if (dev < 0 || dev >= kMaxDevices) return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "device index above 63");
    cudaError_t e = cudaSetDevice(dev);
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    void* static_ptr = g_static_scratch_ptr[dev].load(std::memory_order_relaxed);
    if (!static_ptr) {  // first call on this device: resolve the symbol's device address
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_beam_static_scratch);
        if (e != cudaSuccess || !ptr)
            return fctc::handle_fail(h, FCTC_E_CUDA, e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned NULL");
        g_static_scratch_ptr[dev].store(ptr, std::memory_order_relaxed);
        static_ptr = ptr;
    }

    // Fit the launch into the fixed array: fewer resident warps (pool slots) if need be, never more bytes.
    fctc::BeamParams p{};
    const int sm_count = fctc::handle_sm_count(h);
    fctc::beam_scratch_bytes(sm_count, n_reads, n_rows, n_labels, beam_size, nodes_per_row, 0, &p.nodes_per_slot, &p.n_slots);
    const int64_t per_slot = fctc::beam_pool_bytes(p.nodes_per_slot, beam_size, n_labels, n_rows);
    const int wpc = fctc::beam_warps_per_cta(beam_size, n_labels, n_rows);
    int64_t fit = (int64_t(FCTC_BEAM_STATIC_SCRATCH_BYTES) - 256) / per_slot;
    fit = fit / wpc * wpc;
    if (fit < wpc) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "static scratch too small: one trie pool for %lld rows x %d proposals needs %lld bytes and at least %d "
                 "are required, the built-in array holds %lld; use fctc_beam_search_batch_device with caller scratch",
                 (long long)n_rows, int(p.nodes_per_slot / (n_rows > 0 ? n_rows : 1)), (long long)per_slot,
                 wpc, (long long)FCTC_BEAM_STATIC_SCRATCH_BYTES);
        return fctc::handle_fail(h, FCTC_E_CAPACITY, msg);
    }
    if (p.n_slots > fit) p.n_slots = int(fit);

    p.post = post_dev; p.N = n_reads; p.T = n_rows; p.C = n_labels; p.beam_size = beam_size;
    p.threshold = beam_cut_threshold; p.collapse = collapse_repeats; p.labels = labels_dev; p.path = path_dev;
    p.pool_cap = pool_capacity; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.cursor = pool_used_dev;
    p.scratch = static_ptr;
    p.scratch_bytes = FCTC_BEAM_STATIC_SCRATCH_BYTES;
    p.ticket = static_cast<unsigned long long*>(p.scratch);  // cleared by launch_beam on `stream`
    e = fctc::launch_beam(p, sm_count, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    return 0;
}

