// beam_small.cu — CTC prefix beam search for narrow beams (beam_size * n_labels <= 32), sm_100a.
//
// Same algorithm and results as beam.cu (search::beam_search, src/search.rs:159-300; suffix trie,
// src/tree.rs:47-214; crf variant src/search.rs:38-157), different mapping.  beam.cu gives every
// hypothesis a lane and loops over its labels, which at the reference's default beam of 5 leaves 27
// lanes idle.  Here every CANDIDATE has a lane: lane b*C + s is hypothesis b's own continuation
// (s = 0) or its extension by label s-1.  One posterior row is then a straight-line sequence of
// warp-wide steps with no loops over the beam:
//   * the hypotheses live in registers, replicated over the C lanes of each hypothesis;
//   * "is my trie parent in the beam?" is one MATCH.ANY: continuation lanes ask with the parent's
//     node id, the s = 1 lane of every hypothesis answers with its own;
//   * the merge flags (a parent's extension that lands on a hypothesis already in the beam) travel
//     in one REDUX.OR bit mask;
//   * new trie nodes get their ids from a ballot prefix count — lane order is creation order;
//   * ranking is beam_size rounds of REDUX.MAX over one key per lane (a second REDUX on the
//     node-id key when scores tie: lowest node id first, the oracle's rule);
//   * the new beam is gathered with shuffles from the winning lanes.
// Trie nodes are 32-byte records in the warp's HBM pool: {parent, row | label << 24, skip, -,
// child[4]}; `skip` points to the nearest proper ancestor whose depth is a multiple of 32, which
// turns the final traceback from one dependent load per label into len/32 + 32 of them (lane j
// walks the j-th block of 32 labels).
#include <cuda_fp16.h>

#include "device_utils.cuh"
#include "input.cuh"
#include "kernels.h"

#include <cstdlib>

#include "../../include/fctc.h"

namespace fctc {

namespace {

constexpr int BS_WARPS = 4;  // warps (= reads in flight) per CTA
constexpr int BS_THREADS = BS_WARPS * 32;
constexpr int BS_ROWBUF = 160;  // floats per row prefetch buffer (32 rows of 5 labels)
constexpr unsigned FULL = 0xffffffffu;
constexpr int ROOT = -1;
constexpr int BS_NODE_BYTES = 32;
constexpr int BS_WARP_SMEM = 2 * BS_ROWBUF * 4 + 16 + 128;  // row buffers, the root's child row, traceback block heads

#ifndef BS_MINBLOCKS
#define BS_MINBLOCKS 8
#endif
// HALF (plain search only): the posteriors are binary16, widened exactly to f32 as they are fetched.
static __constant__ uint64_t c_beam_exp2f[32] = FCTC_EXP2F_TAB;

// Out of line, so that the posterior-input path (the common one) keeps its register allocation.
static __device__ __noinline__ void softmax_staged_rows(float* rows, int n_rows, int C, int mode, const uint64_t* xt, int lane) {
    for (int r = lane; r < n_rows; r += 32) softmax_row(rows + r * C, C, mode, xt);
}

template <bool CRF, int CT, bool HALF>
__global__ void __launch_bounds__(BS_THREADS, BS_MINBLOCKS) beam_small_kernel(const BeamParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint64_t s_xt[32];  // glibc expf's table, for the exact softmax of logits (input.cuh)
    if (threadIdx.x < 32) s_xt[threadIdx.x] = c_beam_exp2f[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(FULL, threadIdx.x >> 5, 0);  // tells the compiler it is warp-uniform
    const int slot = blockIdx.x * BS_WARPS + warp;
    if (slot >= p.n_slots) return;
    unsigned char* ws = smem + (size_t)warp * BS_WARP_SMEM;
    float* rowbuf = reinterpret_cast<float*>(ws);
    int* root_child = reinterpret_cast<int*>(ws + 2 * BS_ROWBUF * 4);
    int* heads = root_child + 4;
    int* sel = heads;  // winning lane of every rank; shares the traceback's block-head array
    const uint32_t sel_base = smem_u32(sel);

    const int B = p.beam_size, C = CT ? CT : p.C, NL = C - 1;
    const long long T = p.T;
    const float thr = p.threshold;
    const bool collapse = !CRF && p.collapse != 0;
    const int S = CRF ? p.S : 1;
    const int myb = lane / C, sl = lane - myb * C;  // my hypothesis and my candidate of it
    const unsigned lt_mask = (1u << lane) - 1u;

    const long long cap = p.nodes_per_slot;
    int4* nodes = reinterpret_cast<int4*>(static_cast<unsigned char*>(p.scratch) + 256 + (size_t)slot * (size_t)cap * BS_NODE_BYTES);

    int R = BS_ROWBUF / C;  // rows per prefetch chunk
    if (R > 32) R = 32;
    const int chunk_elems = R * C;

    for (;;) {
        unsigned long long ticket = 0;
        if (lane == 0) ticket = atomicAdd(p.ticket, 1ull);
        ticket = __shfl_sync(FULL, ticket, 0);
        if (ticket >= (unsigned long long)p.N) break;
        const long long n = (long long)ticket;
        const float* rd = p.post + n * T * C * S;
        const uint16_t* rdh = reinterpret_cast<const uint16_t*>(p.post) + n * T * C * S;  // the same rows when HALF
        const bool bf16 = p.in_bf16 != 0;
        auto fetch = [&](long long e) -> float { return HALF ? widen16(__ldg(rdh + e), bf16) : __ldg(rd + e); };
        // element i of a posterior row: staged rows are float32 in shared memory; the crf search reads its
        // state-dependent rows where they lie, so 16-bit values are widened here
        auto at = [&](const float* row, long long i) -> float {
            if (CRF && HALF) return widen16(__ldg(rdh + (row - rd) + i), bf16);
            return row[i];
        };

        // the hypothesis of my lane group (src/search.rs:170-175: one root hypothesis, gap_prob 1)
        int nd = ROOT, par = -2, dep = 0, tip = -1, skp = ROOT, st = 0;
        float lp = 0.0f, gp = 1.0f;
        int status = FCTC_OK;
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
            lp = best; gp = init[0]; st = arg;
            if (bad || arg >= S) status = FCTC_REFERENCE_PANIC;  // unwrap() on a NaN / slice out of range
        }
        if (lane < NL) root_child[lane] = -1;
        int nbeam = 1;
        long long n_nodes = 0;

        float pre[BS_ROWBUF / 32];
        if (!CRF) {
            const long long total = T * C;
#pragma unroll
            for (int j = 0; j < BS_ROWBUF / 32; ++j) {
                const int e = lane + 32 * j;
                pre[j] = (e < chunk_elems && e < total) ? fetch(e) : 0.0f;
            }
        }
        __syncwarp();

        for (long long t0 = 0; t0 < T && status == FCTC_OK; t0 += R) {
            float* rows = rowbuf + ((t0 / R) & 1) * BS_ROWBUF;
            if (!CRF) {
#pragma unroll
                for (int j = 0; j < BS_ROWBUF / 32; ++j) {
                    const int e = lane + 32 * j;
                    if (e < chunk_elems) rows[e] = pre[j];
                }
                const long long base = (t0 + R) * C, total = T * C;  // the next chunk lands while this one is processed
#pragma unroll
                for (int j = 0; j < BS_ROWBUF / 32; ++j) {
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

            for (int rr = 0; rr < rows_here; ++rr) {
                const long long t = t0 + rr;
                const bool active = myb < nbeam;
                const float* pr = CRF ? rd + t * S * C + (size_t)st * C : rows + rr * C;  // my hypothesis' row
                const float p_mine = active ? at(pr, sl) : 0.0f;                               // blank for s = 0
                const float p_tip = (active && sl == 0 && tip >= 0) ? at(pr, tip + 1) : 0.0f;
                // my node's child for my label, asked for now so that the load overlaps the steps below
                int x_child = -1;
                if (active && sl > 0)
                    x_child = (nd == ROOT) ? root_child[sl - 1] : reinterpret_cast<const int*>(nodes + 2 * (size_t)nd + 1)[sl - 1];

                // ---- my own continuation (s = 0 lanes) -------------------------------------------------
                bool has_self = false;
                float s_lp = 0.0f, s_gp = 0.0f;
                if (active && sl == 0) {
                    if (p_mine > thr) {  // src/search.rs:191-198
                        s_gp = __fmul_rn(__fadd_rn(lp, gp), p_mine);
                        has_self = true;
                    }
                    if (collapse && tip >= 0 && !(p_tip < thr)) {  // src/search.rs:205-211
                        s_lp = __fmul_rn(lp, p_tip);
                        has_self = true;
                    }
                }
                // is my trie parent in the beam?  then its extension by my label lands on me
                const bool query = active && sl == 0 && has_self && nd != ROOT;
                const bool answers = active && sl == 1;
                const int mkey = query ? par : (answers ? nd : -1000 - lane);
                const unsigned same = __match_any_sync(FULL, mkey);
                const unsigned ans = same & __ballot_sync(FULL, answers);
                const bool found = query && ans != 0u;
                const int r_lane = found ? __ffs(ans) - 1 : lane;  // the s = 1 lane of the parent's hypothesis
                const float plp = __shfl_sync(FULL, lp, r_lane), pgp = __shfl_sync(FULL, gp, r_lane);
                const int ptip = __shfl_sync(FULL, tip, r_lane);
                const int pst = CRF ? __shfl_sync(FULL, st, r_lane) : 0;
                unsigned fold_bit = 0u;
                if (found) {
                    const float pt = CRF ? at(rd, t * S * C + (long long)pst * C + tip + 1) : p_tip;  // the parent's row
                    if (!(pt < thr)) {
                        float c;
                        if (collapse && ptip == tip) c = __fmul_rn(pgp, pt);  // src/search.rs:220-226
                        else c = __fmul_rn(__fadd_rn(plp, pgp), pt);          // :233-238
                        s_lp = __fadd_rn(s_lp, c);                            // :251
                        fold_bit = 1u << (r_lane + tip);  // the parent's lane for label `tip`: it must not propose
                    }
                }
                const unsigned fold_mask = __reduce_or_sync(FULL, fold_bit);

                // ---- candidates: s = 0 my continuation, s >= 1 my extension by label s-1 --------------
                bool valid = false, is_new = false;
                int x = nd;
                float c_lp = s_lp, c_gp = s_gp;
                if (active) {
                    if (sl == 0) {
                        valid = has_self;
                    } else {
                        const int l = sl - 1;
                        c_gp = 0.0f;
                        if (!(p_mine < thr)) {  // src/search.rs:201-203
                            const bool repeat = collapse && l == tip;
                            x = x_child;
                            bool take = true;
                            if (x < 0) {
                                if (repeat && !(gp > 0.0f)) take = false;  // src/search.rs:212-218
                                else is_new = true;
                            } else if ((fold_mask >> lane) & 1u) {
                                take = false;  // merged into the child's own continuation above
                            }
                            if (take) {
                                c_lp = repeat ? __fmul_rn(gp, p_mine) : __fmul_rn(__fadd_rn(lp, gp), p_mine);
                                valid = true;
                            }
                        }
                    }
                }
                // new trie nodes: ids in lane order = the reference's creation order (beam order, then label)
                const unsigned new_mask = __ballot_sync(FULL, is_new);
                const int total_new = __popc(new_mask);
                if (n_nodes + total_new > cap) {
                    status = FCTC_NODE_POOL_OVERFLOW;
                    break;
                }
                if (is_new) {
                    const int l = sl - 1;
                    const int id = int(n_nodes) + __popc(new_mask & lt_mask);
                    x = id;
                    nodes[2 * (size_t)id] = make_int4(nd, int(t) | (l << 24), (dep & 31) == 0 ? nd : skp, 0);  // src/tree.rs:138-142
                    nodes[2 * (size_t)id + 1] = make_int4(-1, -1, -1, -1);                                      // src/tree.rs:143
                    if (nd == ROOT) root_child[l] = id;
                    else reinterpret_cast<int*>(nodes + 2 * (size_t)nd + 1)[l] = id;
                }
                n_nodes += total_new;

                // ---- ranking keys --------------------------------------------------------------------
                const float tot = __fadd_rn(__fadd_rn(c_lp, c_gp), 0.0f);  // probability(), -0 -> +0
                const bool is_nan = valid && (tot != tot);
                uint32_t hi = valid ? (is_nan ? 1u : order_key(tot)) : 0u;
                const uint32_t lo = valid ? 0x7FFFFFFFu - uint32_t(x) : 0u;  // lower node id ranks first
                const int n_cand = __popc(__ballot_sync(FULL, valid));
                if (n_cand >= 2 && __any_sync(FULL, is_nan)) {  // src/search.rs:261-272 (see oracle note on n >= 2)
                    status = FCTC_INCOMPARABLE_VALUES;
                    break;
                }
                if (n_cand == 0) {  // src/search.rs:274-277
                    status = FCTC_RAN_OUT_OF_BEAM;
                    break;
                }
                const int n_keep = n_cand < B ? n_cand : B;
                // Each round the lane holding the maximum takes the next rank: one REDUX and a compare.  Two
                // candidates with the same score would both take it; that shows up as a winner count above
                // n_keep and the row is ranked again by the exact loop (score, then lowest node id).
                // (A double win can also exhaust the candidates early: then the last maximum is 0.)
                const uint32_t hi0 = hi;
                uint32_t m1 = 0u;  // maxima never increase from round to round
                uint32_t sel_addr = sel_base;
                for (int r = 0; r < n_keep; ++r, sel_addr += 4) {
                    m1 = __reduce_max_sync(FULL, hi);
                    asm volatile(  // if (hi == m1) { sel[r] = lane; hi = 0; } without a branch
                        "{\n\t.reg .pred p;\n\t"
                        "setp.eq.u32 p, %0, %1;\n\t"
                        "@p st.shared.b32 [%2], %3;\n\t"
                        "@p mov.u32 %0, 0;\n\t}"
                        : "+r"(hi)
                        : "r"(m1), "r"(sel_addr), "r"(lane)
                        : "memory");
                }
                const int won = __popc(__ballot_sync(FULL, hi0 != 0u && hi == 0u));
                if (m1 == 0u || won != n_keep) {
                    hi = hi0;
                    __syncwarp();
                    for (int r = 0; r < n_keep; ++r) {
                        const uint32_t x1 = __reduce_max_sync(FULL, hi);
                        const uint32_t my_lo = (hi == x1) ? lo : 0u;
                        const uint32_t x2 = __reduce_max_sync(FULL, my_lo);
                        if (hi == x1 && my_lo == x2) { sel[r] = lane; hi = 0u; }
                    }
                }
                __syncwarp();
                const int my_src = (myb < n_keep) ? sel[myb] : lane;
                const int src0 = sel[0];

                // ---- new beam, renormalised by the top score (src/search.rs:278-282) -----------------
                const float top = __shfl_sync(FULL, __fadd_rn(c_lp, c_gp), src0);
                const int g_x = __shfl_sync(FULL, x, my_src);
                const float g_lp = __shfl_sync(FULL, c_lp, my_src), g_gp = __shfl_sync(FULL, c_gp, my_src);
                const int g_nd = __shfl_sync(FULL, nd, my_src), g_par = __shfl_sync(FULL, par, my_src);
                const int g_dep = __shfl_sync(FULL, dep, my_src), g_tip = __shfl_sync(FULL, tip, my_src);
                const int g_skp = __shfl_sync(FULL, skp, my_src);
                const int g_st = CRF ? __shfl_sync(FULL, st, my_src) : 0;
                const int s_sl = my_src - (my_src / C) * C;
                nd = g_x;
                lp = div_keep_zero(g_lp, top);
                gp = div_keep_zero(g_gp, top);
                if (s_sl == 0) {
                    par = g_par; dep = g_dep; tip = g_tip; skp = g_skp; st = g_st;
                } else {
                    par = g_nd; dep = g_dep + 1; tip = s_sl - 1;
                    skp = (g_dep & 31) == 0 ? g_nd : g_skp;
                    st = CRF ? int(((long long)g_st * NL) % S) + (s_sl - 1) : 0;  // src/search.rs:97
                }
                nbeam = n_keep;
                if (CRF && __any_sync(FULL, myb < nbeam && st >= S)) {
                    status = FCTC_REFERENCE_PANIC;  // slice out of bounds in the reference
                    break;
                }
                __syncwarp();  // root_child updates visible to the next row's lookups
            }
        }
        __syncwarp();

        // ---- traceback (src/search.rs:285-300) of the best hypothesis (lane group 0) ------------------
        int len = 0;
        long long off = 0;
        if (status == FCTC_OK) {
            len = __shfl_sync(FULL, dep, 0);
            if (lane == 0) {
                const unsigned long long want = (unsigned long long)((len + 31) & ~31);
                const unsigned long long o = atomicAdd(p.cursor, want);
                off = (o + want > (unsigned long long)p.pool_cap) ? -1 : (long long)o;
            }
            off = __shfl_sync(FULL, off, 0);
            if (off < 0) {
                status = FCTC_OUTPUT_OVERFLOW;
                len = 0;
                off = 0;
            } else {
                // block heads: the final node, then its chain of skip ancestors (depths (len-1)/32*32, -32, ...)
                int head = __shfl_sync(FULL, nd, 0);
                int head_depth = len;
                while (head_depth > 0) {  // warp-uniform
                    if (lane == 0) {
                        int hnode = head;
                        for (int j = 0; j < 32 && hnode != ROOT; ++j) {
                            heads[j] = hnode;
                            hnode = nodes[2 * (size_t)hnode].z;
                        }
                        head = hnode;
                    }
                    __syncwarp();
                    // lane j walks block j: from its head down to (not including) the next multiple of 32
                    const int d_hi = (lane == 0) ? head_depth : ((head_depth - 1) & ~31) - 32 * (lane - 1);
                    if (d_hi > 0) {
                        const int d_lo = (d_hi - 1) & ~31;
                        int node = heads[lane];
                        for (int d = d_hi; d > d_lo; --d) {
                            const int4 rc = nodes[2 * (size_t)node];
                            p.labels[off + d - 1] = uint8_t(((rc.y >> 24) & 0xff) + 1);
                            p.path[off + d - 1] = rc.y & 0xffffff;
                            node = rc.x;
                        }
                    }
                    head = __shfl_sync(FULL, head, 0);
                    const int done_to = ((head_depth - 1) & ~31) - 32 * 31;  // depth below which the next round starts
                    head_depth = done_to > 0 ? done_to : 0;
                    __syncwarp();
                }
            }
        }
        if (lane == 0) {
            p.offsets[n] = off;
            p.lens[n] = len;
            p.status[n] = status;
        }
        __syncwarp();
    }
}

}  // namespace

bool beam_small_eligible(int beam_size, int C, int64_t T) {
    static const bool forced_wide = [] {  // tuning knob: FCTC_BEAM_SMALL=0 forces the wide kernel (read once)
        const char* e = getenv("FCTC_BEAM_SMALL");
        return e && atoi(e) == 0;
    }();
    if (forced_wide) return false;
    return C >= 2 && C <= 5 && beam_size >= 1 && beam_size * C <= 32 && T < (1LL << 24);
}

int64_t beam_small_slot_bytes(int64_t cap) { return cap * BS_NODE_BYTES; }

int beam_small_warps_per_cta() { return BS_WARPS; }

int beam_small_warps_per_sm() {
    int per_sm = 32;
    static const int per_sm_knob = env_knob("FCTC_BEAM_SMALL_WARPS_PER_SM");  // tuning knob
    if (per_sm_knob > 0) per_sm = per_sm_knob;
    return per_sm;
}

cudaError_t launch_beam_small(const BeamParams& p, cudaStream_t stream) {
    const bool crf = p.S > 0;
    auto kernel = crf ? (p.in_f16 ? (p.C == 5 ? beam_small_kernel<true, 5, true> : beam_small_kernel<true, 0, true>)
                                  : (p.C == 5 ? beam_small_kernel<true, 5, false> : beam_small_kernel<true, 0, false>))
                      : p.in_f16 ? (p.C == 5 ? beam_small_kernel<false, 5, true> : beam_small_kernel<false, 0, true>)
                                 : (p.C == 5 ? beam_small_kernel<false, 5, false> : beam_small_kernel<false, 0, false>);
    const int smem = BS_WARP_SMEM * BS_WARPS;
    cudaError_t e = cudaMemsetAsync(p.ticket, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.cursor, 0, sizeof(unsigned long long), stream);  // the pool fills from entry 0 on every call
    if (e != cudaSuccess) return e;
    const int grid = (p.n_slots + BS_WARPS - 1) / BS_WARPS;
    kernel<<<grid, BS_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace fctc
