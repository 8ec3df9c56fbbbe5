// duplex.cu — batched 2-D ("duplex", pair-consensus) CTC beam search for sm_100a: one warp per pair.
//
// Replaces duplex::beam_search (src/duplex.rs:443-650) with its support types: LogSpace
// (src/duplex.rs:7-80), ProbPair / SearchPoint / SecondaryProbs (:84-210), build_ /
// extend_secondary_probs (:212-249, :338-387), root_probs (:389-409), and the SuffixTree that
// carries a SecondaryProbs per node.  Network 1 drives a prefix beam search in log space; every
// trie node carries the forward probabilities of its labelling over an envelope window of
// network 2; hypotheses are ranked by P1(node) * max_window P2(node).
//
// Organisation (same skeleton as beam.cu):
//   * one warp per pair, pairs taken from an atomic ticket; lane b owns beam entry b (beam <= 32);
//   * both networks are mapped to log space with a re-evaluation of glibc's logf (phred.cuh), so
//     every value that does not pass through exp / log1p is bit-identical to the CPU oracle; in
//     exp_mode 1 (the reference's default build, whose fastexp() evaluates to 0 and turns
//     LogSpace '+' into max) the whole search is exact; in exp_mode 0 (exact exp, the parity
//     target) CUDA's expf / log1pf differ from glibc's in the last bits: tolerance parity;
//   * a node's SecondaryProbs is a ring buffer of W = max window + 1 cells {label, gap, total} in
//     HBM scratch (the reference's Vec with discard_until becomes offset/end indices); `total`
//     (= label (+) gap) is stored so that update_max and children's recurrences need no log-add;
//   * new nodes of a row are built in parallel, one lane per node (each recurrence is sequential
//     in the network-2 row index); extension of the beam's nodes runs in lockstep over that index,
//     which preserves the reference's parents-before-children order (a cell needs its parent's
//     cell of the previous index only).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <cuda_fp16.h>

#include "device_utils.cuh"
#include "duplex_common.cuh"
#include "kernels.h"
#include "phred.cuh"

#include "../../include/fctc.h"

namespace fctc {

struct DuplexLayout {  // per-warp shared memory plan (bytes from the warp's base)
    int beam_stride;   // one beam buffer: 7 arrays of B words
    int off_newcnt, off_fold, off_key, off_node, off_lab, off_gap, off_p2, off_sel, off_row, off_root, off_newlist;
    int warp_bytes;
    int cta_fixed;
};

// SecondaryProbs::get of the parent at index `at` (src/duplex.rs:167-179); only gap and
// probability() are ever used by the recurrences.
// `pslot` is at mod W, kept incrementally by the caller; `par_off` the parent's first valid index.
__device__ __forceinline__ void dx_parent_get(const DxSlot& S, int par, const float* pv, int at, int pslot, int root_len,
                                              int par_off, int par_end, int W, int mode, const uint64_t* XT, float& pgap,
                                              float& ptotal) {
    pgap = dx_neg_inf();
    ptotal = dx_neg_inf();
    if (par == DXROOT) {
        const int idx = at + 1;  // root offset is -1
        if (idx >= 0 && idx < root_len) {
            pgap = S.rootv[idx];
            ptotal = lg_add(dx_neg_inf(), pgap, mode, XT);
        }
    } else if (at >= par_off && at < par_end) {
        pgap = pv[pslot];
        ptotal = pv[W + pslot];
    }
}

// One step of the forward recurrence (src/duplex.rs:234-246 and :373-385).
__device__ __forceinline__ DxCell dx_step(const DxCell& last, float pgap, float ptotal, float p_blank, float p_label,
                                          bool repeat, int mode, const uint64_t* XT) {
    DxCell c;
    c.gap = __fadd_rn(last.total, p_blank);
    c.label = __fadd_rn(p_label, lg_add(last.label, repeat ? pgap : ptotal, mode, XT));
    c.total = lg_add(c.label, c.gap, mode, XT);
    return c;
}

// CRF = true gives duplex::crf_beam_search (src/duplex.rs:652-834): both networks are [T, S, C], every
// hypothesis carries a transition state, rows are read at [t][state], a label moves the state to
// (state * n_base) % n_state + label, and there is no repeat case.  Network-2 values are mapped to
// log space on the fly there (a [T2, S, C] log table per pair would not fit the scratch).
#ifndef DX_MINBLOCKS
#define DX_MINBLOCKS 6
#endif
// (Loops that run once per row are kept rolled: unrolled, the kernel outgrew the instruction cache; see duplex_packed.cu.)
template <bool CRF>
__global__ void __launch_bounds__(DX_THREADS, DX_MINBLOCKS) duplex_search_kernel(const DuplexParams p, const DuplexLayout L) {
    extern __shared__ __align__(16) unsigned char smem[];
    double* s_invc = reinterpret_cast<double*>(smem);
    double* s_logc = s_invc + 16;
    uint64_t* XT = reinterpret_cast<uint64_t*>(s_logc + 16);  // glibc expf's 2^(i/32) table
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // tells the compiler it is warp-uniform
    if (threadIdx.x < 16) {
        s_invc[threadIdx.x] = c_dx_invc[threadIdx.x];
        s_logc[threadIdx.x] = c_dx_logc[threadIdx.x];
    }
    if (threadIdx.x < 32) XT[threadIdx.x] = c_dx_exp2f[threadIdx.x];
    __syncthreads();
    const int slot_id = blockIdx.x * DX_WARPS + warp;
    if (slot_id >= p.n_slots) return;
    unsigned char* ws = smem + L.cta_fixed + (size_t)warp * L.warp_bytes;
    const int B = p.beam_size, C = p.C, NL = C - 1;
    const int T1 = p.T1, T2 = p.T2, W = p.W;
    const int mode = p.exp_mode;
    const bool collapse = !CRF && p.collapse != 0;
    const int NS = CRF ? p.S : 1;
    const float NEG = dx_neg_inf();

    unsigned char* fold = ws + L.off_fold;
    uint32_t* ckey = reinterpret_cast<uint32_t*>(ws + L.off_key);
    int* cnode = reinterpret_cast<int*>(ws + L.off_node);
    float* clab = reinterpret_cast<float*>(ws + L.off_lab);
    float* cgap = reinterpret_cast<float*>(ws + L.off_gap);
    float* cp2 = reinterpret_cast<float*>(ws + L.off_p2);
    int* sel = reinterpret_cast<int*>(ws + L.off_sel);
    float* lp1 = reinterpret_cast<float*>(ws + L.off_row);
    int* root_child = reinterpret_cast<int*>(ws + L.off_root);
    int* newlist = reinterpret_cast<int*>(ws + L.off_newlist);

    const long long cap = p.nodes_per_slot;
    const DxSlot S = dx_slot(static_cast<unsigned char*>(p.scratch) + 256 + (size_t)slot_id * dx_slot_bytes(cap, T2, C, W),
                             cap, T2, C, W);
    const float thr = glibc_logf(p.threshold, s_invc, s_logc);  // src/duplex.rs:454

#pragma unroll 1
    for (int i = lane; i < B * (NL > 0 ? NL : 1); i += 32) fold[i] = 0;

    for (;;) {
        unsigned long long ticket = 0;
        if (lane == 0) ticket = atomicAdd(p.ticket, 1ull);
        ticket = __shfl_sync(DXFULL, ticket, 0);
        if (ticket >= (unsigned long long)p.N) break;
        const long long n = (long long)ticket;
        const float* rd1 = p.post1 + n * (long long)T1 * C * NS;
        const float* rd2 = p.post2 + n * (long long)T2 * C * NS;
        // binary16 posteriors: the same rows behind half pointers, widened on load
        const __half* rh1 = reinterpret_cast<const __half*>(p.post1) + n * (long long)T1 * C * NS;
        const __half* rh2 = reinterpret_cast<const __half*>(p.post2) + n * (long long)T2 * C * NS;
        const bool half_in = p.in_f16 != 0;
        // network-2 value of row i, state s, column c in log space
        auto lp2 = [&](int i, int s_, int c) -> float {
            if (CRF) {
                const size_t e = ((size_t)i * NS + s_) * C + c;
                return glibc_logf(half_in ? __half2float(rh2[e]) : rd2[e], s_invc, s_logc);
            }
            return S.logp2[(size_t)i * C + c];
        };
        const long long* env = p.env ? p.env + n * (long long)T1 * 2 : nullptr;
        int status = FCTC_OK;

        // network 2 in log space (src/duplex.rs:453)
        if (!CRF)
#pragma unroll 1
            for (int i = lane; i < T2 * C; i += 32)
                S.logp2[i] = glibc_logf(half_in ? __half2float(rh2[i]) : rd2[i], s_invc, s_logc);
        int init_state_1 = 0, init_state_2 = 0;
        if (CRF) {  // argmax of the initial state scores (src/duplex.rs:679, 691); NaN aborts the reference
            bool bad = false;
            if (lane == 0) {
#pragma unroll 1
                for (int which = 0; which < 2; ++which) {
                    const float* init = which ? p.init2 + n * p.n_init2 : p.init1 + n * p.n_init1;
                    const int ni = which ? p.n_init2 : p.n_init1;
                    float best = init[0];
                    int arg = 0;
                    bad = bad || (best != best);
#pragma unroll 1
                    for (int i = 1; i < ni; ++i) {
                        const float v = init[i];
                        bad = bad || (v != v);
                        if (v > best) { best = v; arg = i; }
                    }
                    if (arg >= NS) bad = true;
                    if (which) init_state_2 = arg; else init_state_1 = arg;
                }
            }
            init_state_1 = __shfl_sync(DXFULL, init_state_1, 0);
            init_state_2 = __shfl_sync(DXFULL, init_state_2, 0);
            if (__shfl_sync(DXFULL, int(bad), 0)) status = FCTC_REFERENCE_PANIC;
        }
        // root vector over envelope[(0,1)] rows (src/duplex.rs:476-477, 389-409)
        unsigned long long ub0 = env ? (unsigned long long)env[1] : (unsigned long long)T2;
        if (ub0 > (unsigned long long)T2) status = FCTC_INVALID_ENVELOPE;  // the reference would panic on the slice
        __syncwarp();
        int root_len = 0;
        if (status == FCTC_OK) {
            root_len = int(ub0) + 1;
            if (lane == 0) {
                float cur = 0.0f;
                S.rootv[0] = cur;
                int rs = init_state_2;  // crf_root_probs walks the blank transitions (src/duplex.rs:430-438)
#pragma unroll 1
                for (int i = 0; i < int(ub0); ++i) {
                    cur = __fadd_rn(cur, lp2(i, rs, 0));
                    S.rootv[i + 1] = cur;
                    if (CRF) rs = int(((long long)rs * NL) % NS);
                }
            }
        }
        int cur = 0;
        DxBeam bv = dx_beam(ws, B);
        if (lane == 0) {  // src/duplex.rs:465-473
            bv.nd[0] = DXROOT; bv.par[0] = -2; bv.tip[0] = -1; bv.dep[0] = 0;
            bv.lab[0] = NEG; bv.gap[0] = 0.0f; bv.p2[0] = 0.0f;
            bv.st[0] = init_state_1;
        }
#pragma unroll 1
        for (int l = lane; l < NL; l += 32) root_child[l] = -1;
        int nbeam = 1;
        long long n_nodes = 0;
        int last_hi = 0;
        __syncwarp();

        for (int t = 0; t < T1 && status == FCTC_OK; ++t) {
            // ---- envelope row (src/duplex.rs:485-488) ----------------------------------------------
            const unsigned long long lo_u = env ? (unsigned long long)env[2 * t] : 0ull;
            const unsigned long long hi_u = env ? (unsigned long long)env[2 * t + 1] : (unsigned long long)T2;
            const int hi = hi_u > (unsigned long long)T2 ? T2 : int(hi_u);
            if (lo_u >= (unsigned long long)hi || lo_u > (unsigned long long)last_hi) {
                status = FCTC_INVALID_ENVELOPE;
                break;
            }
            const int lo = int(lo_u);
            if (hi - lo + 1 > W) {  // window wider than the scratch plan (max_window too small)
                status = FCTC_NODE_POOL_OVERFLOW;
                break;
            }
            if (!CRF)
#pragma unroll 1
                for (int c = lane; c < C; c += 32)
                    lp1[c] = glibc_logf(half_in ? __half2float(rh1[(size_t)t * C + c]) : rd1[(size_t)t * C + c], s_invc, s_logc);
            // network-1 value of this row for a hypothesis in state s_, column c
            auto row1 = [&](int s_, int c) -> float {
                if (CRF) {
                    const size_t e = ((size_t)t * NS + s_) * C + c;
                    return glibc_logf(half_in ? __half2float(rh1[e]) : rd1[e], s_invc, s_logc);
                }
                return lp1[c];
            };

            // ---- extend the vectors of the beam's nodes (src/duplex.rs:490-522) ----------------------
            if (hi > last_hi) {
                // beam.sort_by_key(node): parents before children; node ids are unique in the beam
                DxBeam nb = dx_beam(ws + (cur ^ 1) * L.beam_stride, B);
                if (lane < nbeam) {
                    const int me = bv.nd[lane];
                    int r = 0;
#pragma unroll 1
                    for (int k = 0; k < nbeam; ++k) r += (bv.nd[k] < me) ? 1 : 0;
                    nb.nd[r] = me; nb.par[r] = bv.par[lane]; nb.tip[r] = bv.tip[lane]; nb.dep[r] = bv.dep[lane];
                    nb.lab[r] = bv.lab[lane]; nb.gap[r] = bv.gap[lane]; nb.p2[r] = bv.p2[lane];
                    nb.st[r] = bv.st[lane];
                }
                __syncwarp();
                cur ^= 1;
                bv = nb;

                const int node = (lane < nbeam) ? bv.nd[lane] : DXROOT;
                const bool active = node >= 0;
                int my_par = -2, my_lab = 0, v_end = 0, par_end = 0;
                const int my_state = (lane < nbeam) ? bv.st[lane] : 0;  // crf_extend uses the entry's own state (:728)
                bool repeat = false;
                float mx = NEG;
                DxCell last{NEG, NEG, NEG};
                float* vd = nullptr;
                if (active) {
                    my_par = bv.par[lane];
                    my_lab = bv.tip[lane];
                    vd = S.vdata + (size_t)node * 2 * W;
                    int off = S.voff[node];
                    v_end = S.vend[node];
                    mx = S.vmx[node];
                    if (lo > off) {  // src/duplex.rs:351-359
                        const int keep = lo - 1;
                        if (keep > off) {  // discard_until (:181-191)
                            if (keep - off >= v_end - off) v_end = keep;  // cleared
                            off = keep;
                        }
                        if (v_end == off) { off = lo; v_end = lo; }
                        // update_max(lo, hi) (:193-204): max over [max(lo, off), min(hi, end)).  The stored
                        // max covers [vlo, end); it stays valid unless a cell that drops out could have
                        // been the maximum (lg_max never picks up NaN cells, so they cannot matter).
                        const int b0 = lo > off ? lo : off;
                        const int e0 = hi < v_end ? hi : v_end;
                        const int old_lo = S.vlo[node];
                        bool rescan = (old_lo > b0) || (e0 <= b0);
                        if (!rescan) {
                            int sl = dx_mod(old_lo, W);
#pragma unroll 1
                            for (int i = old_lo; i < b0; ++i) {
                                if (vd[W + sl] >= mx) rescan = true;
                                if (++sl == W) sl = 0;
                            }
                        }
                        if (rescan) {
                            mx = NEG;
                            int sl = dx_mod(b0, W);
#pragma unroll 1
                            for (int i = b0; i < e0; ++i) {
                                mx = lg_max(mx, vd[W + sl]);
                                if (++sl == W) sl = 0;
                            }
                        }
                        S.vlo[node] = b0;
                        S.voff[node] = off;
                    }
                    if (v_end > off) {
                        const int sl = dx_mod(v_end - 1, W);
                        last.label = S.vlast[node]; last.gap = vd[sl]; last.total = vd[W + sl];
                    }
                    // the parent's label decides the repeat case, independent of collapse (:511)
                    const int plab = (my_par >= 0) ? int(S.nlab[my_par]) : -1;
                    repeat = !CRF && (plab == my_lab);
                    bool par_in_beam = false;
#pragma unroll 1
                    for (int k = 0; k < nbeam; ++k) par_in_beam |= (bv.nd[k] == my_par);
                    par_end = (my_par >= 0) ? ((par_in_beam) ? hi : S.vend[my_par]) : 0;
                }
                __syncwarp();  // every node's new offset is visible before children read their parents
                const int par_off = (active && my_par >= 0) ? S.voff[my_par] : 0;
                const float* pv = (active && my_par >= 0) ? S.vdata + (size_t)my_par * 2 * W : nullptr;
                int imin = active ? v_end : hi;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) {
                    const int o = __shfl_xor_sync(DXFULL, imin, d);
                    imin = o < imin ? o : imin;
                }
                int sl_w = dx_mod(imin, W), sl_p = dx_mod(imin - 1, W);
#pragma unroll 1
                for (int i = imin; i < hi; ++i) {  // lockstep over the network-2 row index
                    if (active && v_end <= i) {
                        float pgap, ptot;
                        // a parent that is in the beam has (by now) cells up to i-1; others are as stored
                        dx_parent_get(S, my_par, pv, i - 1, sl_p, root_len, par_off, (my_par >= 0 && par_end == hi) ? i : par_end, W,
                                      mode, XT, pgap, ptot);
                        const DxCell c = dx_step(last, pgap, ptot, lp2(i, my_state, 0), lp2(i, my_state, my_lab + 1), repeat, mode, XT);
                        vd[sl_w] = c.gap; vd[W + sl_w] = c.total;
                        mx = lg_max(mx, c.total);
                        last = c;
                    }
                    if (++sl_w == W) sl_w = 0;
                    if (++sl_p == W) sl_p = 0;
                    __syncwarp();
                }
                if (active) {
                    S.vlast[node] = last.label;
                    S.vend[node] = hi;
                    S.vmx[node] = mx;
                }
                __syncwarp();
            }
            last_hi = hi;
            __syncwarp();  // lp1 row visible to every lane

            DxBeam nb = dx_beam(ws + (cur ^ 1) * L.beam_stride, B);

            // ---- phase A: own continuation (+ the parent's extension, folded in) ------------------
            if (lane < nbeam) {
                const int b = lane;
                const int node = bv.nd[b];
                const float lab = bv.lab[b], gap = bv.gap[b];
                const int tip = bv.tip[b];
                const float p_blank = row1(bv.st[b], 0);
                bool has_self = false;
                float s_lab = NEG, s_gap = NEG;
                if (p_blank > thr) {  // src/duplex.rs:529-534
                    s_gap = __fadd_rn(lg_add(lab, gap, mode, XT), p_blank);
                    has_self = true;
                }
                if (collapse && tip >= 0) {  // :540-544
                    const float pt = row1(bv.st[b], tip + 1);
                    if (!(pt < thr)) {
                        s_lab = __fadd_rn(lab, pt);
                        has_self = true;
                    }
                }
                if (has_self && node != DXROOT) {
                    const int mypar = bv.par[b];
                    int r = -1;
#pragma unroll 1
                    for (int k = 0; k < nbeam; ++k)
                        if (bv.nd[k] == mypar) r = k;
                    if (r >= 0) {
                        const float pt = row1(bv.st[r], tip + 1);  // the parent's row
                        if (!(pt < thr)) {
                            float c;
                            if (collapse && bv.tip[r] == tip) c = __fadd_rn(bv.gap[r], pt);           // :563-568
                            else c = __fadd_rn(lg_add(bv.lab[r], bv.gap[r], mode, XT), pt);               // :586-590
                            s_lab = lg_add(s_lab, c, mode, XT);                                            // :604
                            fold[r * NL + tip] = 1;
                        }
                    }
                }
                const int s0 = b * C;
                cnode[s0] = node;
                clab[s0] = s_lab;
                cgap[s0] = s_gap;
                cp2[s0] = bv.p2[b];
                ckey[s0] = has_self ? 1u : 0u;
            }
            __syncwarp();

            // ---- phase B1: child proposals; new nodes get provisional ids -2-k -----------------------
            int n_new = 0;
            if (lane < nbeam) {
                const int b = lane;
                const int node = bv.nd[b];
                const float lab = bv.lab[b], gap = bv.gap[b];
                const int tip = bv.tip[b];
#pragma unroll 1
                for (int l = 0; l < NL; ++l) {
                    const int s = b * C + 1 + l;
                    const float pl = row1(bv.st[b], l + 1);
                    uint32_t key = 0;
                    if (!(pl < thr)) {  // :537-539
                        const bool repeat = collapse && l == tip;
                        int x = (node == DXROOT) ? root_child[l] : S.child[(size_t)node * NL + l];
                        bool take = true;
                        if (x < 0) {
                            if (repeat && !(gap > NEG)) take = false;  // :545-561
                            else x = -2 - (n_new++);
                        } else if (fold[b * NL + l]) {
                            fold[b * NL + l] = 0;
                            take = false;
                        }
                        if (take) {
                            cnode[s] = x;
                            clab[s] = repeat ? __fadd_rn(gap, pl) : __fadd_rn(lg_add(lab, gap, mode, XT), pl);
                            cgap[s] = NEG;
                            cp2[s] = bv.p2[b];
                            key = 1u;
                        }
                    }
                    ckey[s] = key;
                }
            }
            const int incl = warp_inclusive_scan(n_new, lane);
            const int total_new = __shfl_sync(DXFULL, incl, 31);
            const int my_base = incl - n_new;
            if (n_nodes + total_new > cap) {
                status = FCTC_NODE_POOL_OVERFLOW;
                break;
            }
            // list the new nodes in creation order (beam rank major, label minor)
            if (lane < nbeam) {
                int k = 0;
#pragma unroll 1
                for (int l = 0; l < NL; ++l) {
                    const int s = lane * C + 1 + l;
                    if (ckey[s] != 0u && cnode[s] <= -2) {
                        newlist[my_base + k] = (lane << 16) | l;
                        cnode[s] = int(n_nodes) + my_base + k;
                        ++k;
                    }
                }
            }
            __syncwarp();

            // ---- phase B2: build the new nodes' vectors, one lane per node (src/duplex.rs:212-249) ----
#pragma unroll 1
            for (int j0 = 0; j0 < total_new; j0 += 32) {
                const int j = j0 + lane;
                if (j < total_new) {
                    const int b = newlist[j] >> 16, l = newlist[j] & 0xffff;
                    const int id = int(n_nodes) + j;
                    const int pnode = bv.nd[b];
                    const int tstate = bv.st[b];  // crf_build uses the proposing hypothesis' state (:766)
                    const bool repeat = collapse && l == bv.tip[b];
                    const int par_end = (pnode >= 0) ? S.vend[pnode] : 0;
                    const int par_off = (pnode >= 0) ? S.voff[pnode] : 0;
                    const float* pv = (pnode >= 0) ? S.vdata + (size_t)pnode * 2 * W : nullptr;
                    float* vd = S.vdata + (size_t)id * 2 * W;
                    DxCell last{NEG, NEG, NEG};
                    float mx = NEG;
                    int sl = dx_mod(lo, W), sl_p = dx_mod(lo - 1, W);
                    for (int i = lo; i < hi; ++i) {
                        float pgap, ptot;
                        dx_parent_get(S, pnode, pv, i - 1, sl_p, root_len, par_off, par_end, W, mode, XT, pgap, ptot);
                        if (++sl_p == W) sl_p = 0;
                        const DxCell c = dx_step(last, pgap, ptot, lp2(i, tstate, 0), lp2(i, tstate, l + 1), repeat, mode, XT);
                        vd[sl] = c.gap; vd[W + sl] = c.total;
                        mx = lg_max(mx, c.total);
                        last = c;
                        if (++sl == W) sl = 0;
                    }
                    S.parent[id] = pnode;
                    S.nlab[id] = uint8_t(l);
                    S.voff[id] = lo;
                    S.vlo[id] = lo;
                    S.vlast[id] = last.label;
                    S.vend[id] = hi;
                    S.vmx[id] = mx;
#pragma unroll 1
                    for (int q = 0; q < NL; ++q) S.child[(size_t)id * NL + q] = -1;
                    if (pnode == DXROOT) root_child[l] = id;
                    else S.child[(size_t)pnode * NL + l] = id;
                }
            }
            n_nodes += total_new;
            __syncwarp();

            // ---- refresh prob_2_max, scores (src/duplex.rs:613-630) -------------------------------------
            bool any_nan = false;
            int n_cand = 0;
            if (lane < nbeam) {
#pragma unroll 1
                for (int sl = 0; sl < C; ++sl) {
                    const int s = lane * C + sl;
                    if (ckey[s] == 0u) continue;
                    const int node = cnode[s];
                    if (node >= 0) cp2[s] = S.vmx[node];
                    const float tot = __fadd_rn(__fadd_rn(lg_add(clab[s], cgap[s], mode, XT), cp2[s]), 0.0f);
                    if (tot != tot) any_nan = true;
                    ckey[s] = (tot != tot) ? 1u : order_key(tot);
                    ++n_cand;
                }
            }
            n_cand = __reduce_add_sync(DXFULL, n_cand);
            any_nan = __any_sync(DXFULL, any_nan);
            __syncwarp();
            if (n_cand >= 2 && any_nan) { status = FCTC_INCOMPARABLE_VALUES; break; }
            if (n_cand == 0) { status = FCTC_RAN_OUT_OF_BEAM; break; }

            // ---- ranking (as beam.cu) -----------------------------------------------------------------------
            const int n_slots_c = nbeam * C;
            const int n_keep = n_cand < B ? n_cand : B;
            uint32_t best_hi = 0, best_lo = 0;
            int best_i = -1;
            auto rescan = [&]() {
                best_hi = 0; best_lo = 0; best_i = -1;
#pragma unroll 1
                for (int i = lane; i < n_slots_c; i += 32) {
                    const uint32_t k = ckey[i];
                    if (k == 0u) continue;
                    const uint32_t lo2 = 0x7FFFFFFFu - uint32_t(cnode[i]);
                    if (k > best_hi || (k == best_hi && lo2 > best_lo)) { best_hi = k; best_lo = lo2; best_i = i; }
                }
            };
            rescan();
#pragma unroll 1
            for (int r = 0; r < n_keep; ++r) {
                const uint32_t m1 = __reduce_max_sync(DXFULL, best_hi);
                const uint32_t m2 = __reduce_max_sync(DXFULL, best_hi == m1 ? best_lo : 0u);
                if (best_hi == m1 && best_lo == m2) {
                    sel[r] = best_i;
                    ckey[best_i] = 0u;
                    rescan();
                }
            }
            __syncwarp();
            if (lane < n_keep) {
                const int i = sel[lane];
                const int b = i / C, sl = i - b * C;
                nb.nd[lane] = cnode[i];
                nb.lab[lane] = clab[i];
                nb.gap[lane] = cgap[i];
                nb.p2[lane] = cp2[i];
                if (sl == 0) { nb.par[lane] = bv.par[b]; nb.tip[lane] = bv.tip[b]; nb.dep[lane] = bv.dep[b]; nb.st[lane] = bv.st[b]; }
                else {
                    nb.par[lane] = bv.nd[b]; nb.tip[lane] = sl - 1; nb.dep[lane] = bv.dep[b] + 1;
                    nb.st[lane] = CRF ? int(((long long)bv.st[b] * NL) % NS) + (sl - 1) : 0;  // src/duplex.rs:775
                }
            }
            if (CRF) {
                const bool bad_state = (lane < n_keep) && (nb.st[lane] >= NS);
                if (__any_sync(DXFULL, bad_state)) status = FCTC_REFERENCE_PANIC;  // the reference's slice would panic
            }
            __syncwarp();
            cur ^= 1;
            bv = nb;
            nbeam = n_keep;
        }
        if (status != FCTC_OK)
#pragma unroll 1
            for (int i = lane; i < B * (NL > 0 ? NL : 1); i += 32) fold[i] = 0;
        __syncwarp();

        // ---- labels of the best hypothesis (src/duplex.rs:638-649) -----------------------------------------------
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
#pragma unroll 1
                    for (int k = len - 1; k >= 0; --k) {
                        p.labels[off + k] = uint8_t(S.nlab[node] + 1);
                        node = S.parent[node];
                    }
                }
            }
            p.offsets[n] = off;
            p.lens[n] = len;
            p.status[n] = status;
        }
        __syncwarp();
    }
}

static bool plan_duplex(int B, int C, DuplexLayout* L) {
    if (B > DX_MAX_BEAM || B < 1 || C < 1 || C > 256) return false;
    const int NL = C - 1 > 0 ? C - 1 : 1;
    auto up = [](long long v, int a) { return (v + a - 1) / a * a; };
    L->beam_stride = int(up(8LL * 4 * B, 16));
    long long o = 2LL * L->beam_stride;
    L->off_newcnt = int(o);  o = up(o + 4LL * B, 16);
    L->off_fold = int(o);    o = up(o + (long long)B * NL, 16);
    const long long nc = (long long)B * C;
    L->off_key = int(o);     o = up(o + 4 * nc, 16);
    L->off_node = int(o);    o = up(o + 4 * nc, 16);
    L->off_lab = int(o);     o = up(o + 4 * nc, 16);
    L->off_gap = int(o);     o = up(o + 4 * nc, 16);
    L->off_p2 = int(o);      o = up(o + 4 * nc, 16);
    L->off_sel = int(o);     o = up(o + 4LL * B, 16);
    L->off_row = int(o);     o = up(o + 4LL * C, 16);
    L->off_root = int(o);    o = up(o + 4LL * NL, 16);
    L->off_newlist = int(o); o = up(o + 4LL * B * NL, 16);
    L->warp_bytes = int(up(o, 128));
    L->cta_fixed = 512;  // logf tables (2 x 16 doubles) + expf table (32 x 8 bytes)
    return L->cta_fixed + (long long)L->warp_bytes * DX_WARPS <= 200 * 1024;
}

bool duplex_shape_supported(int beam_size, int C) {
    DuplexLayout L;
    return plan_duplex(beam_size, C, &L);
}

static int dx_default_nodes_per_row(int C, int B) {
    const long long bound = (long long)(C - 1) * B;
    long long d = 2LL * B;
    if (d < 12) d = 12;
    if (d > bound) d = bound;
    if (d < 1) d = 1;
    return int(d);
}

int64_t duplex_scratch_bytes(int sm_count, int64_t N, int T1, int T2, int C, int beam_size, int W, int nodes_per_row,
                             int64_t budget_bytes, int64_t* nodes_per_slot, int* n_slots, int crf, int64_t* n_rings) {
    if (nodes_per_row <= 0) nodes_per_row = dx_default_nodes_per_row(C, beam_size);
    const long long bound = (long long)(C - 1) * beam_size;
    if (nodes_per_row > bound) nodes_per_row = int(bound > 0 ? bound : 1);
    const long long cap = (long long)nodes_per_row * (T1 > 0 ? T1 : 1);
    // narrow beams run duplex_packed.cu: its pairs share a small pool of rings, so a slot is ~10x smaller and the
    // register file, not memory, sets the number of pairs in flight
    const bool packed = duplex_packed_shape(beam_size, C, crf);
    const long long rings = packed ? duplex_packed_rings(cap, nodes_per_row, bound > 0 ? bound : 1) : cap;
    if (n_rings) *n_rings = rings;
    const long long per_slot = packed ? (long long)duplex_packed_slot_bytes(cap, rings, T2, C, W)
                                      : (long long)dx_slot_bytes(cap, T2, C, W);
    int per_sm = packed ? duplex_packed_slots_per_sm(-1) : 24;  // latency-bound, sequential per pair: throughput follows the resident warps
    static const int per_sm_knob = env_knob("FCTC_DUPLEX_WARPS_PER_SM");  // tuning knob
    if (per_sm_knob > 0) per_sm = per_sm_knob;
    long long slots = (long long)sm_count * per_sm;
    if (slots > N) slots = N;
    if (slots < 1) slots = 1;
    slots = (slots + DX_WARPS - 1) / DX_WARPS * DX_WARPS;
    if (budget_bytes > 0) {  // fewer pairs in flight rather than more memory
        long long fit = (budget_bytes - 256) / per_slot;
        fit = fit / DX_WARPS * DX_WARPS;
        if (fit < DX_WARPS) fit = DX_WARPS;
        if (slots > fit) slots = fit;
    }
    // every pair of a batch takes about the same time, so the pairs run in rounds of `slots`: even the rounds out
    // (16,384 pairs over 9,472 slots would run 9,472 + 6,912; two rounds of 8,192 keep every warp as busy)
    if (N > slots && !packed) {
        const long long rounds = (N + slots - 1) / slots;
        long long even = (N + rounds - 1) / rounds;
        even = (even + DX_WARPS - 1) / DX_WARPS * DX_WARPS;
        if (even < slots) slots = even;
    }
    if (nodes_per_slot) *nodes_per_slot = cap;
    if (n_slots) *n_slots = int(slots);
    return 256 + slots * per_slot;
}

cudaError_t launch_duplex(const DuplexParams& p, cudaStream_t stream) {
    if (p.N <= 0) return cudaSuccess;
    if (duplex_packed_eligible(p)) return launch_duplex_packed(p, stream);  // narrow beams: four pairs per warp
    DuplexLayout L;
    if (!plan_duplex(p.beam_size, p.C, &L)) return cudaErrorInvalidConfiguration;
    const int smem = L.cta_fixed + L.warp_bytes * DX_WARPS;
    const bool crf = p.S > 0;
    cudaError_t e = crf ? cudaFuncSetAttribute(duplex_search_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                        : cudaFuncSetAttribute(duplex_search_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.ticket, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.cursor, 0, sizeof(unsigned long long), stream);  // the pool fills from entry 0 on every call
    if (e != cudaSuccess) return e;
    if (crf) duplex_search_kernel<true><<<p.n_slots / DX_WARPS, DX_THREADS, smem, stream>>>(p, L);
    else duplex_search_kernel<false><<<p.n_slots / DX_WARPS, DX_THREADS, smem, stream>>>(p, L);
    return cudaGetLastError();
}

}  // namespace fctc

// =============================================================================================
// C ABI (include/fctc.h)
// =============================================================================================
namespace {

int dx_check(fctc_handle* h, int64_t n_pairs, int64_t T1, int64_t T2, int32_t C, int32_t beam, float thr) {
    if (!h) return FCTC_E_ARGUMENT;
    if (n_pairs < 0 || T1 < 0 || T2 < 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "negative size");
    if (C <= 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "Empty alphabet given");
    if (C > 256) return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "alphabets above 256 labels are not supported");
    if (T1 == 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "network_output_1 has no rows (the reference aborts on envelope[(0,1)])");
    if (T1 >= (1 << 30) || T2 >= (1 << 30) || T2 * int64_t(C) >= (int64_t(1) << 31))
        return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "network outputs above 2^30 rows are not supported");
    if (int rc = fctc::check_threshold_call(h, C, beam, thr)) return rc;  // src/lib.rs:433-443
    if (!fctc::duplex_shape_supported(beam, C))
        return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "duplex search supports beam_size <= 32 in this release");
    return 0;
}

// What the trie scratch of a host-memory call may take: three quarters of the device memory that is free
// once the handle's own scratch buffer (kept from the previous call, and reused) is counted as free.
int64_t dx_free_budget(const fctc_handle* h) {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return int64_t(8) << 30;
    return int64_t((free_b + fctc::handle_device_buffer_bytes(h, 8)) / 4 * 3);
}

}  // namespace

extern "C" int64_t fctc_duplex_scratch_bytes(const fctc_handle* h, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                             int32_t n_labels, int32_t beam_size, int32_t max_window,
                                             int32_t nodes_per_row) {
    if (!h || n_labels <= 0 || beam_size <= 0 || n_rows_1 <= 0 || n_rows_2 < 0) return 0;
    const int W = int((max_window > 0 && max_window < n_rows_2 ? max_window : n_rows_2) + 1);
    // recommended: a pool per resident warp, capped at 96 GiB (smaller buffers run fewer pairs in flight)
    return fctc::duplex_scratch_bytes(fctc::handle_sm_count(h), n_pairs, int(n_rows_1), int(n_rows_2), n_labels, beam_size, W,
                                      nodes_per_row, int64_t(96) << 30, nullptr, nullptr, 0, nullptr);
}

extern "C" int64_t fctc_crf_duplex_scratch_bytes(const fctc_handle* h, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                                 int32_t n_labels, int32_t beam_size, int32_t max_window,
                                                 int32_t nodes_per_row) {
    if (!h || n_labels <= 0 || beam_size <= 0 || n_rows_1 <= 0 || n_rows_2 < 0) return 0;
    const int W = int((max_window > 0 && max_window < n_rows_2 ? max_window : n_rows_2) + 1);
    return fctc::duplex_scratch_bytes(fctc::handle_sm_count(h), n_pairs, int(n_rows_1), int(n_rows_2), n_labels, beam_size, W,
                                      nodes_per_row, int64_t(96) << 30, nullptr, nullptr, 1, nullptr);
}

static int dx_batch_device_any(fctc_handle* h, const void* post1_dev, const void* post2_dev, int in_f16,
                               int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                               int32_t n_labels, const int64_t* envelope_dev, int32_t max_window,
                               int32_t beam_size, float beam_cut_threshold,
                               int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels_dev,
                               int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                               int32_t* status_dev, unsigned long long* pool_used_dev,
                               void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row,
                               void* stream) {
    if (int rc = dx_check(h, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (!post1_dev || !post2_dev || !labels_dev || !offsets_dev || !lens_dev || !status_dev || !pool_used_dev || !scratch_dev)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (reinterpret_cast<uintptr_t>(scratch_dev) & 255u)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "scratch_dev must be 256-byte aligned");
    if (exp_mode < 0 || exp_mode > 2) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "exp_mode must be 0 (exact exp), 1 (fastexp as shipped) or 2 (CUDA expf/log1pf)");
    cudaError_t e = cudaSetDevice(fctc::handle_device(h));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    fctc::DuplexParams p{};
    p.W = int((max_window > 0 && max_window < n_rows_2 ? max_window : n_rows_2) + 1);
    // use as many pair slots as the caller's scratch holds
    const int64_t need = fctc::duplex_scratch_bytes(fctc::handle_sm_count(h), n_pairs, int(n_rows_1), int(n_rows_2), n_labels,
                                                    beam_size, p.W, nodes_per_row, scratch_bytes, &p.nodes_per_slot, &p.n_slots, 0,
                                                    &p.n_rings);
    if (need > scratch_bytes) {
        char msg[200];
        snprintf(msg, sizeof msg, "duplex scratch too small: %lld bytes given, %lld needed for %d pair slots",
                 (long long)scratch_bytes, (long long)need, p.n_slots);
        return fctc::handle_fail(h, FCTC_E_CAPACITY, msg);
    }
    p.post1 = static_cast<const float*>(post1_dev); p.post2 = static_cast<const float*>(post2_dev); p.in_f16 = in_f16;
    p.env = reinterpret_cast<const long long*>(envelope_dev);
    p.N = n_pairs; p.T1 = int(n_rows_1); p.T2 = int(n_rows_2); p.C = n_labels; p.beam_size = beam_size;
    p.threshold = beam_cut_threshold; p.collapse = collapse_repeats; p.exp_mode = exp_mode;
    p.labels = labels_dev; p.pool_cap = pool_capacity; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.cursor = pool_used_dev; p.scratch = scratch_dev; p.ticket = static_cast<unsigned long long*>(scratch_dev);
    e = fctc::launch_duplex(p, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    return 0;
}

extern "C" int fctc_beam_search_duplex_batch_device(fctc_handle* h, const float* post1_dev, const float* post2_dev,
                                                    int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                                    int32_t n_labels, const int64_t* envelope_dev, int32_t max_window,
                                                    int32_t beam_size, float beam_cut_threshold,
                                                    int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels_dev,
                                                    int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                                    int32_t* status_dev, unsigned long long* pool_used_dev,
                                                    void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row,
                                                    void* stream) {
    return dx_batch_device_any(h, post1_dev, post2_dev, 0, n_pairs, n_rows_1, n_rows_2, n_labels, envelope_dev, max_window,
                               beam_size, beam_cut_threshold, collapse_repeats, exp_mode, labels_dev, pool_capacity,
                               offsets_dev, lens_dev, status_dev, pool_used_dev, scratch_dev, scratch_bytes, nodes_per_row,
                               stream);
}

extern "C" int fctc_beam_search_duplex_batch_device_f16(fctc_handle* h, const void* post1_dev_f16,
                                                        const void* post2_dev_f16, int64_t n_pairs, int64_t n_rows_1,
                                                        int64_t n_rows_2, int32_t n_labels, const int64_t* envelope_dev,
                                                        int32_t max_window, int32_t beam_size, float beam_cut_threshold,
                                                        int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels_dev,
                                                        int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                                        int32_t* status_dev, unsigned long long* pool_used_dev,
                                                        void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row,
                                                        void* stream) {
    return dx_batch_device_any(h, post1_dev_f16, post2_dev_f16, 1, n_pairs, n_rows_1, n_rows_2, n_labels, envelope_dev,
                               max_window, beam_size, beam_cut_threshold, collapse_repeats, exp_mode, labels_dev,
                               pool_capacity, offsets_dev, lens_dev, status_dev, pool_used_dev, scratch_dev, scratch_bytes,
                               nodes_per_row, stream);
}

namespace {

#define DX_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e__)); \
    } while (0)

// A chunk of pairs goes through three steps, so that the next chunk's inputs can be copied in while this chunk's
// kernel runs: stage (inputs to one of two device sets, on the stream given), launch (kernel and the small result
// arrays, on the handle's stream, after the staging event), finish (wait, labels to the host pool at *used).
struct DxInputs {
    float* d1 = nullptr;
    float* d2 = nullptr;
    long long* de = nullptr;
};
struct DxChunk {
    int64_t cnt = 0, stride = 0;
    uint8_t* dl = nullptr;
    unsigned long long* hcur = nullptr;
};

int dx_stage(fctc_handle* h, int set, cudaStream_t cs, const void* post1, const void* post2, int in_f16, const int64_t* env,
             int64_t cnt, int64_t T1, int64_t T2, int32_t C, DxInputs* in) {
    const size_t eb = in_f16 ? 2 : 4;
    const size_t b1 = size_t(cnt) * T1 * C * eb, b2 = size_t(cnt) * T2 * C * eb, be = size_t(cnt) * T1 * 16;
    const int base = set ? 13 : 0;  // persistent buffers 0-2 and 13-15 of the handle
    in->d1 = static_cast<float*>(fctc::handle_device_buffer(h, base, std::max<size_t>(b1, 16)));
    in->d2 = static_cast<float*>(fctc::handle_device_buffer(h, base + 1, std::max<size_t>(b2, 16)));
    in->de = env ? static_cast<long long*>(fctc::handle_device_buffer(h, base + 2, be)) : nullptr;
    if (!in->d1 || !in->d2 || (env && !in->de)) return FCTC_E_NOMEM;
    if (b1) DX_CUDA(cudaMemcpyAsync(in->d1, post1, b1, cudaMemcpyHostToDevice, cs));
    if (b2) DX_CUDA(cudaMemcpyAsync(in->d2, post2, b2, cudaMemcpyHostToDevice, cs));
    if (env) DX_CUDA(cudaMemcpyAsync(in->de, env, be, cudaMemcpyHostToDevice, cs));
    return 0;
}

int dx_launch(fctc_handle* h, const DxInputs& in, cudaEvent_t staged, int in_f16, int64_t cnt, int64_t T1, int64_t T2, int32_t C,
              int W, int32_t beam, float thr, int32_t collapse, int32_t exp_mode, int32_t nodes_per_row, int64_t* offsets,
              int32_t* lens, int32_t* status, DxChunk* ck) {
    cudaStream_t st = fctc::handle_stream(h);
    const int64_t stride = (T1 + 31) & ~int64_t(31);
    uint8_t* dl = static_cast<uint8_t*>(fctc::handle_device_buffer(h, 3, size_t(cnt * stride)));
    int64_t* doff = static_cast<int64_t*>(fctc::handle_device_buffer(h, 4, size_t(cnt) * 8));
    int32_t* dlen = static_cast<int32_t*>(fctc::handle_device_buffer(h, 5, size_t(cnt) * 4));
    int32_t* dst = static_cast<int32_t*>(fctc::handle_device_buffer(h, 6, size_t(cnt) * 4));
    unsigned long long* dcur = static_cast<unsigned long long*>(fctc::handle_device_buffer(h, 7, 256));
    unsigned long long* hcur = static_cast<unsigned long long*>(fctc::handle_pinned_buffer(h, 0, 64));
    if (!dl || !doff || !dlen || !dst || !dcur || !hcur) return FCTC_E_NOMEM;
    fctc::DuplexParams p{};
    p.W = W;
    const int64_t need = fctc::duplex_scratch_bytes(fctc::handle_sm_count(h), cnt, int(T1), int(T2), C, beam, W, nodes_per_row,
                                                    dx_free_budget(h), &p.nodes_per_slot, &p.n_slots, 0, &p.n_rings);
    void* scratch = fctc::handle_device_buffer(h, 8, size_t(need));
    if (!scratch) return FCTC_E_NOMEM;
    if (staged) DX_CUDA(cudaStreamWaitEvent(st, staged, 0));
    DX_CUDA(cudaMemsetAsync(dcur, 0, 8, st));
    p.post1 = in.d1; p.post2 = in.d2; p.in_f16 = in_f16; p.env = in.de; p.N = cnt; p.T1 = int(T1); p.T2 = int(T2); p.C = C;
    p.beam_size = beam; p.threshold = thr; p.collapse = collapse; p.exp_mode = exp_mode; p.labels = dl; p.pool_cap = cnt * stride;
    p.offsets = doff; p.lens = dlen; p.status = dst; p.cursor = dcur; p.scratch = scratch;
    p.ticket = static_cast<unsigned long long*>(scratch);
    DX_CUDA(fctc::launch_duplex(p, st));
    DX_CUDA(cudaMemcpyAsync(hcur, dcur, 8, cudaMemcpyDeviceToHost, st));
    DX_CUDA(cudaMemcpyAsync(offsets, doff, size_t(cnt) * 8, cudaMemcpyDeviceToHost, st));
    DX_CUDA(cudaMemcpyAsync(lens, dlen, size_t(cnt) * 4, cudaMemcpyDeviceToHost, st));
    DX_CUDA(cudaMemcpyAsync(status, dst, size_t(cnt) * 4, cudaMemcpyDeviceToHost, st));
    ck->cnt = cnt; ck->stride = stride; ck->dl = dl; ck->hcur = hcur;
    return 0;
}

int dx_finish(fctc_handle* h, const DxChunk& ck, uint8_t* labels, int64_t pool_cap, int64_t* offsets, int64_t* used) {
    cudaStream_t st = fctc::handle_stream(h);
    DX_CUDA(cudaStreamSynchronize(st));
    const int64_t got = int64_t(*ck.hcur);
    if (*used + got > pool_cap) return fctc::handle_fail(h, FCTC_E_CAPACITY, "result pool too small (n_pairs*round_up(n_rows_1,32) always suffices)");
    if (got) DX_CUDA(cudaMemcpyAsync(labels + *used, ck.dl, size_t(got), cudaMemcpyDeviceToHost, st));
    DX_CUDA(cudaStreamSynchronize(st));
    for (int64_t i = 0; i < ck.cnt; ++i) offsets[i] += *used;
    *used += got;
    return 0;
}

// One synchronous pass over `cnt` pairs of host arrays.  Results are appended to the host pool at *used.
int dx_host_chunk(fctc_handle* h, const void* post1, const void* post2, int in_f16, const int64_t* env, int64_t cnt, int64_t T1,
                  int64_t T2, int32_t C, int W, int32_t beam, float thr, int32_t collapse, int32_t exp_mode,
                  int32_t nodes_per_row, uint8_t* labels, int64_t pool_cap, int64_t* offsets, int32_t* lens,
                  int32_t* status, int64_t* used) {
    DxInputs in;
    DxChunk ck;
    if (int rc = dx_stage(h, 0, fctc::handle_stream(h), post1, post2, in_f16, env, cnt, T1, T2, C, &in)) return rc;
    if (int rc = dx_launch(h, in, nullptr, in_f16, cnt, T1, T2, C, W, beam, thr, collapse, exp_mode, nodes_per_row, offsets, lens,
                           status, &ck))
        return rc;
    return dx_finish(h, ck, labels, pool_cap, offsets, used);
}

}  // namespace

static int dx_batch_host_any(fctc_handle* h, const void* post1_any, const void* post2_any, int in_f16,
                             int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels,
                             const int64_t* envelope_host, int32_t beam_size, float beam_cut_threshold,
                             int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels,
                             int64_t pool_capacity, int64_t* offsets, int32_t* lens, int32_t* status,
                             int64_t* pool_used) {
    const unsigned char* post1_host = static_cast<const unsigned char*>(post1_any);
    const unsigned char* post2_host = static_cast<const unsigned char*>(post2_any);
    const int64_t eb = in_f16 ? 2 : 4;
    if (int rc = dx_check(h, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (!post1_host || (!post2_host && n_rows_2 > 0) || !labels || !offsets || !lens || !status)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null host pointer");
    if (exp_mode < 0 || exp_mode > 2) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "exp_mode must be 0 (exact exp), 1 (fastexp as shipped) or 2 (CUDA expf/log1pf)");
    cudaError_t e = cudaSetDevice(fctc::handle_device(h));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    if (pool_used) *pool_used = 0;
    if (n_pairs == 0) return 0;
    // ring size of the per-node vectors: the widest window the envelope asks for (clamped as the
    // kernel clamps), or all of network 2 without an envelope
    auto widest_window = [&]() -> int64_t {
    int64_t wmax = n_rows_2;
    if (envelope_host) {
        // (16 bytes per network-1 row: for a large batch the scan is split over a few host threads — one thread took
        // as long as the copy-in of the whole chunk)
        const int64_t total = n_pairs * n_rows_1;
        auto scan = [&](int64_t i0, int64_t i1) {
            int64_t w = 1;
            for (int64_t i = i0; i < i1; ++i) {
                const uint64_t lo = uint64_t(envelope_host[2 * i]), hi_u = uint64_t(envelope_host[2 * i + 1]);
                const int64_t hi = hi_u > uint64_t(n_rows_2) ? n_rows_2 : int64_t(hi_u);
                if (lo < uint64_t(hi) && hi - int64_t(lo) > w) w = hi - int64_t(lo);
            }
            return w;
        };
        int n_threads = int(std::min<int64_t>(8, std::max<int64_t>(1, total >> 20)));
        n_threads = std::min<int>(n_threads, std::max(1u, std::thread::hardware_concurrency()));
        if (n_threads <= 1) {
            wmax = scan(0, total);
        } else {
            std::vector<int64_t> part(n_threads, 1);
            std::vector<std::thread> pool;
            for (int k = 0; k < n_threads; ++k)
                pool.emplace_back([&, k]() { part[k] = scan(total * k / n_threads, total * (k + 1) / n_threads); });
            for (std::thread& th : pool) th.join();
            wmax = *std::max_element(part.begin(), part.end());
        }
    }
    return wmax;
    };
    int W = 0;  // found on the host while the first chunk is being copied in
    const int64_t pair_bytes = (n_rows_1 + n_rows_2) * int64_t(n_labels) * eb;
    // a chunk is a whole number of rounds of the pairs in flight (a partial round costs a full one), as many as 2 GiB
    // of posteriors hold, at least one
    const int64_t in_flight = int64_t(fctc::handle_sm_count(h)) *
                              (fctc::duplex_packed_shape(beam_size, n_labels, 0) ? fctc::duplex_packed_slots_per_sm(exp_mode) : 24);
    int64_t chunk = std::max<int64_t>((int64_t(2) << 30) / std::max<int64_t>(pair_bytes, 1), 1);
    if (chunk > in_flight) chunk = chunk / in_flight * in_flight;
    static const int chunk_knob = fctc::env_knob("FCTC_DUPLEX_CHUNK_PAIRS");  // tuning / test knob
    if (chunk_knob > 0) chunk = chunk_knob;
    chunk = std::max<int64_t>(1, std::min<int64_t>(chunk, n_pairs));
    if (chunk < n_pairs) {  // chunks of equal size: the same kernel time in total, and a shorter first copy to wait for
        const int64_t n_chunks = (n_pairs + chunk - 1) / chunk;
        const int64_t even = ((n_pairs + n_chunks - 1) / n_chunks + 3) & ~int64_t(3);
        chunk = std::min(chunk, even);
    }
    int64_t used = 0;
    {
        // chunk k+1 is copied into the other input set on a second stream while chunk k's kernel runs; the kernel is
        // queued before that copy so that a caller with pageable memory (whose copy blocks the host) loses nothing
        cudaStream_t cs = fctc::handle_aux_stream(h);
        cudaEvent_t staged[2] = {nullptr, nullptr};
        for (cudaEvent_t& e : staged)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess)
                return fctc::handle_fail(h, FCTC_E_CUDA, "cudaEventCreate failed");
        auto stage = [&](int64_t first, int set, DxInputs* in) -> int {
            const int64_t cnt = std::min(chunk, n_pairs - first);
            if (int rc = dx_stage(h, set, cs, post1_host + first * n_rows_1 * n_labels * eb,
                                  post2_host + first * n_rows_2 * n_labels * eb, in_f16,
                                  envelope_host ? envelope_host + first * n_rows_1 * 2 : nullptr, cnt, n_rows_1, n_rows_2,
                                  n_labels, in))
                return rc;
            if (cudaEventRecord(staged[set], cs) != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, "cudaEventRecord failed");
            return 0;
        };
        DxInputs in[2];
        if (chunk < n_pairs) {  // both input sets exist before the scratch takes its share of what is free
            const size_t b1 = size_t(chunk) * n_rows_1 * n_labels * eb, b2 = size_t(chunk) * n_rows_2 * n_labels * eb;
            for (int base : {0, 13})
                if (!fctc::handle_device_buffer(h, base, std::max<size_t>(b1, 16)) ||
                    !fctc::handle_device_buffer(h, base + 1, std::max<size_t>(b2, 16)) ||
                    (envelope_host && !fctc::handle_device_buffer(h, base + 2, size_t(chunk) * n_rows_1 * 16))) {
                    for (cudaEvent_t e : staged) cudaEventDestroy(e);
                    return FCTC_E_NOMEM;
                }
        }
        int rc = stage(0, 0, &in[0]);
        W = int(widest_window() + 1);
        int set = 0;
        for (int64_t first = 0; rc == 0 && first < n_pairs; first += chunk, set ^= 1) {
            const int64_t cnt = std::min(chunk, n_pairs - first);
            DxChunk ck;
            rc = dx_launch(h, in[set], staged[set], in_f16, cnt, n_rows_1, n_rows_2, n_labels, W, beam_size, beam_cut_threshold,
                           collapse_repeats, exp_mode, 0, offsets + first, lens + first, status + first, &ck);
            if (rc == 0 && first + chunk < n_pairs) rc = stage(first + chunk, set ^ 1, &in[set ^ 1]);
            if (rc == 0) rc = dx_finish(h, ck, labels, pool_capacity, offsets + first, &used);
        }
        cudaStreamSynchronize(cs);
        for (cudaEvent_t e : staged) cudaEventDestroy(e);
        if (rc) return rc;
    }
    // pairs whose trie (or ring pool) outgrew the default plan: gathered and decoded again in one pass per group of at
    // most 256 MiB of posteriors, at the hard bound (one ring per node: cannot overflow)
    const int32_t bound = std::max<int32_t>(1, (n_labels - 1) * beam_size);
    std::vector<int64_t> again;
    for (int64_t n = 0; n < n_pairs; ++n)
        if (status[n] == FCTC_NODE_POOL_OVERFLOW) again.push_back(n);
    if (!again.empty()) {
        const int64_t b1 = n_rows_1 * int64_t(n_labels) * eb, b2 = n_rows_2 * int64_t(n_labels) * eb, be = n_rows_1 * 16;
        const int64_t group = std::max<int64_t>(1, (int64_t(256) << 20) / std::max<int64_t>(b1 + b2, 1));
        std::vector<unsigned char> g1, g2;
        std::vector<int64_t> ge, off2;
        std::vector<int32_t> len2, st2;
        for (size_t g0 = 0; g0 < again.size(); g0 += size_t(group)) {
            const int64_t k = std::min<int64_t>(group, int64_t(again.size() - g0));
            g1.resize(size_t(std::max<int64_t>(k * b1, 16)));
            g2.resize(size_t(std::max<int64_t>(k * b2, 16)));
            if (envelope_host) ge.resize(size_t(k * n_rows_1 * 2));
            for (int64_t i = 0; i < k; ++i) {
                const int64_t n = again[g0 + size_t(i)];
                memcpy(g1.data() + i * b1, post1_host + n * b1, size_t(b1));
                if (b2) memcpy(g2.data() + i * b2, post2_host + n * b2, size_t(b2));
                if (envelope_host) memcpy(ge.data() + i * n_rows_1 * 2, envelope_host + n * n_rows_1 * 2, size_t(be));
            }
            off2.assign(size_t(k), 0); len2.assign(size_t(k), 0); st2.assign(size_t(k), 0);
            if (int rc = dx_host_chunk(h, g1.data(), g2.data(), in_f16, envelope_host ? ge.data() : nullptr, k, n_rows_1, n_rows_2,
                                       n_labels, W, beam_size, beam_cut_threshold, collapse_repeats, exp_mode, bound, labels,
                                       pool_capacity, off2.data(), len2.data(), st2.data(), &used))
                return rc;
            for (int64_t i = 0; i < k; ++i) {
                const int64_t n = again[g0 + size_t(i)];
                offsets[n] = off2[size_t(i)]; lens[n] = len2[size_t(i)]; status[n] = st2[size_t(i)];
            }
        }
    }
    if (pool_used) *pool_used = used;
    return 0;
}

extern "C" int fctc_beam_search_duplex_batch(fctc_handle* h, const float* post1_host, const float* post2_host,
                                             int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels,
                                             const int64_t* envelope_host, int32_t beam_size, float beam_cut_threshold,
                                             int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels,
                                             int64_t pool_capacity, int64_t* offsets, int32_t* lens, int32_t* status,
                                             int64_t* pool_used) {
    return dx_batch_host_any(h, post1_host, post2_host, 0, n_pairs, n_rows_1, n_rows_2, n_labels, envelope_host, beam_size,
                             beam_cut_threshold, collapse_repeats, exp_mode, labels, pool_capacity, offsets, lens, status,
                             pool_used);
}

extern "C" int fctc_beam_search_duplex_batch_f16(fctc_handle* h, const void* post1_host_f16, const void* post2_host_f16,
                                                 int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels,
                                                 const int64_t* envelope_host, int32_t beam_size,
                                                 float beam_cut_threshold, int32_t collapse_repeats, int32_t exp_mode,
                                                 uint8_t* labels, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                                                 int32_t* status, int64_t* pool_used) {
    return dx_batch_host_any(h, post1_host_f16, post2_host_f16, 1, n_pairs, n_rows_1, n_rows_2, n_labels, envelope_host,
                             beam_size, beam_cut_threshold, collapse_repeats, exp_mode, labels, pool_capacity, offsets, lens,
                             status, pool_used);
}

extern "C" int fctc_beam_search_duplex(fctc_handle* h, const float* post1_host, int64_t n_rows_1, int64_t stride_row_1,
                                       int64_t stride_label_1, const float* post2_host, int64_t n_rows_2,
                                       int64_t stride_row_2, int64_t stride_label_2, int32_t n_labels,
                                       const uint64_t* envelope, int64_t env_stride_row, int64_t env_stride_col,
                                       int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                                       int32_t exp_mode, uint8_t* labels, int64_t capacity, int64_t* out_len) {
    if (int rc = dx_check(h, 1, n_rows_1, n_rows_2, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (!post1_host || (!post2_host && n_rows_2 > 0) || !out_len) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null pointer");
    const int64_t stride = std::max<int64_t>((n_rows_1 + 31) & ~int64_t(31), 32);
    const size_t o2 = (size_t(n_rows_1) * n_labels * 4 + 255) & ~size_t(255);
    const size_t oe = (o2 + size_t(n_rows_2) * n_labels * 4 + 255) & ~size_t(255);
    const size_t ol = (oe + size_t(n_rows_1) * 16 + 255) & ~size_t(255);
    const size_t om = ol + size_t(stride);
    unsigned char* pin = static_cast<unsigned char*>(fctc::handle_pinned_buffer(h, 1, om + 64));
    if (!pin) return FCTC_E_NOMEM;
    float* p1 = reinterpret_cast<float*>(pin);
    float* p2 = reinterpret_cast<float*>(pin + o2);
    int64_t* env = reinterpret_cast<int64_t*>(pin + oe);
    for (int64_t t = 0; t < n_rows_1; ++t)
        for (int32_t c = 0; c < n_labels; ++c) p1[t * n_labels + c] = post1_host[t * stride_row_1 + c * stride_label_1];
    for (int64_t t = 0; t < n_rows_2; ++t)
        for (int32_t c = 0; c < n_labels; ++c) p2[t * n_labels + c] = post2_host[t * stride_row_2 + c * stride_label_2];
    if (envelope)
        for (int64_t t = 0; t < n_rows_1; ++t) {
            env[2 * t] = int64_t(envelope[t * env_stride_row]);
            env[2 * t + 1] = int64_t(envelope[t * env_stride_row + env_stride_col]);
        }
    int64_t* off1 = reinterpret_cast<int64_t*>(pin + om);
    int32_t* len1 = reinterpret_cast<int32_t*>(pin + om + 8);
    int32_t* st1 = reinterpret_cast<int32_t*>(pin + om + 16);
    int64_t used = 0;
    int rc = fctc_beam_search_duplex_batch(h, p1, p2, 1, n_rows_1, n_rows_2, n_labels, envelope ? env : nullptr, beam_size,
                                           beam_cut_threshold, collapse_repeats, exp_mode, pin + ol, stride, off1, len1, st1,
                                           &used);
    if (rc < 0) return rc;
    *out_len = *len1;
    const int64_t m = std::min<int64_t>(*len1, capacity);
    for (int64_t i = 0; i < m; ++i)
        if (labels) labels[i] = pin[ol + *off1 + i];
    return *st1;
}

// ---- crf variant: replaces duplex::crf_beam_search (src/duplex.rs:652-834) as called from
//      src/lib.rs:563-574.  The reference holds no test vector for it (parity unpinned): the check
//      here is against this repository's CPU restatement. ---------------------------------------------
extern "C" int fctc_crf_beam_search_duplex_batch_device_in(
    fctc_handle* h, const void* post1_dev, const float* init1_dev, int32_t n_init1, const void* post2_dev,
    const float* init2_dev, int32_t n_init2, int32_t input_kind, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
    int32_t n_state, int32_t n_labels, const int64_t* envelope_dev, int32_t max_window, int32_t beam_size,
    float beam_cut_threshold, int32_t exp_mode, uint8_t* labels_dev, int64_t pool_capacity, int64_t* offsets_dev,
    int32_t* lens_dev, int32_t* status_dev, unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
    int32_t nodes_per_row, void* stream) {
    if (!h) return FCTC_E_ARGUMENT;
    if (input_kind != FCTC_IN_F32 && input_kind != FCTC_IN_F16)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "the duplex searches take FCTC_IN_F32 or FCTC_IN_F16 posteriors");
    if (int rc = dx_check(h, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (n_state <= 0 || n_init1 <= 0 || n_init2 <= 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "n_state and the init_state vectors must be non-empty");
    if (n_rows_2 * int64_t(n_labels) * n_state >= (int64_t(1) << 40)) return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "network output too large");
    if (!post1_dev || !post2_dev || !init1_dev || !init2_dev || !labels_dev || !offsets_dev || !lens_dev || !status_dev ||
        !pool_used_dev || !scratch_dev)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (reinterpret_cast<uintptr_t>(scratch_dev) & 255u) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "scratch_dev must be 256-byte aligned");
    if (exp_mode < 0 || exp_mode > 2) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "exp_mode must be 0 (exact exp), 1 (fastexp as shipped) or 2 (CUDA expf/log1pf)");
    cudaError_t e = cudaSetDevice(fctc::handle_device(h));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    fctc::DuplexParams p{};
    p.W = int((max_window > 0 && max_window < n_rows_2 ? max_window : n_rows_2) + 1);
    const int64_t need = fctc::duplex_scratch_bytes(fctc::handle_sm_count(h), n_pairs, int(n_rows_1), int(n_rows_2), n_labels,
                                                    beam_size, p.W, nodes_per_row, scratch_bytes, &p.nodes_per_slot, &p.n_slots, 1,
                                                    &p.n_rings);
    if (need > scratch_bytes) return fctc::handle_fail(h, FCTC_E_CAPACITY, "duplex scratch too small (see fctc_crf_duplex_scratch_bytes)");
    p.post1 = static_cast<const float*>(post1_dev); p.post2 = static_cast<const float*>(post2_dev);
    p.in_f16 = input_kind == FCTC_IN_F16;
    p.env = reinterpret_cast<const long long*>(envelope_dev);
    p.N = n_pairs; p.T1 = int(n_rows_1); p.T2 = int(n_rows_2); p.C = n_labels; p.beam_size = beam_size;
    p.threshold = beam_cut_threshold; p.collapse = 0; p.exp_mode = exp_mode;
    p.labels = labels_dev; p.pool_cap = pool_capacity; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.cursor = pool_used_dev; p.scratch = scratch_dev; p.ticket = static_cast<unsigned long long*>(scratch_dev);
    p.S = n_state; p.init1 = init1_dev; p.init2 = init2_dev; p.n_init1 = n_init1; p.n_init2 = n_init2;
    e = fctc::launch_duplex(p, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    return 0;
}

extern "C" int fctc_crf_beam_search_duplex_batch_device(
    fctc_handle* h, const float* post1_dev, const float* init1_dev, int32_t n_init1, const float* post2_dev,
    const float* init2_dev, int32_t n_init2, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_state,
    int32_t n_labels, const int64_t* envelope_dev, int32_t max_window, int32_t beam_size, float beam_cut_threshold,
    int32_t exp_mode, uint8_t* labels_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
    int32_t* status_dev, unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
    int32_t nodes_per_row, void* stream) {
    return fctc_crf_beam_search_duplex_batch_device_in(h, post1_dev, init1_dev, n_init1, post2_dev, init2_dev, n_init2, FCTC_IN_F32,
                                                       n_pairs, n_rows_1, n_rows_2, n_state, n_labels, envelope_dev, max_window,
                                                       beam_size, beam_cut_threshold, exp_mode, labels_dev, pool_capacity,
                                                       offsets_dev, lens_dev, status_dev, pool_used_dev, scratch_dev,
                                                       scratch_bytes, nodes_per_row, stream);
}

// One pair from host memory; arrays C-contiguous: post [T, S, C], init [n_init], envelope usize [T1, 2] or NULL.
extern "C" int fctc_crf_beam_search_duplex(fctc_handle* h, const float* post1_host, int64_t n_rows_1,
                                           const float* init1_host, int32_t n_init1, const float* post2_host,
                                           int64_t n_rows_2, const float* init2_host, int32_t n_init2, int32_t n_state,
                                           int32_t n_labels, const uint64_t* envelope, int32_t beam_size,
                                           float beam_cut_threshold, int32_t exp_mode, uint8_t* labels, int64_t capacity,
                                           int64_t* out_len) {
    if (int rc = dx_check(h, 1, n_rows_1, n_rows_2, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (!post1_host || !post2_host || !init1_host || !init2_host || !out_len) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null pointer");
    if (n_state <= 0 || n_init1 <= 0 || n_init2 <= 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "n_state and the init_state vectors must be non-empty");
    cudaError_t e = cudaSetDevice(fctc::handle_device(h));
    if (e != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e));
    cudaStream_t st = fctc::handle_stream(h);
    const int64_t stride = std::max<int64_t>((n_rows_1 + 31) & ~int64_t(31), 32);
    const size_t b1 = size_t(n_rows_1) * n_state * n_labels * 4, b2 = size_t(n_rows_2) * n_state * n_labels * 4;
    float* d1 = static_cast<float*>(fctc::handle_device_buffer(h, 0, b1));
    float* d2 = static_cast<float*>(fctc::handle_device_buffer(h, 1, std::max<size_t>(b2, 16)));
    long long* de = static_cast<long long*>(fctc::handle_device_buffer(h, 2, size_t(n_rows_1) * 16));
    uint8_t* dl = static_cast<uint8_t*>(fctc::handle_device_buffer(h, 3, size_t(stride)));
    int64_t* doff = static_cast<int64_t*>(fctc::handle_device_buffer(h, 4, 64));
    int32_t* dlen = static_cast<int32_t*>(fctc::handle_device_buffer(h, 5, 64));
    int32_t* dst = static_cast<int32_t*>(fctc::handle_device_buffer(h, 6, 64));
    unsigned long long* dcur = static_cast<unsigned long long*>(fctc::handle_device_buffer(h, 7, 256));
    float* di1 = static_cast<float*>(fctc::handle_device_buffer(h, 11, size_t(n_init1) * 4));
    float* di2 = static_cast<float*>(fctc::handle_device_buffer(h, 12, size_t(n_init2) * 4));
    unsigned char* pin = static_cast<unsigned char*>(fctc::handle_pinned_buffer(h, 3, size_t(stride) + size_t(n_rows_1) * 16 + 64));
    if (!d1 || !d2 || !de || !dl || !doff || !dlen || !dst || !dcur || !di1 || !di2 || !pin) return FCTC_E_NOMEM;
#define DXC(call)                                                                              \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e__)); \
    } while (0)
    int64_t wmax = n_rows_2;
    long long* henv = reinterpret_cast<long long*>(pin + 64);
    if (envelope) {
        wmax = 1;
        for (int64_t t = 0; t < n_rows_1; ++t) {
            henv[2 * t] = (long long)envelope[2 * t];
            henv[2 * t + 1] = (long long)envelope[2 * t + 1];
            const uint64_t lo = envelope[2 * t], hu = envelope[2 * t + 1];
            const int64_t hi = hu > uint64_t(n_rows_2) ? n_rows_2 : int64_t(hu);
            if (lo < uint64_t(hi) && hi - int64_t(lo) > wmax) wmax = hi - int64_t(lo);
        }
        DXC(cudaMemcpyAsync(de, henv, size_t(n_rows_1) * 16, cudaMemcpyHostToDevice, st));
    }
    DXC(cudaMemcpyAsync(d1, post1_host, b1, cudaMemcpyHostToDevice, st));
    if (b2) DXC(cudaMemcpyAsync(d2, post2_host, b2, cudaMemcpyHostToDevice, st));
    DXC(cudaMemcpyAsync(di1, init1_host, size_t(n_init1) * 4, cudaMemcpyHostToDevice, st));
    DXC(cudaMemcpyAsync(di2, init2_host, size_t(n_init2) * 4, cudaMemcpyHostToDevice, st));
    int32_t* hlen = reinterpret_cast<int32_t*>(pin);
    int32_t* hst = hlen + 1;
    int64_t* hoff = reinterpret_cast<int64_t*>(pin + 8);
    uint8_t* hl = pin + 64 + size_t(n_rows_1) * 16;
    const int32_t tries[2] = {0, std::max<int32_t>(1, (n_labels - 1) * beam_size)};
    for (int attempt = 0; attempt < 2; ++attempt) {
        int64_t nps = 0;
        int nsl = 0;
        const int64_t need = fctc::duplex_scratch_bytes(fctc::handle_sm_count(h), 1, int(n_rows_1), int(n_rows_2), n_labels, beam_size,
                                                        int(wmax + 1), tries[attempt], dx_free_budget(h), &nps, &nsl, 1, nullptr);
        void* scratch = fctc::handle_device_buffer(h, 8, size_t(need));
        if (!scratch) return FCTC_E_NOMEM;
        DXC(cudaMemsetAsync(dcur, 0, 8, st));
        int rc = fctc_crf_beam_search_duplex_batch_device(h, d1, di1, n_init1, d2, di2, n_init2, 1, n_rows_1, n_rows_2, n_state, n_labels,
                                                          envelope ? reinterpret_cast<const int64_t*>(de) : nullptr, int32_t(wmax),
                                                          beam_size, beam_cut_threshold, exp_mode, dl, stride, doff, dlen, dst, dcur,
                                                          scratch, need, tries[attempt], st);
        if (rc < 0) return rc;
        DXC(cudaMemcpyAsync(hlen, dlen, 4, cudaMemcpyDeviceToHost, st));
        DXC(cudaMemcpyAsync(hst, dst, 4, cudaMemcpyDeviceToHost, st));
        DXC(cudaMemcpyAsync(hoff, doff, 8, cudaMemcpyDeviceToHost, st));
        DXC(cudaMemcpyAsync(hl, dl, size_t(stride), cudaMemcpyDeviceToHost, st));
        DXC(cudaStreamSynchronize(st));
        if (*hst != FCTC_NODE_POOL_OVERFLOW) break;
    }
#undef DXC
    *out_len = *hlen;
    const int64_t m = std::min<int64_t>(*hlen, capacity);
    for (int64_t i = 0; i < m; ++i)
        if (labels) labels[i] = hl[*hoff + i];
    return *hst;
}

