// duplex_packed.cu — the 2-D ("duplex") beam search for narrow beams: four pairs per warp.
//
// Same search as duplex.cu (duplex::beam_search, src/duplex.rs:443-650, with LogSpace :7-80, SecondaryProbs
// :152-210, build_/extend_secondary_probs :212-249, :338-387, root_probs :389-409), same
// results bit for bit — another mapping of the work onto the warp.  With one pair per warp and one lane per beam
// entry, the reference's default beam of 5 keeps 5 of 32 lanes busy in the beam steps and, in the step that
// dominates — the forward recurrence that builds a SecondaryProbs vector over the envelope window for every
// newly proposed trie node, some 3 nodes x 65 cells per network-1 row — as many lanes as the row proposes nodes.
// Here, for beam_size <= 8:
//   * a warp carries four pairs, eight lanes each (lane = beam entry of its pair); the four pairs advance through
//     the network-1 rows in lockstep, every beam step (extension, own continuation, child proposals, scores,
//     ranking) runs for all four at once with 8-lane shuffles and reductions;
//   * the new nodes of all four pairs go into one job list; each job is taken by TWO lanes.  The recurrence
//         gap_i   = total_{i-1} (x) blank_i
//         label_i = p2[i][l]    (x) (label_{i-1} (+) parent_{i-1}.{gap | total})
//         total_i = label_i (+) gap_i
//     is two chains with one log-add each: the label chain needs nothing from the other, the gap/total chain needs
//     label_i.  The even lane runs the label chain, the odd lane follows one cell behind with gap and total (one
//     shuffle per cell hands label_i over) and stores the cell.  Sixteen nodes are in flight per pass and every
//     lane executes one log-add per cell, in the reference's operation order, so values are unchanged.
//   * a node's SecondaryProbs cells live in a pool of R rings indexed by node id mod R instead of one ring per node
//     id.  A ring is only ever read for a node whose depth is at least (lowest depth in the beam) - 1: the beam's
//     nodes, their parents (extension reads the parent's cells) and their children; beam depths never decrease, so
//     a node below that depth is dead for good and its ring can be handed to a new node.  The hand-over checks
//     exactly that (depth of the ring's previous owner); if it ever fails the pair is flagged
//     FCTC_NODE_POOL_OVERFLOW, and a call with nodes_per_row at the hard bound (what the host entry points retry
//     with) gets one ring per node, which cannot fail.  A pair then needs ~3 MB of scratch instead of ~27 MB
//     (T = 4000, band 65), so the number of pairs in flight is set by registers, not by memory.
// The crf variant and beams above 8 stay with duplex.cu.
#include <cuda_fp16.h>

#include <cstdlib>

#include "duplex_common.cuh"
#include "kernels.h"

#include "../../include/fctc.h"

namespace fctc {

namespace {

constexpr int DP_GROUP = 8;   // lanes per pair
constexpr int DP_PAIRS = 4;   // pairs per warp
constexpr int DP_SLOTS = 16;  // node-build jobs per pass (two lanes each)
constexpr int DP_FIXED = 512; // logf tables (2 x 16 doubles) + expf table (32 x 8 bytes)
// One-warp CTAs per SM the register allocation aims at, by flavour of LogSpace '+' (measured, 1000-row pairs): the exact
// flavour runs best with 128 registers and 16 CTAs, the cheap ones (max-plus, CUDA libm) with more, lighter CTAs.
#ifndef DP_CTAS_EXACT
#define DP_CTAS_EXACT 16
#endif
#ifndef DP_CTAS_OTHER
#define DP_CTAS_OTHER 24
#endif
constexpr int dp_min_ctas(int mode) { return mode == 0 ? DP_CTAS_EXACT : DP_CTAS_OTHER; }

// Scratch of one pair (all sections 256-byte aligned).
struct DpSlot {
    float* logp2;   // [T2][C]
    float* rootv;   // [T2 + 1] running product of network-2 blank probabilities (src/duplex.rs:389-409)
    float* roott;   // [T2 + 1] probability() of the same cells: zero (+) gap
    int* parent;    // [cap]
    int* child;     // [cap][C-1]
    uint8_t* nlab;  // [cap]
    int* voff;      // [cap]  first valid network-2 index of the node's vector
    int* vend;      // [cap]  one past the last
    float* vmx;     // [cap]  max_prob
    int* vlo;       // [cap]  lower bound that max_prob currently covers (lets update_max skip rescans)
    float* vlast;   // [cap]  label probability of the vector's last cell (older label cells are never read)
    int* rdepth;    // [R]    depth of the node that owns the ring
    float2* rings;  // [R][W] cells {gap, total}; node n uses ring n & ring_mask
};
__host__ __device__ inline size_t dp_slot_bytes(long long cap, long long R, int T2, int C, int W) {
    const int NL = C - 1 > 0 ? C - 1 : 1;
    return dx_up256((size_t)T2 * C * 4) + 2 * dx_up256((size_t)(T2 + 1) * 4) + dx_up256((size_t)cap * 4) +
           dx_up256((size_t)cap * NL * 4) + dx_up256((size_t)cap) + 5 * dx_up256((size_t)cap * 4) +
           dx_up256((size_t)R * 4) + dx_up256((size_t)R * (size_t)W * 8);
}
__device__ __forceinline__ DpSlot dp_slot(unsigned char* base, long long cap, long long R, int T2, int C, int W) {
    const int NL = C - 1 > 0 ? C - 1 : 1;
    DpSlot s;
    unsigned char* p = base;
    s.logp2 = reinterpret_cast<float*>(p); p += dx_up256((size_t)T2 * C * 4);
    s.rootv = reinterpret_cast<float*>(p); p += dx_up256((size_t)(T2 + 1) * 4);
    s.roott = reinterpret_cast<float*>(p); p += dx_up256((size_t)(T2 + 1) * 4);
    s.parent = reinterpret_cast<int*>(p);  p += dx_up256((size_t)cap * 4);
    s.child = reinterpret_cast<int*>(p);   p += dx_up256((size_t)cap * NL * 4);
    s.nlab = p;                            p += dx_up256((size_t)cap);
    s.voff = reinterpret_cast<int*>(p);    p += dx_up256((size_t)cap * 4);
    s.vend = reinterpret_cast<int*>(p);    p += dx_up256((size_t)cap * 4);
    s.vmx = reinterpret_cast<float*>(p);   p += dx_up256((size_t)cap * 4);
    s.vlo = reinterpret_cast<int*>(p);     p += dx_up256((size_t)cap * 4);
    s.vlast = reinterpret_cast<float*>(p); p += dx_up256((size_t)cap * 4);
    s.rdepth = reinterpret_cast<int*>(p);  p += dx_up256((size_t)R * 4);
    s.rings = reinterpret_cast<float2*>(p);
    return s;
}

// Out-of-line copies for the once-per-row steps (a dozen call sites): inlined everywhere the kernel was 100 KB of code
// and a quarter of its instruction fetches missed the cache (ncu: sm__icc_request_hit_rate 74 %, "no instruction" the
// leading stall).  The cell loop of phase B2 keeps its inlined LogSpace '+'.
template <int MODE>
__device__ __noinline__ float lga_t(float a, float b, const uint64_t* exp2f_tab) { return lg_add(a, b, MODE, exp2f_tab); }
__device__ __noinline__ float logf_call(float x, const double* invc, const double* logc) { return glibc_logf(x, invc, logc); }

struct PackedLayout {
    int beam_stride;  // bytes of one beam buffer (8 arrays of B words)
    int off_fold, off_key, off_node, off_lab, off_gap, off_p2, off_sel, off_row, off_root, off_info;
    int group_bytes;
    int off_jobs;     // the warp's job list, after the four group areas
    int warp_bytes;
};

// MODE = exp_mode as a compile-time constant: each instantiation carries one flavour of LogSpace '+' only.
template <int MODE>
__global__ void __launch_bounds__(32, dp_min_ctas(MODE)) duplex_packed_kernel(const DuplexParams p, const PackedLayout L) {
    auto lga = [](float a, float b, int, const uint64_t* tab) { return lga_t<MODE>(a, b, tab); };
    extern __shared__ __align__(16) unsigned char smem[];
    double* s_invc = reinterpret_cast<double*>(smem);
    double* s_logc = s_invc + 16;
    uint64_t* XT = reinterpret_cast<uint64_t*>(s_logc + 16);
    const int lane = threadIdx.x;
    if (lane < 16) {
        s_invc[lane] = c_dx_invc[lane];
        s_logc[lane] = c_dx_logc[lane];
    }
    XT[lane] = c_dx_exp2f[lane];
    __syncwarp();

    const int g = lane >> 3, bl = lane & (DP_GROUP - 1);
    const unsigned gmask = 0xFFu << (g * DP_GROUP);
    const int B = p.beam_size, C = p.C, NL = C - 1;
    const int T1 = p.T1, T2 = p.T2, W = p.W;
    constexpr int mode = MODE;
    const bool collapse = p.collapse != 0;
    const bool half_in = p.in_f16 != 0;
    const float NEG = dx_neg_inf();
    const long long cap = p.nodes_per_slot;

    const int slot0 = blockIdx.x * DP_PAIRS;
    const bool have_slot = slot0 + g < p.n_slots;
    unsigned char* wbase = smem + DP_FIXED;
    unsigned char* gs = wbase + g * L.group_bytes;
    int* jobs = reinterpret_cast<int*>(wbase + L.off_jobs);
    const long long R = p.n_rings;                                  // a power of two, or cap (one ring per node)
    const int ring_mask = (R == cap) ? 0x7fffffff : int(R - 1);
    const size_t slot_bytes = dp_slot_bytes(cap, R, T2, C, W);
    unsigned char* scratch0 = static_cast<unsigned char*>(p.scratch) + 256;
    const DpSlot S = dp_slot(scratch0 + (size_t)(have_slot ? slot0 + g : slot0) * slot_bytes, cap, R, T2, C, W);

    unsigned char* fold = gs + L.off_fold;
    uint32_t* ckey = reinterpret_cast<uint32_t*>(gs + L.off_key);
    int* cnode = reinterpret_cast<int*>(gs + L.off_node);
    float* clab = reinterpret_cast<float*>(gs + L.off_lab);
    float* cgap = reinterpret_cast<float*>(gs + L.off_gap);
    float* cp2 = reinterpret_cast<float*>(gs + L.off_p2);
    int* sel = reinterpret_cast<int*>(gs + L.off_sel);
    float* lp1 = reinterpret_cast<float*>(gs + L.off_row);
    int* root_child = reinterpret_cast<int*>(gs + L.off_root);
    int* info = reinterpret_cast<int*>(gs + L.off_info);

    const float thr = logf_call(p.threshold, s_invc, s_logc);  // src/duplex.rs:454

    for (;;) {
        // ---- every group that owns a scratch slot draws its own pair ---------------------------------------
        long long n = -1;
        if (have_slot && bl == 0) {
            const unsigned long long ticket = atomicAdd(p.ticket, 1ull);
            if (ticket < (unsigned long long)p.N) n = (long long)ticket;
        }
        n = __shfl_sync(DXFULL, n, g * DP_GROUP);
        if (!__any_sync(DXFULL, n >= 0)) break;
        int status = (n >= 0) ? FCTC_OK : -1;  // -1: this group sits the round out
        const long long n0 = n >= 0 ? n : 0;
        const float* rd1 = p.post1 + n0 * (long long)T1 * C;
        const float* rd2 = p.post2 + n0 * (long long)T2 * C;
        const __half* rh1 = reinterpret_cast<const __half*>(p.post1) + n0 * (long long)T1 * C;
        const __half* rh2 = reinterpret_cast<const __half*>(p.post2) + n0 * (long long)T2 * C;
        const long long* env = p.env ? p.env + n0 * (long long)T1 * 2 : nullptr;

#pragma unroll 1
        for (int i = bl; i < B * (NL > 0 ? NL : 1); i += DP_GROUP) fold[i] = 0;
        // network 2 in log space (src/duplex.rs:453)
        if (status == FCTC_OK)
#pragma unroll 1
            for (int i = bl; i < T2 * C; i += DP_GROUP)
                S.logp2[i] = logf_call(half_in ? __half2float(rh2[i]) : rd2[i], s_invc, s_logc);
        // root vector over envelope[(0,1)] rows (src/duplex.rs:476-477, 389-409), and its probability()
        const unsigned long long ub0 = env ? (unsigned long long)env[1] : (unsigned long long)T2;
        if (status == FCTC_OK && ub0 > (unsigned long long)T2) status = FCTC_INVALID_ENVELOPE;  // the reference would panic
        __syncwarp();
        int root_len = 0;
        if (status == FCTC_OK) {
            root_len = int(ub0) + 1;
            if (bl == 0) {
                float run = 0.0f;
                S.rootv[0] = run;
#pragma unroll 1
                for (int i = 0; i < int(ub0); ++i) {
                    run = __fadd_rn(run, S.logp2[(size_t)i * C]);
                    S.rootv[i + 1] = run;
                }
            }
        }
        __syncwarp();
        if (status == FCTC_OK)
#pragma unroll 1
            for (int i = bl; i < root_len; i += DP_GROUP) S.roott[i] = lga(NEG, S.rootv[i], mode, XT);
        int cur = 0;
        {
            DxBeam b0 = dx_beam(gs, B);
            if (bl == 0) {  // src/duplex.rs:465-473
                b0.nd[0] = DXROOT; b0.par[0] = -2; b0.tip[0] = -1; b0.dep[0] = 0;
                b0.lab[0] = NEG; b0.gap[0] = 0.0f; b0.p2[0] = 0.0f; b0.st[0] = 0;
            }
        }
#pragma unroll 1
        for (int l = bl; l < NL; l += DP_GROUP) root_child[l] = -1;
        int nbeam = 1;
        int n_nodes = 0;
        int last_hi = 0;
        __syncwarp();

        for (int t = 0; t < T1; ++t) {
            if (!__any_sync(DXFULL, status == FCTC_OK)) break;
            bool ok = status == FCTC_OK;
            // ---- envelope row (src/duplex.rs:485-488) -----------------------------------------------------
            int lo = 0, hi = 0;
            if (ok) {
                const unsigned long long lo_u = env ? (unsigned long long)env[2 * t] : 0ull;
                const unsigned long long hi_u = env ? (unsigned long long)env[2 * t + 1] : (unsigned long long)T2;
                hi = hi_u > (unsigned long long)T2 ? T2 : int(hi_u);
                if (lo_u >= (unsigned long long)hi || lo_u > (unsigned long long)last_hi) {
                    status = FCTC_INVALID_ENVELOPE;
                } else {
                    lo = int(lo_u);
                    if (hi - lo + 1 > W) status = FCTC_NODE_POOL_OVERFLOW;  // window wider than the scratch plan
                }
                ok = status == FCTC_OK;
            }
            if (ok)
#pragma unroll 1
                for (int c = bl; c < C; c += DP_GROUP)
                    lp1[c] = logf_call(half_in ? __half2float(rh1[(size_t)t * C + c]) : rd1[(size_t)t * C + c], s_invc, s_logc);

            // ---- extend the vectors of the beam's nodes (src/duplex.rs:490-522) ----------------------------
            const bool ext = ok && hi > last_hi;
            DxBeam bv = dx_beam(gs + cur * L.beam_stride, B);
            if (ext) {  // beam.sort_by_key(node): parents before children; node ids are unique in the beam
                DxBeam nb = dx_beam(gs + (cur ^ 1) * L.beam_stride, B);
                if (bl < nbeam) {
                    const int me = bv.nd[bl];
                    int r = 0;
#pragma unroll 1
                    for (int k = 0; k < nbeam; ++k) r += (bv.nd[k] < me) ? 1 : 0;
                    nb.nd[r] = me; nb.par[r] = bv.par[bl]; nb.tip[r] = bv.tip[bl]; nb.dep[r] = bv.dep[bl];
                    nb.lab[r] = bv.lab[bl]; nb.gap[r] = bv.gap[bl]; nb.p2[r] = bv.p2[bl]; nb.st[r] = 0;
                }
                cur ^= 1;
                bv = nb;
            }
            __syncwarp();
            {
                const int node = (ext && bl < nbeam) ? bv.nd[bl] : DXROOT;
                const bool active = ext && node >= 0;
                int my_par = -2, my_lab = 0, v_end = 0, par_end = 0;
                bool repeat = false;
                float mx = NEG;
                DxCell last{NEG, NEG, NEG};
                float2* vd = nullptr;
                if (active) {
                    my_par = bv.par[bl];
                    my_lab = bv.tip[bl];
                    vd = S.rings + (size_t)(node & ring_mask) * W;
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
                        // update_max(lo, hi) (:193-204); the stored max covers [vlo, end) and stays valid unless a
                        // cell that drops out could have been the maximum
                        const int b0 = lo > off ? lo : off;
                        const int e0 = hi < v_end ? hi : v_end;
                        const int old_lo = S.vlo[node];
                        bool rescan = (old_lo > b0) || (e0 <= b0);
                        if (!rescan) {
                            int sl = dx_mod(old_lo, W);
#pragma unroll 1
                            for (int i = old_lo; i < b0; ++i) {
                                if (vd[sl].y >= mx) rescan = true;
                                if (++sl == W) sl = 0;
                            }
                        }
                        if (rescan) {
                            mx = NEG;
                            int sl = dx_mod(b0, W);
#pragma unroll 1
                            for (int i = b0; i < e0; ++i) {
                                mx = lg_max(mx, vd[sl].y);
                                if (++sl == W) sl = 0;
                            }
                        }
                        S.vlo[node] = b0;
                        S.voff[node] = off;
                    }
                    if (v_end > off) {
                        const int sl = dx_mod(v_end - 1, W);
                        const float2 cell = vd[sl];
                        last.label = S.vlast[node]; last.gap = cell.x; last.total = cell.y;
                    }
                    // the parent's label decides the repeat case, independent of collapse (:511)
                    const int plab = (my_par >= 0) ? int(S.nlab[my_par]) : -1;
                    repeat = (plab == my_lab);
                    bool par_in_beam = false;
#pragma unroll 1
                    for (int k = 0; k < nbeam; ++k) par_in_beam |= (bv.nd[k] == my_par);
                    par_end = (my_par >= 0) ? (par_in_beam ? hi : S.vend[my_par]) : 0;
                }
                __syncwarp();  // every node's new offset is visible before children read their parents
                const int par_off = (active && my_par >= 0) ? S.voff[my_par] : 0;
                const float2* pv = (active && my_par >= 0) ? S.rings + (size_t)(my_par & ring_mask) * W : nullptr;
                int imin = hi;
                if (ext) imin = __reduce_min_sync(gmask, active ? v_end : hi);
                const int steps = __reduce_max_sync(DXFULL, ext ? hi - imin : 0);
                int sl_w = dx_mod(imin, W), sl_p = dx_mod(imin - 1, W);
#pragma unroll 1
                for (int k = 0; k < steps; ++k) {  // lockstep over the network-2 row index
                    const int i = imin + k;
                    if (active && i < hi && v_end <= i) {
                        // a parent that is in the beam has (by now) cells up to i-1; others are as stored
                        const int at = i - 1;
                        float pgap = NEG, ptot = NEG;
                        if (my_par == DXROOT) {
                            if (at + 1 < root_len) { pgap = S.rootv[at + 1]; ptot = S.roott[at + 1]; }
                        } else if (at >= par_off && at < ((par_end == hi) ? i : par_end)) {
                            const float2 cell = pv[sl_p];
                            pgap = cell.x; ptot = cell.y;
                        }
                        DxCell c;
                        c.gap = __fadd_rn(last.total, S.logp2[(size_t)i * C]);
                        c.label = __fadd_rn(S.logp2[(size_t)i * C + my_lab + 1], lga(last.label, repeat ? pgap : ptot, mode, XT));
                        c.total = lga(c.label, c.gap, mode, XT);
                        vd[sl_w] = make_float2(c.gap, c.total);
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
            }
            if (ok) last_hi = hi;
            __syncwarp();  // lp1 row and the extended vectors visible to every lane

            // ---- phase A: own continuation (+ the parent's extension, folded in) --------------------------
            if (ok && bl < nbeam) {
                const int b = bl;
                const int node = bv.nd[b];
                const float lab = bv.lab[b], gap = bv.gap[b];
                const int tip = bv.tip[b];
                const float p_blank = lp1[0];
                bool has_self = false;
                float s_lab = NEG, s_gap = NEG;
                if (p_blank > thr) {  // src/duplex.rs:529-534
                    s_gap = __fadd_rn(lga(lab, gap, mode, XT), p_blank);
                    has_self = true;
                }
                if (collapse && tip >= 0) {  // :540-544
                    const float pt = lp1[tip + 1];
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
                        const float pt = lp1[tip + 1];  // the parent's row
                        if (!(pt < thr)) {
                            float c;
                            if (collapse && bv.tip[r] == tip) c = __fadd_rn(bv.gap[r], pt);      // :563-568
                            else c = __fadd_rn(lga(bv.lab[r], bv.gap[r], mode, XT), pt);       // :586-590
                            s_lab = lga(s_lab, c, mode, XT);                                   // :604
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

            // ---- phase B1: child proposals; new nodes get provisional ids -2-k ---------------------------
            int n_new = 0;
            if (ok && bl < nbeam) {
                const int b = bl;
                const int node = bv.nd[b];
                const float lab = bv.lab[b], gap = bv.gap[b];
                const int tip = bv.tip[b];
#pragma unroll 1
                for (int l = 0; l < NL; ++l) {
                    const int s = b * C + 1 + l;
                    const float pl = lp1[l + 1];
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
                            clab[s] = repeat ? __fadd_rn(gap, pl) : __fadd_rn(lga(lab, gap, mode, XT), pl);
                            cgap[s] = NEG;
                            cp2[s] = bv.p2[b];
                            key = 1u;
                        }
                    }
                    ckey[s] = key;
                }
            }
            int incl = n_new;  // creation order: beam rank major, label minor
#pragma unroll
            for (int d = 1; d < DP_GROUP; d <<= 1) {
                const int o = __shfl_up_sync(DXFULL, incl, d, DP_GROUP);
                if (bl >= d) incl += o;
            }
            const int total_new = __shfl_sync(DXFULL, incl, DP_GROUP - 1, DP_GROUP);
            const int my_base = incl - n_new;
            if (ok && (long long)n_nodes + total_new > cap) {
                status = FCTC_NODE_POOL_OVERFLOW;
                ok = false;
            }
            // ring hand-over: node id takes ring id & ring_mask; its previous owner (node id - R) must be dead, i.e.
            // lie more than one level below the shallowest hypothesis of the beam
            const int dmin = __reduce_min_sync(gmask, (ok && bl < nbeam) ? bv.dep[bl] : 0x7fffffff);
            bool ring_busy = false;
            if (ok && bl < nbeam)
#pragma unroll 1
                for (int k = 0; k < n_new; ++k) {
                    const int id = n_nodes + my_base + k;
                    if (id > ring_mask && S.rdepth[id & ring_mask] >= dmin - 1) ring_busy = true;
                }
            if (__any_sync(gmask, ring_busy) && ok) {
                status = FCTC_NODE_POOL_OVERFLOW;
                ok = false;
            }
            // the four pairs' new nodes in one job list
            const int gjobs = ok ? total_new : 0;
            const int c0 = __shfl_sync(DXFULL, gjobs, 0), c1 = __shfl_sync(DXFULL, gjobs, 8);
            const int c2 = __shfl_sync(DXFULL, gjobs, 16), c3 = __shfl_sync(DXFULL, gjobs, 24);
            const int gbase = (g > 0 ? c0 : 0) + (g > 1 ? c1 : 0) + (g > 2 ? c2 : 0);
            const int n_jobs = c0 + c1 + c2 + c3;
            if (ok && bl < nbeam) {
                int k = 0;
#pragma unroll 1
                for (int l = 0; l < NL; ++l) {
                    const int s = bl * C + 1 + l;
                    if (ckey[s] != 0u && cnode[s] <= -2) {
                        jobs[gbase + my_base + k] = (g << 16) | (bl << 8) | l;
                        cnode[s] = n_nodes + my_base + k;
                        S.rdepth[(n_nodes + my_base + k) & ring_mask] = bv.dep[bl] + 1;
                        ++k;
                    }
                }
            }
            if (bl == 0) { info[0] = lo; info[1] = hi; info[2] = root_len; info[3] = cur; }
            __syncwarp();

            // ---- phase B2: the new nodes' vectors (src/duplex.rs:212-249), two lanes per node ----------------
            for (int j0 = 0; j0 < n_jobs; j0 += DP_SLOTS) {
                const int j = j0 + (lane >> 1);
                const int role = lane & 1;  // 0: label chain, 1: gap / total chain, one cell behind
                const bool valid = j < n_jobs;
                int jg = 0, jb = 0, jl = 0;
                if (valid) { const int e = jobs[j]; jg = e >> 16; jb = (e >> 8) & 0xff; jl = e & 0xff; }
                unsigned char* js = wbase + jg * L.group_bytes;
                const int* jinfo = reinterpret_cast<const int*>(js + L.off_info);
                const int jlo = jinfo[0], jhi = jinfo[1], jroot_len = jinfo[2];
                const DxBeam jbv = dx_beam(js + jinfo[3] * L.beam_stride, B);
                const DpSlot JS = dp_slot(scratch0 + (size_t)(slot0 + jg) * slot_bytes, cap, R, T2, C, W);
                int id = 0, pnode = DXROOT, par_end = 0, par_off = 0;
                bool repeat = false;
                if (valid) {
                    id = reinterpret_cast<const int*>(js + L.off_node)[jb * C + 1 + jl];
                    pnode = jbv.nd[jb];
                    repeat = collapse && jl == jbv.tip[jb];
                    if (pnode >= 0) { par_end = JS.vend[pnode]; par_off = JS.voff[pnode]; }
                }
                const float2* pv = JS.rings + (size_t)((pnode >= 0 ? pnode : 0) & ring_mask) * W;
                float2* vd = JS.rings + (size_t)(id & ring_mask) * W;
                const int len = valid ? jhi - jlo : 0;
                const int n_iter = __reduce_max_sync(DXFULL, len) + 1;
                // Iteration k: the label lane (role 0) computes cell jlo + k, the gap/total lane (role 1) cell
                // jlo + k - 1.  What a cell reads from memory does not depend on earlier cells, so the loads of
                // iteration k + 1 are issued before the log-add of iteration k.
                const float* col = JS.logp2 + (size_t)(jlo - role) * C + (role == 0 ? jl + 1 : 0);  // p2[row][label | blank]
                // parent term of cell jlo + k = parent cell jlo + k - 1 (SecondaryProbs::get, src/duplex.rs:167-179):
                // present for k in [k_lo, k_hi), read at index pidx (linear for the root vector; a ring of {gap,
                // total} cells otherwise, gap for the repeat case)
                const float* psrc = (pnode == DXROOT) ? (repeat ? JS.rootv : JS.roott)
                                                      : reinterpret_cast<const float*>(pv) + (repeat ? 0 : 1);
                const int pstride = (pnode == DXROOT) ? 1 : 2;
                int k_lo = 0, k_hi = 0, pidx = 0, pwrap = 0x7fffffff;
                if (valid && role == 0) {
                    if (pnode == DXROOT) { k_hi = jroot_len - jlo; pidx = jlo; }
                    else { k_lo = par_off - jlo + 1; k_hi = par_end - jlo + 1; pidx = dx_mod(jlo - 1, W); pwrap = W; }
                    if (k_lo < 0) k_lo = 0;
                    if (k_hi < k_lo) k_hi = k_lo;
                }
                const unsigned k_span = unsigned(k_hi - k_lo);
                const unsigned ulen = unsigned(len);
                float label_prev = NEG, label_in = NEG, total_prev = NEG, mx = NEG;
                int sl = dx_mod(jlo, W);
                int coff = 0;
                float qa_n = NEG, qb_n = 0.0f;
                if (unsigned(0 - role) < ulen) qb_n = col[0];
                if (unsigned(0 - k_lo) < k_span) qa_n = psrc[pidx * pstride];
                for (int k = 0; k < n_iter; ++k) {
                    const float qa = qa_n, qb = qb_n;
                    const bool act = unsigned(k - role) < ulen;
                    const bool act0 = act && role == 0, act1 = act && role == 1;
                    {  // next iteration's inputs
                        coff += C;
                        pidx = (pidx + 1 == pwrap) ? 0 : pidx + 1;
                        qb_n = (unsigned(k + 1 - role) < ulen) ? col[coff] : 0.0f;
                        qa_n = (unsigned(k + 1 - k_lo) < k_span) ? psrc[pidx * pstride] : NEG;
                    }
                    const float gap = __fadd_rn(total_prev, qb);  // gap_i (role 1)
                    const float a = role == 0 ? label_prev : label_in;
                    const float b2 = role == 0 ? qa : gap;
                    const float r = lg_add(a, b2, mode, XT);
                    const float lab_new = __fadd_rn(qb, r);
                    label_prev = act0 ? lab_new : label_prev;
                    if (act1) vd[sl] = make_float2(gap, r);
                    mx = act1 ? lg_max(mx, r) : mx;
                    total_prev = act1 ? r : total_prev;
                    sl = act1 ? ((sl + 1 == W) ? 0 : sl + 1) : sl;
                    label_in = __shfl_sync(DXFULL, label_prev, lane & ~1);
                }
                if (valid) {
                    if (role == 0) {
                        JS.parent[id] = pnode;
                        JS.nlab[id] = uint8_t(jl);
                        JS.voff[id] = jlo;
                        JS.vlo[id] = jlo;
                        JS.vlast[id] = label_prev;
#pragma unroll 1
                        for (int q = 0; q < NL; ++q) JS.child[(size_t)id * NL + q] = -1;
                        if (pnode == DXROOT) reinterpret_cast<int*>(js + L.off_root)[jl] = id;
                        else JS.child[(size_t)pnode * NL + jl] = id;
                    } else {
                        JS.vend[id] = jhi;
                        JS.vmx[id] = mx;
                    }
                }
            }
            if (ok) n_nodes += total_new;
            __syncwarp();

            // ---- refresh prob_2_max, scores (src/duplex.rs:613-630) ---------------------------------------------
            bool any_nan = false;
            int n_cand = 0;
            if (ok && bl < nbeam) {
#pragma unroll 1
                for (int sl = 0; sl < C; ++sl) {
                    const int s = bl * C + sl;
                    if (ckey[s] == 0u) continue;
                    const int node = cnode[s];
                    if (node >= 0) cp2[s] = S.vmx[node];
                    const float tot = __fadd_rn(__fadd_rn(lga(clab[s], cgap[s], mode, XT), cp2[s]), 0.0f);
                    if (tot != tot) any_nan = true;
                    ckey[s] = (tot != tot) ? 1u : order_key(tot);
                    ++n_cand;
                }
            }
            n_cand = __reduce_add_sync(gmask, n_cand);
            any_nan = __any_sync(gmask, any_nan);
            __syncwarp();
            if (ok) {
                if (n_cand >= 2 && any_nan) status = FCTC_INCOMPARABLE_VALUES;
                else if (n_cand == 0) status = FCTC_RAN_OUT_OF_BEAM;
                ok = status == FCTC_OK;
            }

            // ---- ranking: n_keep rounds of 8-lane maxima, lowest node id first on equal scores ---------------
            const int n_slots_c = nbeam * C;
            const int n_keep = ok ? (n_cand < B ? n_cand : B) : 0;
            const int rounds = __reduce_max_sync(DXFULL, n_keep);
            uint32_t best_hi = 0, best_lo = 0;
            int best_i = -1;
            auto rescan = [&]() {
                best_hi = 0; best_lo = 0; best_i = -1;
#pragma unroll 1
                for (int i = bl; i < n_slots_c; i += DP_GROUP) {
                    const uint32_t k = ckey[i];
                    if (k == 0u) continue;
                    const uint32_t lo2 = 0x7FFFFFFFu - uint32_t(cnode[i]);
                    if (k > best_hi || (k == best_hi && lo2 > best_lo)) { best_hi = k; best_lo = lo2; best_i = i; }
                }
            };
            if (ok) rescan();
#pragma unroll 1
            for (int r = 0; r < rounds; ++r) {
                if (r < n_keep) {
                    const uint32_t m1 = __reduce_max_sync(gmask, best_hi);
                    const uint32_t m2 = __reduce_max_sync(gmask, best_hi == m1 ? best_lo : 0u);
                    if (best_hi == m1 && best_lo == m2 && best_i >= 0) {
                        sel[r] = best_i;
                        ckey[best_i] = 0u;
                        rescan();
                    }
                }
            }
            __syncwarp();
            if (ok) {
                DxBeam nb = dx_beam(gs + (cur ^ 1) * L.beam_stride, B);
                if (bl < n_keep) {
                    const int i = sel[bl];
                    const int b = i / C, sl = i - b * C;
                    nb.nd[bl] = cnode[i];
                    nb.lab[bl] = clab[i];
                    nb.gap[bl] = cgap[i];
                    nb.p2[bl] = cp2[i];
                    nb.st[bl] = 0;
                    if (sl == 0) { nb.par[bl] = bv.par[b]; nb.tip[bl] = bv.tip[b]; nb.dep[bl] = bv.dep[b]; }
                    else { nb.par[bl] = bv.nd[b]; nb.tip[bl] = sl - 1; nb.dep[bl] = bv.dep[b] + 1; }
                }
                cur ^= 1;
                nbeam = n_keep;
            }
            __syncwarp();
        }

        // ---- labels of the best hypothesis (src/duplex.rs:638-649) ------------------------------------------------
        if (status >= 0 && bl == 0) {
            const DxBeam bv = dx_beam(gs + cur * L.beam_stride, B);
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

bool plan_packed(int B, int C, PackedLayout* L) {
    if (B < 1 || B > DP_GROUP || C < 1 || C > 8) return false;
    const int NL = C - 1 > 0 ? C - 1 : 1;
    auto up = [](long long v, int a) { return (v + a - 1) / a * a; };
    L->beam_stride = int(up(8LL * 4 * B, 16));
    long long o = 2LL * L->beam_stride;
    L->off_fold = int(o);  o = up(o + (long long)B * NL, 16);
    const long long nc = (long long)B * C;
    L->off_key = int(o);   o = up(o + 4 * nc, 16);
    L->off_node = int(o);  o = up(o + 4 * nc, 16);
    L->off_lab = int(o);   o = up(o + 4 * nc, 16);
    L->off_gap = int(o);   o = up(o + 4 * nc, 16);
    L->off_p2 = int(o);    o = up(o + 4 * nc, 16);
    L->off_sel = int(o);   o = up(o + 4LL * B, 16);
    L->off_row = int(o);   o = up(o + 4LL * C, 16);
    L->off_root = int(o);  o = up(o + 4LL * NL, 16);
    L->off_info = int(o);  o = up(o + 16, 16);
    L->group_bytes = int(up(o, 32));
    L->off_jobs = DP_PAIRS * L->group_bytes;
    L->warp_bytes = int(up(L->off_jobs + 4LL * DP_PAIRS * B * NL, 128));
    return true;
}

}  // namespace

bool duplex_packed_shape(int beam_size, int C, int crf) {
    static const bool forced_off = [] {  // tuning knob: FCTC_DUPLEX_PACKED=0 keeps every shape on duplex.cu (read once)
        const char* e = getenv("FCTC_DUPLEX_PACKED");
        return e && atoi(e) == 0;
    }();
    PackedLayout L;
    return !forced_off && !crf && plan_packed(beam_size, C, &L);
}

// Rings of one pair: one per node when the caller asks for the hard bound of proposals per row (the retry of the
// host entry points: cannot overflow), else 1024 — hundreds of times what is alive at once on nanopore-like data.
int64_t duplex_packed_rings(int64_t cap, int nodes_per_row, int64_t bound) {
    static const int knob = env_knob("FCTC_DUPLEX_RINGS");  // test / tuning knob: a smaller (power-of-two) pool
    int64_t rings = 1024;
    if (knob > 0 && (knob & (knob - 1)) == 0) rings = knob;
    if (nodes_per_row >= bound || cap <= rings) return cap;
    return rings;
}

int64_t duplex_packed_slot_bytes(int64_t cap, int64_t rings, int T2, int C, int W) {
    return int64_t(dp_slot_bytes(cap, rings, T2, C, W));
}

int duplex_packed_slots_per_sm(int exp_mode) {  // resident warps of four pairs
    if (exp_mode < 0) return DP_PAIRS * (DP_CTAS_EXACT > DP_CTAS_OTHER ? DP_CTAS_EXACT : DP_CTAS_OTHER);
    return DP_PAIRS * dp_min_ctas(exp_mode);
}

bool duplex_packed_eligible(const DuplexParams& p) {
    return duplex_packed_shape(p.beam_size, p.C, p.S > 0);
}

cudaError_t launch_duplex_packed(const DuplexParams& p, cudaStream_t stream) {
    PackedLayout L;
    if (!plan_packed(p.beam_size, p.C, &L)) return cudaErrorInvalidConfiguration;
    const int smem = DP_FIXED + L.warp_bytes;
    auto kernel = p.exp_mode == 0 ? duplex_packed_kernel<0> : p.exp_mode == 1 ? duplex_packed_kernel<1> : duplex_packed_kernel<2>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.ticket, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.cursor, 0, sizeof(unsigned long long), stream);  // the pool fills from entry 0 on every call
    if (e != cudaSuccess) return e;
    // pairs in flight: what the scratch holds, at most what this flavour keeps resident, in even rounds (every pair
    // of a batch takes about the same time: 16,384 pairs over 9,472 slots run as two rounds of 8,192)
    int dev = 0, sms = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    long long slots = p.n_slots;
    const long long resident = (long long)sms * duplex_packed_slots_per_sm(p.exp_mode);
    if (slots > resident) slots = resident;
    if (p.N > slots) {
        const long long rounds = (p.N + slots - 1) / slots;
        long long even = (p.N + rounds - 1) / rounds;
        even = (even + DP_PAIRS - 1) / DP_PAIRS * DP_PAIRS;
        if (even < slots) slots = even;
    }
    DuplexParams q = p;
    q.n_slots = int(slots);
    const int grid = int((slots + DP_PAIRS - 1) / DP_PAIRS);
    kernel<<<grid, 32, smem, stream>>>(q, L);
    return cudaGetLastError();
}

}  // namespace fctc
