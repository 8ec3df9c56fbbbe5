// duplex_common.cuh — pieces shared by the two duplex kernels (duplex.cu: one pair per warp, any beam up to 32 and
// the crf variant; duplex_packed.cu: four pairs per warp for beams up to 8): LogSpace arithmetic, the scratch
// layout of one pair slot, the beam record arrays, ring-index helpers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_utils.cuh"
#include "phred.cuh"

namespace fctc {

constexpr int DX_WARPS = 4;
constexpr int DX_THREADS = DX_WARPS * 32;
constexpr int DX_MAX_BEAM = 32;
constexpr unsigned DXFULL = 0xffffffffu;
constexpr int DXROOT = -1;

static __constant__ double c_dx_invc[16] = FCTC_LOGF_INVC;
static __constant__ double c_dx_logc[16] = FCTC_LOGF_LOGC;

__device__ __forceinline__ float dx_neg_inf() { return __uint_as_float(0xff800000u); }

static __constant__ uint64_t c_dx_exp2f[32] = FCTC_EXP2F_TAB;

// LogSpace '+' (src/duplex.rs:42-63): operands ordered so that a+b == b+a, -inf is an exact unit.
// mode 0: exact exp — glibc's expf and log1pf re-evaluated operation by operation (phred.cuh), so the sum has the
//         bits a glibc host computes; mode 1: the reference's default build, whose fastexp() is 0: big + ln_1p(0);
//         mode 2: CUDA's expf / log1pf (last-bit differences from glibc, about a third of the instructions).
// exp2f_tab: FCTC_EXP2F_TAB in shared memory.
__device__ __forceinline__ float lg_add(float a, float b, int mode, const uint64_t* exp2f_tab) {
    // straight-line: the early return for a -inf operand is a select at the end, so that the lanes of a warp stay
    // together (small - big is <= 0 or NaN, its exp in [0, 1] or NaN: the domains of the branch-free libm forms)
    const bool le = a <= b;
    const float big = le ? b : a, small = le ? a : b;
    float y = 0.0f;  // mode 1: ln_1p(fastexp(..)) = ln_1p(0)
    if (mode != 1) {
        const float d = __fadd_rn(small, -big);
        y = (mode == 2) ? log1pf(expf(d)) : glibc_log1pf_unit(glibc_expf_nonpos(d, exp2f_tab));
    }
    const float r = __fadd_rn(big, y);
    return (small == dx_neg_inf()) ? big : r;
}
__device__ __forceinline__ float lg_max(float a, float b) { return (a < b) ? b : a; }  // src/duplex.rs:33-39

struct DxBeam {
    int* nd;
    int* par;
    int* tip;
    int* dep;
    float* lab;
    float* gap;
    float* p2;
    int* st;  // crf transition state
};
__device__ __forceinline__ DxBeam dx_beam(unsigned char* base, int B) {
    DxBeam v;
    v.nd = reinterpret_cast<int*>(base);
    v.par = v.nd + B;
    v.tip = v.par + B;
    v.dep = v.tip + B;
    v.lab = reinterpret_cast<float*>(v.dep + B);
    v.gap = v.lab + B;
    v.p2 = v.gap + B;
    v.st = reinterpret_cast<int*>(v.p2 + B);
    return v;
}

// Scratch of one warp slot (all sections 256-byte aligned).
struct DxSlot {
    float* logp2;   // [T2][C]
    float* rootv;   // [T2 + 1] running product of network-2 blank probabilities (src/duplex.rs:389-409)
    float* roott;   // [T2 + 1] probability() of the same cells: zero (+) gap (packed kernel only)
    int* parent;    // [cap]
    int* child;     // [cap][C-1]
    uint8_t* nlab;  // [cap]
    int* voff;      // [cap]  first valid network-2 index of the node's vector
    int* vend;      // [cap]  one past the last
    float* vmx;     // [cap]  max_prob
    int* vlo;       // [cap]  lower bound that max_prob currently covers (lets update_max skip rescans)
    float* vlast;   // [cap]  label probability of the vector's last cell (older label cells are never read)
    float* vdata;   // [cap][2][W]: gap, total
};
__host__ __device__ inline size_t dx_up256(size_t v) { return (v + 255) & ~size_t(255); }
__host__ __device__ inline size_t dx_slot_bytes(long long cap, int T2, int C, int W) {
    const int NL = C - 1 > 0 ? C - 1 : 1;
    return dx_up256((size_t)T2 * C * 4) + 2 * dx_up256((size_t)(T2 + 1) * 4) + dx_up256((size_t)cap * 4) +
           dx_up256((size_t)cap * NL * 4) + dx_up256((size_t)cap) + 5 * dx_up256((size_t)cap * 4) +
           dx_up256((size_t)cap * 2 * (size_t)W * 4);
}
__device__ __forceinline__ DxSlot dx_slot(unsigned char* base, long long cap, int T2, int C, int W) {
    const int NL = C - 1 > 0 ? C - 1 : 1;
    DxSlot s;
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
    s.vdata = reinterpret_cast<float*>(p);
    return s;
}

struct DxCell {
    float label, gap, total;
};

__device__ __forceinline__ int dx_mod(int i, int W) {  // i mod W for i >= -1
    return i < 0 ? i + W : i % W;
}

}  // namespace fctc
