// device_utils.cuh — sm_100a building blocks shared by the decode kernels: mbarrier + bulk-async
// (TMA engine) global->shared copies, warp scans, cache-hinted stores.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fctc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
    // make the initialised barriers visible to the async (TMA) proxy before the first bulk copy
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}

// 1-D bulk asynchronous copy global -> shared, completion counted in bytes on an mbarrier
// (SASS: UBLKCP).  dst, src 16-byte aligned; bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// The same operations on addresses already in the shared window (32-bit), for loops that keep
// their ring positions as shared addresses.
__device__ __forceinline__ void mbar_init_a(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra WAIT_%=;\n\t}" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s_a(uint32_t dst, const void* src_gmem, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src_gmem), "r"(bytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void sts_v2(uint32_t addr, uint32_t a, uint32_t b) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void lds_v2(uint32_t addr, uint32_t& a, uint32_t& b) {
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr) : "memory");
}

// IEEE a / b that keeps a zero numerator away from the division's out-of-line special-case code:
// (+-0) / b is the same zero for any finite b > 0, and in the beam searches most quotients are of
// that kind (the gap probability of a freshly extended hypothesis is 0).
__device__ __forceinline__ float div_keep_zero(float a, float b) {
    const bool zero_over_finite = a == 0.0f && b > 0.0f && b <= 3.402823466e+38f;
    float num = zero_over_finite ? 1.0f : a;  // the divider itself never sees the zero (no branch to rely on)
    asm volatile("" : "+f"(num));            // ... and the compiler may not fold the select back into a / b
    const float q = num / b;
    return zero_over_finite ? a : q;
}

// Monotone map f32 -> u32 (bigger float, bigger key; >= 1 for every non-NaN input): ranking by REDUX.MAX.
__device__ __forceinline__ uint32_t order_key(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// The shuffle's own predicate (true where the source lane exists) guards the add: two instructions per step
// instead of shuffle + compare + select + add.
__device__ __forceinline__ int warp_inclusive_scan(int v, int lane) {
    (void)lane;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        asm volatile("{\n\t.reg .pred p;\n\t.reg .s32 t;\n\t"
                     "shfl.sync.up.b32 t|p, %0, %1, 0, 0xffffffff;\n\t"
                     "@p add.s32 %0, %0, t;\n\t}"
                     : "+r"(v) : "r"(d));
    }
    return v;
}

}  // namespace fctc
