// input.cuh — what the decoders accept besides float32 posteriors (the reference
// takes float32 posteriors only, src/lib.rs:171-188, and its caller's previous step is a softmax):
//   * 16-bit storage: IEEE binary16 or bfloat16, widened exactly to f32 as a value is fetched;
//   * logits (or log-probabilities): the softmax over the label axis is applied in the load stage.
// The softmax arithmetic is fixed so that a host can reproduce it:
//     m = first maximum of the row (strict '>' from label 0);  e_c = exp(z_c - m);
//     S = (((e_0 + e_1) + e_2) + ...) in index order;  p_c = e_c / S            — all in f32.
// mode 1 ("exact"): exp is glibc's expf, evaluated operation by operation (phred.cuh), the division is IEEE: p is,
//   bit for bit, what that recipe gives on a glibc host (oracle_softmax), so decoding logits equals decoding the
//   host-side softmax of the same logits.
// mode 2 ("fast"): exp through ex2.approx and one approximate reciprocal of S: a few ulp off, an order of
//   magnitude fewer instructions.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

#include "phred.cuh"

namespace fctc {

__device__ __forceinline__ float widen16(uint16_t bits, bool bf16) {
    return bf16 ? __uint_as_float(uint32_t(bits) << 16) : __half2float(__ushort_as_half(bits));
}
// two packed 16-bit values -> two floats (low half first)
__device__ __forceinline__ float2 widen16x2(uint32_t w, bool bf16) {
    if (bf16) return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
}

__device__ __forceinline__ float softmax_exp(float d, int mode, const uint64_t* exp2f_tab) {  // d = z - max <= 0 (or NaN)
    if (mode == 1) return glibc_expf_nonpos(d, exp2f_tab);
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(d * 1.4426950408889634f));
    return e;
}

// In place, C values at row[0], row[stride], ...
__device__ __forceinline__ void softmax_row(float* row, int C, int mode, const uint64_t* exp2f_tab) {
    float m = row[0];
    for (int c = 1; c < C; ++c)
        if (row[c] > m) m = row[c];
    float S = 0.0f;
    for (int c = 0; c < C; ++c) {
        const float e = softmax_exp(__fadd_rn(row[c], -m), mode, exp2f_tab);
        row[c] = e;
        S = (c == 0) ? e : __fadd_rn(S, e);
    }
    if (mode == 1) {
        for (int c = 0; c < C; ++c) row[c] = __fdiv_rn(row[c], S);
    } else {
        const float inv = __frcp_rn(S);
        for (int c = 0; c < C; ++c) row[c] = __fmul_rn(row[c], inv);
    }
}

}  // namespace fctc
