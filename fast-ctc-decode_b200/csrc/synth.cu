// synth.cu — synthetic posterior generator (bench + tests input).  Not part of the
// decode path.  Counter-based: every row is a pure function of (seed, read, row), so any shard
// of any size regenerates the same bytes.
#include "duplex_common.cuh"
#include "kernels.h"
#include "phred.cuh"

namespace fctc {

__device__ __forceinline__ uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ float u24(uint64_t w, int half) {  // uniform in [0,1), 24 bits
    uint32_t bits = half ? uint32_t(w >> 40) : uint32_t(w >> 8) & 0xFFFFFFu;
    return float(bits) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(256) synth_kernel(float* post, long long N, long long T, int C, int kind,
                                                    uint64_t seed, long long first_read) {
    const long long rows = N * T;
    for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < rows;
         g += (long long)gridDim.x * blockDim.x) {
        const long long n = g / T, t = g - n * T;
        // kind 2 = kind 1's hidden label track with an independent noise stream (second read of a duplex pair)
        const uint64_t noise_seed = (kind == 2) ? (seed ^ 0x5BD1E9955BD1E995ull) : seed;
        const uint64_t key = mix64(noise_seed ^ mix64(uint64_t(first_read + n) * 0x100000001B3ull + 0x51ull)) + uint64_t(t) * 0xD1342543DE82EF95ull;
        float* row = post + g * C;
        if (kind == 0) {
            // softmax(3 * (2u - 1)) in f32
            float mx = -1e30f;
            for (int c = 0; c < C; ++c) {
                const uint64_t w = mix64(key + uint64_t(c >> 1) * 0x632BE59BD9B4E019ull);
                const float z = 3.0f * (2.0f * u24(w, c & 1) - 1.0f);
                row[c] = z;
                mx = fmaxf(mx, z);
            }
            float sum = 0.0f;
            for (int c = 0; c < C; ++c) {
                const float e = expf(row[c] - mx);
                row[c] = e;
                sum += e;
            }
            const float inv = 1.0f / sum;
            for (int c = 0; c < C; ++c) row[c] *= inv;
        } else {
            // peaked, nanopore-like: the dominant label is shared by row pairs (so repeats occur),
            // blank with probability 0.65, everything else 0.05*u^2
            const uint64_t pair_key = mix64(seed ^ mix64(uint64_t(first_read + n) * 0x100000001B3ull + 0x51ull)) +
                                      uint64_t(t >> 1) * 0xD1342543DE82EF95ull + 0xABCDull;
            const uint64_t wd = mix64(pair_key);
            int dom = 0;
            if (u24(wd, 0) >= 0.65f && C > 1) {
                dom = 1 + int(u24(wd, 1) * float(C - 1));
                if (dom > C - 1) dom = C - 1;
            }
            float sum = 0.0f;
            for (int c = 0; c < C; ++c) {
                const uint64_t w = mix64(key + uint64_t(c >> 1) * 0x632BE59BD9B4E019ull);
                const float u = u24(w, c & 1);
                const float v = (c == dom) ? 1.0f : 0.05f * u * u;
                row[c] = v;
                sum += v;
            }
            for (int c = 0; c < C; ++c) row[c] = row[c] / sum;
        }
    }
}

cudaError_t launch_synth(float* post, int64_t N, int64_t T, int C, int kind, uint64_t seed, int64_t first_read,
                         cudaStream_t stream) {
    const long long rows = N * T;
    if (rows <= 0) return cudaSuccess;
    long long blocks = (rows + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    synth_kernel<<<dim3((unsigned)blocks), 256, 0, stream>>>(post, N, T, C, kind, seed, first_read);
    return cudaGetLastError();
}

// Test hook: leaves `byte` in every byte of every SM's shared memory (one CTA per SM, each taking the whole
// 227 KB), so that a kernel that reads shared memory it has not written meets a chosen pattern instead of
// whatever the previous kernel left.  0xFF gives -1 / NaN words.
__global__ void __launch_bounds__(1024) poison_smem_kernel(unsigned int word, int bytes, unsigned int* sink) {
    extern __shared__ unsigned int pool[];
    for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) pool[i] = word;
    __syncthreads();
    if (pool[(threadIdx.x * 131) % (bytes / 4)] != word) sink[0] = 1;  // keeps the stores alive
}

cudaError_t launch_poison_smem(int sm_count, int byte, unsigned int* sink, cudaStream_t stream) {
    const int bytes = 227 * 1024;
    cudaError_t e = cudaFuncSetAttribute(poison_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    const unsigned int b = (unsigned int)(byte & 0xff);
    poison_smem_kernel<<<sm_count * 2, 1024, bytes, stream>>>(b * 0x01010101u, bytes, sink);
    return cudaGetLastError();
}


// Test hook: the DEVICE builds of the libm restatements over an array (which: 0 expf on x <= 0, 1 log1pf on [0, 1],
// 2 logf, 3 log10f, 4 LogSpace '+' of (x, y) in exact-exp mode), so that the arithmetic the kernels run — fused
// operations, the branch-free division, the table in shared memory — is compared with the host's libm directly,
// not only through decode results.
__global__ void __launch_bounds__(256) libm_probe_kernel(const float* x, const float* y, float* out, long long n, int which) {
    __shared__ uint64_t xt[32];
    __shared__ double invc[16], logc[16];
    static const uint64_t tab[32] = FCTC_EXP2F_TAB;
    static const double ic[16] = FCTC_LOGF_INVC, lc[16] = FCTC_LOGF_LOGC;
    if (threadIdx.x < 32) xt[threadIdx.x] = tab[threadIdx.x];
    if (threadIdx.x < 16) { invc[threadIdx.x] = ic[threadIdx.x]; logc[threadIdx.x] = lc[threadIdx.x]; }
    __syncthreads();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = x[i];
        float r;
        if (which == 0) r = glibc_expf_nonpos(v, xt);
        else if (which == 1) r = glibc_log1pf_unit(v);
        else if (which == 2) r = glibc_logf(v, invc, logc);
        else if (which == 3) r = glibc_log10f(v, invc, logc);
        else if (which == 5) {  // fast minus exact quality value for the miss probability v; y = {qscale, qbias}
            const QualityParams qp = make_quality_params(y[0], y[1]);
            r = phred_fast_q(v, qp) - phred_exact_q(v, y[0], y[1], invc, logc);
        } else if (which == 6) {  // the eps the fast path allows at that value
            const QualityParams qp = make_quality_params(y[0], y[1]);
            r = phred_fast_eps(phred_fast_q(v, qp), qp);
        } else r = lg_add(v, y[i], 0, xt);
        out[i] = r;
    }
}

cudaError_t launch_libm_probe(const float* x, const float* y, float* out, long long n, int which, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    libm_probe_kernel<<<(unsigned)blocks, 256, 0, stream>>>(x, y, out, n, which);
    return cudaGetLastError();
}

}  // namespace fctc
