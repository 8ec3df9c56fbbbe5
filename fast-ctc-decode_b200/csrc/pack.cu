// pack.cu — slot-form Viterbi results squeezed into a packed pool on the device.
//
// The Viterbi kernel gives every read a fixed slot of round_up(n_rows, 32) entries (deterministic, no
// atomics), of which a read fills only its emitted labels — about two thirds on softmax rows, a third on
// nanopore-like rows.  A caller in host memory pays PCIe and host-DRAM bandwidth for every byte that
// comes back, so fctc_viterbi_batch_packed copies the used prefixes together before the read-back:
//   1. one block scans round_up(lens[n], 32) over the chunk's reads -> packed offsets, total in *cursor
//      (placement is read order, so the packed pool is as deterministic as the slot form);
//   2. one warp per read moves labels / quality characters / rows in 16-byte pieces (slot and packed
//      starts are both multiples of 32 entries); entries past a read's length are written as 0.
// No reference counterpart: src/search.rs:320-383 returns one read at a time in host memory.
#include "device_utils.cuh"
#include "kernels.h"

namespace fctc {

namespace {

constexpr int PK_SCAN_THREADS = 1024;

__global__ void __launch_bounds__(PK_SCAN_THREADS) pack_scan_kernel(const int32_t* __restrict__ lens, long long n,
                                                                    long long* __restrict__ out_offsets,
                                                                    unsigned long long* __restrict__ cursor) {
    __shared__ long long warp_sums[PK_SCAN_THREADS / 32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long c0 = 0; c0 < n; c0 += PK_SCAN_THREADS) {
        const long long i = c0 + threadIdx.x;
        const long long v = (i < n) ? ((long long)(lens[i] + 31) & ~31ll) : 0;
        long long incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        long long wbase = 0, total = 0;
#pragma unroll
        for (int w = 0; w < PK_SCAN_THREADS / 32; ++w) {
            const long long s = warp_sums[w];
            if (w < warp) wbase += s;
            total += s;
        }
        const long long c = carry;
        if (i < n) out_offsets[i] = c + wbase + incl - v;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *cursor = (unsigned long long)carry;
}

// Copies `len` elements of ES bytes from src to dst (both 16-byte aligned), zero-filling up to the next
// multiple of 32 elements; the lanes of one warp take consecutive 16-byte pieces.
template <int ES>
__device__ __forceinline__ void pack_copy_array(const unsigned char* __restrict__ src, unsigned char* __restrict__ dst,
                                                int len, int lane) {
    constexpr int PER = 16 / ES;  // elements per piece
    const int padded = (len + 31) & ~31;
    for (int e = lane * PER; e < padded; e += 32 * PER) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (e < len) {
            v = *reinterpret_cast<const uint4*>(src + (size_t)e * ES);
            if (e + PER > len) {  // the piece that holds the read's last entry: clear what follows it
                unsigned char* b = reinterpret_cast<unsigned char*>(&v);
                for (int k = (len - e) * ES; k < 16; ++k) b[k] = 0;
            }
        }
        *reinterpret_cast<uint4*>(dst + (size_t)e * ES) = v;
    }
}

__global__ void __launch_bounds__(256) pack_copy_kernel(const PackParams p) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    const unsigned long long total = *p.cursor;
    if (total > (unsigned long long)p.out_capacity) return;  // cannot happen for buffers of N * round_up(T, 32) entries
    for (long long n = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps) {
        const int len = p.lens[n];
        const long long s = p.offsets[n] - p.offset_base, d = p.out_offsets[n];
        pack_copy_array<1>(p.labels + s, p.out_labels + d, len, lane);
        pack_copy_array<4>(reinterpret_cast<const unsigned char*>(p.path + s), reinterpret_cast<unsigned char*>(p.out_path + d),
                           len, lane);
        if (p.quals) {
            if (p.qual_bytes == 1)
                pack_copy_array<1>(static_cast<const unsigned char*>(p.quals) + s, static_cast<unsigned char*>(p.out_quals) + d,
                                   len, lane);
            else
                pack_copy_array<4>(static_cast<const unsigned char*>(p.quals) + s * 4,
                                   static_cast<unsigned char*>(p.out_quals) + d * 4, len, lane);
        }
    }
}

}  // namespace

cudaError_t launch_pack(const PackParams& p, int sm_count, cudaStream_t stream) {
    if (p.N <= 0) return cudaMemsetAsync(p.cursor, 0, 8, stream);
    pack_scan_kernel<<<1, PK_SCAN_THREADS, 0, stream>>>(p.lens, p.N, reinterpret_cast<long long*>(p.out_offsets), p.cursor);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    long long blocks = (p.N + 7) / 8;
    const long long cap = (long long)sm_count * 8;
    if (blocks > cap) blocks = cap;
    pack_copy_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace fctc
