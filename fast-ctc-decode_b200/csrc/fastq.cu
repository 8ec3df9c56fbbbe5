// fastq.cu — FASTQ record assembly on the device.
//
// The reference returns a read as one string, sequence then quality characters
// (src/search.rs:378-380), and leaves it to the caller to split it and to wrap it into a record.
// With batched decoding on the GPU that host pass would touch every label again, so the slot-form
// result of the Viterbi kernel (labels, quality characters, lens, status) is turned into FASTQ text
// where it lies:
//     @<prefix><first_index + n>\n<sequence>\n+\n<quality>\n
// records concatenated in read order; reads whose status is not FCTC_OK produce no record.
// Three small kernels: record lengths, an exclusive scan over them (block partials, one block for
// the partials, add back), and the writer (one warp per read, lanes copy bytes in strides of 32 so
// that both the loads from the slot and the stores into the text are contiguous).
#include "device_utils.cuh"
#include "kernels.h"

#include "../../include/fctc.h"

namespace fctc {

namespace {

constexpr int FQ_SCAN_THREADS = 256;
constexpr int FQ_SCAN_ITEMS = 4;  // per thread
constexpr int FQ_SCAN_TILE = FQ_SCAN_THREADS * FQ_SCAN_ITEMS;

__device__ __forceinline__ int decimal_digits(unsigned long long v) {
    int d = 1;
    while (v >= 10ull) { v /= 10ull; ++d; }
    return d;
}

__global__ void __launch_bounds__(256) fq_lengths_kernel(const FastqParams p, long long* rec_len) {
    for (long long n = blockIdx.x * (long long)blockDim.x + threadIdx.x; n < p.N; n += (long long)gridDim.x * blockDim.x) {
        long long len = 0;
        if (p.status[n] == FCTC_OK) {
            const long long L = p.lens[n];
            len = 1 + p.prefix_len + decimal_digits((unsigned long long)(p.first_index + n)) + 1 + L + 3 + L + 1;
        }
        rec_len[n] = len;
    }
}

// block-wise exclusive scan: out[i] = sum of in[0..i) within the tile, tile totals to block_sums
__global__ void __launch_bounds__(FQ_SCAN_THREADS) fq_scan_tiles_kernel(const long long* in, long long* out, long long n,
                                                                         long long* block_sums) {
    __shared__ long long warp_sums[FQ_SCAN_THREADS / 32];
    const long long base = (long long)blockIdx.x * FQ_SCAN_TILE + (long long)threadIdx.x * FQ_SCAN_ITEMS;
    long long v[FQ_SCAN_ITEMS], tsum = 0;
#pragma unroll
    for (int k = 0; k < FQ_SCAN_ITEMS; ++k) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        tsum += v[k];
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    long long incl = tsum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const long long o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    long long wbase = 0, total = 0;
#pragma unroll
    for (int w = 0; w < FQ_SCAN_THREADS / 32; ++w) {
        const long long s = warp_sums[w];
        if (w < warp) wbase += s;
        total += s;
    }
    long long run = wbase + incl - tsum;
#pragma unroll
    for (int k = 0; k < FQ_SCAN_ITEMS; ++k) {
        if (base + k < n) out[base + k] = run;
        run += v[k];
    }
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// one block: exclusive scan of the tile totals in place (sequential over chunks of 256; at most a few thousand tiles)
__global__ void __launch_bounds__(FQ_SCAN_THREADS) fq_scan_sums_kernel(long long* sums, long long n_tiles, long long* grand_total) {
    __shared__ long long warp_sums[FQ_SCAN_THREADS / 32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long c0 = 0; c0 < n_tiles; c0 += FQ_SCAN_THREADS) {
        const long long i = c0 + threadIdx.x;
        const long long v = (i < n_tiles) ? sums[i] : 0;
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
        for (int w = 0; w < FQ_SCAN_THREADS / 32; ++w) {
            const long long s = warp_sums[w];
            if (w < warp) wbase += s;
            total += s;
        }
        const long long c = carry;
        if (i < n_tiles) sums[i] = c + wbase + incl - v;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *grand_total = carry;
}

__global__ void __launch_bounds__(256) fq_add_sums_kernel(long long* offs, long long n, const long long* sums,
                                                          const long long* grand_total) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i <= n; i += (long long)gridDim.x * blockDim.x) {
        if (i == n) offs[n] = *grand_total;
        else offs[i] += sums[i / FQ_SCAN_TILE];
    }
}

__global__ void __launch_bounds__(256) fq_write_kernel(const FastqParams p, const long long* rec_off) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    const long long total = rec_off[p.N];
    if (total > p.out_capacity) return;  // the host reports FCTC_E_CAPACITY from the same number
    for (long long n = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps) {
        if (p.status[n] != FCTC_OK) continue;
        const long long L = p.lens[n];
        const long long src = p.offsets[n];
        unsigned char* o = p.out + rec_off[n];
        const unsigned long long idx = (unsigned long long)(p.first_index + n);
        const int digits = decimal_digits(idx);
        const int head = 1 + p.prefix_len + digits + 1;
        if (lane == 0) {
            o[0] = '@';
            for (int k = 0; k < p.prefix_len; ++k) o[1 + k] = p.prefix[k];
            unsigned long long v = idx;
            for (int k = digits - 1; k >= 0; --k) { o[1 + p.prefix_len + k] = (unsigned char)('0' + int(v % 10ull)); v /= 10ull; }
            o[head - 1] = '\n';
            o[head + L] = '\n';
            o[head + L + 1] = '+';
            o[head + L + 2] = '\n';
            o[head + L + 3 + L] = '\n';
        }
        unsigned char* seq = o + head;
        unsigned char* qual = o + head + L + 3;
        for (long long k = lane; k < L; k += 32) {
            seq[k] = p.alphabet[p.labels[src + k]];
            qual[k] = p.quals[src + k];
        }
    }
}

}  // namespace

int64_t fastq_scratch_bytes(int64_t N) {
    const int64_t tiles = (N + FQ_SCAN_TILE - 1) / FQ_SCAN_TILE;
    return 8 * (N + tiles + 2) + 256;
}

cudaError_t launch_fastq(const FastqParams& p, long long* rec_off, void* scratch, cudaStream_t stream) {
    // scratch: [N] record lengths, [tiles] tile sums, [1] grand total
    long long* rec_len = static_cast<long long*>(scratch);
    const long long tiles = (p.N + FQ_SCAN_TILE - 1) / FQ_SCAN_TILE;
    long long* sums = rec_len + p.N;
    long long* grand = sums + tiles;
    if (p.N <= 0) return cudaMemsetAsync(rec_off, 0, 8, stream);
    const unsigned blocks = (unsigned)((p.N + 255) / 256 < 148 * 8 ? (p.N + 255) / 256 : 148 * 8);
    fq_lengths_kernel<<<blocks, 256, 0, stream>>>(p, rec_len);
    fq_scan_tiles_kernel<<<(unsigned)tiles, FQ_SCAN_THREADS, 0, stream>>>(rec_len, rec_off, p.N, sums);
    fq_scan_sums_kernel<<<1, FQ_SCAN_THREADS, 0, stream>>>(sums, tiles, grand);
    fq_add_sums_kernel<<<blocks, 256, 0, stream>>>(rec_off, p.N, sums, grand);
    const long long want = (p.N + 7) / 8;  // 8 warps per block
    fq_write_kernel<<<(unsigned)(want < 148 * 16 ? want : 148 * 16), 256, 0, stream>>>(p, rec_off);
    return cudaGetLastError();
}

}  // namespace fctc

// ---------------------------------------------------------------------------------------------
// Duplex envelope from the two reads' Viterbi paths.
//
// The reference leaves the envelope to the caller (src/lib.rs:376-378).  This constructor needs no
// aligner: base k of read 1 is anchored to base (k * L2) / L1 of read 2 — the two calls of one
// molecule have nearly the same length, while their translocation speeds (rows per base) differ
// locally, which a band that is diagonal in ROWS cannot follow.  For network-1 row t, with k(t) bases
// of read 1 emitted at or before t:
//     c(t) = path2[((k-1) * L2) / L1]                      k >= 1
//          = min(path2[0], (t * T2) / T1)                  k == 0
//          = (t * T2) / T1                                 either read empty
//     hi(t) = min(T2, c(t) + pad + 1)
//     lo(t) = min(max(0, c(t) - pad), hi(t-1))             (lo(0) without the second term)
// c is non-decreasing, so hi is, and lo(t) <= hi(t-1) < ... keeps the reference's validity rule
// (lo < hi, lo <= previous hi; src/duplex.rs:484-488).  One thread per (pair, row): a binary search
// in read 1's emission rows, two loads from read 2's.
// ---------------------------------------------------------------------------------------------
namespace fctc {

namespace {

__device__ __forceinline__ long long env_center(const EnvelopeParams& p, long long n, long long t) {
    const int L1 = p.lens1[n], L2 = p.lens2[n];
    const long long prop = (t * p.T2) / p.T1;
    if (L1 <= 0 || L2 <= 0) return prop < p.T2 ? prop : p.T2 - 1;
    const int* a = p.path1 + p.offsets1[n];
    const int* b = p.path2 + p.offsets2[n];
    int lo = 0, hi = L1;  // k = number of entries of a[] that are <= t
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] <= t) lo = mid + 1; else hi = mid;
    }
    const int k = lo;
    if (k == 0) {
        const long long first = b[0];
        return prop < first ? prop : first;
    }
    if (p.partner) {  // aligned anchor (launch_align), kept inside read 2
        int j = p.partner[p.offsets1[n] + k - 1];
        j = j < 0 ? 0 : (j >= L2 ? L2 - 1 : j);
        return b[j];
    }
    return b[(long long)(k - 1) * L2 / L1];
}

__global__ void __launch_bounds__(256) envelope_kernel(const EnvelopeParams p) {
    const long long total = p.N * p.T1;
    int widest = 0;
    for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long n = g / p.T1, t = g - n * p.T1;
        const long long c = env_center(p, n, t);
        long long hi = c + p.pad + 1;
        if (hi > p.T2) hi = p.T2;
        long long lo = c - p.pad;
        if (lo < 0) lo = 0;
        if (t > 0) {
            long long ph = env_center(p, n, t - 1) + p.pad + 1;
            if (ph > p.T2) ph = p.T2;
            if (lo > ph) lo = ph;
        }
        p.env[2 * g] = lo;
        p.env[2 * g + 1] = hi;
        if (p.max_window && int(hi - lo) > widest) widest = int(hi - lo);
    }
    if (p.max_window) {  // one atomic per warp: the widest window is what the duplex kernel sizes its rings by
        widest = __reduce_max_sync(0xffffffffu, widest);
        if ((threadIdx.x & 31) == 0 && widest > 0) atomicMax(p.max_window, widest);
    }
}

}  // namespace

cudaError_t launch_envelope(const EnvelopeParams& p, cudaStream_t stream) {
    const long long total = p.N * p.T1;
    if (total <= 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (p.max_window) {
        cudaError_t e = cudaMemsetAsync(p.max_window, 0, sizeof(int), stream);
        if (e != cudaSuccess) return e;
    }
    envelope_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace fctc
