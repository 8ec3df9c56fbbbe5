// align.cu — what sits on either side of the decode path, on the device:
//
//   * banded alignment of the two Viterbi label strings of a duplex pair, so that the envelope of the 2-D search
//     follows the pair through insertions and deletions.  The reference leaves the envelope to the caller
//     (src/lib.rs:376-378) and only checks it (src/duplex.rs:484-488); the proportional anchor of
//     fctc_duplex_envelope_device (base k <-> base k*L2/L1) drifts wherever one read has lost or gained bases, and
//     the search's cost is linear in the window width that drift forces.
//   * the move table and the dwell (run length in rows) of every emitted base, from the emission rows that
//     search::viterbi_search returns as `path` (src/search.rs:378-380 hands back only the rows).
//
// Alignment (one warp per pair).  a = read 1's labels (length L1), b = read 2's (L2).  Unit-cost edit distance
// over the band  j in [i*L2/L1 - band, i*L2/L1 + band]  of row i (integer division; cells outside are infinite):
//     D[0][j] = j,   D[i][j] = min(D[i-1][j-1] + (a[i-1] != b[j-1]),  D[i-1][j] + 1,  D[i][j-1] + 1)
// with ties resolved diagonal, then up, then left.  The left dependency is a running minimum: with
// A[j] = min(diagonal, up), D[i][j] = j + min over j' <= j of (A[j'] - j'), which the lanes compute as a prefix
// minimum (a few cells per lane, one warp scan per row).  One byte per cell records the move; lane 0 walks back
// from (L1, L2): a diagonal move pairs base i-1 with base j-1, an up move (a base read 2 does not have) anchors base
// i-1 to the read-2 base before the gap.  If the band does not connect the corners (very different lengths), the
// pair keeps the proportional anchor.
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_utils.cuh"
#include "kernels.h"

#include "../../include/fctc.h"

namespace fctc {

namespace {

constexpr int AL_WARPS = 4;
constexpr int AL_MAX_CPL = 8;           // band cells per lane: band half-width up to 127 bases
constexpr int AL_INF = 0x3fffffff;
enum : uint8_t { AL_DIAG = 0, AL_UP = 1, AL_LEFT = 2 };

__global__ void __launch_bounds__(AL_WARPS * 32) align_kernel(const AlignParams p) {
    extern __shared__ int al_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = blockIdx.x * AL_WARPS + warp;
    if (slot >= p.n_slots) return;
    const int bw = p.band, width = 2 * bw + 1;
    const int cpl = (width + 31) / 32;  // cells per lane, contiguous: lane l owns cells [l*cpl, (l+1)*cpl)
    int* rowbuf = al_smem + warp * 2 * (width + 1);  // two rows of D, ping-pong
    uint8_t* dirs = p.dirs + (size_t)slot * p.dir_bytes_per_slot;

    for (long long n = slot; n < p.N; n += p.n_slots) {
        const int L1 = p.lens1[n], L2 = p.lens2[n];
        int* partner = p.partner + p.offsets1[n];
        if (L1 <= 0 || L2 <= 0) continue;  // nothing to anchor: the envelope falls back to the diagonal band
        const uint8_t* a = p.labels1 + p.offsets1[n];
        const uint8_t* b = p.labels2 + p.offsets2[n];
        int* prev = rowbuf;
        int* cur = rowbuf + (width + 1);
        // row 0: D[0][j] = j inside the band
        for (int q = lane; q < width; q += 32) {
            const int j = q - bw;
            prev[q] = (j >= 0 && j <= L2) ? j : AL_INF;
        }
        __syncwarp();
        int off_prev = -bw;
        for (int i = 1; i <= L1; ++i) {
            const int off = int(((long long)i * L2) / L1) - bw;
            const int shift = off - off_prev;
            const int ai = a[i - 1];
            int A[AL_MAX_CPL], dmove[AL_MAX_CPL];
            int run = AL_INF;  // min over this lane's cells so far of A[q] - q
#pragma unroll
            for (int c = 0; c < AL_MAX_CPL; ++c) {
                A[c] = AL_INF;
                dmove[c] = AL_UP;
                const int q = lane * cpl + c;
                if (c < cpl && q < width) {
                    const int j = off + q;
                    if (j >= 0 && j <= L2) {
                        const int qp = q + shift;  // the same column in the previous row's band
                        const int up = (qp >= 0 && qp < width) ? prev[qp] : AL_INF;
                        const int dg = (j >= 1 && qp - 1 >= 0 && qp - 1 < width) ? prev[qp - 1] : AL_INF;
                        const int dcost = (dg < AL_INF) ? dg + (ai != int(b[j - 1]) ? 1 : 0) : AL_INF;
                        const int ucost = (up < AL_INF) ? up + 1 : AL_INF;
                        A[c] = dcost <= ucost ? dcost : ucost;
                        dmove[c] = dcost <= ucost ? AL_DIAG : AL_UP;
                    }
                    const int v = (A[c] < AL_INF) ? A[c] - q : AL_INF;
                    run = v < run ? v : run;
                }
            }
            // exclusive prefix minimum of the lanes' totals
            int incl = run;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d && o < incl) incl = o;
            }
            int carry = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) carry = AL_INF;
#pragma unroll
            for (int c = 0; c < AL_MAX_CPL; ++c) {
                const int q = lane * cpl + c;
                if (c < cpl && q < width) {
                    const int v = (A[c] < AL_INF) ? A[c] - q : AL_INF;
                    // best of the cells to the left (their value + distance) against this cell's own A
                    const int left = (carry < AL_INF) ? carry + q : AL_INF;
                    int d = A[c];
                    uint8_t mv = uint8_t(dmove[c]);
                    if (left < d) { d = left; mv = AL_LEFT; }
                    cur[q] = d;
                    dirs[(size_t)i * width + q] = mv;
                    carry = v < carry ? v : carry;
                }
            }
            __syncwarp();
            int* t = prev; prev = cur; cur = t;
            off_prev = off;
        }
        // traceback from (L1, L2); the last row's band is centred on L2, so that cell is index bw
        if (lane == 0) {
            const bool connected = prev[bw] < AL_INF;
            int i = L1, j = L2;
            while (i > 0) {
                if (!connected) {  // the band lost the path: proportional anchor, as fctc_duplex_envelope_device
                    partner[i - 1] = int(((long long)(i - 1) * L2) / L1);
                    --i;
                    continue;
                }
                const int off = int(((long long)i * L2) / L1) - bw;
                const uint8_t mv = dirs[(size_t)i * width + (j - off)];
                if (mv == AL_DIAG) { partner[i - 1] = j - 1; --i; --j; }
                else if (mv == AL_UP) { partner[i - 1] = j > 0 ? j - 1 : 0; --i; }
                else --j;
            }
        }
        __syncwarp();
    }
}

// Move table and dwells (one warp per read): moves[n][t] = 1 where row t emitted a base, dwell[k] = rows from the
// k-th emission to the next one (to the end of the read for the last).
__global__ void __launch_bounds__(256) moves_kernel(const MovesParams p) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long n = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps) {
        const int len = p.status[n] == FCTC_OK ? p.lens[n] : 0;
        const int* path = p.path + p.offsets[n];
        if (p.moves) {
            uint8_t* mv = p.moves + n * p.T;
            for (long long t = lane; t < p.T; t += 32) mv[t] = 0;
            __syncwarp();
            for (int k = lane; k < len; k += 32) mv[path[k]] = 1;
        }
        if (p.dwell) {
            int* dw = p.dwell + p.offsets[n];
            for (int k = lane; k < len; k += 32) dw[k] = (k + 1 < len ? path[k + 1] : int(p.T)) - path[k];
        }
    }
}

}  // namespace

int64_t align_scratch_bytes(int sm_count, int64_t N, int64_t max_len, int band, int64_t budget_bytes, int* n_slots,
                            int64_t* dir_bytes_per_slot) {
    const int64_t per_slot = ((max_len + 1) * int64_t(2 * band + 1) + 255) & ~int64_t(255);
    int64_t slots = int64_t(sm_count) * 32;
    if (slots > N) slots = N;
    if (slots < 1) slots = 1;
    if (budget_bytes > 0 && slots * per_slot > budget_bytes) slots = budget_bytes / per_slot;
    if (n_slots) *n_slots = int(slots);
    if (dir_bytes_per_slot) *dir_bytes_per_slot = per_slot;
    return slots * per_slot;
}

cudaError_t launch_align(const AlignParams& p, cudaStream_t stream) {
    if (p.N <= 0 || p.n_slots <= 0) return cudaSuccess;
    if (p.band < 1 || (2 * p.band + 1 + 31) / 32 > AL_MAX_CPL) return cudaErrorInvalidValue;
    const int smem = AL_WARPS * 2 * (2 * p.band + 2) * int(sizeof(int));
    const int grid = (p.n_slots + AL_WARPS - 1) / AL_WARPS;
    align_kernel<<<grid, AL_WARPS * 32, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_moves(const MovesParams& p, int sm_count, cudaStream_t stream) {
    if (p.N <= 0) return cudaSuccess;
    long long blocks = (p.N + 7) / 8;
    if (blocks > (long long)sm_count * 8) blocks = (long long)sm_count * 8;
    moves_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace fctc
