// kernels.h — host-side launch interface of the CUDA kernels (internal to libfctc_b200.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct fctc_handle;

namespace fctc {

// Handle internals needed by entry points that live outside api.cu.
int handle_device(const fctc_handle* h);
int handle_sm_count(const fctc_handle* h);
int handle_fail(fctc_handle* h, int code, const char* message);
int check_beam_call(fctc_handle* h, int64_t n_reads, int64_t n_rows, int32_t n_labels, int32_t beam_size, float thr);
int check_threshold_call(fctc_handle* h, int32_t n_labels, int32_t beam_size, float thr);  // src/lib.rs:433-443
// Persistent, growable device / pinned-host buffers owned by the handle (index 0..15 each).
void* handle_device_buffer(fctc_handle* h, int index, size_t bytes);  // NULL on failure (error text set)
size_t handle_device_buffer_bytes(const fctc_handle* h, int index);   // current size of a persistent buffer
void* handle_pinned_buffer(fctc_handle* h, int index, size_t bytes);
cudaStream_t handle_stream(fctc_handle* h);
cudaStream_t handle_aux_stream(fctc_handle* h);  // a second stream of the handle, for copies that run under a kernel

// Integer tuning knob from the environment (development aid), 0 when unset or not positive.  Each call site keeps
// the value in a function-local static, so the environment is read once per process, not once per launch.
int env_knob(const char* name);

struct ViterbiParams {
    const float* post;  // [N, T, C] contiguous, device (binary16 values behind the same pointer when in_f16)
    int in_f16;         // 1: the values are 16 bits wide (IEEE binary16, or bfloat16 with in_bf16); widened exactly to f32 before use
    int in_bf16;        // with in_f16: the 16-bit values are bfloat16
    int softmax;        // 0: the input holds posteriors; 1 / 2: logits, softmax in the load stage (input.cuh: exact / fast)
    int64_t N, T;
    int C;
    int qstring;
    float qscale, qbias;
    int collapse;
    uint8_t* labels;  // result pool in slot form: read n owns [n*slot_stride, n*slot_stride + lens[n])
    void* quals;      // uint8_t* or uint32_t* (qual_bytes)
    int qual_bytes;
    int32_t* path;
    int64_t slot_stride;  // round_up(T, 32)
    int64_t offset_base;  // offsets[n] = offset_base + n*slot_stride (lets a chunk report global offsets)
    int64_t* offsets;     // [N]
    int32_t* lens;        // [N]
    int32_t* status;      // [N]
};

// Chooses the streaming (bulk-async staged) kernel when the shape allows it, else the generic
// one; returns cudaSuccess or the launch error.  `which` (optional) receives 0 = streaming,
// 1 = generic.
cudaError_t launch_viterbi(const ViterbiParams& p, int sm_count, cudaStream_t stream, int* which);

// Slot form -> packed pool (pack.cu): read n moves from [offsets[n] - offset_base, +lens[n]) of the slot arrays to
// [out_offsets[n], +lens[n]) of the packed ones, out_offsets = exclusive scan of round_up(lens, 32) in read order,
// *cursor = the total.
struct PackParams {
    const uint8_t* labels;
    const void* quals;  // may be null
    int qual_bytes;
    const int32_t* path;
    const int64_t* offsets;
    int64_t offset_base;
    const int32_t* lens;
    int64_t N;
    uint8_t* out_labels;
    void* out_quals;
    int32_t* out_path;
    int64_t* out_offsets;
    int64_t out_capacity;
    unsigned long long* cursor;
};
cudaError_t launch_pack(const PackParams& p, int sm_count, cudaStream_t stream);

struct BeamParams {
    const float* post;  // [N, T, C] (binary16 values behind the same pointer when in_f16)
    int in_f16;         // 1: 16-bit values (binary16, or bfloat16 with in_bf16), widened exactly to f32 as they are fetched (plain search only)
    int in_bf16;
    int softmax;        // 0 posteriors; 1 / 2: logits, softmax applied to every staged row (input.cuh: exact / fast)
    int64_t N, T;
    int C;
    int beam_size;
    float threshold;
    int collapse;
    uint8_t* labels;
    int32_t* path;
    int64_t pool_cap;
    int64_t* offsets;
    int32_t* lens;
    int32_t* status;
    unsigned long long* cursor;
    void* scratch;          // trie pools, one slot per resident warp
    int64_t scratch_bytes;
    int64_t nodes_per_slot;  // node capacity of one slot
    int n_slots;             // slots available in scratch
    unsigned long long* ticket;  // read counter for dynamic scheduling (inside scratch)
    // Launches that share one scratch and run at the same time (the host pipeline's consecutive chunks, wide beams):
    // a warp takes its trie pool from a ring of free slot numbers and puts it back when it has no read left, so the
    // next launch's CTAs fill the SMs as this launch's last reads end.  Null: slot = position in the grid.
    unsigned int* slot_ring;     // [ring_size] slot number + 1, 0 = taken
    unsigned int* ring_ctr;      // [0] takes so far, [1] returns so far
    int ring_size;
    // crf variant (search::crf_beam_search): S > 0 selects it; post is [N, T, S, C]
    int S;
    const float* init;  // [N, n_init] initial state scores
    int n_init;
};

// budget_bytes > 0: use at most that much (fewer pool slots = fewer reads in flight); 0: the full plan.
int64_t beam_scratch_bytes(int sm_count, int64_t N, int64_t T, int C, int beam_size, int nodes_per_row,
                           int64_t budget_bytes, int64_t* nodes_per_slot, int* n_slots);
cudaError_t launch_beam(const BeamParams& p, int sm_count, cudaStream_t stream);
// fills a slot ring with every slot number and clears its two counters (ring_ctr points at two words)
cudaError_t launch_beam_ring_init(unsigned int* slot_ring, unsigned int* ring_ctr, int ring_size, cudaStream_t stream);
// the wide kernel's resident warps per SM for this shape (0 when the narrow-beam kernel takes it)
int beam_wide_warps_per_sm(int beam_size, int C, int64_t T);
bool beam_shape_supported(int beam_size, int C);
// bytes of one warp's trie pool for `cap` nodes, for the kernel that (beam_size, C, T) selects
int64_t beam_pool_bytes(int64_t cap, int beam_size, int C, int64_t T);
int beam_warps_per_cta(int beam_size, int C, int64_t T);
// narrow-beam kernel (beam_small.cu): one lane per candidate, beam_size * C <= 32
bool beam_small_eligible(int beam_size, int C, int64_t T);
int64_t beam_small_slot_bytes(int64_t cap);
int beam_small_warps_per_cta();
int beam_small_warps_per_sm();
cudaError_t launch_beam_small(const BeamParams& p, cudaStream_t stream);

struct DuplexParams {
    const float* post1;    // [N, T1, C]
    const float* post2;    // [N, T2, C]
    int in_f16;            // 1: both are binary16 behind the same pointers, widened exactly to f32 on load (plain search only)
    const long long* env;  // [N, T1, 2] network-2 row bounds per network-1 row, or NULL (= [0, T2) everywhere)
    int64_t N;
    int T1, T2, C;
    int beam_size;
    float threshold;
    int collapse;
    int exp_mode;  // 0 exact exp (parity target, glibc's expf/log1pf), 1 the reference's default build (fastexp == 0: '+' is max), 2 CUDA's expf/log1pf
    int W;         // ring size of a node's vector: max window + 1
    uint8_t* labels;
    int64_t pool_cap;
    int64_t* offsets;
    int32_t* lens;
    int32_t* status;
    unsigned long long* cursor;
    void* scratch;
    int64_t nodes_per_slot;
    int64_t n_rings;  // packed kernel: rings of SecondaryProbs cells per pair slot (a power of two, or nodes_per_slot)
    int n_slots;
    unsigned long long* ticket;
    // crf variant (duplex::crf_beam_search): S > 0 selects it; posts are [N, T, S, C]
    int S;
    const float* init1;  // [N, n_init1]
    const float* init2;  // [N, n_init2]
    int n_init1, n_init2;
};
bool duplex_shape_supported(int beam_size, int C);
// crf: the state-conditioned search (always duplex.cu); n_rings (optional) receives DuplexParams::n_rings
int64_t duplex_scratch_bytes(int sm_count, int64_t N, int T1, int T2, int C, int beam_size, int W, int nodes_per_row,
                             int64_t budget_bytes, int64_t* nodes_per_slot, int* n_slots, int crf = 0,
                             int64_t* n_rings = nullptr);
cudaError_t launch_duplex(const DuplexParams& p, cudaStream_t stream);
// duplex_packed.cu: four pairs per warp, beam_size <= 8, at most 8 labels, plain (non-crf) search; same scratch plan
bool duplex_packed_eligible(const DuplexParams& p);
bool duplex_packed_shape(int beam_size, int C, int crf);
int64_t duplex_packed_rings(int64_t cap, int nodes_per_row, int64_t bound);
int64_t duplex_packed_slot_bytes(int64_t cap, int64_t rings, int T2, int C, int W);
int duplex_packed_slots_per_sm(int exp_mode);  // exp_mode < 0: the largest over the modes (scratch planning)
cudaError_t launch_duplex_packed(const DuplexParams& p, cudaStream_t stream);

// FASTQ record assembly from the Viterbi slot form (fastq.cu)
struct FastqParams {
    const uint8_t* labels;   // slot-form labels (alphabet indices)
    const uint8_t* quals;    // slot-form quality characters (qual_bytes == 1)
    const int64_t* offsets;  // [N]
    const int32_t* lens;     // [N]
    const int32_t* status;   // [N]
    int64_t N;
    int64_t first_index;     // number of read 0 in the record names
    unsigned char* out;
    int64_t out_capacity;
    int prefix_len;
    char prefix[64];
    unsigned char alphabet[256];  // label index -> character
};
int64_t fastq_scratch_bytes(int64_t N);
cudaError_t launch_fastq(const FastqParams& p, long long* rec_off, void* scratch, cudaStream_t stream);

// duplex envelope from the two reads' Viterbi paths (fastq.cu)
struct EnvelopeParams {
    const int32_t* path1;    // slot-form emission rows of read 1 / read 2
    const int64_t* offsets1;
    const int32_t* lens1;
    const int32_t* path2;
    const int64_t* offsets2;
    const int32_t* lens2;
    int64_t N, T1, T2;
    int64_t pad;             // half-width of the window, in network-2 rows
    long long* env;          // [N, T1, 2]
    int* max_window;         // optional: receives the widest window (hi - lo) of the whole batch
    const int32_t* partner;  // optional, read 1's slot layout: base k of read 1 is anchored to base partner[k] of read 2
                             // (launch_align); null = the proportional anchor k * L2 / L1
};

// banded alignment of the pairs' label strings (align.cu)
struct AlignParams {
    const uint8_t* labels1;   // slot-form labels of read 1 / read 2
    const int64_t* offsets1;
    const int32_t* lens1;
    const uint8_t* labels2;
    const int64_t* offsets2;
    const int32_t* lens2;
    int64_t N;
    int band;                 // half-width of the band, in bases
    int32_t* partner;         // out, read 1's slot layout
    uint8_t* dirs;            // scratch: n_slots x dir_bytes_per_slot move bytes
    int64_t dir_bytes_per_slot;
    int n_slots;
};
int64_t align_scratch_bytes(int sm_count, int64_t N, int64_t max_len, int band, int64_t budget_bytes, int* n_slots,
                            int64_t* dir_bytes_per_slot);
cudaError_t launch_align(const AlignParams& p, cudaStream_t stream);

// move table / dwells from the emission rows (align.cu)
struct MovesParams {
    const int32_t* path;
    const int64_t* offsets;
    const int32_t* lens;
    const int32_t* status;
    int64_t N, T;
    uint8_t* moves;  // optional [N, T]
    int32_t* dwell;  // optional, same layout as path
};
cudaError_t launch_moves(const MovesParams& p, int sm_count, cudaStream_t stream);
cudaError_t launch_envelope(const EnvelopeParams& p, cudaStream_t stream);
cudaError_t launch_libm_probe(const float* x, const float* y, float* out, long long n, int which, cudaStream_t stream);
cudaError_t launch_poison_smem(int sm_count, int byte, unsigned int* sink, cudaStream_t stream);
cudaError_t launch_synth(float* post, int64_t N, int64_t T, int C, int kind, uint64_t seed, int64_t first_read,
                         cudaStream_t stream);

}  // namespace fctc
