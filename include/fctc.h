/* fctc.h — C ABI of libfctc_b200.so, the B200 (sm_100a) batched CTC decoder.
 *
 * This is the drop-in boundary for the decode path of nanoporetech/fast-ctc-decode v0.3.7: each
 * entry point below is what the reference's `src/lib.rs` would bind over `extern "C"` in place of
 * its calls into `src/search.rs` / `src/duplex.rs`.  The signatures use plain pointers and sizes
 * only (no C++/torch types), so they are bindable from Rust (`extern "C"` + build.rs linking the
 * library), Python (ctypes / the bundled `fast_ctc_decode` extension) or C.
 * Reference citations are file:line relative to the reference checkout.
 *
 * Conventions
 *   - Posteriors are float32, label axis innermost, blank label first (src/lib.rs:604-616).
 *   - Decoders return *alphabet indices* (1..n_labels-1, blank never appears); the caller maps
 *     them to label strings exactly as src/search.rs:291,358 do (labels may be multi-character).
 *   - Batch results are a pool: read n owns entries [offsets[n], offsets[n]+lens[n]) of `labels` /
 *     `quals` / `path`.  Viterbi uses fixed slots (offsets[n] = n * round_up(n_rows, 32), so the
 *     layout is fully deterministic and needs no atomics); beam search packs reads with one atomic
 *     add per read (rounded up to 32 entries), so there the *placement* of a read depends on
 *     scheduling while the *content* per read is deterministic.
 *   - *pool_used_dev (device entry points with a packed pool): one 64-bit counter in device memory.  Every call
 *     clears it on `stream` before its kernel runs and leaves the number of pool entries handed out, so the
 *     caller need not initialise it; results of successive calls go to separate pools (or the caller copies them
 *     out in between).
 *   - `path` entries are row indices into the read (int32 on the device, the reference's
 *     Vec<usize>): emission row for Viterbi, first-proposal row for beam search.
 *   - Return value: >= 0 is a per-call decode status for single-read calls (FCTC_OK or one of the
 *     reference's SearchError values); < 0 is a call-level failure, text in fctc_last_error().
 *   - Every call that takes a handle must be issued from one host thread at a time per handle
 *     (one handle per device / per host thread; src/lib.rs:199 releases the GIL the same way).
 *   - There is no CPU fallback: without a usable CUDA device fctc_create fails.
 */
#ifndef FCTC_H
#define FCTC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Per-read decode status — mirrors enum SearchError, src/lib.rs:36-58. */
#define FCTC_OK 0
#define FCTC_RAN_OUT_OF_BEAM 1      /* "Ran out of search space (beam_cut_threshold too high)" */
#define FCTC_INCOMPARABLE_VALUES 2  /* "Failed to compare values (NaNs in input?)" */
#define FCTC_INVALID_ENVELOPE 3     /* "Invalid envelope values" */
#define FCTC_NODE_POOL_OVERFLOW 4   /* device trie pool too small for this read (batch-device API only;
                                       the host entry points retry with a larger pool) */
#define FCTC_OUTPUT_OVERFLOW 5      /* result pool (pool_capacity) too small for this read */
#define FCTC_INVALID_QUALITY_CHAR 6 /* phred code is not a Unicode scalar: the reference's
                                       char::from_u32(..).unwrap() (src/search.rs:35) would abort */

#define FCTC_REFERENCE_PANIC 7      /* a condition on which the reference aborts the process: NaN met by a crf
                                       search's argmax().unwrap() (src/search.rs:404-412), or a crf state that
                                       leaves [0, n_state) */

/* Call-level failures. */
#define FCTC_E_CUDA (-1)
#define FCTC_E_ARGUMENT (-2)
#define FCTC_E_NOMEM (-3)
#define FCTC_E_UNSUPPORTED (-4)
#define FCTC_E_CAPACITY (-5)

typedef struct fctc_handle fctc_handle;

/* Library / device management (no reference counterpart: the reference is a stateless CPU
 * library; the handle owns the CUDA stream, pinned staging and device scratch of one device). */
int fctc_version(void);                      /* 10000*major + 100*minor + patch */
int fctc_device_count(void);                 /* 0 if no usable CUDA device */
int fctc_create(int device, fctc_handle** out);
void fctc_destroy(fctc_handle* h);
const char* fctc_last_error(const fctc_handle* h); /* h may be NULL: last fctc_create error */
/* SearchError Display text, src/lib.rs:43-56 (NULL for statuses that have no reference text). */
const char* fctc_status_message(int status);
int fctc_device_sm_count(const fctc_handle* h);

/* ------------------------------------------------------------------------------------------
 * Viterbi (greedy) decode — replaces search::viterbi_search, src/search.rs:320-383 (+ find_max
 * :303-318, phred :31-36) as called from src/lib.rs:200-207.
 *
 * fctc_viterbi_batch_device: posteriors already in HBM, C-contiguous [n_reads, n_rows, n_labels];
 *   results stay in HBM (pool form, see above).  Asynchronous on `stream` (a cudaStream_t; NULL is
 *   the legacy default stream, as everywhere in CUDA).  qual_bytes selects the quality element type: 1 = uint8
 *   (enough whenever 40*|qscale|+|qbias|+33 < 256), 4 = uint32 (any Unicode scalar, as the
 *   reference's `char`); 1 is refused (FCTC_E_ARGUMENT) when a code could exceed 255.  quals_dev
 *   may be NULL when qstring == 0.  pool_capacity (entries of labels_dev / quals_dev / path_dev)
 *   must be at least n_reads * round_up(n_rows, 32), else FCTC_E_CAPACITY.
 * ------------------------------------------------------------------------------------------ */
int fctc_viterbi_batch_device(fctc_handle* h, const float* post_dev, int64_t n_reads, int64_t n_rows,
                              int32_t n_labels, int32_t qstring, float qscale, float qbias,
                              int32_t collapse_repeats, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                              int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                              int32_t* status_dev, void* stream);

/* fctc_viterbi_batch: the same over HOST memory (new batched entry point of the north star):
 *   post_host is C-contiguous [n_reads, n_rows, n_labels]; outputs are host arrays in pool form
 *   (quals: qual_bytes 1 = uint8, 4 = uint32 Unicode scalars; path is int32).  Reads are streamed
 *   through the device in chunks, host<->device copies overlapped with decoding; synchronous.
 *   Pinned (page-locked) host buffers give full copy overlap; pageable ones work, slower.
 *   pool_capacity must be >= n_reads * round_up(n_rows, 32) (else FCTC_E_CAPACITY); *pool_used
 *   receives that product.  qual_bytes == 1 is
 *   refused (FCTC_E_ARGUMENT) when a quality code could exceed 255 for the given qscale/qbias. */
int fctc_viterbi_batch(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                       int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                       void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                       int32_t* lens, int32_t* status, int64_t* pool_used);

/* fctc_viterbi_batch_device_f16 / fctc_viterbi_batch_f16: the same two calls for posteriors stored as
 *   IEEE binary16 (the precision basecallers compute their scores in) — SURVEY.md section 8(f) rank 2,
 *   no counterpart in the reference, whose functions take float32 arrays only (src/lib.rs:171).  Every
 *   value is widened exactly to float32 before use, so the result is, bit for bit, what the reference
 *   returns for the widened array; what changes is the traffic: 10 instead of 20 bytes per row over
 *   PCIe and HBM.  Five labels with n_rows % 8 == 0 take the streaming kernel, other shapes a
 *   slower generic one. */
int fctc_viterbi_batch_device_f16(fctc_handle* h, const void* post_dev_f16, int64_t n_reads, int64_t n_rows,
                                  int32_t n_labels, int32_t qstring, float qscale, float qbias,
                                  int32_t collapse_repeats, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                                  int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                  int32_t* status_dev, void* stream);
int fctc_viterbi_batch_f16(fctc_handle* h, const void* post_host_f16, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                           int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                           void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                           int32_t* lens, int32_t* status, int64_t* pool_used);

/* fctc_viterbi_batch_packed / fctc_viterbi_batch_packed_f16: fctc_viterbi_batch(_f16) with the results packed
 *   before they cross PCIe — no counterpart in the reference (src/search.rs:320-383 decodes one read in host
 *   memory).  The slot form gives every read round_up(n_rows, 32) entries although it fills only its emitted
 *   labels (a third to two thirds of the rows), and fctc_viterbi_batch copies whole slots back.  Here the used
 *   prefixes are moved together on the device first: offsets[n] is the exclusive sum of round_up(lens[m], 32)
 *   over m < n (read order, deterministic; entries between a read's end and the next start are 0), *pool_used
 *   the total, and only that many entries of labels / quals / path are written.  pool_capacity may be anything
 *   >= the total (n_reads * round_up(n_rows, 32) always suffices); FCTC_E_CAPACITY otherwise.  Same arguments
 *   and per-read results as fctc_viterbi_batch; about 1.5x less device-to-host traffic. */
int fctc_viterbi_batch_packed(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                              int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                              void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                              int32_t* lens, int32_t* status, int64_t* pool_used);
int fctc_viterbi_batch_packed_f16(fctc_handle* h, const void* post_host_f16, int64_t n_reads, int64_t n_rows,
                                  int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                                  uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity,
                                  int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used);

/* ------------------------------------------------------------------------------------------
 * Other input kinds — SURVEY.md section 8(f) rank 2; no counterpart in the reference, whose functions take float32
 * posteriors only (src/lib.rs:171-188) although the caller's previous step is a softmax over the labels.
 * input_kind = element type | what the values are:
 *   FCTC_IN_F32 / FCTC_IN_F16 (IEEE binary16) / FCTC_IN_BF16 (bfloat16): 16-bit values are widened exactly to
 *     float32 as they are fetched, so results equal the reference's on the widened array at half the traffic;
 *   | FCTC_IN_LOGITS_EXACT or | FCTC_IN_LOGITS_FAST: the array holds logits (or log-probabilities), and the softmax
 *     over the label axis is applied in the kernels' load stage instead of by a kernel of the caller's.  The recipe
 *     is fixed — m = first maximum of the row, e_c = exp(z_c - m), S = e_0 + e_1 + ... in index order, p_c = e_c / S,
 *     all float32.  EXACT evaluates exp as glibc's expf does and divides with IEEE rounding, so p is bit for bit
 *     what the recipe gives on a glibc host and the decode equals the reference's on that host-side softmax; FAST
 *     uses the hardware's 2^x and one reciprocal (a few ulp off, far cheaper; a Viterbi decode without quality
 *     strings then needs no softmax at all, since it keeps the order of the values).
 * The four calls below are fctc_viterbi_batch_device, fctc_viterbi_batch_packed (packed results),
 * fctc_beam_search_batch_device and fctc_beam_search_batch with that extra argument; everything else is unchanged.
 * The streaming Viterbi kernel takes five labels with n_rows % 4 == 0 (float32) or % 8 == 0 (16-bit); other shapes
 * run a slower generic kernel.
 * ------------------------------------------------------------------------------------------ */
#define FCTC_IN_F32 0
#define FCTC_IN_F16 1
#define FCTC_IN_BF16 2
#define FCTC_IN_LOGITS_EXACT 0x10
#define FCTC_IN_LOGITS_FAST 0x20
int fctc_viterbi_batch_device_in(fctc_handle* h, const void* in_dev, int32_t input_kind, int64_t n_reads, int64_t n_rows,
                                 int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                                 uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes, int32_t* path_dev,
                                 int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                 void* stream);
int fctc_viterbi_batch_in(fctc_handle* h, const void* in_host, int32_t input_kind, int64_t n_reads, int64_t n_rows,
                          int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                          uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity,
                          int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used);
int fctc_beam_search_batch_device_in(fctc_handle* h, const void* in_dev, int32_t input_kind, int64_t n_reads,
                                     int64_t n_rows, int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                     int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                     int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                     unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
                                     int32_t nodes_per_row, void* stream);
int fctc_beam_search_batch_in(fctc_handle* h, const void* in_host, int32_t input_kind, int64_t n_reads, int64_t n_rows,
                              int32_t n_labels, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                              uint8_t* labels, int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                              int32_t* status, int64_t* pool_used);

/* fctc_viterbi_search: one read, any strides (in elements) — the shape of the pyo3 entry point
 *   src/lib.rs:171-212.  labels/quals/path hold `capacity` entries (n_rows always suffices).
 *   Returns the decode status; n_rows == 0 or n_labels == 0 gives FCTC_E_ARGUMENT (the reference
 *   aborts on its asserts, src/search.rs:328-331). */
int fctc_viterbi_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_labels,
                        int64_t stride_row, int64_t stride_label, int32_t qstring, float qscale, float qbias,
                        int32_t collapse_repeats, uint8_t* labels, uint32_t* quals, int64_t* path, int64_t capacity,
                        int64_t* out_len);

/* ------------------------------------------------------------------------------------------
 * 1-D CTC prefix beam search — replaces search::beam_search, src/search.rs:159-301, and its
 * support types SearchPoint (:7-28), tree::SuffixTree (src/tree.rs:17-194) and vec2d::Vec2D
 * (src/vec2d.rs:5-41), as called from src/lib.rs:354-360.  Argument validation (beam_size >= 1,
 * 0 <= threshold < 1/n_labels; src/lib.rs:333-349) is the caller's, as in the reference, but is
 * re-checked here and reported as FCTC_E_ARGUMENT.
 *
 * The device API needs trie scratch at scratch_dev (256-byte aligned): one pool of nodes_per_row
 * proposals per posterior row for every read in flight (hard bound (n_labels-1)*beam_size; 0
 * selects a default that the host entry points grow on FCTC_NODE_POOL_OVERFLOW).
 * fctc_beam_scratch_bytes() is the recommended size (a pool per resident warp, capped at 96 GiB or 70 % of the device's memory, whichever is more);
 * a smaller buffer is accepted — the launch then keeps fewer reads in flight (at least 4 pools
 * must fit, else FCTC_E_CAPACITY).
 * ------------------------------------------------------------------------------------------ */
int64_t fctc_beam_scratch_bytes(const fctc_handle* h, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                                int32_t beam_size, int32_t nodes_per_row);
int fctc_beam_search_batch_device(fctc_handle* h, const float* post_dev, int64_t n_reads, int64_t n_rows,
                                  int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                  int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                  int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                  unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
                                  int32_t nodes_per_row, void* stream);
/* fctc_beam_search_batch_device_f16 / fctc_beam_search_batch_f16: fctc_beam_search_batch_device and
 *   fctc_beam_search_batch for binary16 posteriors (no counterpart in the
 *   reference, which takes float32 only).  Values are widened exactly to float32 as they are fetched,
 *   so sequences and paths equal the reference's on the widened array. */
int fctc_beam_search_batch_device_f16(fctc_handle* h, const void* post_dev_f16, int64_t n_reads, int64_t n_rows,
                                      int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                      int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                      int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                      int32_t* status_dev, unsigned long long* pool_used_dev, void* scratch_dev,
                                      int64_t scratch_bytes, int32_t nodes_per_row, void* stream);
int fctc_beam_search_batch_f16(fctc_handle* h, const void* post_host_f16, int64_t n_reads, int64_t n_rows,
                               int32_t n_labels, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                               uint8_t* labels, int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                               int32_t* status, int64_t* pool_used);

/* fctc_beam_search_batch_device_static — no counterpart in the reference.  The same decode as
 * fctc_beam_search_batch_device for callers that cannot allocate device memory when they call
 * (all HBM pre-assigned to their own arena, allocator calls forbidden on the calling thread, ...):
 * the trie scratch is a fixed FCTC_BEAM_STATIC_SCRATCH_BYTES array inside the library's device
 * image, so the call performs no cudaMalloc/cudaFree.  That array occupies HBM on every GPU that
 * loads the library, until the process exits.  The launch uses as many trie pools as fit (fewer
 * resident warps, never more memory); if not even 4 pools of nodes_per_row * n_rows nodes fit, the
 * call fails with FCTC_E_CAPACITY and a message naming the sizes.  The first call on each device
 * looks the array's device address up (cudaGetSymbolAddress) — issue it before any stream capture.
 * All calls on a device share the array: the launch itself clears the only part that must start at
 * zero, but calls that can overlap on the GPU corrupt each other and must be serialised. */
#define FCTC_BEAM_STATIC_SCRATCH_BYTES (64ll << 20)
int fctc_beam_search_batch_device_static(fctc_handle* h, const float* post_dev, int64_t n_reads, int64_t n_rows,
                                         int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                         int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                         int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                         int32_t* status_dev, unsigned long long* pool_used_dev,
                                         int32_t nodes_per_row, void* stream);
int fctc_beam_search_batch(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                           int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats, uint8_t* labels,
                           int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens, int32_t* status,
                           int64_t* pool_used);
int fctc_beam_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_labels, int64_t stride_row,
                     int64_t stride_label, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                     uint8_t* labels, int64_t* path, int64_t capacity, int64_t* out_len);

/* ------------------------------------------------------------------------------------------
 * Several GPUs in one call — the batched form of the reference's threading affordance: its functions release
 * the GIL (src/lib.rs:199, :353) so that callers decode reads on several threads; here the reads of ONE host array
 * are cut into n_devices contiguous ranges [g*N/G, (g+1)*N/G), range g is decoded on devices[g] by its own host
 * thread (the library keeps one handle per device ordinal for these calls, created on first use), and the
 * results land in the caller's arrays in read order.  Reads are independent: no data moves between devices.
 * A device may be listed more than once (its ranges then run one after the other).  Arguments and per-read
 * results as fctc_viterbi_batch / fctc_beam_search_batch; pool_capacity must be n_reads * round_up(n_rows, 32).
 * Viterbi results are in slot form (offsets[n] = n * round_up(n_rows, 32)); beam-search range g fills the pool
 * from entry lo_g * round_up(n_rows, 32) on, *pool_used is the number of entries handed out over all ranges.
 * Failures: < 0 with the text in fctc_multi_last_error() (per calling thread).  fctc_multi_release() destroys
 * the cached handles (optional; call it when no multi call is running).
 * ------------------------------------------------------------------------------------------ */
int fctc_viterbi_batch_multi(const int32_t* devices, int32_t n_devices, const float* post_host, int64_t n_reads,
                             int64_t n_rows, int32_t n_labels, int32_t qstring, float qscale, float qbias,
                             int32_t collapse_repeats, uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path,
                             int64_t pool_capacity, int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used);
int fctc_beam_search_batch_multi(const int32_t* devices, int32_t n_devices, const float* post_host, int64_t n_reads,
                                 int64_t n_rows, int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                 int32_t collapse_repeats, uint8_t* labels, int32_t* path, int64_t pool_capacity,
                                 int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used);
const char* fctc_multi_last_error(void);
void fctc_multi_release(void);

/* ------------------------------------------------------------------------------------------
 * 2-D ("duplex") beam search — replaces duplex::beam_search, src/duplex.rs:443-650, with LogSpace
 * (:7-80; fastexp src/fastexp.rs), ProbPair/SearchPoint/SecondaryProbs (:84-210), build_/
 * extend_secondary_probs (:212-249, :338-387) and root_probs (:389-409), as called from
 * src/lib.rs:475-483.  Two posterior arrays per pair, [n_rows_1, C] and [n_rows_2, C]; `envelope`
 * gives, per network-1 row, the [lower, upper) network-2 rows to align against ([n_pairs,
 * n_rows_1, 2]; NULL = all of network 2, src/lib.rs:457-469).  Output: labels only (the reference
 * returns just the sequence).  exp_mode 0 = exact exp (the --no-default-features build, the parity
 * target; glibc's expf and log1pf are evaluated operation by operation, so results are bit-identical to
 * a glibc host), 1 = the reference's default build (its fastexp() evaluates to 0, so '+' is max),
 * 2 = exact exp through CUDA's own expf / log1pf (faster; last-bit differences from a glibc host can flip
 * a near-tie).
 * beam_size <= 32 in this release.  max_window (device API): the widest upper-lower of the
 * envelope, 0 = n_rows_2; a row wider than that gets FCTC_NODE_POOL_OVERFLOW.  The device API
 * uses as many pair slots as scratch_bytes holds (at least 4 must fit).
 * ------------------------------------------------------------------------------------------ */
int64_t fctc_duplex_scratch_bytes(const fctc_handle* h, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                  int32_t n_labels, int32_t beam_size, int32_t max_window, int32_t nodes_per_row);
int fctc_beam_search_duplex_batch_device(fctc_handle* h, const float* post1_dev, const float* post2_dev,
                                         int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels,
                                         const int64_t* envelope_dev, int32_t max_window, int32_t beam_size,
                                         float beam_cut_threshold, int32_t collapse_repeats, int32_t exp_mode,
                                         uint8_t* labels_dev, int64_t pool_capacity, int64_t* offsets_dev,
                                         int32_t* lens_dev, int32_t* status_dev, unsigned long long* pool_used_dev,
                                         void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row, void* stream);

/* fctc_beam_search_duplex_batch_device_f16: the device entry point above for binary16 posteriors
 * (no counterpart in the reference).  Both networks are widened
 * exactly to float32 as they are read, before the log mapping. */
int fctc_beam_search_duplex_batch_device_f16(fctc_handle* h, const void* post1_dev_f16, const void* post2_dev_f16,
                                             int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels,
                                             const int64_t* envelope_dev, int32_t max_window, int32_t beam_size,
                                             float beam_cut_threshold, int32_t collapse_repeats, int32_t exp_mode,
                                             uint8_t* labels_dev, int64_t pool_capacity, int64_t* offsets_dev,
                                             int32_t* lens_dev, int32_t* status_dev, unsigned long long* pool_used_dev,
                                             void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row,
                                             void* stream);


int fctc_beam_search_duplex_batch(fctc_handle* h, const float* post1_host, const float* post2_host, int64_t n_pairs,
                                  int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels, const int64_t* envelope_host,
                                  int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                                  int32_t exp_mode, uint8_t* labels, int64_t pool_capacity, int64_t* offsets,
                                  int32_t* lens, int32_t* status, int64_t* pool_used);
/* The host batch call for binary16 posteriors (both networks; widened exactly as they are read): half the PCIe bytes.
 * No counterpart in the reference. */
int fctc_beam_search_duplex_batch_f16(fctc_handle* h, const void* post1_host_f16, const void* post2_host_f16,
                                      int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t n_labels,
                                      const int64_t* envelope_host, int32_t beam_size, float beam_cut_threshold,
                                      int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels, int64_t pool_capacity,
                                      int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used);
/* One pair, any strides (elements) — the shape of the pyo3 entry point src/lib.rs:402-488;
 * envelope is the reference's usize array [n_rows_1, 2] (strides in elements) or NULL. */
int fctc_beam_search_duplex(fctc_handle* h, const float* post1_host, int64_t n_rows_1, int64_t stride_row_1,
                            int64_t stride_label_1, const float* post2_host, int64_t n_rows_2, int64_t stride_row_2,
                            int64_t stride_label_2, int32_t n_labels, const uint64_t* envelope,
                            int64_t env_stride_row, int64_t env_stride_col, int32_t beam_size,
                            float beam_cut_threshold, int32_t collapse_repeats, int32_t exp_mode, uint8_t* labels,
                            int64_t capacity, int64_t* out_len);

/* ------------------------------------------------------------------------------------------
 * State-conditioned ("crf") decoders over [n_rows, n_state, n_labels] posteriors plus an initial
 * state score vector — replace search::crf_greedy_search (src/search.rs:385-423) and
 * search::crf_beam_search (src/search.rs:38-157) as called from src/lib.rs:238-245, 275-281.
 * A label moves the state to (state * (n_labels-1)) % n_state + label.  NaNs (the reference's
 * argmax().unwrap()) and states outside [0, n_state) abort the reference; here the read gets
 * FCTC_REFERENCE_PANIC.  Greedy results are in slot form (as Viterbi), beam results in the packed
 * pool (as beam search; scratch from fctc_beam_scratch_bytes).  Host arrays of the single-read
 * calls must be C-contiguous.  crf_beam_search_duplex (src/duplex.rs:652-834) is the last pair of declarations
 * below: built on the duplex kernel; the reference holds no test vector for it, so its results are checked
 * against this repository's CPU restatement only (parity unpinned by the reference).
 * ------------------------------------------------------------------------------------------ */
int fctc_crf_greedy_batch_device(fctc_handle* h, const float* post_dev, const float* init_dev, int64_t n_reads,
                                 int64_t n_rows, int32_t n_state, int32_t n_labels, int32_t n_init, int32_t qstring,
                                 float qscale, float qbias, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                                 int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                 int32_t* status_dev, void* stream);
int fctc_crf_beam_search_batch_device(fctc_handle* h, const float* post_dev, const float* init_dev, int64_t n_reads,
                                      int64_t n_rows, int32_t n_state, int32_t n_labels, int32_t n_init,
                                      int32_t beam_size, float beam_cut_threshold, uint8_t* labels_dev,
                                      int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev,
                                      int32_t* lens_dev, int32_t* status_dev, unsigned long long* pool_used_dev,
                                      void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row, void* stream);
/* The same two searches over 16-bit posteriors as the network's last layer leaves them: input_kind is FCTC_IN_F32,
 * FCTC_IN_F16 or FCTC_IN_BF16 (no logits kinds here), every value is widened exactly to float32 where it is read
 * (the searches read one state's row per step, in place), so the results equal those of the float32 call on the
 * widened array.  init_dev stays float32 (n_init values per read). */
int fctc_crf_greedy_batch_device_in(fctc_handle* h, const void* post_dev, int32_t input_kind, const float* init_dev,
                                    int64_t n_reads, int64_t n_rows, int32_t n_state, int32_t n_labels, int32_t n_init,
                                    int32_t qstring, float qscale, float qbias, uint8_t* labels_dev, void* quals_dev,
                                    int32_t qual_bytes, int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev,
                                    int32_t* lens_dev, int32_t* status_dev, void* stream);
int fctc_crf_beam_search_batch_device_in(fctc_handle* h, const void* post_dev, int32_t input_kind, const float* init_dev,
                                         int64_t n_reads, int64_t n_rows, int32_t n_state, int32_t n_labels,
                                         int32_t n_init, int32_t beam_size, float beam_cut_threshold, uint8_t* labels_dev,
                                         int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev,
                                         int32_t* lens_dev, int32_t* status_dev, unsigned long long* pool_used_dev,
                                         void* scratch_dev, int64_t scratch_bytes, int32_t nodes_per_row, void* stream);
int fctc_crf_greedy_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_state, int32_t n_labels,
                           const float* init_host, int32_t n_init, int32_t qstring, float qscale, float qbias,
                           uint8_t* labels, uint32_t* quals, int64_t* path, int64_t capacity, int64_t* out_len);
int fctc_crf_beam_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_state, int32_t n_labels,
                         const float* init_host, int32_t n_init, int32_t beam_size, float beam_cut_threshold,
                         uint8_t* labels, int64_t* path, int64_t capacity, int64_t* out_len);

/* Scratch for the crf duplex device call (one ring per trie node: larger than fctc_duplex_scratch_bytes' plan). */
int64_t fctc_crf_duplex_scratch_bytes(const fctc_handle* h, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                      int32_t n_labels, int32_t beam_size, int32_t max_window, int32_t nodes_per_row);
int fctc_crf_beam_search_duplex_batch_device(fctc_handle* h, const float* post1_dev, const float* init1_dev,
                                             int32_t n_init1, const float* post2_dev, const float* init2_dev,
                                             int32_t n_init2, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                             int32_t n_state, int32_t n_labels, const int64_t* envelope_dev,
                                             int32_t max_window, int32_t beam_size, float beam_cut_threshold,
                                             int32_t exp_mode, uint8_t* labels_dev, int64_t pool_capacity,
                                             int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                             unsigned long long* pool_used_dev, void* scratch_dev,
                                             int64_t scratch_bytes, int32_t nodes_per_row, void* stream);
/* input_kind FCTC_IN_F32 or FCTC_IN_F16 (both networks alike), as fctc_beam_search_duplex_batch_device_f16. */
int fctc_crf_beam_search_duplex_batch_device_in(fctc_handle* h, const void* post1_dev, const float* init1_dev,
                                                int32_t n_init1, const void* post2_dev, const float* init2_dev,
                                                int32_t n_init2, int32_t input_kind, int64_t n_pairs, int64_t n_rows_1,
                                                int64_t n_rows_2, int32_t n_state, int32_t n_labels,
                                                const int64_t* envelope_dev, int32_t max_window, int32_t beam_size,
                                                float beam_cut_threshold, int32_t exp_mode, uint8_t* labels_dev,
                                                int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                                int32_t* status_dev, unsigned long long* pool_used_dev, void* scratch_dev,
                                                int64_t scratch_bytes, int32_t nodes_per_row, void* stream);
int fctc_crf_beam_search_duplex(fctc_handle* h, const float* post1_host, int64_t n_rows_1, const float* init1_host,
                                int32_t n_init1, const float* post2_host, int64_t n_rows_2, const float* init2_host,
                                int32_t n_init2, int32_t n_state, int32_t n_labels, const uint64_t* envelope,
                                int32_t beam_size, float beam_cut_threshold, int32_t exp_mode, uint8_t* labels,
                                int64_t capacity, int64_t* out_len);

/* ------------------------------------------------------------------------------------------
 * FASTQ record assembly on the device — SURVEY.md section 8(f) rank 3; no counterpart in the
 * reference, which returns sequence and quality as one concatenated string
 * (src/search.rs:378-380) and leaves records to the caller.  Input is the slot-form result of
 * fctc_viterbi_batch_device with qstring = 1 and qual_bytes = 1; output is the text
 *     @<name_prefix><first_index + n>\n<sequence>\n+\n<quality>\n
 * for every read whose status is FCTC_OK (others produce no record), concatenated in read order
 * in out_dev.  alphabet gives one character per label index (n_labels of them; index 0, the
 * blank, is never emitted).  record_offsets_dev[n_reads + 1] receives each record's first byte
 * and, last, the total length; when that total exceeds out_capacity nothing is written and the
 * caller learns the size it needs from that last entry (sum of 2*len + 7 + strlen(prefix) +
 * digits per read bounds it).  scratch_dev: fctc_fastq_scratch_bytes(n_reads) bytes, 8-byte
 * aligned.  Asynchronous on `stream`.
 * ------------------------------------------------------------------------------------------ */
int64_t fctc_fastq_scratch_bytes(const fctc_handle* h, int64_t n_reads);
int fctc_fastq_records_device(fctc_handle* h, const uint8_t* labels_dev, const uint8_t* quals_dev,
                              const int64_t* offsets_dev, const int32_t* lens_dev, const int32_t* status_dev,
                              int64_t n_reads, const char* alphabet, int32_t n_labels, const char* name_prefix,
                              int64_t first_index, uint8_t* out_dev, int64_t out_capacity, int64_t* record_offsets_dev,
                              void* scratch_dev, int64_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Duplex envelope from the two reads' Viterbi paths — SURVEY.md section 8(f) rank 4; no counterpart
 * in the reference, which leaves the envelope to the caller (src/lib.rs:376-378).  Inputs are the
 * slot-form path / offsets / lens arrays of fctc_viterbi_batch_device run on both networks of every
 * pair.  No aligner: base k of read 1 is anchored to base (k * len2) / len1 of read 2, and row t of
 * network 1 gets the window [c - pad_rows, c + pad_rows + 1) around the network-2 row c of its
 * anchor (clamped to [0, n_rows_2), lower bound pulled down to the previous row's upper bound so
 * that the envelope passes the reference's validity rule, src/duplex.rs:484-488; pairs with an
 * empty read get the plain diagonal band).  Where the anchor jumps (a stretch of read 2 without
 * bases) a window is wider than 2 * pad_rows + 1; max_window_dev (optional, one int32 on the
 * device) receives the widest window of the batch, the max_window to hand to
 * fctc_beam_search_duplex_batch_device.
 * envelope_dev: int64 [n_pairs, n_rows_1, 2], ready for fctc_beam_search_duplex_batch_device.
 * ------------------------------------------------------------------------------------------ */
int fctc_duplex_envelope_device(fctc_handle* h, const int32_t* path1_dev, const int64_t* offsets1_dev,
                                const int32_t* lens1_dev, const int32_t* path2_dev, const int64_t* offsets2_dev,
                                const int32_t* lens2_dev, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                int32_t pad_rows, int64_t* envelope_dev, int32_t* max_window_dev, void* stream);

/* ------------------------------------------------------------------------------------------
 * Duplex envelope from a banded alignment of the two reads' Viterbi label strings — SURVEY.md section 8(f) rank 4 as
 * written; no counterpart in the reference (src/lib.rs:376-378 leaves the envelope to the caller, src/duplex.rs:484-488
 * only checks it).  fctc_duplex_envelope_device above anchors base k of read 1 to base k*L2/L1 of read 2, which
 * drifts by as many bases as one read has lost or gained up to that point and forces wide windows (the 2-D
 * search's cost is linear in the window).  Here the label strings are aligned first (unit-cost edit distance inside
 * a band of +-band_bases around the proportional diagonal, ties diagonal > up > left), base k is anchored to its
 * aligned partner, and the window rule is the same as above.  Inputs: the slot-form labels / path / offsets / lens
 * of fctc_viterbi_batch_device on both networks.  partner_dev: int32, read 1's pool layout (entry offsets1[n] + k =
 * index in read 2 of the partner of base k), also an output.  scratch_dev: fctc_align_scratch_bytes(h, n_pairs,
 * n_rows_1, band_bases) bytes for all pairs in flight; less is accepted (fewer pairs at a time), below one pair's
 * worth FCTC_E_CAPACITY.  Pairs whose lengths differ by more than the band keep the proportional anchor.
 * ------------------------------------------------------------------------------------------ */
int64_t fctc_align_scratch_bytes(const fctc_handle* h, int64_t n_pairs, int64_t max_bases, int32_t band_bases);
int fctc_duplex_envelope_aligned_device(fctc_handle* h, const uint8_t* labels1_dev, const int32_t* path1_dev,
                                        const int64_t* offsets1_dev, const int32_t* lens1_dev, const uint8_t* labels2_dev,
                                        const int32_t* path2_dev, const int64_t* offsets2_dev, const int32_t* lens2_dev,
                                        int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t band_bases,
                                        int32_t pad_rows, int64_t* envelope_dev, int32_t* max_window_dev,
                                        int32_t* partner_dev, void* scratch_dev, int64_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Move table and dwells from the emission rows — SURVEY.md section 8(f) rank 3, second half; no counterpart in the
 * reference, which returns the rows (`path`, src/search.rs:378-380) and leaves derived formats to the caller.
 * moves_dev (optional): uint8 [n_reads, n_rows], 1 where the row emitted a base (the basecallers' "move table");
 * dwell_dev (optional): int32 in the layout of path_dev, rows from a base's emission to the next one (to the end of
 * the read for the last base).  Reads whose status is not FCTC_OK give an all-zero move row and no dwells.
 * ------------------------------------------------------------------------------------------ */
int fctc_move_table_device(fctc_handle* h, const int32_t* path_dev, const int64_t* offsets_dev, const int32_t* lens_dev,
                           const int32_t* status_dev, int64_t n_reads, int64_t n_rows, uint8_t* moves_dev,
                           int32_t* dwell_dev, void* stream);

/* ------------------------------------------------------------------------------------------
 * Synthetic posterior generator used by bench.py and the tests (no reference counterpart; it
 * implements SURVEY.md §8d's input definition): fills post_dev[n_reads, n_rows, n_labels] with
 * kind 0: softmax(3*(2u-1)) rows, kind 1: peaked rows (dominant label, blank w.p. 0.65), kind 2: kind
 * 1's dominant-label track with an independent noise stream (the second read of a duplex pair),
 * u from a counter-based generator keyed (seed, first_read + read, row).  Asynchronous on stream.
 * ------------------------------------------------------------------------------------------ */
int fctc_synth_posteriors_device(fctc_handle* h, float* post_dev, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                                 int32_t kind, uint64_t seed, int64_t first_read, void* stream);

/* Test hook (replaces nothing): fills every SM's shared memory with `byte` and waits, so that the next
 * kernels start from a chosen pattern (0xFF = -1 / NaN words) instead of what earlier kernels left.
 * The GPU test-suite runs every test behind it: no kernel may depend on stale shared memory. */
int fctc_test_poison_shared_memory(fctc_handle* h, int32_t byte);

/* Test hook (replaces nothing): runs the DEVICE builds of the libm restatements over an array — which = 0 expf (x <= 0),
 * 1 log1pf (0 <= x <= 1), 2 logf, 3 log10f, 4 LogSpace '+' of (x[i], y[i]) in exact-exp mode — so that tests compare the
 * arithmetic the kernels execute with the host's libm directly; 5: the quality value of the Viterbi kernel's fast path
 * (hardware lg2) minus the reference's (glibc log10f and three rounded operations) for the miss probability x[i], with
 * y = {qscale, qbias}; 6: the distance from a rounding boundary below which the fast path defers to the exact one at
 * that value.  y_dev may be NULL unless which >= 4. */
int fctc_test_libm_device(fctc_handle* h, const float* x_dev, const float* y_dev, float* out_dev, int64_t n, int32_t which,
                          void* stream);

/* Host instantiation of the device phred()/log10f code (same source, compiled for the host), so
 * that CPU tests can compare it with glibc without a GPU.  Replaces nothing; test hook. */
uint32_t fctc_host_phred(float prob, float qscale, float qbias);
void fctc_host_log10f(const float* x, float* y, int64_t n);
void fctc_host_logf(const float* x, float* y, int64_t n);
void fctc_host_expf(const float* x, float* y, int64_t n);   /* the LogSpace '+' pair of the duplex search */
void fctc_host_log1pf(const float* x, float* y, int64_t n);
void fctc_host_expf_nonpos(const float* x, float* y, int64_t n);  /* the branch-free forms the kernels run: x <= 0 or NaN */
void fctc_host_log1pf_unit(const float* x, float* y, int64_t n);  /* ... 0 <= x <= 1 or NaN */

#ifdef __cplusplus
}
#endif
#endif /* FCTC_H */
