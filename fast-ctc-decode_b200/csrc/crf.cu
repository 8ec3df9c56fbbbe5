// crf.cu — the state-conditioned ("crf") decoders: search::crf_greedy_search
// (src/search.rs:385-423) as a kernel of its own, and the C ABI of search::crf_beam_search
// (src/search.rs:38-157), whose kernel is the CRF instantiation of beam.cu's.
//
// Posteriors are [T, n_state, C]; a decode walks one state per time step:
//     label = argmax(probs[t][state][:]);  if label > 0: state = (state * n_base) % n_state + label - 1
// so each step's 4*C-byte row depends on the previous step's result — a pointer chase through
// time with no work to share inside a read.  The greedy kernel therefore gives every read one
// *thread* (thousands of independent chains in flight hide the HBM latency); quality characters
// are per emitted label (no run mean), from the exact phred path of phred.cuh.
// argmax / max are ndarray-stats' (first maximum wins; any NaN makes the reference abort through
// unwrap(), reported here as FCTC_REFERENCE_PANIC, as is a state that leaves [0, n_state)).
#include <algorithm>
#include <cstdio>

#include "device_utils.cuh"
#include "input.cuh"
#include "kernels.h"
#include "phred.cuh"

#include "../../include/fctc.h"

namespace fctc {

__constant__ double c_crf_invc[16] = FCTC_LOGF_INVC;
__constant__ double c_crf_logc[16] = FCTC_LOGF_LOGC;

struct CrfGreedyParams {
    const float* post;  // [N, T, S, C]; 16-bit values behind the same pointer when in_kind != 0
    int in_kind;        // FCTC_IN_F32 / F16 / BF16: 16-bit values are widened exactly to f32 as they are read
    const float* init;  // [N, n_init]
    int64_t N, T;
    int S, C, n_init;
    int qstring;
    float qscale, qbias;
    uint8_t* labels;
    void* quals;
    int qual_bytes;
    int32_t* path;
    int64_t slot_stride;
    int64_t* offsets;
    int32_t* lens;
    int32_t* status;
};

// first maximum; *nan set when any element is NaN (the reference's unwrap() would panic)
__device__ __forceinline__ void first_argmax(const float* v, int n, int& arg, float& best, bool& nan) {
    best = v[0];
    arg = 0;
    nan = nan || (best != best);
    for (int i = 1; i < n; ++i) {
        const float x = v[i];
        nan = nan || (x != x);
        if (x > best) { best = x; arg = i; }
    }
}

// the same over 16-bit values (binary16 or bfloat16), each widened exactly before it is compared
__device__ __forceinline__ void first_argmax16(const uint16_t* v, int n, bool bf16, int& arg, float& best, bool& nan) {
    best = widen16(v[0], bf16);
    arg = 0;
    nan = nan || (best != best);
    for (int i = 1; i < n; ++i) {
        const float x = widen16(v[i], bf16);
        nan = nan || (x != x);
        if (x > best) { best = x; arg = i; }
    }
}

template <typename QT>
__global__ void __launch_bounds__(128) crf_greedy_kernel(const CrfGreedyParams p) {
    __shared__ double s_invc[16], s_logc[16];
    if (threadIdx.x < 16) {
        s_invc[threadIdx.x] = c_crf_invc[threadIdx.x];
        s_logc[threadIdx.x] = c_crf_logc[threadIdx.x];
    }
    __syncthreads();
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const int S = p.S, C = p.C, n_base = C - 1;
    const float* rd = p.post + n * p.T * S * C;
    const uint16_t* rdh = reinterpret_cast<const uint16_t*>(p.post) + n * p.T * S * C;
    const bool wide = p.in_kind == 0, bf16 = p.in_kind == 2;
    uint8_t* out_lab = p.labels + n * p.slot_stride;
    QT* out_qual = reinterpret_cast<QT*>(p.quals) + n * p.slot_stride;
    int* out_path = p.path + n * p.slot_stride;
    bool panic = false, bad_char = false;
    int state;
    float dummy;
    first_argmax(p.init + n * p.n_init, p.n_init, state, dummy, panic);  // src/search.rs:404
    int count = 0;
    for (long long t = 0; t < p.T && !panic; ++t) {
        if (state < 0 || state >= S) { panic = true; break; }  // slice index out of range
        int label;
        float prob;
        if (wide) first_argmax(rd + (t * S + state) * C, C, label, prob, panic);  // :407, :412
        else first_argmax16(rdh + (t * S + state) * C, C, bf16, label, prob, panic);
        if (panic) break;
        if (label > 0) {
            out_path[count] = int(t);
            out_lab[count] = uint8_t(label);
            if (p.qstring) {
                const uint32_t code = phred_code(prob, p.qscale, p.qbias, s_invc, s_logc);  // :413
                if (sizeof(QT) == 4 && !is_unicode_scalar(code)) bad_char = true;
                out_qual[count] = QT(code);
            }
            ++count;
            state = int(((long long)state * n_base) % S) + (label - 1);  // :414
        }
    }
    p.offsets[n] = n * p.slot_stride;
    p.lens[n] = panic ? 0 : count;
    p.status[n] = panic ? FCTC_REFERENCE_PANIC : (bad_char ? FCTC_INVALID_QUALITY_CHAR : FCTC_OK);
}

static cudaError_t launch_crf_greedy(const CrfGreedyParams& p, cudaStream_t stream) {
    if (p.N <= 0) return cudaSuccess;
    const unsigned grid = unsigned((p.N + 127) / 128);
    if (p.qual_bytes == 4) crf_greedy_kernel<uint32_t><<<grid, 128, 0, stream>>>(p);
    else crf_greedy_kernel<uint8_t><<<grid, 128, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace fctc

// =============================================================================================
// C ABI
// =============================================================================================
namespace {

int crf_check(fctc_handle* h, int64_t N, int64_t T, int32_t S, int32_t C, int32_t n_init) {
    if (!h) return FCTC_E_ARGUMENT;
    if (C <= 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "Empty alphabet given");  // src/lib.rs:227-228
    if (C > 256) return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "alphabets above 256 labels are not supported");
    if (N < 0 || T <= 0 || S <= 0 || n_init <= 0)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "network_output and init_state must be non-empty (the reference asserts it)");
    if (T >= (1 << 30)) return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "reads above 2^30 rows are not supported");
    return 0;
}

bool crf_quality_fits_u8(float qscale, float qbias) {
    if (!(qscale == qscale) || !(qbias == qbias)) return true;
    return std::fabs(40.0 * double(qscale)) + std::fabs(double(qbias)) + 34.0 < 255.0;
}

// state-conditioned posteriors come as float32, binary16 or bfloat16; logits are not defined for them here
int crf_input_kind(fctc_handle* h, int kind) {
    if (!h) return FCTC_E_ARGUMENT;
    if (kind != FCTC_IN_F32 && kind != FCTC_IN_F16 && kind != FCTC_IN_BF16)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "crf input_kind must be FCTC_IN_F32, FCTC_IN_F16 or FCTC_IN_BF16");
    return 0;
}

#define CRF_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) return fctc::handle_fail(h, FCTC_E_CUDA, cudaGetErrorString(e__)); \
    } while (0)

}  // namespace

extern "C" int fctc_crf_greedy_batch_device_in(fctc_handle* h, const void* post_dev, int32_t input_kind, const float* init_dev,
                                               int64_t n_reads, int64_t n_rows, int32_t n_state, int32_t n_labels,
                                               int32_t n_init, int32_t qstring, float qscale, float qbias,
                                               uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes, int32_t* path_dev,
                                               int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                               int32_t* status_dev, void* stream) {
    if (int rc = crf_input_kind(h, input_kind)) return rc;
    if (int rc = crf_check(h, n_reads, n_rows, n_state, n_labels, n_init)) return rc;
    if (!post_dev || !init_dev || !labels_dev || !path_dev || !offsets_dev || !lens_dev || !status_dev)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (qstring && !quals_dev) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "qstring requested but quals_dev is NULL");
    if (qual_bytes != 1 && qual_bytes != 4) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "qual_bytes must be 1 or 4");
    if (qstring && qual_bytes == 1 && !crf_quality_fits_u8(qscale, qbias))
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "qual_bytes=1 cannot hold the quality codes of this qscale/qbias; use 4");
    const int64_t stride = (n_rows + 31) & ~int64_t(31);
    if (pool_capacity < n_reads * stride) return fctc::handle_fail(h, FCTC_E_CAPACITY, "result pool too small: n_reads*round_up(n_rows,32) entries needed");
    CRF_CUDA(cudaSetDevice(fctc::handle_device(h)));
    fctc::CrfGreedyParams p{};
    p.post = static_cast<const float*>(post_dev); p.in_kind = input_kind; p.init = init_dev; p.N = n_reads; p.T = n_rows;
    p.S = n_state; p.C = n_labels; p.n_init = n_init;
    p.qstring = qstring; p.qscale = qscale; p.qbias = qbias; p.labels = labels_dev; p.quals = quals_dev;
    p.qual_bytes = qual_bytes; p.path = path_dev; p.slot_stride = stride; p.offsets = offsets_dev; p.lens = lens_dev;
    p.status = status_dev;
    CRF_CUDA(fctc::launch_crf_greedy(p, static_cast<cudaStream_t>(stream)));
    return 0;
}

extern "C" int fctc_crf_greedy_batch_device(fctc_handle* h, const float* post_dev, const float* init_dev, int64_t n_reads,
                                            int64_t n_rows, int32_t n_state, int32_t n_labels, int32_t n_init,
                                            int32_t qstring, float qscale, float qbias, uint8_t* labels_dev,
                                            void* quals_dev, int32_t qual_bytes, int32_t* path_dev, int64_t pool_capacity,
                                            int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev, void* stream) {
    return fctc_crf_greedy_batch_device_in(h, post_dev, FCTC_IN_F32, init_dev, n_reads, n_rows, n_state, n_labels, n_init, qstring,
                                           qscale, qbias, labels_dev, quals_dev, qual_bytes, path_dev, pool_capacity, offsets_dev,
                                           lens_dev, status_dev, stream);
}

extern "C" int fctc_crf_beam_search_batch_device_in(fctc_handle* h, const void* post_dev, int32_t input_kind,
                                                    const float* init_dev, int64_t n_reads, int64_t n_rows, int32_t n_state,
                                                    int32_t n_labels, int32_t n_init, int32_t beam_size,
                                                    float beam_cut_threshold, uint8_t* labels_dev, int32_t* path_dev,
                                                    int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                                    int32_t* status_dev, unsigned long long* pool_used_dev, void* scratch_dev,
                                                    int64_t scratch_bytes, int32_t nodes_per_row, void* stream) {
    if (int rc = crf_input_kind(h, input_kind)) return rc;
    if (int rc = crf_check(h, n_reads, n_rows, n_state, n_labels, n_init)) return rc;
    // the reference's wrapper validates only the alphabet (src/lib.rs:263-270); beam_size 0 would
    // empty the beam at once (RanOutOfBeam), a negative value cannot be passed as usize
    if (beam_size <= 0) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "beam_size cannot be 0");
    if (!fctc::beam_shape_supported(beam_size, n_labels))
        return fctc::handle_fail(h, FCTC_E_UNSUPPORTED, "beam_size x labels exceeds the per-warp shared-memory plan");
    if (!post_dev || !init_dev || !labels_dev || !path_dev || !offsets_dev || !lens_dev || !status_dev || !pool_used_dev || !scratch_dev)
        return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (reinterpret_cast<uintptr_t>(scratch_dev) & 255u) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "scratch_dev must be 256-byte aligned");
    CRF_CUDA(cudaSetDevice(fctc::handle_device(h)));
    fctc::BeamParams p{};
    const int64_t need = fctc::beam_scratch_bytes(fctc::handle_sm_count(h), n_reads, n_rows, n_labels, beam_size, nodes_per_row,
                                                  scratch_bytes, &p.nodes_per_slot, &p.n_slots);
    if (scratch_bytes < need) return fctc::handle_fail(h, FCTC_E_CAPACITY, "beam scratch too small (see fctc_beam_scratch_bytes)");
    p.post = static_cast<const float*>(post_dev); p.in_f16 = input_kind != FCTC_IN_F32; p.in_bf16 = input_kind == FCTC_IN_BF16;
    p.N = n_reads; p.T = n_rows; p.C = n_labels; p.beam_size = beam_size; p.threshold = beam_cut_threshold;
    p.collapse = 0; p.labels = labels_dev; p.path = path_dev; p.pool_cap = pool_capacity; p.offsets = offsets_dev;
    p.lens = lens_dev; p.status = status_dev; p.cursor = pool_used_dev; p.scratch = scratch_dev; p.scratch_bytes = scratch_bytes;
    p.ticket = static_cast<unsigned long long*>(scratch_dev);
    p.S = n_state; p.init = init_dev; p.n_init = n_init;
    CRF_CUDA(fctc::launch_beam(p, fctc::handle_sm_count(h), static_cast<cudaStream_t>(stream)));
    return 0;
}

extern "C" int fctc_crf_beam_search_batch_device(fctc_handle* h, const float* post_dev, const float* init_dev,
                                                 int64_t n_reads, int64_t n_rows, int32_t n_state, int32_t n_labels,
                                                 int32_t n_init, int32_t beam_size, float beam_cut_threshold,
                                                 uint8_t* labels_dev, int32_t* path_dev, int64_t pool_capacity,
                                                 int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                                 unsigned long long* pool_used_dev, void* scratch_dev,
                                                 int64_t scratch_bytes, int32_t nodes_per_row, void* stream) {
    return fctc_crf_beam_search_batch_device_in(h, post_dev, FCTC_IN_F32, init_dev, n_reads, n_rows, n_state, n_labels, n_init,
                                                beam_size, beam_cut_threshold, labels_dev, path_dev, pool_capacity, offsets_dev,
                                                lens_dev, status_dev, pool_used_dev, scratch_dev, scratch_bytes, nodes_per_row,
                                                stream);
}

// ---- single-read host entry points: the shape of the pyo3 functions src/lib.rs:214-286 ---------------
extern "C" int fctc_crf_greedy_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_state,
                                      int32_t n_labels, const float* init_host, int32_t n_init, int32_t qstring,
                                      float qscale, float qbias, uint8_t* labels, uint32_t* quals, int64_t* path,
                                      int64_t capacity, int64_t* out_len) {
    if (int rc = crf_check(h, 1, n_rows, n_state, n_labels, n_init)) return rc;
    if (!post_host || !init_host || !out_len) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null pointer");
    CRF_CUDA(cudaSetDevice(fctc::handle_device(h)));
    cudaStream_t st = fctc::handle_stream(h);
    const int64_t stride = (n_rows + 31) & ~int64_t(31);
    const size_t pb = size_t(n_rows) * n_state * n_labels * 4;
    float* dpost = static_cast<float*>(fctc::handle_device_buffer(h, 0, pb));
    float* dinit = static_cast<float*>(fctc::handle_device_buffer(h, 1, size_t(n_init) * 4));
    uint8_t* dl = static_cast<uint8_t*>(fctc::handle_device_buffer(h, 3, size_t(stride)));
    uint32_t* dq = static_cast<uint32_t*>(fctc::handle_device_buffer(h, 9, size_t(stride) * 4));
    int32_t* dp = static_cast<int32_t*>(fctc::handle_device_buffer(h, 10, size_t(stride) * 4));
    int64_t* doff = static_cast<int64_t*>(fctc::handle_device_buffer(h, 4, 64));
    int32_t* dlen = static_cast<int32_t*>(fctc::handle_device_buffer(h, 5, 64));
    int32_t* dst = static_cast<int32_t*>(fctc::handle_device_buffer(h, 6, 64));
    unsigned char* pin = static_cast<unsigned char*>(fctc::handle_pinned_buffer(h, 2, size_t(stride) * 9 + 64));
    if (!dpost || !dinit || !dl || !dq || !dp || !doff || !dlen || !dst || !pin) return FCTC_E_NOMEM;
    CRF_CUDA(cudaMemcpyAsync(dpost, post_host, pb, cudaMemcpyHostToDevice, st));
    CRF_CUDA(cudaMemcpyAsync(dinit, init_host, size_t(n_init) * 4, cudaMemcpyHostToDevice, st));
    int rc = fctc_crf_greedy_batch_device(h, dpost, dinit, 1, n_rows, n_state, n_labels, n_init, qstring, qscale, qbias, dl, dq,
                                          4, dp, stride, doff, dlen, dst, st);
    if (rc < 0) return rc;
    int32_t* hlen = reinterpret_cast<int32_t*>(pin);
    int32_t* hst = hlen + 1;
    uint8_t* hl = pin + 64;
    uint32_t* hq = reinterpret_cast<uint32_t*>(pin + 64 + stride);
    int32_t* hp = reinterpret_cast<int32_t*>(pin + 64 + stride * 5);
    CRF_CUDA(cudaMemcpyAsync(hlen, dlen, 4, cudaMemcpyDeviceToHost, st));
    CRF_CUDA(cudaMemcpyAsync(hst, dst, 4, cudaMemcpyDeviceToHost, st));
    CRF_CUDA(cudaMemcpyAsync(hl, dl, size_t(stride), cudaMemcpyDeviceToHost, st));
    CRF_CUDA(cudaMemcpyAsync(hq, dq, size_t(stride) * 4, cudaMemcpyDeviceToHost, st));
    CRF_CUDA(cudaMemcpyAsync(hp, dp, size_t(stride) * 4, cudaMemcpyDeviceToHost, st));
    CRF_CUDA(cudaStreamSynchronize(st));
    *out_len = *hlen;
    const int64_t m = std::min<int64_t>(*hlen, capacity);
    for (int64_t i = 0; i < m; ++i) {
        if (labels) labels[i] = hl[i];
        if (path) path[i] = hp[i];
        if (quals && qstring) quals[i] = hq[i];
    }
    return *hst;
}

extern "C" int fctc_crf_beam_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_state,
                                    int32_t n_labels, const float* init_host, int32_t n_init, int32_t beam_size,
                                    float beam_cut_threshold, uint8_t* labels, int64_t* path, int64_t capacity,
                                    int64_t* out_len) {
    if (int rc = crf_check(h, 1, n_rows, n_state, n_labels, n_init)) return rc;
    if (!post_host || !init_host || !out_len) return fctc::handle_fail(h, FCTC_E_ARGUMENT, "null pointer");
    CRF_CUDA(cudaSetDevice(fctc::handle_device(h)));
    cudaStream_t st = fctc::handle_stream(h);
    const int64_t stride = std::max<int64_t>((n_rows + 31) & ~int64_t(31), 32);
    const size_t pb = size_t(n_rows) * n_state * n_labels * 4;
    float* dpost = static_cast<float*>(fctc::handle_device_buffer(h, 0, pb));
    float* dinit = static_cast<float*>(fctc::handle_device_buffer(h, 1, size_t(n_init) * 4));
    uint8_t* dl = static_cast<uint8_t*>(fctc::handle_device_buffer(h, 3, size_t(stride)));
    int32_t* dp = static_cast<int32_t*>(fctc::handle_device_buffer(h, 10, size_t(stride) * 4));
    int64_t* doff = static_cast<int64_t*>(fctc::handle_device_buffer(h, 4, 64));
    int32_t* dlen = static_cast<int32_t*>(fctc::handle_device_buffer(h, 5, 64));
    int32_t* dst = static_cast<int32_t*>(fctc::handle_device_buffer(h, 6, 64));
    unsigned long long* dcur = static_cast<unsigned long long*>(fctc::handle_device_buffer(h, 7, 256));
    unsigned char* pin = static_cast<unsigned char*>(fctc::handle_pinned_buffer(h, 2, size_t(stride) * 9 + 64));
    if (!dpost || !dinit || !dl || !dp || !doff || !dlen || !dst || !dcur || !pin) return FCTC_E_NOMEM;
    CRF_CUDA(cudaMemcpyAsync(dpost, post_host, pb, cudaMemcpyHostToDevice, st));
    CRF_CUDA(cudaMemcpyAsync(dinit, init_host, size_t(n_init) * 4, cudaMemcpyHostToDevice, st));
    int32_t* hlen = reinterpret_cast<int32_t*>(pin);
    int32_t* hst = hlen + 1;
    int64_t* hoff = reinterpret_cast<int64_t*>(pin + 8);
    uint8_t* hl = pin + 64;
    int32_t* hp = reinterpret_cast<int32_t*>(pin + 64 + stride * 5);
    // default trie pool first, the hard bound of (C-1)*beam proposals per row if that overflows
    const int32_t tries[2] = {0, std::max<int32_t>(1, (n_labels - 1) * beam_size)};
    for (int attempt = 0; attempt < 2; ++attempt) {
        const int64_t need = fctc::beam_scratch_bytes(fctc::handle_sm_count(h), 1, n_rows, n_labels, beam_size, tries[attempt],
                                                      0, nullptr, nullptr);
        void* scratch = fctc::handle_device_buffer(h, 8, size_t(need));
        if (!scratch) return FCTC_E_NOMEM;
        CRF_CUDA(cudaMemsetAsync(dcur, 0, 8, st));
        int rc = fctc_crf_beam_search_batch_device(h, dpost, dinit, 1, n_rows, n_state, n_labels, n_init, beam_size,
                                                   beam_cut_threshold, dl, dp, stride, doff, dlen, dst, dcur, scratch, need,
                                                   tries[attempt], st);
        if (rc < 0) return rc;
        CRF_CUDA(cudaMemcpyAsync(hlen, dlen, 4, cudaMemcpyDeviceToHost, st));
        CRF_CUDA(cudaMemcpyAsync(hst, dst, 4, cudaMemcpyDeviceToHost, st));
        CRF_CUDA(cudaMemcpyAsync(hoff, doff, 8, cudaMemcpyDeviceToHost, st));
        CRF_CUDA(cudaMemcpyAsync(hl, dl, size_t(stride), cudaMemcpyDeviceToHost, st));
        CRF_CUDA(cudaMemcpyAsync(hp, dp, size_t(stride) * 4, cudaMemcpyDeviceToHost, st));
        CRF_CUDA(cudaStreamSynchronize(st));
        if (*hst != FCTC_NODE_POOL_OVERFLOW) break;
    }
    *out_len = *hlen;
    const int64_t m = std::min<int64_t>(*hlen, capacity), o = *hoff;
    for (int64_t i = 0; i < m; ++i) {
        if (labels) labels[i] = hl[o + i];
        if (path) path[i] = hp[o + i];
    }
    return *hst;
}
