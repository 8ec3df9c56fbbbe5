// api.cu — the extern "C" boundary (include/fctc.h): handle, argument checks, host<->device
// streaming for the host-memory entry points.  No decode arithmetic happens on the host.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/fctc.h"
#include "kernels.h"
#include "phred.cuh"

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
};

struct Slot {  // one in-flight chunk of the host-memory pipeline
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    cudaEvent_t kernel_done = nullptr;  // the chunk's kernel alone (chunks that share one trie scratch run in turn)
    DevBuf post, labels, quals, path, offsets, lens, status, cursor, scratch;
    DevBuf slot_labels, slot_quals, slot_path, slot_offsets;  // slot-form results ahead of the packing step
    unsigned long long* cursor_host = nullptr;  // pinned
    int64_t dst_first = 0, count = 0;           // result rows of the chunk in flight
    bool busy = false;
};

}  // namespace

struct fctc_handle {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    Slot slot[2];
    DevBuf misc;
    void* pinned = nullptr;  // staging for single-read calls
    size_t pinned_bytes = 0;
    DevBuf extra_dev[16];    // buffers of entry points that live outside api.cu (duplex, crf)
    void* extra_pin[16] = {};
    size_t extra_pin_bytes[16] = {};
};

namespace {

int fail(fctc_handle* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    else g_create_error = buf;
    return code;
}

#define FCTC_CUDA(h, call)                                                                         \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(h, e__ == cudaErrorMemoryAllocation ? FCTC_E_NOMEM : FCTC_E_CUDA, "%s: %s", #call, \
                        cudaGetErrorString(e__));                                                  \
    } while (0)

int ensure(fctc_handle* h, DevBuf& b, size_t bytes) {
    if (bytes <= b.bytes) return 0;
    if (b.p) FCTC_CUDA(h, cudaFree(b.p));
    b.p = nullptr;
    b.bytes = 0;
    size_t want = std::max<size_t>(bytes, 256);
    FCTC_CUDA(h, cudaMalloc(&b.p, want));
    b.bytes = want;
    return 0;
}

int ensure_pinned(fctc_handle* h, size_t bytes) {
    if (bytes <= h->pinned_bytes) return 0;
    if (h->pinned) FCTC_CUDA(h, cudaFreeHost(h->pinned));
    h->pinned = nullptr;
    h->pinned_bytes = 0;
    FCTC_CUDA(h, cudaMallocHost(&h->pinned, bytes));
    h->pinned_bytes = bytes;
    return 0;
}

void release(DevBuf& b) {
    if (b.p) cudaFree(b.p);
    b = DevBuf{};
}

// Largest quality code phred() can produce for these parameters: p is clamped to [1e-4, 1] for
// prob <= 1, but prob < 0 (garbage input) gives p > 1 and a negative q, which saturates to 0.
bool quality_fits_u8(float qscale, float qbias) {
    if (!(qscale == qscale) || !(qbias == qbias)) return true;  // NaN -> code 33
    double hi = std::max(std::fabs(40.0 * double(qscale)) + std::fabs(double(qbias)), 0.0);
    return hi + 33.0 + 1.0 < 255.0;
}

int check_common(fctc_handle* h, int64_t n_reads, int64_t n_rows, int32_t n_labels) {
    if (!h) return FCTC_E_ARGUMENT;
    if (n_reads < 0 || n_rows < 0) return fail(h, FCTC_E_ARGUMENT, "n_reads and n_rows must be >= 0");
    if (n_labels <= 0) return fail(h, FCTC_E_ARGUMENT, "Empty alphabet given");  // src/lib.rs:191
    if (n_labels > 256) return fail(h, FCTC_E_UNSUPPORTED, "alphabets above 256 labels are not supported (labels are uint8)");
    if (n_rows > (int64_t(1) << 30)) return fail(h, FCTC_E_UNSUPPORTED, "reads above 2^30 rows are not supported");
    return 0;
}

int roundup32(int64_t v) { return int((v + 31) & ~int64_t(31)); }

// The internal "in_f16" argument of the batch paths carries a whole FCTC_IN_* input kind: element type in the low
// four bits (0 f32, 1 binary16, 2 bfloat16), softmax mode above (FCTC_IN_LOGITS_EXACT / FCTC_IN_LOGITS_FAST).
template <class Params>
void set_input(Params& p, int kind) {
    p.in_f16 = (kind & 0xf) != 0;
    p.in_bf16 = (kind & 0xf) == 2;
    p.softmax = (kind >> 4) & 3;
}

int check_input_kind(fctc_handle* h, int kind) {
    if (!h) return FCTC_E_ARGUMENT;
    if ((kind & 0xf) > 2 || (kind >> 4) > 2 || kind < 0)
        return fail(h, FCTC_E_ARGUMENT, "input_kind must be FCTC_IN_F32 / _F16 / _BF16, optionally or-ed with FCTC_IN_LOGITS_EXACT or _FAST");
    return 0;
}

}  // namespace

// All fctc_* definitions below take C linkage from their declarations in include/fctc.h.

int fctc_version(void) { return 100; }

int fctc_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char* fctc_status_message(int status) {
    switch (status) {  // src/lib.rs:43-56
        case FCTC_RAN_OUT_OF_BEAM: return "Ran out of search space (beam_cut_threshold too high)";
        case FCTC_INCOMPARABLE_VALUES: return "Failed to compare values (NaNs in input?)";
        case FCTC_INVALID_ENVELOPE: return "Invalid envelope values";
        default: return nullptr;
    }
}

const char* fctc_last_error(const fctc_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int fctc_device_sm_count(const fctc_handle* h) { return h ? h->sm_count : 0; }

int fctc_create(int device, fctc_handle** out) {
    if (!out) return FCTC_E_ARGUMENT;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, FCTC_E_CUDA, "no usable CUDA device (%s); libfctc_b200 has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= n) return fail(nullptr, FCTC_E_ARGUMENT, "device %d out of range (0..%d)", device, n - 1);
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return fail(nullptr, FCTC_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return fail(nullptr, FCTC_E_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a (B200) only",
                    device, prop.major, prop.minor);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, FCTC_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    fctc_handle* h = new fctc_handle();
    h->device = device;
    h->sm_count = prop.multiProcessorCount;
    bool ok = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; i < 2 && ok; ++i) {
        ok = ok && cudaStreamCreateWithFlags(&h->slot[i].stream, cudaStreamNonBlocking) == cudaSuccess;
        ok = ok && cudaEventCreateWithFlags(&h->slot[i].done, cudaEventDisableTiming) == cudaSuccess;
        ok = ok && cudaEventCreateWithFlags(&h->slot[i].kernel_done, cudaEventDisableTiming) == cudaSuccess;
        ok = ok && cudaMallocHost(reinterpret_cast<void**>(&h->slot[i].cursor_host), 64) == cudaSuccess;
    }
    if (!ok) {
        int rc = fail(nullptr, FCTC_E_CUDA, "stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
        fctc_destroy(h);
        return rc;
    }
    *out = h;
    return 0;
}

void fctc_destroy(fctc_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    for (Slot& s : h->slot) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        for (DevBuf* b : {&s.post, &s.labels, &s.quals, &s.path, &s.offsets, &s.lens, &s.status, &s.cursor, &s.scratch,
                          &s.slot_labels, &s.slot_quals, &s.slot_path, &s.slot_offsets})
            release(*b);
        if (s.cursor_host) cudaFreeHost(s.cursor_host);
        if (s.done) cudaEventDestroy(s.done);
        if (s.kernel_done) cudaEventDestroy(s.kernel_done);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    release(h->misc);
    for (DevBuf& b : h->extra_dev) release(b);
    for (void* q : h->extra_pin)
        if (q) cudaFreeHost(q);
    if (h->pinned) cudaFreeHost(h->pinned);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

// ---------------------------------------------------------------------------------------------
// Viterbi
// ---------------------------------------------------------------------------------------------
namespace {
int viterbi_batch_device_any(fctc_handle* h, const void* post_dev, int in_f16, int64_t n_reads, int64_t n_rows,
                             int32_t n_labels, int32_t qstring, float qscale, float qbias,
                             int32_t collapse_repeats, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                             int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                             int32_t* status_dev, void* stream) {
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (n_rows == 0) return fail(h, FCTC_E_ARGUMENT, "n_rows must be >= 1 (the reference asserts a non-empty matrix)");
    if (!post_dev || !labels_dev || !path_dev || !offsets_dev || !lens_dev || !status_dev)
        return fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (qstring && !quals_dev) return fail(h, FCTC_E_ARGUMENT, "qstring requested but quals_dev is NULL");
    if (qual_bytes != 1 && qual_bytes != 4) return fail(h, FCTC_E_ARGUMENT, "qual_bytes must be 1 or 4");
    if (qstring && qual_bytes == 1 && !quality_fits_u8(qscale, qbias))
        return fail(h, FCTC_E_ARGUMENT, "qual_bytes=1 cannot hold the quality codes of qscale=%g qbias=%g; use 4", qscale,
                    qbias);
    const int64_t stride = roundup32(n_rows);
    if (pool_capacity < n_reads * stride)
        return fail(h, FCTC_E_CAPACITY, "result pool too small: %lld entries given, n_reads*round_up(n_rows,32) = %lld needed",
                    (long long)pool_capacity, (long long)(n_reads * stride));
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::ViterbiParams p{};
    p.post = static_cast<const float*>(post_dev); set_input(p, in_f16); p.N = n_reads; p.T = n_rows; p.C = n_labels;
    p.qstring = qstring; p.qscale = qscale; p.qbias = qbias; p.collapse = collapse_repeats;
    p.labels = labels_dev; p.quals = quals_dev; p.qual_bytes = qual_bytes; p.path = path_dev;
    p.slot_stride = stride; p.offset_base = 0; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    FCTC_CUDA(h, fctc::launch_viterbi(p, h->sm_count, static_cast<cudaStream_t>(stream), nullptr));
    return 0;
}
}  // namespace

int fctc_viterbi_batch_device(fctc_handle* h, const float* post_dev, int64_t n_reads, int64_t n_rows,
                              int32_t n_labels, int32_t qstring, float qscale, float qbias,
                              int32_t collapse_repeats, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                              int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                              int32_t* status_dev, void* stream) {
    return viterbi_batch_device_any(h, post_dev, 0, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats,
                                    labels_dev, quals_dev, qual_bytes, path_dev, pool_capacity, offsets_dev, lens_dev,
                                    status_dev, stream);
}

int fctc_viterbi_batch_device_f16(fctc_handle* h, const void* post_dev_f16, int64_t n_reads, int64_t n_rows,
                                  int32_t n_labels, int32_t qstring, float qscale, float qbias,
                                  int32_t collapse_repeats, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                                  int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                  int32_t* status_dev, void* stream) {
    return viterbi_batch_device_any(h, post_dev_f16, 1, n_reads, n_rows, n_labels, qstring, qscale, qbias,
                                    collapse_repeats, labels_dev, quals_dev, qual_bytes, path_dev, pool_capacity,
                                    offsets_dev, lens_dev, status_dev, stream);
}

namespace {

// Host-memory pipeline shared by the batch entry points: reads are cut into chunks, each chunk
// runs H2D -> kernel -> D2H on one of two streams so that the copies of one chunk overlap the
// other chunk's work.  The device-side result pool is per chunk (cursor reset to 0), the host
// pool is the concatenation of the chunks' used prefixes.
struct HostOut {
    uint8_t* labels;
    void* quals;  // may be null
    int qual_bytes;
    int32_t* path;
    int64_t cap;
    int64_t* offsets;
    int32_t* lens;
    int32_t* status;
    int64_t used = 0;
    bool overflow = false;
};

int drain_slot(fctc_handle* h, Slot& s, HostOut& out) {
    if (!s.busy) return 0;
    FCTC_CUDA(h, cudaEventSynchronize(s.done));
    const int64_t used = int64_t(*s.cursor_host);
    if (out.used + used > out.cap) {
        out.overflow = true;
    } else if (used > 0) {
        FCTC_CUDA(h, cudaMemcpyAsync(out.labels + out.used, s.labels.p, size_t(used), cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(out.path + out.used, s.path.p, size_t(used) * 4, cudaMemcpyDeviceToHost, s.stream));
        if (out.quals)
            FCTC_CUDA(h, cudaMemcpyAsync(static_cast<uint8_t*>(out.quals) + out.used * out.qual_bytes, s.quals.p,
                                         size_t(used) * out.qual_bytes, cudaMemcpyDeviceToHost, s.stream));
    }
    // offsets of this chunk are relative to the chunk's device pool: rebase (host, one add per read)
    if (!out.overflow)
        for (int64_t i = 0; i < s.count; ++i) out.offsets[s.dst_first + i] += out.used;
    if (!out.overflow) out.used += used;
    s.busy = false;
    return 0;
}

// launch(slot, count) enqueues the decode kernel for `count` reads resident in slot.post on
// slot.stream, writing slot.{labels,quals,path,offsets,lens,status,cursor}.
// Reads [src_first, src_first + n) of post_host go to result rows [dst_first, ...).
template <class Launch>
int host_pipeline(fctc_handle* h, const void* post_host_any, int elem_bytes, int64_t n_reads, int64_t n_rows,
                  int32_t n_labels, HostOut& out, int64_t chunk, Launch&& launch) {
    const unsigned char* post_host = static_cast<const unsigned char*>(post_host_any);
    const int64_t read_bytes = n_rows * int64_t(n_labels) * elem_bytes;
    const int64_t slot_entries = roundup32(n_rows);
    chunk = std::max<int64_t>(1, std::min(chunk, n_reads));
    for (int i = 0; i < 2; ++i) {
        Slot& s = h->slot[i];
        if (int rc = ensure(h, s.post, size_t(std::max<int64_t>(chunk * read_bytes, 16)))) return rc;
        if (int rc = ensure(h, s.labels, size_t(chunk * slot_entries))) return rc;
        if (int rc = ensure(h, s.path, size_t(chunk * slot_entries) * 4)) return rc;
        if (out.quals)
            if (int rc = ensure(h, s.quals, size_t(chunk * slot_entries) * out.qual_bytes)) return rc;
        if (int rc = ensure(h, s.offsets, size_t(chunk) * 8)) return rc;
        if (int rc = ensure(h, s.lens, size_t(chunk) * 4)) return rc;
        if (int rc = ensure(h, s.status, size_t(chunk) * 4)) return rc;
        if (int rc = ensure(h, s.cursor, 256)) return rc;
        s.busy = false;
    }
    // FCTC_TRACE_PIPELINE=1 prints, per chunk, when its copy-in started / ended and when its kernel ended
    // (milliseconds on the device clock since the first chunk was queued): a development aid.
    static const bool trace = getenv("FCTC_TRACE_PIPELINE") != nullptr;
    std::vector<cudaEvent_t> tev;
    auto mark = [&](cudaStream_t st) {
        if (!trace) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        tev.push_back(e);
    };
    // Queue order matters: the copy engines take work in submission order, so the read-back of chunk i
    // (which has to wait for kernel i) is only queued after the copy-in of chunk i+1 — otherwise that
    // copy-in sits behind it and the transfer never overlaps the kernel (measured: fully serial).
    auto queue_readback = [&](Slot& s, int64_t first, int64_t cnt) -> int {
        FCTC_CUDA(h, cudaMemcpyAsync(s.cursor_host, s.cursor.p, 8, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(out.offsets + first, s.offsets.p, size_t(cnt) * 8, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(out.lens + first, s.lens.p, size_t(cnt) * 4, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(out.status + first, s.status.p, size_t(cnt) * 4, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaEventRecord(s.done, s.stream));
        return 0;
    };
    int which = 0;
    int64_t prev_first = -1, prev_cnt = 0;
    for (int64_t first = 0; first < n_reads; first += chunk, which ^= 1) {
        Slot& s = h->slot[which];
        if (int rc = drain_slot(h, s, out)) return rc;
        const int64_t cnt = std::min(chunk, n_reads - first);
        s.dst_first = first;
        s.count = cnt;
        mark(s.stream);
        if (read_bytes > 0)
            FCTC_CUDA(h, cudaMemcpyAsync(s.post.p, post_host + first * read_bytes, size_t(cnt * read_bytes),
                                         cudaMemcpyHostToDevice, s.stream));
        mark(s.stream);
        FCTC_CUDA(h, cudaMemsetAsync(s.cursor.p, 0, 8, s.stream));
        if (int rc = launch(s, cnt)) return rc;
        mark(s.stream);
        s.busy = true;
        if (prev_first >= 0)
            if (int rc = queue_readback(h->slot[which ^ 1], prev_first, prev_cnt)) return rc;
        prev_first = first;
        prev_cnt = cnt;
    }
    if (prev_first >= 0)
        if (int rc = queue_readback(h->slot[which ^ 1], prev_first, prev_cnt)) return rc;
    // drain in submission order: the slot used before the last one first
    if (int rc = drain_slot(h, h->slot[which], out)) return rc;
    if (int rc = drain_slot(h, h->slot[which ^ 1], out)) return rc;
    FCTC_CUDA(h, cudaStreamSynchronize(h->slot[0].stream));
    FCTC_CUDA(h, cudaStreamSynchronize(h->slot[1].stream));
    for (size_t i = 0; i + 2 < tev.size() + 0 && trace; i += 3) {
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, tev[0], tev[i]);
        cudaEventElapsedTime(&b, tev[0], tev[i + 1]);
        cudaEventElapsedTime(&c, tev[0], tev[i + 2]);
        fprintf(stderr, "[fctc pipeline] chunk %zu: copy-in %.2f..%.2f ms, kernel done %.2f ms\n", i / 3, a, b, c);
    }
    for (cudaEvent_t e : tev) cudaEventDestroy(e);
    return 0;
}

int64_t default_chunk_reads(int64_t read_bytes) {  // ~64 MiB of posteriors per chunk
    return std::max<int64_t>(1, (int64_t(64) << 20) / std::max<int64_t>(read_bytes, 1));
}

}  // namespace

namespace {
int viterbi_batch_any(fctc_handle* h, const void* post_host_any, int in_f16, int64_t n_reads, int64_t n_rows,
                      int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                      uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity,
                      int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used) {
    const unsigned char* post_host = static_cast<const unsigned char*>(post_host_any);
    const int elem_bytes = (in_f16 & 0xf) ? 2 : 4;
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (n_rows == 0) return fail(h, FCTC_E_ARGUMENT, "n_rows must be >= 1 (the reference asserts a non-empty matrix)");
    if (!post_host || !labels || !path || !offsets || !lens || !status)
        return fail(h, FCTC_E_ARGUMENT, "null host pointer");
    if (qstring && !quals) return fail(h, FCTC_E_ARGUMENT, "qstring requested but quals is NULL");
    if (qual_bytes != 1 && qual_bytes != 4) return fail(h, FCTC_E_ARGUMENT, "qual_bytes must be 1 or 4");
    if (qstring && qual_bytes == 1 && !quality_fits_u8(qscale, qbias))
        return fail(h, FCTC_E_ARGUMENT, "qual_bytes=1 cannot hold the quality codes of qscale=%g qbias=%g; use 4", qscale,
                    qbias);
    const int64_t stride = roundup32(n_rows);
    if (pool_used) *pool_used = 0;
    if (pool_capacity < n_reads * stride)
        return fail(h, FCTC_E_CAPACITY, "result pool too small: %lld entries given, n_reads*round_up(n_rows,32) = %lld needed",
                    (long long)pool_capacity, (long long)(n_reads * stride));
    FCTC_CUDA(h, cudaSetDevice(h->device));
    if (n_reads == 0) return 0;

    // Slot-form results need no cursor, so the whole pipeline is asynchronous: per chunk
    // H2D posteriors -> kernel -> D2H of the chunk's slots, alternating between two streams.
    const int64_t read_bytes = n_rows * int64_t(n_labels) * elem_bytes;
    const int64_t chunk = std::min(default_chunk_reads(read_bytes), n_reads);
    for (int i = 0; i < 2; ++i) {
        Slot& s = h->slot[i];
        if (int rc = ensure(h, s.post, size_t(chunk * read_bytes))) return rc;
        if (int rc = ensure(h, s.labels, size_t(chunk * stride))) return rc;
        if (int rc = ensure(h, s.path, size_t(chunk * stride) * 4)) return rc;
        if (qstring)
            if (int rc = ensure(h, s.quals, size_t(chunk * stride) * qual_bytes)) return rc;
        if (int rc = ensure(h, s.offsets, size_t(chunk) * 8)) return rc;
        if (int rc = ensure(h, s.lens, size_t(chunk) * 4)) return rc;
        if (int rc = ensure(h, s.status, size_t(chunk) * 4)) return rc;
    }
    int which = 0;
    for (int64_t first = 0; first < n_reads; first += chunk, which ^= 1) {
        Slot& s = h->slot[which];
        const int64_t cnt = std::min(chunk, n_reads - first);
        FCTC_CUDA(h, cudaMemcpyAsync(s.post.p, post_host + first * read_bytes, size_t(cnt * read_bytes),
                                     cudaMemcpyHostToDevice, s.stream));
        fctc::ViterbiParams p{};
        p.post = static_cast<const float*>(s.post.p); set_input(p, in_f16); p.N = cnt; p.T = n_rows; p.C = n_labels;
        p.qstring = qstring; p.qscale = qscale; p.qbias = qbias; p.collapse = collapse_repeats;
        p.labels = static_cast<uint8_t*>(s.labels.p); p.quals = s.quals.p; p.qual_bytes = qual_bytes;
        p.path = static_cast<int32_t*>(s.path.p); p.slot_stride = stride; p.offset_base = first * stride;
        p.offsets = static_cast<int64_t*>(s.offsets.p); p.lens = static_cast<int32_t*>(s.lens.p);
        p.status = static_cast<int32_t*>(s.status.p);
        FCTC_CUDA(h, fctc::launch_viterbi(p, h->sm_count, s.stream, nullptr));
        const size_t ent = size_t(cnt * stride), o = size_t(first * stride);
        FCTC_CUDA(h, cudaMemcpyAsync(labels + o, s.labels.p, ent, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(path + o, s.path.p, ent * 4, cudaMemcpyDeviceToHost, s.stream));
        if (qstring)
            FCTC_CUDA(h, cudaMemcpyAsync(static_cast<uint8_t*>(quals) + o * qual_bytes, s.quals.p, ent * qual_bytes,
                                         cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(offsets + first, s.offsets.p, size_t(cnt) * 8, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(lens + first, s.lens.p, size_t(cnt) * 4, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(status + first, s.status.p, size_t(cnt) * 4, cudaMemcpyDeviceToHost, s.stream));
    }
    FCTC_CUDA(h, cudaStreamSynchronize(h->slot[0].stream));
    FCTC_CUDA(h, cudaStreamSynchronize(h->slot[1].stream));
    if (pool_used) *pool_used = n_reads * stride;
    return 0;
}
}  // namespace

namespace {
// fctc_viterbi_batch_packed: the chunk's kernel writes slot form into side buffers, pack.cu squeezes the used
// prefixes into the slot's result buffers, and the shared host pipeline reads back just those (its per-chunk
// cursor tells the host how many entries that is).
int viterbi_batch_packed_any(fctc_handle* h, const void* post_host, int in_f16, int64_t n_reads, int64_t n_rows,
                             int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                             uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity,
                             int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used) {
    const int elem_bytes = (in_f16 & 0xf) ? 2 : 4;
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (n_rows == 0) return fail(h, FCTC_E_ARGUMENT, "n_rows must be >= 1 (the reference asserts a non-empty matrix)");
    if (!post_host || !labels || !path || !offsets || !lens || !status)
        return fail(h, FCTC_E_ARGUMENT, "null host pointer");
    if (qstring && !quals) return fail(h, FCTC_E_ARGUMENT, "qstring requested but quals is NULL");
    if (qual_bytes != 1 && qual_bytes != 4) return fail(h, FCTC_E_ARGUMENT, "qual_bytes must be 1 or 4");
    if (qstring && qual_bytes == 1 && !quality_fits_u8(qscale, qbias))
        return fail(h, FCTC_E_ARGUMENT, "qual_bytes=1 cannot hold the quality codes of qscale=%g qbias=%g; use 4", qscale,
                    qbias);
    if (pool_used) *pool_used = 0;
    FCTC_CUDA(h, cudaSetDevice(h->device));
    if (n_reads == 0) return 0;
    const int64_t stride = roundup32(n_rows);
    const int64_t read_bytes = n_rows * int64_t(n_labels) * elem_bytes;
    const int64_t chunk = std::min(default_chunk_reads(read_bytes), n_reads);
    for (int i = 0; i < 2; ++i) {
        Slot& s = h->slot[i];
        if (int rc = ensure(h, s.slot_labels, size_t(chunk * stride))) return rc;
        if (int rc = ensure(h, s.slot_path, size_t(chunk * stride) * 4)) return rc;
        if (qstring)
            if (int rc = ensure(h, s.slot_quals, size_t(chunk * stride) * qual_bytes)) return rc;
        if (int rc = ensure(h, s.slot_offsets, size_t(chunk) * 8)) return rc;
    }
    HostOut out{labels, qstring ? quals : nullptr, qual_bytes, path, pool_capacity, offsets, lens, status};
    auto launch = [&](Slot& s, int64_t cnt) -> int {
        fctc::ViterbiParams p{};
        p.post = static_cast<const float*>(s.post.p); set_input(p, in_f16); p.N = cnt; p.T = n_rows; p.C = n_labels;
        p.qstring = qstring; p.qscale = qscale; p.qbias = qbias; p.collapse = collapse_repeats;
        p.labels = static_cast<uint8_t*>(s.slot_labels.p); p.quals = s.slot_quals.p; p.qual_bytes = qual_bytes;
        p.path = static_cast<int32_t*>(s.slot_path.p); p.slot_stride = stride; p.offset_base = 0;
        p.offsets = static_cast<int64_t*>(s.slot_offsets.p); p.lens = static_cast<int32_t*>(s.lens.p);
        p.status = static_cast<int32_t*>(s.status.p);
        FCTC_CUDA(h, fctc::launch_viterbi(p, h->sm_count, s.stream, nullptr));
        fctc::PackParams k{};
        k.labels = p.labels; k.quals = qstring ? s.slot_quals.p : nullptr; k.qual_bytes = qual_bytes; k.path = p.path;
        k.offsets = p.offsets; k.offset_base = 0; k.lens = p.lens; k.N = cnt;
        k.out_labels = static_cast<uint8_t*>(s.labels.p); k.out_quals = s.quals.p;
        k.out_path = static_cast<int32_t*>(s.path.p); k.out_offsets = static_cast<int64_t*>(s.offsets.p);
        k.out_capacity = cnt * stride; k.cursor = static_cast<unsigned long long*>(s.cursor.p);
        FCTC_CUDA(h, fctc::launch_pack(k, h->sm_count, s.stream));
        return 0;
    };
    if (int rc = host_pipeline(h, post_host, elem_bytes, n_reads, n_rows, n_labels, out, chunk, launch)) return rc;
    if (pool_used) *pool_used = out.used;
    if (out.overflow)
        return fail(h, FCTC_E_CAPACITY, "result pool too small: %lld entries given (n_reads*round_up(n_rows,32) always suffices)",
                    (long long)pool_capacity);
    return 0;
}
}  // namespace

int fctc_viterbi_batch_packed(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                              int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                              void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                              int32_t* lens, int32_t* status, int64_t* pool_used) {
    return viterbi_batch_packed_any(h, post_host, 0, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats,
                                    labels, quals, qual_bytes, path, pool_capacity, offsets, lens, status, pool_used);
}

int fctc_viterbi_batch_packed_f16(fctc_handle* h, const void* post_host_f16, int64_t n_reads, int64_t n_rows,
                                  int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                                  uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity,
                                  int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used) {
    return viterbi_batch_packed_any(h, post_host_f16, 1, n_reads, n_rows, n_labels, qstring, qscale, qbias,
                                    collapse_repeats, labels, quals, qual_bytes, path, pool_capacity, offsets, lens, status,
                                    pool_used);
}

int fctc_viterbi_batch(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                       int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                       void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                       int32_t* lens, int32_t* status, int64_t* pool_used) {
    return viterbi_batch_any(h, post_host, 0, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats, labels,
                             quals, qual_bytes, path, pool_capacity, offsets, lens, status, pool_used);
}

int fctc_viterbi_batch_f16(fctc_handle* h, const void* post_host_f16, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                           int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                           void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                           int32_t* lens, int32_t* status, int64_t* pool_used) {
    return viterbi_batch_any(h, post_host_f16, 1, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats,
                             labels, quals, qual_bytes, path, pool_capacity, offsets, lens, status, pool_used);
}

namespace {

// Packs a strided [T, C] host view into the handle's pinned staging buffer, followed by room for
// one read's results; returns the layout.
struct SingleStage {
    float* packed;
    uint8_t* labels;
    uint32_t* quals;
    int32_t* path;
    int64_t* off;
    int32_t* len;
    int32_t* st;
    int64_t slot_entries;
};

int stage_single(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_labels, int64_t stride_row,
                 int64_t stride_label, SingleStage* S) {
    const int64_t slot_entries = std::max<int64_t>(roundup32(n_rows), 32);
    const size_t post_bytes = size_t(n_rows) * n_labels * 4;
    const size_t lab_off = (post_bytes + 255) & ~size_t(255);
    const size_t qual_off = lab_off + size_t(slot_entries);
    const size_t path_off = qual_off + size_t(slot_entries) * 4;
    const size_t meta_off = path_off + size_t(slot_entries) * 4;
    if (int rc = ensure_pinned(h, meta_off + 64)) return rc;
    unsigned char* pin = static_cast<unsigned char*>(h->pinned);
    S->packed = reinterpret_cast<float*>(pin);
    for (int64_t t = 0; t < n_rows; ++t)
        for (int32_t c = 0; c < n_labels; ++c) S->packed[t * n_labels + c] = post_host[t * stride_row + c * stride_label];
    S->labels = pin + lab_off;
    S->quals = reinterpret_cast<uint32_t*>(pin + qual_off);
    S->path = reinterpret_cast<int32_t*>(pin + path_off);
    S->off = reinterpret_cast<int64_t*>(pin + meta_off);
    S->len = reinterpret_cast<int32_t*>(pin + meta_off + 8);
    S->st = reinterpret_cast<int32_t*>(pin + meta_off + 16);
    S->slot_entries = slot_entries;
    return 0;
}

}  // namespace

int fctc_viterbi_batch_device_in(fctc_handle* h, const void* in_dev, int32_t input_kind, int64_t n_reads, int64_t n_rows,
                                 int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                                 uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes, int32_t* path_dev,
                                 int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                 void* stream) {
    if (int rc = check_input_kind(h, input_kind)) return rc;
    return viterbi_batch_device_any(h, in_dev, input_kind, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats,
                                    labels_dev, quals_dev, qual_bytes, path_dev, pool_capacity, offsets_dev, lens_dev,
                                    status_dev, stream);
}

int fctc_viterbi_batch_in(fctc_handle* h, const void* in_host, int32_t input_kind, int64_t n_reads, int64_t n_rows,
                          int32_t n_labels, int32_t qstring, float qscale, float qbias, int32_t collapse_repeats,
                          uint8_t* labels, void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity,
                          int64_t* offsets, int32_t* lens, int32_t* status, int64_t* pool_used) {
    if (int rc = check_input_kind(h, input_kind)) return rc;
    return viterbi_batch_packed_any(h, in_host, input_kind, n_reads, n_rows, n_labels, qstring, qscale, qbias,
                                    collapse_repeats, labels, quals, qual_bytes, path, pool_capacity, offsets, lens, status,
                                    pool_used);
}

int fctc_viterbi_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_labels,
                        int64_t stride_row, int64_t stride_label, int32_t qstring, float qscale, float qbias,
                        int32_t collapse_repeats, uint8_t* labels, uint32_t* quals, int64_t* path, int64_t capacity,
                        int64_t* out_len) {
    if (int rc = check_common(h, 1, n_rows, n_labels)) return rc;
    if (n_rows == 0) return fail(h, FCTC_E_ARGUMENT, "n_rows must be >= 1 (the reference asserts a non-empty matrix)");
    if (!post_host || !out_len) return fail(h, FCTC_E_ARGUMENT, "null pointer");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    SingleStage S;
    if (int rc = stage_single(h, post_host, n_rows, n_labels, stride_row, stride_label, &S)) return rc;
    int64_t used = 0;
    int rc = fctc_viterbi_batch(h, S.packed, 1, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats, S.labels,
                                S.quals, 4, S.path, S.slot_entries, S.off, S.len, S.st, &used);
    if (rc < 0) return rc;
    const int64_t len = *S.len, o = *S.off;
    *out_len = len;
    const int64_t m = std::min<int64_t>(len, capacity);
    for (int64_t i = 0; i < m; ++i) {
        if (labels) labels[i] = S.labels[o + i];
        if (path) path[i] = S.path[o + i];
        if (quals && qstring) quals[i] = S.quals[o + i];
    }
    return *S.st;
}

// ---------------------------------------------------------------------------------------------
// 1-D beam search
// ---------------------------------------------------------------------------------------------
namespace fctc { bool beam_shape_supported(int beam_size, int C); }

namespace {

int check_beam(fctc_handle* h, int32_t n_labels, int32_t beam_size, float thr) {
    if (beam_size <= 0) return fail(h, FCTC_E_ARGUMENT, "beam_size cannot be 0");                       // src/lib.rs:339-340
    if (thr < -0.0f) return fail(h, FCTC_E_ARGUMENT, "beam_cut_threshold must be at least 0.0");         // :341-344
    const float max_cut = 1.0f / float(n_labels);
    if (thr >= max_cut) return fail(h, FCTC_E_ARGUMENT, "beam_cut_threshold cannot be more than %g", max_cut);  // :345-349
    if (!fctc::beam_shape_supported(beam_size, n_labels))
        return fail(h, FCTC_E_UNSUPPORTED, "beam_size %d x %d labels exceeds the per-warp shared-memory plan", beam_size,
                    n_labels);
    return 0;
}

}  // namespace

namespace fctc {
int env_knob(const char* name) {
    const char* e = getenv(name);
    const long long v = e ? atoll(e) : 0;
    return v > 0 && v < (1ll << 31) ? int(v) : 0;
}
int handle_device(const fctc_handle* h) { return h->device; }
int handle_sm_count(const fctc_handle* h) { return h->sm_count; }
int handle_fail(fctc_handle* h, int code, const char* message) { return fail(h, code, "%s", message); }
int check_beam_call(fctc_handle* h, int64_t n_reads, int64_t n_rows, int32_t n_labels, int32_t beam_size, float thr) {
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    return check_beam(h, n_labels, beam_size, thr);
}
int check_threshold_call(fctc_handle* h, int32_t n_labels, int32_t beam_size, float thr) {
    if (beam_size <= 0) return fail(h, FCTC_E_ARGUMENT, "beam_size cannot be 0");
    if (thr < -0.0f) return fail(h, FCTC_E_ARGUMENT, "beam_cut_threshold must be at least 0.0");
    const float max_cut = 1.0f / float(n_labels);
    if (thr >= max_cut) return fail(h, FCTC_E_ARGUMENT, "beam_cut_threshold cannot be more than %g", max_cut);
    return 0;
}
void* handle_device_buffer(fctc_handle* h, int index, size_t bytes) {
    if (index < 0 || index >= 16) return nullptr;
    if (ensure(h, h->extra_dev[index], bytes)) return nullptr;
    return h->extra_dev[index].p;
}
size_t handle_device_buffer_bytes(const fctc_handle* h, int index) {
    return (index < 0 || index >= 16) ? 0 : h->extra_dev[index].bytes;
}
void* handle_pinned_buffer(fctc_handle* h, int index, size_t bytes) {
    if (index < 0 || index >= 16) return nullptr;
    if (bytes <= h->extra_pin_bytes[index]) return h->extra_pin[index];
    if (h->extra_pin[index]) cudaFreeHost(h->extra_pin[index]);
    h->extra_pin[index] = nullptr;
    h->extra_pin_bytes[index] = 0;
    const size_t want = std::max<size_t>(bytes, 256);
    cudaError_t e = cudaMallocHost(&h->extra_pin[index], want);
    if (e != cudaSuccess) {
        fail(h, FCTC_E_NOMEM, "cudaMallocHost(%zu): %s", want, cudaGetErrorString(e));
        return nullptr;
    }
    h->extra_pin_bytes[index] = want;
    return h->extra_pin[index];
}
cudaStream_t handle_stream(fctc_handle* h) { return h->stream; }
cudaStream_t handle_aux_stream(fctc_handle* h) { return h->slot[1].stream; }
}  // namespace fctc

int64_t fctc_beam_scratch_bytes(const fctc_handle* h, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                                int32_t beam_size, int32_t nodes_per_row) {
    if (!h || n_labels <= 0 || beam_size <= 0) return 0;
    // recommended size: one trie pool per resident warp, capped at 96 GiB or 70 % of the device's memory, whichever is
    // more (a smaller buffer is accepted by the device entry points: they run as many pools as fit, at least 4).
    // At beam 32 and 10,000 rows a pool is 24 MB and the 4,736 resident warps of a B200 want 106 GiB: under a flat
    // 96 GiB cap 4,292 of them ran and the rest of a launch's first round waited for a pool.
    int64_t cap = int64_t(96) << 30;
    size_t free_b = 0, total_b = 0;
    int dev = -1;
    if (cudaGetDevice(&dev) == cudaSuccess) {
        if (dev != h->device) cudaSetDevice(h->device);
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) cap = std::max<int64_t>(cap, int64_t(total_b / 10 * 7));
        else cudaGetLastError();
        if (dev != h->device) cudaSetDevice(dev);
    }
    return fctc::beam_scratch_bytes(h->sm_count, n_reads, n_rows, n_labels, beam_size, nodes_per_row, cap, nullptr, nullptr);
}

namespace {
int beam_search_batch_device_any(fctc_handle* h, const void* post_dev, int in_f16, int64_t n_reads, int64_t n_rows,
                                 int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                 int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                 int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                 unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
                                 int32_t nodes_per_row, void* stream) {
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (int rc = check_beam(h, n_labels, beam_size, beam_cut_threshold)) return rc;
    if (!post_dev || !labels_dev || !path_dev || !offsets_dev || !lens_dev || !status_dev || !pool_used_dev || !scratch_dev)
        return fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (reinterpret_cast<uintptr_t>(scratch_dev) & 255u) return fail(h, FCTC_E_ARGUMENT, "scratch_dev must be 256-byte aligned");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::BeamParams p{};
    int64_t need = fctc::beam_scratch_bytes(h->sm_count, n_reads, n_rows, n_labels, beam_size, nodes_per_row,
                                            scratch_bytes, &p.nodes_per_slot, &p.n_slots);
    if (scratch_bytes < need)
        return fail(h, FCTC_E_CAPACITY, "beam scratch too small: %lld bytes given, %lld needed", (long long)scratch_bytes,
                    (long long)need);
    p.post = static_cast<const float*>(post_dev); set_input(p, in_f16); p.N = n_reads; p.T = n_rows; p.C = n_labels;
    p.beam_size = beam_size;
    p.threshold = beam_cut_threshold; p.collapse = collapse_repeats; p.labels = labels_dev; p.path = path_dev;
    p.pool_cap = pool_capacity; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.cursor = pool_used_dev; p.scratch = scratch_dev; p.scratch_bytes = scratch_bytes;
    p.ticket = static_cast<unsigned long long*>(scratch_dev);
    FCTC_CUDA(h, fctc::launch_beam(p, h->sm_count, static_cast<cudaStream_t>(stream)));
    return 0;
}
}  // namespace

int fctc_beam_search_batch_device(fctc_handle* h, const float* post_dev, int64_t n_reads, int64_t n_rows,
                                  int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                  int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                  int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                  unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
                                  int32_t nodes_per_row, void* stream) {
    return beam_search_batch_device_any(h, post_dev, 0, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                        collapse_repeats, labels_dev, path_dev, pool_capacity, offsets_dev, lens_dev,
                                        status_dev, pool_used_dev, scratch_dev, scratch_bytes, nodes_per_row, stream);
}

int fctc_beam_search_batch_device_f16(fctc_handle* h, const void* post_dev_f16, int64_t n_reads, int64_t n_rows,
                                      int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                      int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                      int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                                      int32_t* status_dev, unsigned long long* pool_used_dev, void* scratch_dev,
                                      int64_t scratch_bytes, int32_t nodes_per_row, void* stream) {
    return beam_search_batch_device_any(h, post_dev_f16, 1, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                        collapse_repeats, labels_dev, path_dev, pool_capacity, offsets_dev, lens_dev,
                                        status_dev, pool_used_dev, scratch_dev, scratch_bytes, nodes_per_row, stream);
}

namespace {

int beam_host_pass(fctc_handle* h, const void* post_host, int in_f16, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                   int32_t beam_size, float thr, int32_t collapse, int32_t nodes_per_row, HostOut& out) {
    const int elem_bytes = (in_f16 & 0xf) ? 2 : 4;
    const int64_t slot_entries = std::max<int64_t>(roundup32(n_rows), 32);
    // The search keeps one read per resident warp busy for the read's whole length, so a chunk is a whole number of
    // rounds of the resident warps (a 64 MiB chunk would leave most of the GPU idle), capped at 2 GiB.  How many
    // rounds: the narrow-beam kernel on float32 rows is about as fast as the link (copy-bound end to end), and small
    // chunks keep more of the copies under the kernels — one round (beam 5, 65,536 reads: 2.08e9 -> 2.23e9 rows/s);
    // every other case is kernel-bound and loses time between kernels — three rounds (beam 32: 4.9e8 vs 5.5e8).
    const int64_t read_bytes = std::max<int64_t>(n_rows * int64_t(n_labels) * elem_bytes, 1);
    int round_reads = 0;
    fctc::beam_scratch_bytes(h->sm_count, n_reads, n_rows, n_labels, beam_size, nodes_per_row, 0, nullptr, &round_reads);
    const int rounds = (elem_bytes == 4 && fctc::beam_small_eligible(beam_size, n_labels, n_rows)) ? 1 : 3;
    int64_t chunk = std::max<int64_t>(default_chunk_reads(read_bytes), int64_t(std::max(round_reads, 1)) * rounds);
    chunk = std::min<int64_t>(chunk, std::max<int64_t>((int64_t(2) << 30) / read_bytes, 1));
    static const int chunk_knob = fctc::env_knob("FCTC_BEAM_CHUNK_READS");  // tuning knob
    if (chunk_knob > 0) chunk = chunk_knob;
    chunk = std::min<int64_t>(chunk, n_reads);
    int64_t nodes_per_slot = 0;
    int n_slots = 0;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) free_b = size_t(16) << 30;
    // One trie scratch serves both pipeline slots: every kernel fills the GPU on its own, so the two
    // chunks' kernels are made to run in turn (an event per kernel) and each gets the whole budget —
    // with a scratch per slot, wide beams on long reads ran at half the resident warps.  The budget is
    // three quarters of what is free once the handle's reusable scratch is counted as free.
    Slot& owner = h->slot[0];
    const int64_t budget = std::max<int64_t>(int64_t((free_b + owner.scratch.bytes) / 4 * 3), int64_t(1) << 28);
    const int64_t need = fctc::beam_scratch_bytes(h->sm_count, chunk, n_rows, n_labels, beam_size, nodes_per_row, budget,
                                                  &nodes_per_slot, &n_slots);
    // Wide beams with a pool for every warp the SMs can hold: the chunks' kernels are not made to run in turn but
    // share the pools through a ring of free slot numbers (kernels.h), so that a chunk's first CTAs start as the
    // chunk before it runs out of reads.  A read takes 0.7x to 1.7x the median time, and a kernel alone on the GPU
    // spends that spread with most warps idle (one round of beam-32 reads: 112 ms alone, 57-65 ms in a long launch).
    const int wide_per_sm = fctc::beam_wide_warps_per_sm(beam_size, n_labels, n_rows);
    static const int overlap_knob = fctc::env_knob("FCTC_BEAM_OVERLAP");  // tuning knob: 1 = never share (kernels in turn)
    const bool shared_pools = wide_per_sm > 0 && n_slots == h->sm_count * wide_per_sm && chunk < n_reads && overlap_knob != 1;
    const size_t ring_off = (size_t(need) + 255) & ~size_t(255);
    if (int rc = ensure(h, owner.scratch, shared_pools ? ring_off + size_t(n_slots) * 4 + 256 : size_t(need))) return rc;
    unsigned int* ring = reinterpret_cast<unsigned int*>(static_cast<unsigned char*>(owner.scratch.p) + ring_off);
    if (shared_pools) {
        FCTC_CUDA(h, fctc::launch_beam_ring_init(ring + 64, ring, n_slots, owner.stream));
        FCTC_CUDA(h, cudaStreamSynchronize(owner.stream));
    }
    // every resident warp decodes one read at a time: a chunk that is not a whole number of such rounds
    // pays a full round for its leftover reads
    if (chunk > n_slots && chunk < n_reads) chunk = chunk / n_slots * n_slots;
    bool first_launch = true;
    auto launch = [&](Slot& s, int64_t cnt) -> int {
        fctc::BeamParams p{};
        fctc::beam_scratch_bytes(h->sm_count, cnt, n_rows, n_labels, beam_size, nodes_per_row, need, &p.nodes_per_slot, &p.n_slots);
        p.post = static_cast<const float*>(s.post.p); set_input(p, in_f16); p.N = cnt; p.T = n_rows; p.C = n_labels;
        p.beam_size = beam_size; p.threshold = thr; p.collapse = collapse;
        p.labels = static_cast<uint8_t*>(s.labels.p); p.path = static_cast<int32_t*>(s.path.p);
        p.pool_cap = cnt * slot_entries; p.offsets = static_cast<int64_t*>(s.offsets.p);
        p.lens = static_cast<int32_t*>(s.lens.p); p.status = static_cast<int32_t*>(s.status.p);
        p.cursor = static_cast<unsigned long long*>(s.cursor.p); p.scratch = owner.scratch.p;
        p.scratch_bytes = int64_t(need); p.ticket = static_cast<unsigned long long*>(owner.scratch.p);
        Slot& other = (&s == &h->slot[0]) ? h->slot[1] : h->slot[0];
        if (shared_pools) {
            p.ticket = static_cast<unsigned long long*>(s.cursor.p) + 1;  // a counter per chunk: two kernels are in flight
            p.slot_ring = ring + 64; p.ring_ctr = ring; p.ring_size = n_slots;
        } else if (!first_launch) {
            FCTC_CUDA(h, cudaStreamWaitEvent(s.stream, other.kernel_done, 0));  // the scratch is free again
        }
        first_launch = false;
        FCTC_CUDA(h, fctc::launch_beam(p, h->sm_count, s.stream));
        FCTC_CUDA(h, cudaEventRecord(s.kernel_done, s.stream));
        return 0;
    };
    return host_pipeline(h, post_host, elem_bytes, n_reads, n_rows, n_labels, out, chunk, launch);
}

int beam_search_batch_any(fctc_handle* h, const void* post_host_any, int in_f16, int64_t n_reads, int64_t n_rows,
                          int32_t n_labels, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                          uint8_t* labels, int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                          int32_t* status, int64_t* pool_used) {
    const unsigned char* post_host = static_cast<const unsigned char*>(post_host_any);
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (int rc = check_beam(h, n_labels, beam_size, beam_cut_threshold)) return rc;
    if ((!post_host && n_rows > 0) || !labels || !path || !offsets || !lens || !status)
        return fail(h, FCTC_E_ARGUMENT, "null host pointer");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    if (pool_used) *pool_used = 0;
    if (n_reads == 0) return 0;
    HostOut out{labels, nullptr, 1, path, pool_capacity, offsets, lens, status};
    if (int rc = beam_host_pass(h, post_host, in_f16, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                collapse_repeats, 0, out))
        return rc;
    if (!out.overflow) {
        // reads whose trie outgrew the default pool: gathered into one array and decoded again in one pass (groups of
        // at most 256 MiB of posteriors) with the hard bound of (n_labels-1)*beam_size proposals per row
        const int32_t bound = std::max<int32_t>(1, (n_labels - 1) * beam_size);
        const int64_t read_bytes = n_rows * int64_t(n_labels) * ((in_f16 & 0xf) ? 2 : 4);
        std::vector<int64_t> again;
        for (int64_t n = 0; n < n_reads; ++n)
            if (status[n] == FCTC_NODE_POOL_OVERFLOW) again.push_back(n);
        const int64_t group = std::max<int64_t>(1, (int64_t(256) << 20) / std::max<int64_t>(read_bytes, 1));
        std::vector<unsigned char> gathered;
        std::vector<int64_t> off2;
        std::vector<int32_t> len2, st2;
        for (size_t g0 = 0; g0 < again.size() && !out.overflow; g0 += size_t(group)) {
            const int64_t k = std::min<int64_t>(group, int64_t(again.size() - g0));
            gathered.resize(size_t(std::max<int64_t>(k * read_bytes, 16)));
            for (int64_t i = 0; i < k; ++i)
                memcpy(gathered.data() + i * read_bytes, post_host + again[g0 + i] * read_bytes, size_t(read_bytes));
            off2.assign(size_t(k), 0); len2.assign(size_t(k), 0); st2.assign(size_t(k), 0);
            HostOut two{labels, nullptr, 1, path, pool_capacity, off2.data(), len2.data(), st2.data()};
            two.used = out.used;
            if (int rc = beam_host_pass(h, gathered.data(), in_f16, k, n_rows, n_labels, beam_size, beam_cut_threshold,
                                        collapse_repeats, bound, two))
                return rc;
            out.used = two.used;
            out.overflow = two.overflow;
            if (!two.overflow)
                for (int64_t i = 0; i < k; ++i) {
                    const int64_t n = again[g0 + i];
                    offsets[n] = off2[size_t(i)]; lens[n] = len2[size_t(i)]; status[n] = st2[size_t(i)];
                }
        }
    }
    if (pool_used) *pool_used = out.used;
    if (out.overflow)
        return fail(h, FCTC_E_CAPACITY, "result pool too small: %lld entries given (n_reads*round_up(n_rows,32) always suffices)",
                    (long long)pool_capacity);
    return 0;
}

}  // namespace

int fctc_beam_search_batch(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                           int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats, uint8_t* labels,
                           int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens, int32_t* status,
                           int64_t* pool_used) {
    return beam_search_batch_any(h, post_host, 0, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold, collapse_repeats,
                                 labels, path, pool_capacity, offsets, lens, status, pool_used);
}

int fctc_beam_search_batch_f16(fctc_handle* h, const void* post_host_f16, int64_t n_reads, int64_t n_rows,
                               int32_t n_labels, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                               uint8_t* labels, int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                               int32_t* status, int64_t* pool_used) {
    return beam_search_batch_any(h, post_host_f16, 1, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                 collapse_repeats, labels, path, pool_capacity, offsets, lens, status, pool_used);
}

int fctc_beam_search_batch_device_in(fctc_handle* h, const void* in_dev, int32_t input_kind, int64_t n_reads,
                                     int64_t n_rows, int32_t n_labels, int32_t beam_size, float beam_cut_threshold,
                                     int32_t collapse_repeats, uint8_t* labels_dev, int32_t* path_dev,
                                     int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev, int32_t* status_dev,
                                     unsigned long long* pool_used_dev, void* scratch_dev, int64_t scratch_bytes,
                                     int32_t nodes_per_row, void* stream) {
    if (int rc = check_input_kind(h, input_kind)) return rc;
    return beam_search_batch_device_any(h, in_dev, input_kind, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                        collapse_repeats, labels_dev, path_dev, pool_capacity, offsets_dev, lens_dev,
                                        status_dev, pool_used_dev, scratch_dev, scratch_bytes, nodes_per_row, stream);
}

int fctc_beam_search_batch_in(fctc_handle* h, const void* in_host, int32_t input_kind, int64_t n_reads, int64_t n_rows,
                              int32_t n_labels, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                              uint8_t* labels, int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                              int32_t* status, int64_t* pool_used) {
    if (int rc = check_input_kind(h, input_kind)) return rc;
    return beam_search_batch_any(h, in_host, input_kind, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                 collapse_repeats, labels, path, pool_capacity, offsets, lens, status, pool_used);
}

int fctc_beam_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_labels, int64_t stride_row,
                     int64_t stride_label, int32_t beam_size, float beam_cut_threshold, int32_t collapse_repeats,
                     uint8_t* labels, int64_t* path, int64_t capacity, int64_t* out_len) {
    if (int rc = check_common(h, 1, n_rows, n_labels)) return rc;
    if (!out_len || (!post_host && n_rows > 0)) return fail(h, FCTC_E_ARGUMENT, "null pointer");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    SingleStage S;
    if (int rc = stage_single(h, post_host, n_rows, n_labels, stride_row, stride_label, &S)) return rc;
    int64_t used = 0;
    int rc = fctc_beam_search_batch(h, S.packed, 1, n_rows, n_labels, beam_size, beam_cut_threshold, collapse_repeats,
                                    S.labels, S.path, S.slot_entries, S.off, S.len, S.st, &used);
    if (rc < 0) return rc;
    const int64_t len = *S.len, o = *S.off;
    *out_len = len;
    const int64_t m = std::min<int64_t>(len, capacity);
    for (int64_t i = 0; i < m; ++i) {
        if (labels) labels[i] = S.labels[o + i];
        if (path) path[i] = S.path[o + i];
    }
    return *S.st;
}

// ---------------------------------------------------------------------------------------------
// FASTQ record assembly
// ---------------------------------------------------------------------------------------------
int64_t fctc_fastq_scratch_bytes(const fctc_handle* h, int64_t n_reads) {
    (void)h;
    return n_reads < 0 ? 0 : fctc::fastq_scratch_bytes(n_reads);
}

int fctc_fastq_records_device(fctc_handle* h, const uint8_t* labels_dev, const uint8_t* quals_dev,
                              const int64_t* offsets_dev, const int32_t* lens_dev, const int32_t* status_dev,
                              int64_t n_reads, const char* alphabet, int32_t n_labels, const char* name_prefix,
                              int64_t first_index, uint8_t* out_dev, int64_t out_capacity, int64_t* record_offsets_dev,
                              void* scratch_dev, int64_t scratch_bytes, void* stream) {
    if (!h) return FCTC_E_ARGUMENT;
    if (n_reads < 0 || first_index < 0) return fail(h, FCTC_E_ARGUMENT, "negative read count or first index");
    if (!labels_dev || !quals_dev || !offsets_dev || !lens_dev || !status_dev || !out_dev || !record_offsets_dev || !scratch_dev)
        return fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (!alphabet || n_labels < 1 || n_labels > 256)
        return fail(h, FCTC_E_ARGUMENT, "alphabet must give one character for each of 1..256 labels");
    const size_t plen = name_prefix ? strlen(name_prefix) : 0;
    if (plen > 63) return fail(h, FCTC_E_ARGUMENT, "name_prefix longer than 63 characters");
    if (scratch_bytes < fctc::fastq_scratch_bytes(n_reads))
        return fail(h, FCTC_E_CAPACITY, "fastq scratch too small: %lld bytes given, %lld needed", (long long)scratch_bytes,
                    (long long)fctc::fastq_scratch_bytes(n_reads));
    if (reinterpret_cast<uintptr_t>(scratch_dev) & 7u) return fail(h, FCTC_E_ARGUMENT, "scratch_dev must be 8-byte aligned");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::FastqParams p{};
    p.labels = labels_dev; p.quals = quals_dev; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.N = n_reads; p.first_index = first_index; p.out = out_dev; p.out_capacity = out_capacity;
    p.prefix_len = int(plen);
    memcpy(p.prefix, name_prefix ? name_prefix : "", plen);
    for (int i = 0; i < 256; ++i) p.alphabet[i] = (unsigned char)(i < n_labels ? alphabet[i] : '?');
    FCTC_CUDA(h, fctc::launch_fastq(p, reinterpret_cast<long long*>(record_offsets_dev), scratch_dev,
                                    static_cast<cudaStream_t>(stream)));
    return 0;
}

int fctc_duplex_envelope_device(fctc_handle* h, const int32_t* path1_dev, const int64_t* offsets1_dev,
                                const int32_t* lens1_dev, const int32_t* path2_dev, const int64_t* offsets2_dev,
                                const int32_t* lens2_dev, int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2,
                                int32_t pad_rows, int64_t* envelope_dev, int32_t* max_window_dev, void* stream) {
    if (!h) return FCTC_E_ARGUMENT;
    if (n_pairs < 0 || n_rows_1 <= 0 || n_rows_2 <= 0 || pad_rows < 0)
        return fail(h, FCTC_E_ARGUMENT, "n_pairs >= 0, n_rows_1 > 0, n_rows_2 > 0 and pad_rows >= 0 are required");
    if (!path1_dev || !offsets1_dev || !lens1_dev || !path2_dev || !offsets2_dev || !lens2_dev || !envelope_dev)
        return fail(h, FCTC_E_ARGUMENT, "null device pointer");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::EnvelopeParams p{};
    p.path1 = path1_dev; p.offsets1 = offsets1_dev; p.lens1 = lens1_dev;
    p.path2 = path2_dev; p.offsets2 = offsets2_dev; p.lens2 = lens2_dev;
    p.N = n_pairs; p.T1 = n_rows_1; p.T2 = n_rows_2; p.pad = pad_rows;
    p.env = reinterpret_cast<long long*>(envelope_dev);
    p.max_window = max_window_dev;
    FCTC_CUDA(h, fctc::launch_envelope(p, static_cast<cudaStream_t>(stream)));
    return 0;
}

int64_t fctc_align_scratch_bytes(const fctc_handle* h, int64_t n_pairs, int64_t max_bases, int32_t band_bases) {
    if (!h || n_pairs < 0 || max_bases < 0 || band_bases < 1) return 0;
    return fctc::align_scratch_bytes(h->sm_count, n_pairs, max_bases, band_bases, 0, nullptr, nullptr);
}

int fctc_duplex_envelope_aligned_device(fctc_handle* h, const uint8_t* labels1_dev, const int32_t* path1_dev,
                                        const int64_t* offsets1_dev, const int32_t* lens1_dev, const uint8_t* labels2_dev,
                                        const int32_t* path2_dev, const int64_t* offsets2_dev, const int32_t* lens2_dev,
                                        int64_t n_pairs, int64_t n_rows_1, int64_t n_rows_2, int32_t band_bases,
                                        int32_t pad_rows, int64_t* envelope_dev, int32_t* max_window_dev,
                                        int32_t* partner_dev, void* scratch_dev, int64_t scratch_bytes, void* stream) {
    if (!h) return FCTC_E_ARGUMENT;
    if (n_pairs < 0 || n_rows_1 <= 0 || n_rows_2 <= 0 || pad_rows < 0)
        return fail(h, FCTC_E_ARGUMENT, "n_pairs >= 0, n_rows_1 > 0, n_rows_2 > 0 and pad_rows >= 0 are required");
    if (band_bases < 1 || band_bases > 127) return fail(h, FCTC_E_ARGUMENT, "band_bases must lie in 1..127");
    if (!labels1_dev || !path1_dev || !offsets1_dev || !lens1_dev || !labels2_dev || !path2_dev || !offsets2_dev || !lens2_dev ||
        !envelope_dev || !partner_dev || !scratch_dev)
        return fail(h, FCTC_E_ARGUMENT, "null device pointer");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::AlignParams a{};
    a.labels1 = labels1_dev; a.offsets1 = offsets1_dev; a.lens1 = lens1_dev;
    a.labels2 = labels2_dev; a.offsets2 = offsets2_dev; a.lens2 = lens2_dev;
    a.N = n_pairs; a.band = band_bases; a.partner = partner_dev; a.dirs = static_cast<uint8_t*>(scratch_dev);
    // a read emits at most one base per row: n_rows_1 bounds read 1's length
    fctc::align_scratch_bytes(h->sm_count, n_pairs, n_rows_1, band_bases, scratch_bytes, &a.n_slots, &a.dir_bytes_per_slot);
    if (n_pairs > 0 && a.n_slots < 1)
        return fail(h, FCTC_E_CAPACITY, "alignment scratch too small: %lld bytes given, one pair needs %lld (fctc_align_scratch_bytes)",
                    (long long)scratch_bytes, (long long)a.dir_bytes_per_slot);
    FCTC_CUDA(h, fctc::launch_align(a, static_cast<cudaStream_t>(stream)));
    fctc::EnvelopeParams p{};
    p.path1 = path1_dev; p.offsets1 = offsets1_dev; p.lens1 = lens1_dev;
    p.path2 = path2_dev; p.offsets2 = offsets2_dev; p.lens2 = lens2_dev;
    p.N = n_pairs; p.T1 = n_rows_1; p.T2 = n_rows_2; p.pad = pad_rows;
    p.env = reinterpret_cast<long long*>(envelope_dev);
    p.max_window = max_window_dev;
    p.partner = partner_dev;
    FCTC_CUDA(h, fctc::launch_envelope(p, static_cast<cudaStream_t>(stream)));
    return 0;
}

int fctc_move_table_device(fctc_handle* h, const int32_t* path_dev, const int64_t* offsets_dev, const int32_t* lens_dev,
                           const int32_t* status_dev, int64_t n_reads, int64_t n_rows, uint8_t* moves_dev,
                           int32_t* dwell_dev, void* stream) {
    if (!h) return FCTC_E_ARGUMENT;
    if (n_reads < 0 || n_rows <= 0) return fail(h, FCTC_E_ARGUMENT, "n_reads >= 0 and n_rows > 0 are required");
    if (!path_dev || !offsets_dev || !lens_dev || !status_dev || (!moves_dev && !dwell_dev))
        return fail(h, FCTC_E_ARGUMENT, "null device pointer (at least one of moves_dev / dwell_dev must be given)");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::MovesParams p{};
    p.path = path_dev; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.N = n_reads; p.T = n_rows; p.moves = moves_dev; p.dwell = dwell_dev;
    FCTC_CUDA(h, fctc::launch_moves(p, h->sm_count, static_cast<cudaStream_t>(stream)));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Synthetic input + host test hooks
// ---------------------------------------------------------------------------------------------
int fctc_synth_posteriors_device(fctc_handle* h, float* post_dev, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                                 int32_t kind, uint64_t seed, int64_t first_read, void* stream) {
    if (!h || !post_dev) return FCTC_E_ARGUMENT;
    FCTC_CUDA(h, cudaSetDevice(h->device));
    FCTC_CUDA(h, fctc::launch_synth(post_dev, n_reads, n_rows, n_labels, kind, seed, first_read,
                                    static_cast<cudaStream_t>(stream)));
    return 0;
}

int fctc_test_poison_shared_memory(fctc_handle* h, int32_t byte) {
    if (!h) return FCTC_E_ARGUMENT;
    FCTC_CUDA(h, cudaSetDevice(h->device));
    if (int rc = ensure(h, h->misc, 256)) return rc;
    FCTC_CUDA(h, fctc::launch_poison_smem(h->sm_count, byte, static_cast<unsigned int*>(h->misc.p), h->stream));
    FCTC_CUDA(h, cudaStreamSynchronize(h->stream));
    return 0;
}

int fctc_test_libm_device(fctc_handle* h, const float* x_dev, const float* y_dev, float* out_dev, int64_t n, int32_t which,
                          void* stream) {
    if (!h || !x_dev || !out_dev || which < 0 || which > 6 || (which >= 4 && !y_dev)) return FCTC_E_ARGUMENT;
    FCTC_CUDA(h, cudaSetDevice(h->device));
    FCTC_CUDA(h, fctc::launch_libm_probe(x_dev, y_dev, out_dev, n, which, static_cast<cudaStream_t>(stream)));
    return 0;
}

uint32_t fctc_host_phred(float prob, float qscale, float qbias) {
    static const double invc[16] = FCTC_LOGF_INVC;
    static const double logc[16] = FCTC_LOGF_LOGC;
    return fctc::phred_code(prob, qscale, qbias, invc, logc);
}

void fctc_host_logf(const float* x, float* y, int64_t n) {
    static const double invc[16] = FCTC_LOGF_INVC;
    static const double logc[16] = FCTC_LOGF_LOGC;
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_logf(x[i], invc, logc);
}

void fctc_host_expf(const float* x, float* y, int64_t n) {
    static const uint64_t tab[32] = FCTC_EXP2F_TAB;
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_expf(x[i], tab);
}

void fctc_host_log1pf(const float* x, float* y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_log1pf(x[i]);
}

void fctc_host_expf_nonpos(const float* x, float* y, int64_t n) {
    static const uint64_t tab[32] = FCTC_EXP2F_TAB;
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_expf_nonpos(x[i], tab);
}

void fctc_host_log1pf_unit(const float* x, float* y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_log1pf_unit(x[i]);
}

void fctc_host_log10f(const float* x, float* y, int64_t n) {
    static const double invc[16] = FCTC_LOGF_INVC;
    static const double logc[16] = FCTC_LOGF_LOGC;
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_log10f(x[i], invc, logc);
}

