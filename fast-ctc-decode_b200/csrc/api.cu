// api.cu — the extern "C" boundary (include/fctc.h): handle, argument checks, host<->device
// streaming for the host-memory entry points.  No decode arithmetic happens on the host.
#include <cuda_runtime.h>

#include <algorithm>
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
    DevBuf post, labels, quals, path, offsets, lens, status, cursor, scratch;
    unsigned long long* cursor_host = nullptr;  // pinned
    int64_t first = 0, count = 0;               // reads of the chunk in flight
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
    if (n_reads < 0 || n_rows <= 0) return fail(h, FCTC_E_ARGUMENT, "n_reads must be >= 0 and n_rows >= 1");
    if (n_labels <= 0) return fail(h, FCTC_E_ARGUMENT, "Empty alphabet given");  // src/lib.rs:191
    if (n_labels > 256) return fail(h, FCTC_E_UNSUPPORTED, "alphabets above 256 labels are not supported (labels are uint8)");
    if (n_rows > (int64_t(1) << 30)) return fail(h, FCTC_E_UNSUPPORTED, "reads above 2^30 rows are not supported");
    return 0;
}

int roundup32(int64_t v) { return int((v + 31) & ~int64_t(31)); }

}  // namespace

extern "C" {

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
        for (DevBuf* b : {&s.post, &s.labels, &s.quals, &s.path, &s.offsets, &s.lens, &s.status, &s.cursor, &s.scratch})
            release(*b);
        if (s.cursor_host) cudaFreeHost(s.cursor_host);
        if (s.done) cudaEventDestroy(s.done);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    release(h->misc);
    if (h->pinned) cudaFreeHost(h->pinned);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

// ---------------------------------------------------------------------------------------------
// Viterbi
// ---------------------------------------------------------------------------------------------
int fctc_viterbi_batch_device(fctc_handle* h, const float* post_dev, int64_t n_reads, int64_t n_rows,
                              int32_t n_labels, int32_t qstring, float qscale, float qbias,
                              int32_t collapse_repeats, uint8_t* labels_dev, void* quals_dev, int32_t qual_bytes,
                              int32_t* path_dev, int64_t pool_capacity, int64_t* offsets_dev, int32_t* lens_dev,
                              int32_t* status_dev, unsigned long long* pool_used_dev, void* stream) {
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (!post_dev || !labels_dev || !path_dev || !offsets_dev || !lens_dev || !status_dev || !pool_used_dev)
        return fail(h, FCTC_E_ARGUMENT, "null device pointer");
    if (qstring && !quals_dev) return fail(h, FCTC_E_ARGUMENT, "qstring requested but quals_dev is NULL");
    if (qual_bytes != 1 && qual_bytes != 4) return fail(h, FCTC_E_ARGUMENT, "qual_bytes must be 1 or 4");
    if (qstring && qual_bytes == 1 && !quality_fits_u8(qscale, qbias))
        return fail(h, FCTC_E_ARGUMENT, "qual_bytes=1 cannot hold the quality codes of qscale=%g qbias=%g; use 4", qscale,
                    qbias);
    if ((reinterpret_cast<uintptr_t>(labels_dev) | reinterpret_cast<uintptr_t>(path_dev) |
         reinterpret_cast<uintptr_t>(quals_dev)) & 15u)
        return fail(h, FCTC_E_ARGUMENT, "result pool arrays must be 16-byte aligned");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    fctc::ViterbiParams p{};
    p.post = post_dev; p.N = n_reads; p.T = n_rows; p.C = n_labels;
    p.qstring = qstring; p.qscale = qscale; p.qbias = qbias; p.collapse = collapse_repeats;
    p.labels = labels_dev; p.quals = quals_dev; p.qual_bytes = qual_bytes; p.path = path_dev;
    p.pool_cap = pool_capacity; p.offsets = offsets_dev; p.lens = lens_dev; p.status = status_dev;
    p.cursor = pool_used_dev;
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : h->stream;
    FCTC_CUDA(h, fctc::launch_viterbi(p, h->sm_count, st, nullptr));
    return 0;
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
    for (int64_t i = 0; i < s.count; ++i) out.offsets[s.first + i] += out.used;
    out.used += used;
    s.busy = false;
    return 0;
}

}  // namespace

int fctc_viterbi_batch(fctc_handle* h, const float* post_host, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                       int32_t qstring, float qscale, float qbias, int32_t collapse_repeats, uint8_t* labels,
                       void* quals, int32_t qual_bytes, int32_t* path, int64_t pool_capacity, int64_t* offsets,
                       int32_t* lens, int32_t* status, int64_t* pool_used) {
    if (int rc = check_common(h, n_reads, n_rows, n_labels)) return rc;
    if (!post_host || !labels || !path || !offsets || !lens || !status)
        return fail(h, FCTC_E_ARGUMENT, "null host pointer");
    if (qstring && !quals) return fail(h, FCTC_E_ARGUMENT, "qstring requested but quals is NULL");
    if (qual_bytes != 1 && qual_bytes != 4) return fail(h, FCTC_E_ARGUMENT, "qual_bytes must be 1 or 4");
    if (qstring && qual_bytes == 1 && !quality_fits_u8(qscale, qbias))
        return fail(h, FCTC_E_ARGUMENT, "qual_bytes=1 cannot hold the quality codes of qscale=%g qbias=%g; use 4", qscale,
                    qbias);
    FCTC_CUDA(h, cudaSetDevice(h->device));
    if (pool_used) *pool_used = 0;
    if (n_reads == 0) return 0;

    const int64_t read_bytes = n_rows * int64_t(n_labels) * 4;
    const int64_t slot_entries = roundup32(n_rows);
    // chunk: ~64 MiB of posteriors, at least one read, at most what keeps 2 x the SM count busy
    int64_t chunk = std::max<int64_t>(1, (int64_t(64) << 20) / read_bytes);
    chunk = std::min(chunk, n_reads);
    HostOut out{labels, qstring ? quals : nullptr, qual_bytes, path, pool_capacity, offsets, lens, status};

    for (int i = 0; i < 2; ++i) {
        Slot& s = h->slot[i];
        if (int rc = ensure(h, s.post, size_t(chunk * read_bytes))) return rc;
        if (int rc = ensure(h, s.labels, size_t(chunk * slot_entries))) return rc;
        if (int rc = ensure(h, s.path, size_t(chunk * slot_entries) * 4)) return rc;
        if (qstring)
            if (int rc = ensure(h, s.quals, size_t(chunk * slot_entries) * qual_bytes)) return rc;
        if (int rc = ensure(h, s.offsets, size_t(chunk) * 8)) return rc;
        if (int rc = ensure(h, s.lens, size_t(chunk) * 4)) return rc;
        if (int rc = ensure(h, s.status, size_t(chunk) * 4)) return rc;
        if (int rc = ensure(h, s.cursor, 256)) return rc;
        s.busy = false;
    }

    int which = 0;
    for (int64_t first = 0; first < n_reads; first += chunk, which ^= 1) {
        Slot& s = h->slot[which];
        if (int rc = drain_slot(h, s, out)) return rc;
        const int64_t cnt = std::min(chunk, n_reads - first);
        s.first = first;
        s.count = cnt;
        FCTC_CUDA(h, cudaMemcpyAsync(s.post.p, post_host + first * n_rows * n_labels, size_t(cnt * read_bytes),
                                     cudaMemcpyHostToDevice, s.stream));
        FCTC_CUDA(h, cudaMemsetAsync(s.cursor.p, 0, 8, s.stream));
        fctc::ViterbiParams p{};
        p.post = static_cast<const float*>(s.post.p); p.N = cnt; p.T = n_rows; p.C = n_labels;
        p.qstring = qstring; p.qscale = qscale; p.qbias = qbias; p.collapse = collapse_repeats;
        p.labels = static_cast<uint8_t*>(s.labels.p); p.quals = s.quals.p; p.qual_bytes = qual_bytes;
        p.path = static_cast<int32_t*>(s.path.p); p.pool_cap = cnt * slot_entries;
        p.offsets = static_cast<int64_t*>(s.offsets.p); p.lens = static_cast<int32_t*>(s.lens.p);
        p.status = static_cast<int32_t*>(s.status.p); p.cursor = static_cast<unsigned long long*>(s.cursor.p);
        FCTC_CUDA(h, fctc::launch_viterbi(p, h->sm_count, s.stream, nullptr));
        FCTC_CUDA(h, cudaMemcpyAsync(s.cursor_host, s.cursor.p, 8, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(offsets + first, s.offsets.p, size_t(cnt) * 8, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(lens + first, s.lens.p, size_t(cnt) * 4, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaMemcpyAsync(status + first, s.status.p, size_t(cnt) * 4, cudaMemcpyDeviceToHost, s.stream));
        FCTC_CUDA(h, cudaEventRecord(s.done, s.stream));
        s.busy = true;
    }
    // drain in submission order: the slot used before the last one first
    if (int rc = drain_slot(h, h->slot[which], out)) return rc;
    if (int rc = drain_slot(h, h->slot[which ^ 1], out)) return rc;
    FCTC_CUDA(h, cudaStreamSynchronize(h->slot[0].stream));
    FCTC_CUDA(h, cudaStreamSynchronize(h->slot[1].stream));
    if (pool_used) *pool_used = out.used;
    if (out.overflow)
        return fail(h, FCTC_E_CAPACITY, "result pool too small: %lld entries given (n_reads*round_up(n_rows,32) always suffices)",
                    (long long)pool_capacity);
    return 0;
}

int fctc_viterbi_search(fctc_handle* h, const float* post_host, int64_t n_rows, int32_t n_labels,
                        int64_t stride_row, int64_t stride_label, int32_t qstring, float qscale, float qbias,
                        int32_t collapse_repeats, uint8_t* labels, uint32_t* quals, int64_t* path, int64_t capacity,
                        int64_t* out_len) {
    if (int rc = check_common(h, 1, n_rows, n_labels)) return rc;
    if (!post_host || !out_len) return fail(h, FCTC_E_ARGUMENT, "null pointer");
    FCTC_CUDA(h, cudaSetDevice(h->device));
    const int64_t slot_entries = roundup32(n_rows);
    const size_t post_bytes = size_t(n_rows) * n_labels * 4;
    const size_t lab_off = (post_bytes + 255) & ~size_t(255);
    const size_t qual_off = lab_off + size_t(slot_entries);
    const size_t path_off = qual_off + size_t(slot_entries) * 4;
    const size_t meta_off = path_off + size_t(slot_entries) * 4;
    if (int rc = ensure_pinned(h, meta_off + 64)) return rc;
    unsigned char* pin = static_cast<unsigned char*>(h->pinned);
    float* packed = reinterpret_cast<float*>(pin);
    for (int64_t t = 0; t < n_rows; ++t)
        for (int32_t c = 0; c < n_labels; ++c) packed[t * n_labels + c] = post_host[t * stride_row + c * stride_label];
    int64_t* off1 = reinterpret_cast<int64_t*>(pin + meta_off);
    int32_t* len1 = reinterpret_cast<int32_t*>(pin + meta_off + 8);
    int32_t* st1 = reinterpret_cast<int32_t*>(pin + meta_off + 16);
    int64_t used = 0;
    int rc = fctc_viterbi_batch(h, packed, 1, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats, pin + lab_off,
                                pin + qual_off, 4, reinterpret_cast<int32_t*>(pin + path_off), slot_entries, off1, len1,
                                st1, &used);
    if (rc < 0) return rc;
    const int64_t len = *len1;
    *out_len = len;
    const int64_t m = std::min<int64_t>(len, capacity);
    const int64_t o = *off1;
    for (int64_t i = 0; i < m; ++i) {
        if (labels) labels[i] = pin[lab_off + o + i];
        if (path) path[i] = reinterpret_cast<int32_t*>(pin + path_off)[o + i];
        if (quals && qstring) quals[i] = reinterpret_cast<uint32_t*>(pin + qual_off)[o + i];
    }
    return *st1;
}

// ---------------------------------------------------------------------------------------------
// Synthetic input + host test hooks
// ---------------------------------------------------------------------------------------------
int fctc_synth_posteriors_device(fctc_handle* h, float* post_dev, int64_t n_reads, int64_t n_rows, int32_t n_labels,
                                 int32_t kind, uint64_t seed, int64_t first_read, void* stream) {
    if (!h || !post_dev) return FCTC_E_ARGUMENT;
    FCTC_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : h->stream;
    FCTC_CUDA(h, fctc::launch_synth(post_dev, n_reads, n_rows, n_labels, kind, seed, first_read, st));
    return 0;
}

uint32_t fctc_host_phred(float prob, float qscale, float qbias) {
    static const double invc[16] = FCTC_LOGF_INVC;
    static const double logc[16] = FCTC_LOGF_LOGC;
    return fctc::phred_code(prob, qscale, qbias, invc, logc);
}

void fctc_host_log10f(const float* x, float* y, int64_t n) {
    static const double invc[16] = FCTC_LOGF_INVC;
    static const double logc[16] = FCTC_LOGF_LOGC;
    for (int64_t i = 0; i < n; ++i) y[i] = fctc::glibc_log10f(x[i], invc, logc);
}

}  // extern "C"
