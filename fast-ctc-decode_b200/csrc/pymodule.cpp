// pymodule.cpp — CPython extension `fast_ctc_decode`: the drop-in for the reference's pyo3 module
// (src/lib.rs:142-628), written in C++ over the C ABI (include/fctc.h) because no Rust toolchain
// exists in this image.  Same function names, parameter names, order and defaults
// (src/lib.rs:171-179, 215-216, 253-254, 319-322, 402-410, 491-494), same exception classes and
// message texts; the decode itself runs in libfctc_b200.so on the GPU — there is no CPU path.
//
// Additions (the north star's batched entry points): viterbi_search_batch, beam_search_batch,
// beam_search_duplex_batch over [N, T, C] float32 arrays, returning one result per read.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/fctc.h"

namespace py = pybind11;

namespace {

int g_exp_mode = 0;  // 0 = exact exp (--no-default-features semantics), 1 = default build (fastexp == 0), 2 = CUDA libm

// One handle per host thread at a time (src/lib.rs releases the GIL the same way).  A thread that ends hands its handle
// (streams, pinned staging, device buffers) to a pool from which the next new thread takes it, so that programs
// which decode from short-lived threads hold as many handles as they ever had threads at once, not one per thread
// ever started.  Nothing is destroyed at exit (the CUDA runtime may be gone by then); the pool itself is never freed.
struct HandlePool {
    std::mutex lock;
    std::vector<fctc_handle*> idle;
};
HandlePool& handle_pool() {
    static HandlePool* pool = new HandlePool();
    return *pool;
}
struct Handle {
    fctc_handle* h = nullptr;
    ~Handle() {
        if (!h) return;
        HandlePool& pool = handle_pool();
        std::lock_guard<std::mutex> lk(pool.lock);
        pool.idle.push_back(h);
    }
};

// Rust's Display for f32 prints the shortest decimal that round-trips (src/lib.rs:346-349 formats
// max_beam_cut with it): "0.2", "0.33333334".
std::string rust_f32(float v) {
    char buf[64];
    for (int prec = 1; prec <= 9; ++prec) {
        snprintf(buf, sizeof buf, "%.*g", prec, double(v));
        if (strtof(buf, nullptr) == v) break;
    }
    return buf;
}

fctc_handle* handle() {
    thread_local Handle th;
    if (!th.h) {
        {
            HandlePool& pool = handle_pool();
            std::lock_guard<std::mutex> lk(pool.lock);
            if (!pool.idle.empty()) {
                th.h = pool.idle.back();
                pool.idle.pop_back();
                return th.h;
            }
        }
        const char* dev = std::getenv("FCTC_DEVICE");
        int rc = fctc_create(dev ? std::atoi(dev) : 0, &th.h);
        if (rc != 0) throw std::runtime_error(std::string("fast_ctc_decode (B200): ") + fctc_last_error(nullptr));
    }
    return th.h;
}

// seq_to_vec, src/lib.rs:144-146: tuple(alphabet), str() of every element
std::vector<std::string> alphabet_of(const py::object& alphabet) {
    py::tuple t = py::tuple(alphabet);
    std::vector<std::string> out;
    out.reserve(t.size());
    for (const auto& item : t) out.push_back(py::str(item).cast<std::string>());
    return out;
}

using F32 = py::array_t<float, 0>;  // no forcecast: a non-float32 argument raises TypeError, as pyo3 does

const py::buffer_info need_dims(const py::array& a, int nd, const char* name) {
    if (a.ndim() != nd) throw py::type_error(std::string(name) + ": expected a " + std::to_string(nd) + "-dimensional float32 array");
    return a.request();
}

void append_utf8(std::string& s, uint32_t cp) {
    if (cp < 0x80) s.push_back(char(cp));
    else if (cp < 0x800) { s.push_back(char(0xC0 | (cp >> 6))); s.push_back(char(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) { s.push_back(char(0xE0 | (cp >> 12))); s.push_back(char(0x80 | ((cp >> 6) & 0x3F))); s.push_back(char(0x80 | (cp & 0x3F))); }
    else { s.push_back(char(0xF0 | (cp >> 18))); s.push_back(char(0x80 | ((cp >> 12) & 0x3F))); s.push_back(char(0x80 | ((cp >> 6) & 0x3F))); s.push_back(char(0x80 | (cp & 0x3F))); }
}

[[noreturn]] void raise_call_error(int rc) {
    const std::string msg = fctc_last_error(handle());
    if (rc == FCTC_E_ARGUMENT) throw py::value_error(msg);
    throw std::runtime_error(msg);
}

void raise_status(int status) {
    if (status == FCTC_OK) return;
    if (const char* m = fctc_status_message(status)) throw std::runtime_error(m);  // PyRuntimeError(format!("{}", e)), src/lib.rs:210
    if (status == FCTC_INVALID_QUALITY_CHAR)
        throw std::runtime_error("quality score is not a valid character (the reference would abort here)");
    if (status == FCTC_REFERENCE_PANIC)
        throw std::runtime_error("NaN or out-of-range state met by a crf search (the reference would abort here)");
    throw std::runtime_error("decode failed with status " + std::to_string(status));
}

std::string join_labels(const std::vector<std::string>& alphabet, const uint8_t* labels, int64_t n) {
    std::string s;
    for (int64_t i = 0; i < n; ++i) s += alphabet[labels[i]];
    return s;
}

int64_t stride_of(const py::buffer_info& b, int axis) { return b.strides[axis] / int64_t(sizeof(float)); }

// ---- viterbi_search, src/lib.rs:171-212 ------------------------------------------------------------
py::tuple viterbi_search(const F32& network_output, const py::object& alphabet, bool qstring, float qscale, float qbias,
                         bool collapse_repeats) {
    const auto names = alphabet_of(alphabet);
    const py::buffer_info b = need_dims(network_output, 2, "network_output");
    if (names.empty()) throw py::value_error("Empty alphabet given");
    if (int64_t(names.size()) != b.shape[1]) throw py::value_error("alphabet size does not match probability matrix dimensions");
    const int64_t T = b.shape[0];
    std::vector<uint8_t> labels(size_t(std::max<int64_t>(T, 1)));
    std::vector<uint32_t> quals(labels.size());
    std::vector<int64_t> path(labels.size());
    int64_t len = 0;
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = fctc_viterbi_search(handle(), static_cast<const float*>(b.ptr), T, int32_t(b.shape[1]), stride_of(b, 0),
                                 stride_of(b, 1), qstring, qscale, qbias, collapse_repeats, labels.data(), quals.data(),
                                 path.data(), int64_t(labels.size()), &len);
    }
    if (rc < 0) raise_call_error(rc);
    raise_status(rc);
    std::string seq = join_labels(names, labels.data(), len);
    if (qstring)
        for (int64_t i = 0; i < len; ++i) append_utf8(seq, quals[size_t(i)]);  // src/search.rs:378-380
    path.resize(size_t(len));
    return py::make_tuple(py::str(seq), py::cast(path));
}

// ---- beam_search, src/lib.rs:319-365 ------------------------------------------------------------------
py::tuple beam_search(const F32& network_output, const py::object& alphabet, int64_t beam_size, float beam_cut_threshold,
                      bool collapse_repeats) {
    const auto names = alphabet_of(alphabet);
    const py::buffer_info b = need_dims(network_output, 2, "network_output");
    const float max_beam_cut = 1.0f / float(names.size());
    if (int64_t(names.size()) != b.shape[1])
        throw py::value_error("alphabet size " + std::to_string(names.size()) + " does not match probability matrix inner dimension " +
                              std::to_string(b.shape[1]));
    if (beam_size < 0) throw py::type_error("beam_size must be a non-negative integer");  // usize extraction
    if (beam_size == 0) throw py::value_error("beam_size cannot be 0");
    if (beam_cut_threshold < -0.0f) throw py::value_error("beam_cut_threshold must be at least 0.0");
    if (beam_cut_threshold >= max_beam_cut) throw py::value_error("beam_cut_threshold cannot be more than " + rust_f32(max_beam_cut));
    const int64_t T = b.shape[0];
    std::vector<uint8_t> labels(size_t(std::max<int64_t>(T, 1)));
    std::vector<int64_t> path(labels.size());
    int64_t len = 0;
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = fctc_beam_search(handle(), static_cast<const float*>(b.ptr), T, int32_t(b.shape[1]), stride_of(b, 0),
                              stride_of(b, 1), int32_t(std::min<int64_t>(beam_size, 1 << 30)), beam_cut_threshold,
                              collapse_repeats, labels.data(), path.data(), int64_t(labels.size()), &len);
    }
    if (rc < 0) raise_call_error(rc);
    raise_status(rc);
    path.resize(size_t(len));
    return py::make_tuple(py::str(join_labels(names, labels.data(), len)), py::cast(path));
}

// ---- beam_search_duplex, src/lib.rs:402-488 --------------------------------------------------------------
py::str beam_search_duplex(const F32& network_output_1, const F32& network_output_2, const py::object& alphabet,
                           const py::object& envelope, int64_t beam_size, float beam_cut_threshold, bool collapse_repeats) {
    const auto names = alphabet_of(alphabet);
    const py::buffer_info b1 = need_dims(network_output_1, 2, "network_output_1");
    const py::buffer_info b2 = need_dims(network_output_2, 2, "network_output_2");
    const float max_beam_cut = 1.0f / float(names.size());
    if (b1.shape[1] != b2.shape[1]) throw py::value_error("inner axes of the network outputs do not match");
    if (int64_t(names.size()) != b1.shape[1])
        throw py::value_error("alphabet size " + std::to_string(names.size()) + " does not match probability matrix inner dimension " +
                              std::to_string(b1.shape[1]));
    if (beam_size < 0) throw py::type_error("beam_size must be a non-negative integer");
    if (beam_size == 0) throw py::value_error("beam_size cannot be 0");
    if (beam_cut_threshold < -0.0f) throw py::value_error("beam_cut_threshold must be at least 0.0");
    if (beam_cut_threshold >= max_beam_cut) throw py::value_error("beam_cut_threshold cannot be more than " + rust_f32(max_beam_cut));
    const uint64_t* env = nullptr;
    int64_t er = 0, ec = 0;
    py::array_t<uint64_t, 0> env_arr;
    if (!envelope.is_none()) {
        if (!py::isinstance<py::array_t<uint64_t, 0>>(envelope)) throw py::type_error("envelope: expected a 2-dimensional uint64 array");
        env_arr = envelope.cast<py::array_t<uint64_t, 0>>();
        if (env_arr.ndim() != 2) throw py::type_error("envelope: expected a 2-dimensional uint64 array");
        if (env_arr.shape(0) != b1.shape[0]) throw py::value_error("the lengths of network_output_1 and envelope do not match");
        if (env_arr.shape(1) != 2) throw py::value_error("the inner axis of envelope must have size 2");
        env = env_arr.data();
        er = env_arr.strides(0) / 8;
        ec = env_arr.strides(1) / 8;
    }
    const int64_t T1 = b1.shape[0];
    std::vector<uint8_t> labels(size_t(std::max<int64_t>(T1, 1)));
    int64_t len = 0;
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = fctc_beam_search_duplex(handle(), static_cast<const float*>(b1.ptr), T1, stride_of(b1, 0), stride_of(b1, 1),
                                     static_cast<const float*>(b2.ptr), b2.shape[0], stride_of(b2, 0), stride_of(b2, 1),
                                     int32_t(b1.shape[1]), env, er, ec, int32_t(std::min<int64_t>(beam_size, 1 << 30)),
                                     beam_cut_threshold, collapse_repeats, g_exp_mode, labels.data(), int64_t(labels.size()), &len);
    }
    if (rc < 0) raise_call_error(rc);
    raise_status(rc);
    return py::str(join_labels(names, labels.data(), len));
}

// ---- batched entry points (new) -----------------------------------------------------------------------------
py::list viterbi_search_batch(const py::array_t<float, py::array::c_style>& network_outputs, const py::object& alphabet,
                              bool qstring, float qscale, float qbias, bool collapse_repeats) {
    const auto names = alphabet_of(alphabet);
    if (network_outputs.ndim() != 3) throw py::type_error("network_outputs: expected a C-contiguous float32 array [N, T, C]");
    const int64_t N = network_outputs.shape(0), T = network_outputs.shape(1), C = network_outputs.shape(2);
    if (names.empty()) throw py::value_error("Empty alphabet given");
    if (int64_t(names.size()) != C) throw py::value_error("alphabet size does not match probability matrix dimensions");
    const int64_t stride = (T + 31) & ~int64_t(31);
    std::vector<uint8_t> labels(size_t(N * stride));
    std::vector<uint32_t> quals(qstring ? labels.size() : 0);
    std::vector<int32_t> path(labels.size());
    std::vector<int64_t> offsets(static_cast<size_t>(N));
    std::vector<int32_t> lens(static_cast<size_t>(N)), status(static_cast<size_t>(N));
    int64_t used = 0;
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = fctc_viterbi_batch(handle(), network_outputs.data(), N, T, int32_t(C), qstring, qscale, qbias, collapse_repeats,
                                labels.data(), qstring ? quals.data() : nullptr, 4, path.data(), int64_t(labels.size()),
                                offsets.data(), lens.data(), status.data(), &used);
    }
    if (rc < 0) raise_call_error(rc);
    py::list out;
    for (int64_t n = 0; n < N; ++n) {
        raise_status(status[size_t(n)]);
        const int64_t o = offsets[size_t(n)], len = lens[size_t(n)];
        std::string seq = join_labels(names, labels.data() + o, len);
        if (qstring)
            for (int64_t i = 0; i < len; ++i) append_utf8(seq, quals[size_t(o + i)]);
        py::list p;
        for (int64_t i = 0; i < len; ++i) p.append(int64_t(path[size_t(o + i)]));
        out.append(py::make_tuple(py::str(seq), p));
    }
    return out;
}

py::list beam_search_batch(const py::array_t<float, py::array::c_style>& network_outputs, const py::object& alphabet,
                           int64_t beam_size, float beam_cut_threshold, bool collapse_repeats) {
    const auto names = alphabet_of(alphabet);
    if (network_outputs.ndim() != 3) throw py::type_error("network_outputs: expected a C-contiguous float32 array [N, T, C]");
    const int64_t N = network_outputs.shape(0), T = network_outputs.shape(1), C = network_outputs.shape(2);
    const float max_beam_cut = 1.0f / float(names.size());
    if (int64_t(names.size()) != C)
        throw py::value_error("alphabet size " + std::to_string(names.size()) + " does not match probability matrix inner dimension " +
                              std::to_string(C));
    if (beam_size <= 0) throw py::value_error("beam_size cannot be 0");
    if (beam_cut_threshold < -0.0f) throw py::value_error("beam_cut_threshold must be at least 0.0");
    if (beam_cut_threshold >= max_beam_cut) throw py::value_error("beam_cut_threshold cannot be more than " + rust_f32(max_beam_cut));
    const int64_t stride = std::max<int64_t>((T + 31) & ~int64_t(31), 32);
    std::vector<uint8_t> labels(size_t(N * stride));
    std::vector<int32_t> path(labels.size());
    std::vector<int64_t> offsets(static_cast<size_t>(N));
    std::vector<int32_t> lens(static_cast<size_t>(N)), status(static_cast<size_t>(N));
    int64_t used = 0;
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = fctc_beam_search_batch(handle(), network_outputs.data(), N, T, int32_t(C), int32_t(std::min<int64_t>(beam_size, 1 << 30)),
                                    beam_cut_threshold, collapse_repeats, labels.data(), path.data(), int64_t(labels.size()),
                                    offsets.data(), lens.data(), status.data(), &used);
    }
    if (rc < 0) raise_call_error(rc);
    py::list out;
    for (int64_t n = 0; n < N; ++n) {
        raise_status(status[size_t(n)]);
        const int64_t o = offsets[size_t(n)], len = lens[size_t(n)];
        py::list p;
        for (int64_t i = 0; i < len; ++i) p.append(int64_t(path[size_t(o + i)]));
        out.append(py::make_tuple(py::str(join_labels(names, labels.data() + o, len)), p));
    }
    return out;
}

py::list beam_search_duplex_batch(const py::array_t<float, py::array::c_style>& network_outputs_1,
                                  const py::array_t<float, py::array::c_style>& network_outputs_2, const py::object& alphabet,
                                  const py::object& envelope, int64_t beam_size, float beam_cut_threshold, bool collapse_repeats) {
    const auto names = alphabet_of(alphabet);
    if (network_outputs_1.ndim() != 3 || network_outputs_2.ndim() != 3)
        throw py::type_error("network_outputs: expected C-contiguous float32 arrays [N, T, C]");
    const int64_t N = network_outputs_1.shape(0), T1 = network_outputs_1.shape(1), C = network_outputs_1.shape(2);
    const int64_t T2 = network_outputs_2.shape(1);
    const float max_beam_cut = 1.0f / float(names.size());
    if (network_outputs_2.shape(2) != C || network_outputs_2.shape(0) != N) throw py::value_error("inner axes of the network outputs do not match");
    if (int64_t(names.size()) != C)
        throw py::value_error("alphabet size " + std::to_string(names.size()) + " does not match probability matrix inner dimension " +
                              std::to_string(C));
    if (beam_size <= 0) throw py::value_error("beam_size cannot be 0");
    if (beam_cut_threshold < -0.0f) throw py::value_error("beam_cut_threshold must be at least 0.0");
    if (beam_cut_threshold >= max_beam_cut) throw py::value_error("beam_cut_threshold cannot be more than " + rust_f32(max_beam_cut));
    py::array_t<int64_t, py::array::c_style | py::array::forcecast> env_arr;
    const int64_t* env = nullptr;
    if (!envelope.is_none()) {
        env_arr = envelope.cast<py::array_t<int64_t, py::array::c_style | py::array::forcecast>>();
        if (env_arr.ndim() != 3 || env_arr.shape(0) != N || env_arr.shape(1) != T1)
            throw py::value_error("the lengths of network_output_1 and envelope do not match");
        if (env_arr.shape(2) != 2) throw py::value_error("the inner axis of envelope must have size 2");
        env = env_arr.data();
    }
    const int64_t stride = std::max<int64_t>((T1 + 31) & ~int64_t(31), 32);
    std::vector<uint8_t> labels(size_t(N * stride));
    std::vector<int64_t> offsets(static_cast<size_t>(N));
    std::vector<int32_t> lens(static_cast<size_t>(N)), status(static_cast<size_t>(N));
    int64_t used = 0;
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = fctc_beam_search_duplex_batch(handle(), network_outputs_1.data(), network_outputs_2.data(), N, T1, T2, int32_t(C), env,
                                           int32_t(std::min<int64_t>(beam_size, 1 << 30)), beam_cut_threshold, collapse_repeats,
                                           g_exp_mode, labels.data(), int64_t(labels.size()), offsets.data(), lens.data(),
                                           status.data(), &used);
    }
    if (rc < 0) raise_call_error(rc);
    py::list out;
    for (int64_t n = 0; n < N; ++n) {
        raise_status(status[size_t(n)]);
        out.append(py::str(join_labels(names, labels.data() + offsets[size_t(n)], lens[size_t(n)])));
    }
    return out;
}

#include "pymodule_crf.inc"

}  // namespace

PYBIND11_MODULE(fast_ctc_decode, m) {
    m.doc() = "Methods for labelling RNN results using CTC decoding — B200 (CUDA) build of the fast_ctc_decode API.";
    using namespace pybind11::literals;
    m.def("viterbi_search", &viterbi_search, "network_output"_a, "alphabet"_a, "qstring"_a = false, "qscale"_a = 1.0f,
          "qbias"_a = 0.0f, "collapse_repeats"_a = true);
    m.def("beam_search", &beam_search, "network_output"_a, "alphabet"_a, "beam_size"_a = 5, "beam_cut_threshold"_a = 0.0f,
          "collapse_repeats"_a = true);
    m.def("beam_search_duplex", &beam_search_duplex, "network_output_1"_a, "network_output_2"_a, "alphabet"_a,
          "envelope"_a = py::none(), "beam_size"_a = 5, "beam_cut_threshold"_a = 0.0f, "collapse_repeats"_a = true);
    register_crf(m);
    m.def("viterbi_search_batch", &viterbi_search_batch, "network_outputs"_a, "alphabet"_a, "qstring"_a = false,
          "qscale"_a = 1.0f, "qbias"_a = 0.0f, "collapse_repeats"_a = true);
    m.def("beam_search_batch", &beam_search_batch, "network_outputs"_a, "alphabet"_a, "beam_size"_a = 5,
          "beam_cut_threshold"_a = 0.0f, "collapse_repeats"_a = true);
    m.def("beam_search_duplex_batch", &beam_search_duplex_batch, "network_outputs_1"_a, "network_outputs_2"_a, "alphabet"_a,
          "envelope"_a = py::none(), "beam_size"_a = 5, "beam_cut_threshold"_a = 0.0f, "collapse_repeats"_a = true);
    m.def("set_exp_mode", [](const std::string& mode) {
        if (mode == "exact") g_exp_mode = 0;
        else if (mode == "fastexp") g_exp_mode = 1;
        else if (mode == "cuda") g_exp_mode = 2;
        else throw py::value_error("mode must be 'exact' (the --no-default-features build), 'fastexp' (the default build) or "
                                   "'cuda' (CUDA's expf/log1pf: faster, last-bit differences from a glibc host)");
    }, "mode"_a, "LogSpace '+' of the duplex searches: 'exact' exp (default here, bit-identical to a glibc host), the "
                 "reference's default 'fastexp' build, or 'cuda'.");
    m.attr("__version__") = "0.3.7";  // CARGO_PKG_VERSION, src/lib.rs:626
}
