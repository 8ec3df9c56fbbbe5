"""ctypes front end of the CPU oracle (oracle/ctc_oracle.cpp).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.

Every function returns alphabet *indices*; `to_text` applies the alphabet the way the reference
does (src/lib.rs:144-146 + src/search.rs:291,358,378-380).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libctc_oracle.so")

OK, RAN_OUT_OF_BEAM, INCOMPARABLE, INVALID_ENVELOPE, PANIC = 0, 1, 2, 3, 100

_lib = None


def build(force=False):
    """Compile the oracle with the committed Makefile (g++ -O2 -ffp-contract=off)."""
    src = os.path.join(_HERE, "ctc_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libctc_oracle.so"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        fp, i64, i32, u32, u64 = (C.POINTER(C.c_float), C.c_int64, C.c_int32, C.c_uint32, C.c_uint64)
        vp = C.c_void_p
        L.oracle_phred.restype = u32
        L.oracle_phred.argtypes = [C.c_float, C.c_float, C.c_float, C.POINTER(C.c_int)]
        L.oracle_log10f_n.restype = None
        L.oracle_log10f_n.argtypes = [vp, vp, i64]
        L.oracle_logf_n.restype = None
        L.oracle_logf_n.argtypes = [vp, vp, i64]
        for name in ("oracle_expf_n", "oracle_log1pf_n"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [vp, vp, i64]
        L.oracle_softmax_rows.restype = None
        L.oracle_softmax_rows.argtypes = [vp, vp, i64, i64]
        L.oracle_fastexp_as_shipped.restype = C.c_float
        L.oracle_fastexp_as_shipped.argtypes = [C.c_float]
        L.oracle_logadd.restype = C.c_float
        L.oracle_logadd.argtypes = [C.c_float, C.c_float, C.c_int]
        L.oracle_viterbi.argtypes = [vp, i64, i64, i64, i64, C.c_int, C.c_float, C.c_float, C.c_int, vp, vp, vp, i64, vp]
        L.oracle_beam_search.argtypes = [vp, i64, i64, i64, i64, i64, C.c_float, C.c_int, vp, vp, i64, vp, vp]
        L.oracle_crf_greedy.argtypes = [vp, i64, i64, i64, vp, i64, C.c_int, C.c_float, C.c_float, vp, vp, vp, i64, vp]
        L.oracle_crf_beam_search.argtypes = [vp, i64, i64, i64, vp, i64, i64, C.c_float, vp, vp, i64, vp]
        L.oracle_beam_search_duplex.argtypes = [vp, i64, vp, i64, i64, vp, i64, C.c_float, C.c_int, C.c_int, vp, i64, vp, vp]
        L.oracle_crf_beam_search_duplex.argtypes = [vp, i64, vp, i64, vp, i64, vp, i64, i64, i64, vp, i64, C.c_float,
                                                    C.c_int, vp, i64, vp]
        L.oracle_viterbi_batch.argtypes = [vp, i64, i64, i64, C.c_int, C.c_float, C.c_float, C.c_int, vp, vp, vp, i64,
                                           vp, vp, C.c_int]
        L.oracle_beam_search_batch.argtypes = [vp, i64, i64, i64, i64, C.c_float, C.c_int, vp, vp, i64, vp, vp, vp,
                                               C.c_int]
        L.oracle_beam_search_duplex_batch.argtypes = [vp, vp, i64, i64, i64, i64, vp, i64, C.c_float, C.c_int, C.c_int,
                                                      vp, i64, vp, vp, C.c_int]
        L.oracle_trie_new.restype = vp
        L.oracle_trie_new.argtypes = [i64]
        L.oracle_trie_free.argtypes = [vp]
        L.oracle_trie_add.restype = i32
        L.oracle_trie_add.argtypes = [vp, i32, i64, i64]
        L.oracle_trie_child.restype = i32
        L.oracle_trie_child.argtypes = [vp, i32, i64]
        L.oracle_trie_label.restype = i64
        L.oracle_trie_label.argtypes = [vp, i32]
        L.oracle_trie_data.argtypes = [vp, i32, vp]
        L.oracle_trie_ancestors.restype = i64
        L.oracle_trie_ancestors.argtypes = [vp, i32, vp, vp, i64]
        L.oracle_secondary_get_gap.restype = C.c_float
        L.oracle_secondary_get_gap.argtypes = [i64, vp, vp, i64, i64]
        L.oracle_secondary_update_max.restype = C.c_float
        L.oracle_secondary_update_max.argtypes = [i64, vp, vp, i64, i64, i64, C.c_int]
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    a = np.asarray(a)
    if a.dtype != np.float32:
        raise TypeError("posteriors must be float32")
    return a


def softmax(logits):
    """float32 softmax over the last axis by the recipe of the logits entry points (input.cuh), with the system libm."""
    z = np.ascontiguousarray(_f32(logits))
    out = np.empty_like(z)
    lib().oracle_softmax_rows(_ptr(z), _ptr(out), z.size // z.shape[-1], z.shape[-1])
    return out


def hardware_threads():
    return int(lib().oracle_hardware_threads())


def phred(prob, qscale=1.0, qbias=0.0):
    valid = C.c_int(0)
    code = lib().oracle_phred(float(prob), float(qscale), float(qbias), C.byref(valid))
    return int(code), bool(valid.value)


def viterbi(post, qstring=False, qscale=1.0, qbias=0.0, collapse_repeats=True):
    post = _f32(post)
    T, Cn = post.shape
    st, sc = (s // 4 for s in post.strides)
    cap = max(T, 1)
    labels = np.zeros(cap, np.int32)
    quals = np.zeros(cap, np.uint32)
    path = np.zeros(cap, np.int64)
    n = C.c_int64(0)
    status = lib().oracle_viterbi(_ptr(post), T, Cn, st, sc, int(qstring), qscale, qbias, int(collapse_repeats),
                                  _ptr(labels), _ptr(quals), _ptr(path), cap, C.byref(n))
    k = n.value
    return status, labels[:k].copy(), (quals[:k].copy() if qstring else None), path[:k].copy()


def beam_search(post, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True, return_nodes=False):
    post = _f32(post)
    T, Cn = post.shape
    st, sc = (s // 4 for s in post.strides)
    cap = max(T, 1)
    labels = np.zeros(cap, np.int32)
    path = np.zeros(cap, np.int64)
    n = C.c_int64(0)
    nodes = C.c_int64(0)
    status = lib().oracle_beam_search(_ptr(post), T, Cn, st, sc, beam_size, beam_cut_threshold, int(collapse_repeats),
                                      _ptr(labels), _ptr(path), cap, C.byref(n), C.byref(nodes))
    k = n.value
    out = (status, labels[:k].copy(), path[:k].copy())
    return out + (nodes.value,) if return_nodes else out


def crf_greedy(post, init_state, qstring=False, qscale=1.0, qbias=0.0):
    post = np.ascontiguousarray(_f32(post))
    init_state = np.ascontiguousarray(_f32(init_state))
    T, S, Cn = post.shape
    cap = max(T, 1)
    labels = np.zeros(cap, np.int32)
    quals = np.zeros(cap, np.uint32)
    path = np.zeros(cap, np.int64)
    n = C.c_int64(0)
    status = lib().oracle_crf_greedy(_ptr(post), T, S, Cn, _ptr(init_state), init_state.size, int(qstring), qscale,
                                     qbias, _ptr(labels), _ptr(quals), _ptr(path), cap, C.byref(n))
    k = n.value
    return status, labels[:k].copy(), (quals[:k].copy() if qstring else None), path[:k].copy()


def crf_beam_search(post, init_state, beam_size=5, beam_cut_threshold=0.0):
    post = np.ascontiguousarray(_f32(post))
    init_state = np.ascontiguousarray(_f32(init_state))
    T, S, Cn = post.shape
    cap = max(T, 1)
    labels = np.zeros(cap, np.int32)
    path = np.zeros(cap, np.int64)
    n = C.c_int64(0)
    status = lib().oracle_crf_beam_search(_ptr(post), T, S, Cn, _ptr(init_state), init_state.size, beam_size,
                                          beam_cut_threshold, _ptr(labels), _ptr(path), cap, C.byref(n))
    k = n.value
    return status, labels[:k].copy(), path[:k].copy()


def beam_search_duplex(post1, post2, envelope=None, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True,
                       exp_mode=0, return_cells=False):
    post1 = np.ascontiguousarray(_f32(post1))
    post2 = np.ascontiguousarray(_f32(post2))
    T1, Cn = post1.shape
    T2 = post2.shape[0]
    env = None if envelope is None else np.ascontiguousarray(envelope, dtype=np.uint64)
    cap = max(T1, 1)
    labels = np.zeros(cap, np.int32)
    n = C.c_int64(0)
    cells = C.c_int64(0)
    status = lib().oracle_beam_search_duplex(_ptr(post1), T1, _ptr(post2), T2, Cn, _ptr(env), beam_size,
                                             beam_cut_threshold, int(collapse_repeats), exp_mode, _ptr(labels), cap,
                                             C.byref(n), C.byref(cells))
    out = (status, labels[:n.value].copy())
    return out + (cells.value,) if return_cells else out


def crf_beam_search_duplex(post1, init1, post2, init2, envelope=None, beam_size=5, beam_cut_threshold=0.0, exp_mode=0):
    post1 = np.ascontiguousarray(_f32(post1))
    post2 = np.ascontiguousarray(_f32(post2))
    init1 = np.ascontiguousarray(_f32(init1))
    init2 = np.ascontiguousarray(_f32(init2))
    T1, S, Cn = post1.shape
    T2 = post2.shape[0]
    env = None if envelope is None else np.ascontiguousarray(envelope, dtype=np.uint64)
    cap = max(T1, 1)
    labels = np.zeros(cap, np.int32)
    n = C.c_int64(0)
    status = lib().oracle_crf_beam_search_duplex(_ptr(post1), T1, _ptr(init1), init1.size, _ptr(post2), T2,
                                                 _ptr(init2), init2.size, S, Cn, _ptr(env), beam_size,
                                                 beam_cut_threshold, exp_mode, _ptr(labels), cap, C.byref(n))
    return status, labels[:n.value].copy()


def viterbi_batch(post, qstring=False, qscale=1.0, qbias=0.0, collapse_repeats=True, n_threads=0, keep=True):
    """post: C-contiguous [N,T,C] float32.  Returns (labels[N,T], quals[N,T] | None, path[N,T], lens[N], status[N])
    or, with keep=False (timing), only (lens, status)."""
    post = _f32(post)
    assert post.flags.c_contiguous and post.ndim == 3
    N, T, Cn = post.shape
    lens = np.zeros(N, np.int64)
    status = np.zeros(N, np.int32)
    labels = np.zeros((N, T), np.int32) if keep else None
    quals = np.zeros((N, T), np.uint32) if (keep and qstring) else None
    path = np.zeros((N, T), np.int64) if keep else None
    lib().oracle_viterbi_batch(_ptr(post), N, T, Cn, int(qstring), qscale, qbias, int(collapse_repeats), _ptr(labels),
                               _ptr(quals), _ptr(path), T, _ptr(lens), _ptr(status), n_threads)
    return (labels, quals, path, lens, status) if keep else (lens, status)


def beam_search_batch(post, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True, n_threads=0, keep=True):
    post = _f32(post)
    assert post.flags.c_contiguous and post.ndim == 3
    N, T, Cn = post.shape
    lens = np.zeros(N, np.int64)
    status = np.zeros(N, np.int32)
    nodes = np.zeros(N, np.int64)
    labels = np.zeros((N, T), np.int32) if keep else None
    path = np.zeros((N, T), np.int64) if keep else None
    lib().oracle_beam_search_batch(_ptr(post), N, T, Cn, beam_size, beam_cut_threshold, int(collapse_repeats),
                                   _ptr(labels), _ptr(path), T, _ptr(lens), _ptr(status), _ptr(nodes), n_threads)
    return (labels, path, lens, status, nodes) if keep else (lens, status, nodes)


def beam_search_duplex_batch(post1, post2, envelope=None, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True,
                             exp_mode=0, n_threads=0):
    post1, post2 = _f32(post1), _f32(post2)
    assert post1.flags.c_contiguous and post2.flags.c_contiguous
    N, T1, Cn = post1.shape
    T2 = post2.shape[1]
    env = None if envelope is None else np.ascontiguousarray(envelope, dtype=np.uint64)
    lens = np.zeros(N, np.int64)
    status = np.zeros(N, np.int32)
    labels = np.zeros((N, T1), np.int32)
    lib().oracle_beam_search_duplex_batch(_ptr(post1), _ptr(post2), N, T1, T2, Cn, _ptr(env), beam_size,
                                          beam_cut_threshold, int(collapse_repeats), exp_mode, _ptr(labels), T1,
                                          _ptr(lens), _ptr(status), n_threads)
    return labels, lens, status


def to_text(alphabet, labels, quals=None):
    """alphabet: str (one label per character) or a sequence of str; labels: alphabet indices."""
    names = [str(a) for a in alphabet]
    s = "".join(names[int(i)] for i in labels)
    if quals is not None:
        s += "".join(chr(int(q)) for q in quals)
    return s
