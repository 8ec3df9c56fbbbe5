"""fctc_b200 — thin ctypes binding of the C ABI in include/fctc.h (libfctc_b200.so).

This module only marshals pointers; every decode runs in the CUDA library.  It raises if the
library is missing or no B200 is present — there is no CPU fallback.
Device-resident entry points take raw device pointers (e.g. ``tensor.data_ptr()``) so that any
allocator (torch, cuda-python, the caller's own) can own the memory.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FCTC_LIB") or os.path.join(os.path.dirname(_HERE), "libfctc_b200.so")

OK, RAN_OUT_OF_BEAM, INCOMPARABLE_VALUES, INVALID_ENVELOPE, NODE_POOL_OVERFLOW, OUTPUT_OVERFLOW, INVALID_QUALITY_CHAR, \
    REFERENCE_PANIC = range(8)
E_CUDA, E_ARGUMENT, E_NOMEM, E_UNSUPPORTED, E_CAPACITY = -1, -2, -3, -4, -5
IN_F32, IN_F16, IN_BF16, IN_LOGITS_EXACT, IN_LOGITS_FAST = 0, 1, 2, 0x10, 0x20


def to_bf16_bits(a):
    """float32 array -> bfloat16 bit patterns (uint16), round to nearest even; NaNs stay NaNs."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    nan = np.isnan(a)
    if nan.any():
        r = np.where(nan, np.uint16(0x7FC0), r)
    return r


def from_bf16_bits(b):
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def _elem_kind(arr, kind):
    want = {1: np.float16, 2: np.uint16}.get(kind & 0xF, np.float32)  # an unknown kind is the library's to refuse
    if not isinstance(arr, np.ndarray) or arr.dtype != want or arr.ndim != 3 or not arr.flags.c_contiguous:
        raise TypeError(f"input_kind {kind:#x} needs a C-contiguous {np.dtype(want).name} array of shape [N, T, C]")
    return arr

_lib = None

_i64, _i32, _f32, _vp, _u64 = C.c_int64, C.c_int32, C.c_float, C.c_void_p, C.c_uint64

_SIGNATURES = {
    "fctc_version": (C.c_int, []),
    "fctc_device_count": (C.c_int, []),
    "fctc_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "fctc_destroy": (None, [_vp]),
    "fctc_last_error": (C.c_char_p, [_vp]),
    "fctc_status_message": (C.c_char_p, [C.c_int]),
    "fctc_device_sm_count": (C.c_int, [_vp]),
    "fctc_viterbi_batch_device": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp,
                                            _i64, _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp, _i64,
                                     _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch_device_f16": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32,
                                                _vp, _i64, _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch_f16": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp, _i64,
                                         _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch_packed": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp, _i64,
                                            _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch_packed_f16": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp,
                                                _i64, _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch_device_in": (C.c_int, [_vp, _vp, _i32, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp,
                                               _i64, _vp, _vp, _vp, _vp]),
    "fctc_viterbi_batch_in": (C.c_int, [_vp, _vp, _i32, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp, _i64,
                                        _vp, _vp, _vp, _vp]),
    "fctc_beam_search_batch_device_in": (C.c_int, [_vp, _vp, _i32, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp,
                                                   _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_beam_search_batch_in": (C.c_int, [_vp, _vp, _i32, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp, _vp,
                                            _vp, _vp]),
    "fctc_viterbi_search": (C.c_int, [_vp, _vp, _i64, _i32, _i64, _i64, _i32, _f32, _f32, _i32, _vp, _vp, _vp, _i64,
                                      _vp]),
    "fctc_beam_scratch_bytes": (_i64, [_vp, _i64, _i64, _i32, _i32, _i32]),
    "fctc_beam_search_batch_device": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp,
                                                _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_beam_search_batch_device_static": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64,
                                                       _vp, _vp, _vp, _vp, _i32, _vp]),
    "fctc_beam_search_batch": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp, _vp, _vp,
                                         _vp]),
    "fctc_beam_search_batch_device_f16": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp,
                                                    _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_beam_search_batch_f16": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp, _vp,
                                             _vp, _vp]),
    "fctc_beam_search": (C.c_int, [_vp, _vp, _i64, _i32, _i64, _i64, _i32, _f32, _i32, _vp, _vp, _i64, _vp]),
    "fctc_viterbi_batch_multi": (C.c_int, [_vp, _i32, _vp, _i64, _i64, _i32, _i32, _f32, _f32, _i32, _vp, _vp, _i32, _vp, _i64,
                                           _vp, _vp, _vp, _vp]),
    "fctc_beam_search_batch_multi": (C.c_int, [_vp, _i32, _vp, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _i64, _vp, _vp,
                                               _vp, _vp]),
    "fctc_multi_last_error": (C.c_char_p, []),
    "fctc_multi_release": (None, []),
    "fctc_duplex_scratch_bytes": (_i64, [_vp, _i64, _i64, _i64, _i32, _i32, _i32, _i32]),
    "fctc_crf_duplex_scratch_bytes": (_i64, [_vp, _i64, _i64, _i64, _i32, _i32, _i32, _i32]),
    "fctc_beam_search_duplex_batch_device": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _i32, _i32, _f32,
                                                       _i32, _i32, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _i32,
                                                       _vp]),
    "fctc_beam_search_duplex_batch_device_f16": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _i32, _i32, _f32,
                                                           _i32, _i32, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _i32,
                                                           _vp]),
    "fctc_beam_search_duplex_batch": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _i32, _f32, _i32, _i32,
                                                _vp, _i64, _vp, _vp, _vp, _vp]),
    "fctc_beam_search_duplex_batch_f16": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _i32, _f32, _i32, _i32,
                                                    _vp, _i64, _vp, _vp, _vp, _vp]),
    "fctc_beam_search_duplex": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _i64, _i64, _i64, _i32, _vp, _i64, _i64,
                                          _i32, _f32, _i32, _i32, _vp, _i64, _vp]),
    "fctc_crf_greedy_batch_device": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _f32, _f32, _vp, _vp,
                                               _i32, _vp, _i64, _vp, _vp, _vp, _vp]),
    "fctc_crf_beam_search_batch_device": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _vp,
                                                    _i64, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_crf_greedy_batch_device_in": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _f32, _f32, _vp,
                                                  _vp, _i32, _vp, _i64, _vp, _vp, _vp, _vp]),
    "fctc_crf_beam_search_batch_device_in": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _f32, _vp,
                                                       _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_crf_beam_search_duplex_batch_device_in": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _i64, _i64, _i64,
                                                              _i32, _i32, _vp, _i32, _i32, _f32, _i32, _vp, _i64, _vp, _vp,
                                                              _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_crf_greedy_search": (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _i32, _i32, _f32, _f32, _vp, _vp, _vp, _i64,
                                         _vp]),
    "fctc_crf_beam_search": (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _i32, _i32, _f32, _vp, _vp, _i64, _vp]),
    "fctc_crf_beam_search_duplex_batch_device": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp, _i32, _i64, _i64, _i64, _i32,
                                                           _i32, _vp, _i32, _i32, _f32, _i32, _vp, _i64, _vp, _vp, _vp,
                                                           _vp, _vp, _i64, _i32, _vp]),
    "fctc_crf_beam_search_duplex": (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _i64, _vp, _i32, _i32, _i32, _vp, _i32,
                                              _f32, _i32, _vp, _i64, _vp]),
    "fctc_fastq_scratch_bytes": (_i64, [_vp, _i64]),
    "fctc_fastq_records_device": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, C.c_char_p, _i32, C.c_char_p, _i64, _vp,
                                            _i64, _vp, _vp, _i64, _vp]),
    "fctc_duplex_envelope_device": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp]),
    "fctc_align_scratch_bytes": (_i64, [_vp, _i64, _i64, _i32]),
    "fctc_duplex_envelope_aligned_device": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32,
                                                      _i32, _vp, _vp, _vp, _vp, _i64, _vp]),
    "fctc_move_table_device": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp, _vp]),
    "fctc_synth_posteriors_device": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i32, _u64, _i64, _vp]),
    "fctc_test_poison_shared_memory": (C.c_int, [_vp, _i32]),
    "fctc_test_libm_device": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "fctc_host_phred": (C.c_uint32, [_f32, _f32, _f32]),
    "fctc_host_log10f": (None, [_vp, _vp, _i64]),
    "fctc_host_logf": (None, [_vp, _vp, _i64]),
    "fctc_host_expf": (None, [_vp, _vp, _i64]),
    "fctc_host_log1pf": (None, [_vp, _vp, _i64]),
    "fctc_host_expf_nonpos": (None, [_vp, _vp, _i64]),
    "fctc_host_log1pf_unit": (None, [_vp, _vp, _i64]),
}


class FctcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libfctc_b200 error {code}: {message}")
        self.code = code


def lib():
    """Load libfctc_b200.so (built by build.py).  Raises if it is missing: no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built; run `python fast-ctc-decode_b200/build.py`")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def declared_symbols():
    return sorted(_SIGNATURES)


def _np_ptr(a):
    return None if a is None else a.ctypes.data


def round_up32(v):
    return (int(v) + 31) & ~31


class Decoder:
    """One handle = one device + its streams and staging buffers (include/fctc.h: fctc_create)."""

    def __init__(self, device=0):
        self._L = lib()
        h = _vp()
        rc = self._L.fctc_create(int(device), C.byref(h))
        if rc != 0:
            raise FctcError(rc, (self._L.fctc_last_error(None) or b"").decode())
        self._h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._L.fctc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc < 0:
            raise FctcError(rc, (self._L.fctc_last_error(self._h) or b"").decode())
        return rc

    @property
    def sm_count(self):
        return int(self._L.fctc_device_sm_count(self._h))

    # ---- device-resident API (raw pointers) ---------------------------------------------------
    def synth_posteriors_device(self, post_ptr, n_reads, n_rows, n_labels, kind=0, seed=0, first_read=0, stream=0):
        self._check(self._L.fctc_synth_posteriors_device(self._h, post_ptr, n_reads, n_rows, n_labels, kind, seed,
                                                         first_read, stream or None))

    def viterbi_batch_device(self, post_ptr, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats,
                             labels_ptr, quals_ptr, qual_bytes, path_ptr, pool_capacity, offsets_ptr, lens_ptr,
                             status_ptr, stream=0, f16=False):
        """f16=True: post_ptr addresses binary16 posteriors (fctc_viterbi_batch_device_f16)."""
        fn = self._L.fctc_viterbi_batch_device_f16 if f16 else self._L.fctc_viterbi_batch_device
        self._check(fn(self._h, post_ptr, n_reads, n_rows, n_labels, int(qstring), qscale, qbias, int(collapse_repeats),
                       labels_ptr, quals_ptr or None, qual_bytes, path_ptr, pool_capacity, offsets_ptr, lens_ptr,
                       status_ptr, stream or None))

    def duplex_envelope_device(self, path1_ptr, offsets1_ptr, lens1_ptr, path2_ptr, offsets2_ptr, lens2_ptr, n_pairs,
                               n_rows_1, n_rows_2, pad_rows, envelope_ptr, max_window_ptr=0, stream=0):
        """int64 [n_pairs, n_rows_1, 2] envelope from the two reads' slot-form Viterbi paths."""
        self._check(self._L.fctc_duplex_envelope_device(self._h, path1_ptr, offsets1_ptr, lens1_ptr, path2_ptr,
                                                        offsets2_ptr, lens2_ptr, n_pairs, n_rows_1, n_rows_2, pad_rows,
                                                        envelope_ptr, max_window_ptr or None, stream or None))

    def align_scratch_bytes(self, n_pairs, max_bases, band_bases):
        return int(self._L.fctc_align_scratch_bytes(self._h, n_pairs, max_bases, band_bases))

    def duplex_envelope_aligned_device(self, labels1_ptr, path1_ptr, offsets1_ptr, lens1_ptr, labels2_ptr, path2_ptr,
                                       offsets2_ptr, lens2_ptr, n_pairs, n_rows_1, n_rows_2, band_bases, pad_rows,
                                       envelope_ptr, max_window_ptr, partner_ptr, scratch_ptr, scratch_bytes, stream=0):
        """Envelope from a banded alignment of the two reads' Viterbi label strings (fctc_duplex_envelope_aligned_device)."""
        self._check(self._L.fctc_duplex_envelope_aligned_device(self._h, labels1_ptr, path1_ptr, offsets1_ptr, lens1_ptr,
                                                                labels2_ptr, path2_ptr, offsets2_ptr, lens2_ptr, n_pairs,
                                                                n_rows_1, n_rows_2, band_bases, pad_rows, envelope_ptr,
                                                                max_window_ptr or None, partner_ptr, scratch_ptr,
                                                                scratch_bytes, stream or None))

    def move_table_device(self, path_ptr, offsets_ptr, lens_ptr, status_ptr, n_reads, n_rows, moves_ptr=0, dwell_ptr=0,
                          stream=0):
        """uint8 move table [n_reads, n_rows] and / or int32 dwells (path layout) from the emission rows."""
        self._check(self._L.fctc_move_table_device(self._h, path_ptr, offsets_ptr, lens_ptr, status_ptr, n_reads, n_rows,
                                                   moves_ptr or None, dwell_ptr or None, stream or None))

    def libm_device(self, x_ptr, y_ptr, out_ptr, n, which, stream=0):
        """Test hook: the device builds of the libm restatements over an array (fctc_test_libm_device)."""
        self._check(self._L.fctc_test_libm_device(self._h, x_ptr, y_ptr or None, out_ptr, n, which, stream or None))

    def poison_shared_memory(self, byte=0xFF):
        """Test hook: every SM's shared memory is filled with `byte` before the next kernel starts."""
        self._check(self._L.fctc_test_poison_shared_memory(self._h, byte))

    def fastq_scratch_bytes(self, n_reads):
        return int(self._L.fctc_fastq_scratch_bytes(self._h, n_reads))

    def fastq_records_device(self, labels_ptr, quals_ptr, offsets_ptr, lens_ptr, status_ptr, n_reads, alphabet,
                             name_prefix, first_index, out_ptr, out_capacity, record_offsets_ptr, scratch_ptr,
                             scratch_bytes, stream=0):
        """FASTQ text from the slot-form Viterbi result (qstring, qual_bytes=1), assembled on the device."""
        alpha = alphabet.encode("ascii") if isinstance(alphabet, str) else bytes(alphabet)
        self._check(self._L.fctc_fastq_records_device(self._h, labels_ptr, quals_ptr, offsets_ptr, lens_ptr, status_ptr,
                                                      n_reads, alpha, len(alpha), name_prefix.encode("ascii"), first_index,
                                                      out_ptr, out_capacity, record_offsets_ptr, scratch_ptr,
                                                      scratch_bytes, stream or None))

    def beam_scratch_bytes(self, n_reads, n_rows, n_labels, beam_size, nodes_per_row=0):
        return int(self._L.fctc_beam_scratch_bytes(self._h, n_reads, n_rows, n_labels, beam_size, nodes_per_row))

    def beam_search_batch_device(self, post_ptr, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                 collapse_repeats, labels_ptr, path_ptr, pool_capacity, offsets_ptr, lens_ptr,
                                 status_ptr, pool_used_ptr, scratch_ptr, scratch_bytes, nodes_per_row=0, stream=0,
                                 f16=False):
        """f16=True: post_ptr addresses binary16 posteriors (fctc_beam_search_batch_device_f16)."""
        fn = self._L.fctc_beam_search_batch_device_f16 if f16 else self._L.fctc_beam_search_batch_device
        self._check(fn(self._h, post_ptr, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold, int(collapse_repeats),
                       labels_ptr, path_ptr, pool_capacity, offsets_ptr, lens_ptr, status_ptr, pool_used_ptr, scratch_ptr,
                       scratch_bytes, nodes_per_row, stream or None))

    def duplex_scratch_bytes(self, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size, max_window=0, nodes_per_row=0):
        return int(self._L.fctc_duplex_scratch_bytes(self._h, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size,
                                                     max_window, nodes_per_row))

    def crf_duplex_scratch_bytes(self, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size, max_window=0, nodes_per_row=0):
        return int(self._L.fctc_crf_duplex_scratch_bytes(self._h, n_pairs, n_rows_1, n_rows_2, n_labels, beam_size,
                                                         max_window, nodes_per_row))

    def last_error(self):
        return (self._L.fctc_last_error(self._h) or b"").decode()

    def beam_search_duplex_batch_device(self, post1_ptr, post2_ptr, n_pairs, n_rows_1, n_rows_2, n_labels, envelope_ptr,
                                        max_window, beam_size, beam_cut_threshold, collapse_repeats, exp_mode,
                                        labels_ptr, pool_capacity, offsets_ptr, lens_ptr, status_ptr, pool_used_ptr,
                                        scratch_ptr, scratch_bytes, nodes_per_row=0, stream=0, f16=False):
        fn = self._L.fctc_beam_search_duplex_batch_device_f16 if f16 else self._L.fctc_beam_search_duplex_batch_device
        self._check(fn(self._h, post1_ptr, post2_ptr, n_pairs, n_rows_1,
                                                                 n_rows_2, n_labels, envelope_ptr or None, max_window,
                                                                 beam_size, beam_cut_threshold, int(collapse_repeats),
                                                                 exp_mode, labels_ptr, pool_capacity, offsets_ptr,
                                                                 lens_ptr, status_ptr, pool_used_ptr, scratch_ptr,
                                                                 scratch_bytes, nodes_per_row, stream or None))

    def beam_search_batch_device_static(self, post_ptr, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                        collapse_repeats, labels_ptr, path_ptr, pool_capacity, offsets_ptr, lens_ptr,
                                        status_ptr, pool_used_ptr, nodes_per_row=0, stream=0):
        """Beam search with the library's built-in scratch array: for callers that must not allocate
        device memory.  Limited to FCTC_BEAM_STATIC_SCRATCH_BYTES (64 MiB) of trie pools; calls on one
        device share that array and must not overlap on the GPU."""
        self._check(self._L.fctc_beam_search_batch_device_static(self._h, post_ptr, n_reads, n_rows, n_labels,
                                                                 beam_size, beam_cut_threshold, int(collapse_repeats),
                                                                 labels_ptr, path_ptr, pool_capacity, offsets_ptr,
                                                                 lens_ptr, status_ptr, pool_used_ptr, nodes_per_row,
                                                                 stream or None))

    # ---- host-memory batch API ------------------------------------------------------------------
    def viterbi_batch(self, post, qstring=False, qscale=1.0, qbias=0.0, collapse_repeats=True, qual_bytes=1,
                      out=None, packed=False):
        """post: C-contiguous float32 — or float16, decoded as the exactly widened float32 values —
        [N, T, C] (numpy; pinned memory gives full copy overlap).
        packed=True: fctc_viterbi_batch_packed — results squeezed together on the device before the read-back
        (offsets are then the running sum of round_up(len, 32), not n * round_up(T, 32)).
        Returns dict(labels, quals, path, offsets, lens, status, used, d2h_bytes)."""
        f16 = getattr(post, "dtype", None) == np.float16
        post = np.ascontiguousarray(post) if f16 else _as_f32_3d(post)
        if post.ndim != 3:
            raise ValueError("posteriors must be [N, T, C]")
        N, T, Cn = post.shape
        cap = N * round_up32(T)
        o = out or {}
        labels = o.get("labels", None)
        labels = np.empty(cap, np.uint8) if labels is None else labels
        quals = o.get("quals", None)
        if qstring and quals is None:
            quals = np.empty(cap, np.uint8 if qual_bytes == 1 else np.uint32)
        path = o.get("path", None)
        path = np.empty(cap, np.int32) if path is None else path
        offsets = o.get("offsets", None)
        offsets = np.empty(N, np.int64) if offsets is None else offsets
        lens = o.get("lens", None)
        lens = np.empty(N, np.int32) if lens is None else lens
        status = o.get("status", None)
        status = np.empty(N, np.int32) if status is None else status
        used = _i64(0)
        if packed:
            fn = self._L.fctc_viterbi_batch_packed_f16 if f16 else self._L.fctc_viterbi_batch_packed
        else:
            fn = self._L.fctc_viterbi_batch_f16 if f16 else self._L.fctc_viterbi_batch
        self._check(fn(self._h, post.ctypes.data, N, T, Cn, int(qstring), qscale, qbias, int(collapse_repeats),
                       labels.ctypes.data, _np_ptr(quals) if qstring else None, qual_bytes, path.ctypes.data,
                       labels.size, offsets.ctypes.data, lens.ctypes.data, status.ctypes.data, C.byref(used)))
        # bytes the call moved device -> host: result entries (label + row + quality) and 16 B per read
        d2h = used.value * (1 + 4 + (qual_bytes if qstring else 0)) + N * 16
        return dict(labels=labels, quals=quals if qstring else None, path=path, offsets=offsets, lens=lens,
                    status=status, used=used.value, d2h_bytes=d2h)

    def beam_search_batch(self, post, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True, out=None):
        """post: C-contiguous float32 — or float16, decoded as the exactly widened float32 values — [N, T, C].
        `out` may carry preallocated (ideally pinned) labels / path / offsets / lens / status arrays."""
        f16 = getattr(post, "dtype", None) == np.float16
        post = np.ascontiguousarray(post) if f16 else _as_f32_3d(post)
        if post.ndim != 3:
            raise ValueError("posteriors must be [N, T, C]")
        N, T, Cn = post.shape
        cap = N * round_up32(T)
        o = out or {}

        def take(name, n, dtype):
            a = o.get(name, None)
            return np.empty(n, dtype) if a is None else a
        labels, path = take("labels", cap, np.uint8), take("path", cap, np.int32)
        offsets, lens, status = take("offsets", N, np.int64), take("lens", N, np.int32), take("status", N, np.int32)
        used = _i64(0)
        fn = self._L.fctc_beam_search_batch_f16 if f16 else self._L.fctc_beam_search_batch
        self._check(fn(self._h, post.ctypes.data, N, T, Cn, beam_size, beam_cut_threshold, int(collapse_repeats),
                       labels.ctypes.data, path.ctypes.data, labels.size, offsets.ctypes.data, lens.ctypes.data,
                       status.ctypes.data, C.byref(used)))
        return dict(labels=labels, path=path, offsets=offsets, lens=lens, status=status, used=used.value)

    def viterbi_batch_in(self, arr, input_kind, qstring=False, qscale=1.0, qbias=0.0, collapse_repeats=True, qual_bytes=1):
        """fctc_viterbi_batch_in: [N, T, C] float32 / float16 / bfloat16 bits (uint16), posteriors or logits
        (IN_* | IN_LOGITS_*); packed results as viterbi_batch(packed=True)."""
        arr = _elem_kind(arr, input_kind)
        N, T, Cn = arr.shape
        cap = N * round_up32(T)
        labels, path = np.empty(cap, np.uint8), np.empty(cap, np.int32)
        quals = np.empty(cap, np.uint8 if qual_bytes == 1 else np.uint32) if qstring else None
        offsets, lens, status = np.empty(N, np.int64), np.empty(N, np.int32), np.empty(N, np.int32)
        used = _i64(0)
        self._check(self._L.fctc_viterbi_batch_in(self._h, arr.ctypes.data, input_kind, N, T, Cn, int(qstring), qscale, qbias,
                                                  int(collapse_repeats), labels.ctypes.data, _np_ptr(quals), qual_bytes,
                                                  path.ctypes.data, cap, offsets.ctypes.data, lens.ctypes.data,
                                                  status.ctypes.data, C.byref(used)))
        return dict(labels=labels, quals=quals, path=path, offsets=offsets, lens=lens, status=status, used=used.value)

    def beam_search_batch_in(self, arr, input_kind, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True):
        """fctc_beam_search_batch_in: as beam_search_batch for any input kind."""
        arr = _elem_kind(arr, input_kind)
        N, T, Cn = arr.shape
        cap = N * max(round_up32(T), 32)
        labels, path = np.empty(cap, np.uint8), np.empty(cap, np.int32)
        offsets, lens, status = np.empty(N, np.int64), np.empty(N, np.int32), np.empty(N, np.int32)
        used = _i64(0)
        self._check(self._L.fctc_beam_search_batch_in(self._h, arr.ctypes.data, input_kind, N, T, Cn, beam_size,
                                                      beam_cut_threshold, int(collapse_repeats), labels.ctypes.data,
                                                      path.ctypes.data, cap, offsets.ctypes.data, lens.ctypes.data,
                                                      status.ctypes.data, C.byref(used)))
        return dict(labels=labels, path=path, offsets=offsets, lens=lens, status=status, used=used.value)

    def viterbi_batch_device_in(self, in_ptr, input_kind, n_reads, n_rows, n_labels, qstring, qscale, qbias, collapse_repeats,
                                labels_ptr, quals_ptr, qual_bytes, path_ptr, pool_capacity, offsets_ptr, lens_ptr,
                                status_ptr, stream=0):
        self._check(self._L.fctc_viterbi_batch_device_in(self._h, in_ptr, input_kind, n_reads, n_rows, n_labels, int(qstring),
                                                         qscale, qbias, int(collapse_repeats), labels_ptr, quals_ptr or None,
                                                         qual_bytes, path_ptr, pool_capacity, offsets_ptr, lens_ptr,
                                                         status_ptr, stream or None))

    def beam_search_batch_device_in(self, in_ptr, input_kind, n_reads, n_rows, n_labels, beam_size, beam_cut_threshold,
                                    collapse_repeats, labels_ptr, path_ptr, pool_capacity, offsets_ptr, lens_ptr, status_ptr,
                                    pool_used_ptr, scratch_ptr, scratch_bytes, nodes_per_row=0, stream=0):
        self._check(self._L.fctc_beam_search_batch_device_in(self._h, in_ptr, input_kind, n_reads, n_rows, n_labels, beam_size,
                                                             beam_cut_threshold, int(collapse_repeats), labels_ptr, path_ptr,
                                                             pool_capacity, offsets_ptr, lens_ptr, status_ptr, pool_used_ptr,
                                                             scratch_ptr, scratch_bytes, nodes_per_row, stream or None))

    def beam_search_duplex_batch(self, post1, post2, envelope=None, beam_size=5, beam_cut_threshold=0.0,
                                 collapse_repeats=True, exp_mode=0):
        """post1 [N,T1,C], post2 [N,T2,C] float32 (or both float16) C-contiguous; envelope int64 [N,T1,2] or None."""
        f16 = getattr(post1, "dtype", None) == np.float16 and getattr(post2, "dtype", None) == np.float16
        if f16:
            post1, post2 = np.ascontiguousarray(post1), np.ascontiguousarray(post2)
        else:
            post1, post2 = _as_f32_3d(post1), _as_f32_3d(post2)
        N, T1, Cn = post1.shape
        T2 = post2.shape[1]
        if post2.shape[0] != N or post2.shape[2] != Cn:
            raise ValueError("inner axes of the network outputs do not match")
        env = None
        if envelope is not None:
            env = np.ascontiguousarray(envelope, dtype=np.int64)
            if env.shape != (N, T1, 2):
                raise ValueError("envelope must have shape [N, T1, 2]")
        cap = N * max(round_up32(T1), 32)
        labels = np.empty(cap, np.uint8)
        offsets = np.empty(N, np.int64)
        lens = np.empty(N, np.int32)
        status = np.empty(N, np.int32)
        used = _i64(0)
        fn = self._L.fctc_beam_search_duplex_batch_f16 if f16 else self._L.fctc_beam_search_duplex_batch
        self._check(fn(self._h, post1.ctypes.data, post2.ctypes.data, N, T1, T2, Cn,
                                                          _np_ptr(env), beam_size, beam_cut_threshold,
                                                          int(collapse_repeats), exp_mode, labels.ctypes.data, cap,
                                                          offsets.ctypes.data, lens.ctypes.data, status.ctypes.data,
                                                          C.byref(used)))
        return dict(labels=labels, offsets=offsets, lens=lens, status=status, used=used.value)

    def beam_search_duplex(self, post1, post2, envelope=None, beam_size=5, beam_cut_threshold=0.0,
                           collapse_repeats=True, exp_mode=0):
        post1, post2 = _as_f32_2d(post1), _as_f32_2d(post2)
        T1, Cn = post1.shape
        T2 = post2.shape[0]
        if post2.shape[1] != Cn:
            raise ValueError("inner axes of the network outputs do not match")
        s1r, s1c = (s // 4 for s in post1.strides)
        s2r, s2c = (s // 4 for s in post2.strides)
        env, er, ec = None, 0, 0
        if envelope is not None:
            env = np.asarray(envelope)
            if env.dtype != np.uint64:
                raise TypeError("envelope must be a uint64 (usize) array")
            if env.shape[0] != T1:
                raise ValueError("the lengths of network_output_1 and envelope do not match")
            if env.ndim != 2 or env.shape[1] != 2:
                raise ValueError("the inner axis of envelope must have size 2")
            er, ec = (s // 8 for s in env.strides)
        cap = max(T1, 1)
        labels = np.empty(cap, np.uint8)
        n = _i64(0)
        status = self._check(self._L.fctc_beam_search_duplex(self._h, post1.ctypes.data, T1, s1r, s1c,
                                                             post2.ctypes.data, T2, s2r, s2c, Cn, _np_ptr(env), er, ec,
                                                             beam_size, beam_cut_threshold, int(collapse_repeats),
                                                             exp_mode, labels.ctypes.data, cap, C.byref(n)))
        return status, labels[:n.value].copy()

    # ---- single-read API (shape of the pyo3 functions; any strides) --------------------------------
    def viterbi_search(self, post, qstring=False, qscale=1.0, qbias=0.0, collapse_repeats=True):
        post = _as_f32_2d(post)
        T, Cn = post.shape
        st, sc = (s // 4 for s in post.strides)
        cap = max(T, 1)
        labels = np.empty(cap, np.uint8)
        quals = np.empty(cap, np.uint32)
        path = np.empty(cap, np.int64)
        n = _i64(0)
        status = self._check(self._L.fctc_viterbi_search(self._h, post.ctypes.data, T, Cn, st, sc, int(qstring),
                                                         qscale, qbias, int(collapse_repeats), labels.ctypes.data,
                                                         quals.ctypes.data, path.ctypes.data, cap, C.byref(n)))
        k = n.value
        return status, labels[:k].copy(), (quals[:k].copy() if qstring else None), path[:k].copy()

    def crf_greedy_search(self, post, init_state, qstring=False, qscale=1.0, qbias=0.0):
        post = np.ascontiguousarray(post)
        init_state = np.ascontiguousarray(init_state)
        if post.dtype != np.float32 or post.ndim != 3 or init_state.dtype != np.float32 or init_state.ndim != 1:
            raise TypeError("expected float32 arrays [T, S, C] and [n_init]")
        T, S, Cn = post.shape
        cap = max(T, 1)
        labels = np.empty(cap, np.uint8)
        quals = np.empty(cap, np.uint32)
        path = np.empty(cap, np.int64)
        n = _i64(0)
        status = self._check(self._L.fctc_crf_greedy_search(self._h, post.ctypes.data, T, S, Cn, init_state.ctypes.data,
                                                            init_state.size, int(qstring), qscale, qbias,
                                                            labels.ctypes.data, quals.ctypes.data, path.ctypes.data, cap,
                                                            C.byref(n)))
        k = n.value
        return status, labels[:k].copy(), (quals[:k].copy() if qstring else None), path[:k].copy()

    def crf_beam_search(self, post, init_state, beam_size=5, beam_cut_threshold=0.0):
        post = np.ascontiguousarray(post)
        init_state = np.ascontiguousarray(init_state)
        if post.dtype != np.float32 or post.ndim != 3 or init_state.dtype != np.float32 or init_state.ndim != 1:
            raise TypeError("expected float32 arrays [T, S, C] and [n_init]")
        T, S, Cn = post.shape
        cap = max(T, 1)
        labels = np.empty(cap, np.uint8)
        path = np.empty(cap, np.int64)
        n = _i64(0)
        status = self._check(self._L.fctc_crf_beam_search(self._h, post.ctypes.data, T, S, Cn, init_state.ctypes.data,
                                                          init_state.size, beam_size, beam_cut_threshold,
                                                          labels.ctypes.data, path.ctypes.data, cap, C.byref(n)))
        k = n.value
        return status, labels[:k].copy(), path[:k].copy()

    def crf_beam_search_duplex(self, post1, init1, post2, init2, envelope=None, beam_size=5, beam_cut_threshold=0.0,
                               exp_mode=0):
        post1, post2 = np.ascontiguousarray(post1), np.ascontiguousarray(post2)
        init1, init2 = np.ascontiguousarray(init1), np.ascontiguousarray(init2)
        for a, nd in ((post1, 3), (post2, 3), (init1, 1), (init2, 1)):
            if a.dtype != np.float32 or a.ndim != nd:
                raise TypeError("expected float32 arrays [T, S, C] and [n_init]")
        T1, S, Cn = post1.shape
        if post2.shape[2] != Cn or post2.shape[1] != S:
            raise ValueError("inner axes of the network outputs do not match")
        env = None
        if envelope is not None:
            env = np.ascontiguousarray(envelope, dtype=np.uint64)
            if env.shape[0] != T1:
                raise ValueError("the lengths of network_output_1 and envelope do not match")
            if env.ndim != 2 or env.shape[1] != 2:
                raise ValueError("the inner axis of envelope must have size 2")
        cap = max(T1, 1)
        labels = np.empty(cap, np.uint8)
        n = _i64(0)
        status = self._check(self._L.fctc_crf_beam_search_duplex(self._h, post1.ctypes.data, T1, init1.ctypes.data,
                                                                 init1.size, post2.ctypes.data, post2.shape[0],
                                                                 init2.ctypes.data, init2.size, S, Cn, _np_ptr(env),
                                                                 beam_size, beam_cut_threshold, exp_mode,
                                                                 labels.ctypes.data, cap, C.byref(n)))
        return status, labels[:n.value].copy()

    def beam_search(self, post, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True):
        post = _as_f32_2d(post)
        T, Cn = post.shape
        st, sc = (s // 4 for s in post.strides)
        cap = max(T, 1)
        labels = np.empty(cap, np.uint8)
        path = np.empty(cap, np.int64)
        n = _i64(0)
        status = self._check(self._L.fctc_beam_search(self._h, post.ctypes.data, T, Cn, st, sc, beam_size,
                                                      beam_cut_threshold, int(collapse_repeats), labels.ctypes.data,
                                                      path.ctypes.data, cap, C.byref(n)))
        k = n.value
        return status, labels[:k].copy(), path[:k].copy()


def _multi_check(rc):
    if rc < 0:
        raise FctcError(rc, (lib().fctc_multi_last_error() or b"").decode())


def viterbi_batch_multi(devices, post, qstring=False, qscale=1.0, qbias=0.0, collapse_repeats=True, qual_bytes=1):
    """fctc_viterbi_batch_multi: reads [g*N/G, (g+1)*N/G) of post [N, T, C] on devices[g], one host thread each;
    results in read order (slot form).  Returns dict(labels, quals, path, offsets, lens, status, used)."""
    post = _as_f32_3d(post)
    N, T, Cn = post.shape
    cap = N * round_up32(T)
    dev = np.ascontiguousarray(devices, dtype=np.int32)
    labels, path = np.empty(cap, np.uint8), np.empty(cap, np.int32)
    quals = np.empty(cap, np.uint8 if qual_bytes == 1 else np.uint32) if qstring else None
    offsets, lens, status = np.empty(N, np.int64), np.empty(N, np.int32), np.empty(N, np.int32)
    used = _i64(0)
    _multi_check(lib().fctc_viterbi_batch_multi(dev.ctypes.data, dev.size, post.ctypes.data, N, T, Cn, int(qstring), qscale,
                                                qbias, int(collapse_repeats), labels.ctypes.data, _np_ptr(quals), qual_bytes,
                                                path.ctypes.data, cap, offsets.ctypes.data, lens.ctypes.data,
                                                status.ctypes.data, C.byref(used)))
    return dict(labels=labels, quals=quals, path=path, offsets=offsets, lens=lens, status=status, used=used.value)


def beam_search_batch_multi(devices, post, beam_size=5, beam_cut_threshold=0.0, collapse_repeats=True):
    """fctc_beam_search_batch_multi: as viterbi_batch_multi for the beam search (packed pool per device range)."""
    post = _as_f32_3d(post)
    N, T, Cn = post.shape
    cap = N * max(round_up32(T), 32)
    dev = np.ascontiguousarray(devices, dtype=np.int32)
    labels, path = np.empty(cap, np.uint8), np.empty(cap, np.int32)
    offsets, lens, status = np.empty(N, np.int64), np.empty(N, np.int32), np.empty(N, np.int32)
    used = _i64(0)
    _multi_check(lib().fctc_beam_search_batch_multi(dev.ctypes.data, dev.size, post.ctypes.data, N, T, Cn, beam_size,
                                                    beam_cut_threshold, int(collapse_repeats), labels.ctypes.data,
                                                    path.ctypes.data, cap, offsets.ctypes.data, lens.ctypes.data,
                                                    status.ctypes.data, C.byref(used)))
    return dict(labels=labels, path=path, offsets=offsets, lens=lens, status=status, used=used.value)


def _as_f32_3d(a):
    if not isinstance(a, np.ndarray) or a.dtype != np.float32 or a.ndim != 3 or not a.flags.c_contiguous:
        raise TypeError("expected a C-contiguous float32 numpy array of shape [N, T, C]")
    return a


def _as_f32_2d(a):
    if not isinstance(a, np.ndarray) or a.dtype != np.float32 or a.ndim != 2:
        raise TypeError("expected a float32 numpy array of shape [T, C]")
    return a
