"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/fctc.h declares,
its host build of the phred/log10f code matches glibc bit for bit, and it refuses to run without a
GPU (no CPU fallback).  No compute call is made on a device here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fctc():
    import fctc_b200
    if not os.path.exists(fctc_b200.LIB_PATH):
        import subprocess
        import sys
        subprocess.check_call([sys.executable, os.path.join(ROOT, "fast-ctc-decode_b200", "build.py")])
    return fctc_b200


def test_exports_every_declared_symbol(fctc):
    header = open(os.path.join(ROOT, "include", "fctc.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(fctc_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 15
    L = C.CDLL(fctc.LIB_PATH)
    missing = [name for name in sorted(declared) if not hasattr(L, name)]
    assert not missing, f"libfctc_b200.so lacks {missing}"
    # the python binding declares exactly the header's symbols
    assert set(fctc.declared_symbols()) == declared


def test_no_cpu_fallback_without_gpu(fctc):
    L = fctc.lib()
    if L.fctc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(fctc.FctcError) as e:
        fctc.Decoder(0)
    assert "no usable CUDA device" in str(e.value)


def test_status_messages_match_reference(fctc):
    """src/lib.rs:43-56"""
    L = fctc.lib()
    assert L.fctc_status_message(1) == b"Ran out of search space (beam_cut_threshold too high)"
    assert L.fctc_status_message(2) == b"Failed to compare values (NaNs in input?)"
    assert L.fctc_status_message(3) == b"Invalid envelope values"
    assert L.fctc_status_message(0) is None


def test_log10f_replica_matches_glibc(fctc, oracle):
    """The device quality path re-evaluates glibc's log10f (phred.cuh).  Its host build must agree
    with the libm the oracle (and Rust's f32::log10) calls, bit for bit: strided sweep over every
    binade of the positive normal range plus a dense sweep of phred's domain [1e-4, 1]."""
    L, O = fctc.lib(), oracle.lib()
    bits = np.concatenate([
        np.arange(0x00800000, 0x7F800000, 4099, dtype=np.uint32),
        np.arange(np.float32(1e-4).view(np.uint32) - 64, np.float32(1.0).view(np.uint32) + 64, 37, dtype=np.uint32),
        np.array([0x7F800000, 0x7FC00000, 0x3F800000, 0x3F7FFFFF, 0x3F800001], dtype=np.uint32),
    ])
    x = bits.view(np.float32).copy()
    mine = np.empty_like(x)
    ref = np.empty_like(x)
    L.fctc_host_log10f(x.ctypes.data, mine.ctypes.data, x.size)
    O.oracle_log10f_n(x.ctypes.data, ref.ctypes.data, x.size)
    assert x.size > 3_000_000
    assert np.array_equal(mine.view(np.uint32), ref.view(np.uint32))


def test_logf_replica_matches_glibc(fctc, oracle):
    """The duplex kernel maps posteriors to log space with the same re-evaluation of glibc's logf;
    sweep every binade incl. subnormals, plus the special values."""
    L, O = fctc.lib(), oracle.lib()
    bits = np.concatenate([
        np.arange(0x00000001, 0x7F800000, 4099, dtype=np.uint32),
        np.array([0, 0x80000000, 0x7F800000, 0xFF800000, 0x7FC00000, 0xBF800000, 0x3F800000, 0x00800000, 0x007FFFFF, 1],
                 dtype=np.uint32),
    ])
    x = bits.view(np.float32).copy()
    mine = np.empty_like(x)
    ref = np.empty_like(x)
    L.fctc_host_logf(x.ctypes.data, mine.ctypes.data, x.size)
    O.oracle_logf_n(x.ctypes.data, ref.ctypes.data, x.size)
    nan = np.isnan(ref)
    assert np.array_equal(np.isnan(mine), nan)
    assert np.array_equal(mine.view(np.uint32)[~nan], ref.view(np.uint32)[~nan])


def test_expf_log1pf_replicas_match_glibc(fctc, oracle):
    """LogSpace '+' (src/duplex.rs:45-52) is big + ln_1p(exp(small - big)) through the system libm.  The duplex
    kernels evaluate glibc's expf and log1pf themselves (phred.cuh); the host build of that code must agree with libm
    bit for bit: strided sweeps over all bit patterns, dense over the domains the search uses (expf of (-104, 0],
    log1pf of [0, 1]), plus special values; the branch-free forms the kernels actually run are swept over their own
    domains.  tools/libm_replica_sweep.c is the exhaustive version (every bit pattern)."""
    L, O = fctc.lib(), oracle.lib()

    def check(mine_fn, ref_fn, bits):
        x = bits.view(np.float32).copy()
        mine, ref = np.empty_like(x), np.empty_like(x)
        mine_fn(x.ctypes.data, mine.ctypes.data, x.size)
        ref_fn(x.ctypes.data, ref.ctypes.data, x.size)
        nan = np.isnan(ref)
        assert np.array_equal(np.isnan(mine), nan)
        bad = np.flatnonzero(mine.view(np.uint32)[~nan] != ref.view(np.uint32)[~nan])
        assert bad.size == 0, f"{bad.size} differ, first x={x[~nan][bad[0]]!r}"
        return x.size

    special = np.array([0, 0x80000000, 0x7F800000, 0xFF800000, 0x7FC00000, 0xBF800000, 0x3F800000, 0x00800000, 0x007FFFFF,
                        1, 0x4202422F, 0xC27C65D9, 0xC2CFF1B4, 0xC2CFF1B5, 0x42B17218, 0xBE95F61A, 0xBE95F61F, 0x3ED413D7,
                        0x31000000, 0x24800000], dtype=np.uint32)
    strided = np.arange(0, 2**32, 1999, dtype=np.uint64).astype(np.uint32)
    neg = np.arange(0xB0000000, 0xC2D20000, 61, dtype=np.uint32)      # -4.7e-10 .. -105: the search's small - big
    unit = np.arange(0x00000000, 0x3F800001, 53, dtype=np.uint32)     # [0, 1]: exp of a non-positive number
    n = check(L.fctc_host_expf, O.oracle_expf_n, np.concatenate([strided, neg, special]))
    n += check(L.fctc_host_log1pf, O.oracle_log1pf_n, np.concatenate([strided, unit, special]))
    nonpos = np.concatenate([np.arange(0x80000000, 2**32, 499, dtype=np.uint64).astype(np.uint32), neg,
                             np.array([0, 0x80000000, 0xFF800000, 0x7FC00000, 0xFFC00000, 0xC27C65D9, 0xC2CFF1B4, 0xC2CFF1B5],
                                      dtype=np.uint32)])
    n += check(L.fctc_host_expf_nonpos, O.oracle_expf_n, nonpos)
    n += check(L.fctc_host_log1pf_unit, O.oracle_log1pf_n,
               np.concatenate([unit, np.arange(0x3ED41000, 0x3ED42000, dtype=np.uint32),      # around 0.41422
                               np.arange(0x3F7FF000, 0x3F800001, dtype=np.uint32),            # up to 1
                               np.array([0, 0x7FC00000, 0x30FFFFFF, 0x31000000, 0x247FFFFF, 0x24800000], dtype=np.uint32)]))
    assert n > 40_000_000


def test_phred_host_build_matches_oracle(fctc, oracle):
    L = fctc.lib()
    rng = np.random.default_rng(3)
    probs = np.concatenate([rng.random(20000, dtype=np.float32), np.float32([0, 1, 0.5, 0.9999, 0.99995, -1, 2, np.nan])])
    for qscale, qbias in [(1.0, 0.0), (0.7, 2.0), (-1.0, 40.0), (10.0, 100.0), (1e9, 0.0), (3000.0, 55000.0)]:
        for p in probs[:: (1 if (qscale, qbias) == (1.0, 0.0) else 20)]:
            assert L.fctc_host_phred(p, qscale, qbias) == oracle.phred(p, qscale, qbias)[0]
