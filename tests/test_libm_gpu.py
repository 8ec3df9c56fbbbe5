"""The libm restatements as the DEVICE executes them (phred.cuh compiled for sm_100a: fused operations, the branch-free
division, tables in shared memory) against the host's libm, value by value: expf on non-positive arguments, log1pf on
[0, 1], logf, log10f, and LogSpace '+' (src/duplex.rs:42-63) itself in exact-exp mode.  Tens of millions of values per
function, dense where the searches live; tools/libm_replica_sweep.c covers every bit pattern for the host builds."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def on_device(dec, which, x, y=None):
    import torch
    dx = torch.from_numpy(x).cuda()
    dy = torch.from_numpy(y).cuda() if y is not None else None
    out = torch.empty_like(dx)
    dec.libm_device(dx.data_ptr(), dy.data_ptr() if dy is not None else 0, out.data_ptr(), x.size, which,
                    torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def same_bits(a, b):
    nan = np.isnan(b)
    assert np.array_equal(np.isnan(a), nan)
    bad = np.flatnonzero(a.view(np.uint32)[~nan] != b.view(np.uint32)[~nan])
    assert bad.size == 0, f"{bad.size} of {a.size} differ"


def host(oracle, name, x):
    ref = np.empty_like(x)
    getattr(oracle.lib(), name)(x.ctypes.data, ref.ctypes.data, x.size)
    return ref


def test_expf_nonpos_on_device(dec, oracle):
    bits = np.concatenate([np.arange(0x80000000, 2**32, 127, dtype=np.uint64).astype(np.uint32),
                           np.arange(0xB0000000, 0xC2D20000, 17, dtype=np.uint32),
                           np.array([0, 0x80000000, 0xFF800000, 0x7FC00000, 0xC27C65D9, 0xC2CFF1B4, 0xC2CFF1B5], np.uint32)])
    x = bits.view(np.float32).copy()
    same_bits(on_device(dec, 0, x), host(oracle, "oracle_expf_n", x))
    assert x.size > 30_000_000


def test_log1pf_unit_on_device(dec, oracle):
    bits = np.concatenate([np.arange(0, 0x3F800001, 23, dtype=np.uint32), np.arange(0x3ED40000, 0x3ED50000, dtype=np.uint32),
                           np.arange(0x3F7F0000, 0x3F800001, dtype=np.uint32),
                           np.array([0, 0x7FC00000, 0x30FFFFFF, 0x31000000, 0x247FFFFF, 0x24800000], np.uint32)])
    x = bits.view(np.float32).copy()
    same_bits(on_device(dec, 1, x), host(oracle, "oracle_log1pf_n", x))
    assert x.size > 40_000_000


def test_logf_log10f_on_device(dec, oracle):
    bits = np.concatenate([np.arange(1, 0x7F800000, 97, dtype=np.uint32),
                           np.array([0, 0x80000000, 0x7F800000, 0xFF800000, 0x7FC00000, 0xBF800000, 0x3F800000], np.uint32)])
    x = bits.view(np.float32).copy()
    same_bits(on_device(dec, 2, x), host(oracle, "oracle_logf_n", x))
    pos = np.arange(0x00800000, 0x7F800001, 97, dtype=np.uint32).view(np.float32).copy()
    same_bits(on_device(dec, 3, pos), host(oracle, "oracle_log10f_n", pos))


def test_logspace_add_on_device(dec, oracle):
    """LogSpace '+' on pairs of log-probabilities at every distance from 0 to far beyond where the smaller one stops
    mattering, plus -inf / +inf / NaN operands in both positions."""
    rng = np.random.default_rng(1)
    n = 8_000_000
    a = (-rng.random(n, dtype=np.float32) * np.float32(900.0)).astype(np.float32)
    gap = (rng.random(n, dtype=np.float32) ** 3 * np.float32(120.0)).astype(np.float32)
    b = (a - gap).astype(np.float32)
    swap = rng.random(n) < 0.5
    x, y = np.where(swap, b, a).astype(np.float32), np.where(swap, a, b).astype(np.float32)
    special = np.array([-np.inf, np.inf, np.nan, 0.0, -0.0, -1.0, -1e-30, -104.0, -88.0], np.float32)
    sx, sy = np.meshgrid(special, special)
    x, y = np.concatenate([x, sx.ravel(), a[:1000]]), np.concatenate([y, sy.ravel(), a[:1000]])
    got = on_device(dec, 4, np.ascontiguousarray(x), np.ascontiguousarray(y))
    want = np.array([oracle.lib().oracle_logadd(float(p), float(q), 0) for p, q in zip(x[-1081:], y[-1081:])], np.float32)
    same_bits(got[-1081:], want)
    # the bulk through the vectorised libm calls: big + log1pf(expf(small - big)), -inf an exact unit
    big, small = np.maximum(x[:n], y[:n]), np.minimum(x[:n], y[:n])
    d = (small - big).astype(np.float32)
    e = host(oracle, "oracle_expf_n", d)
    l = host(oracle, "oracle_log1pf_n", e)
    same_bits(got[:n], (big + l).astype(np.float32))


@pytest.mark.parametrize("qscale,qbias", [(1.0, 0.0), (0.7, 2.0), (2.5, -3.0), (1.0, 33.0), (0.1, 0.0), (40.0, 100.0), (-1.0, 50.0),
                                          (0.013, -0.5), (3.3, 1e3)])
def test_fast_quality_value_stays_inside_its_margin(dec, qscale, qbias):
    """The Viterbi kernel takes the quality character from a hardware lg2 whenever the value lies farther than eps
    from a rounding boundary (csrc/phred.cuh, phred_fast).  Here the distance between that value and the reference's
    (glibc log10f, three rounded float32 operations) is measured on the device for EVERY float32 miss probability from
    1e-4 to 1 and a stride through (1, 2^40]: it must stay below a quarter of eps."""
    import torch
    lo = int(np.float32(1e-4).view(np.uint32))
    bits = np.concatenate([np.arange(lo, 0x3F800001, dtype=np.uint32), np.arange(0x3F800000, 0x53800000, 4099, dtype=np.uint32)])
    x = bits.view(np.float32).copy()
    y = np.array([qscale, qbias], np.float32)
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    diff, eps = torch.empty_like(dx), torch.empty_like(dx)
    st = torch.cuda.current_stream().cuda_stream
    dec.libm_device(dx.data_ptr(), dy.data_ptr(), diff.data_ptr(), x.size, 5, st)
    dec.libm_device(dx.data_ptr(), dy.data_ptr(), eps.data_ptr(), x.size, 6, st)
    torch.cuda.synchronize()
    ratio = float((diff.abs() / eps).max())
    assert x.size > 100_000_000 and ratio < 0.25, ratio
