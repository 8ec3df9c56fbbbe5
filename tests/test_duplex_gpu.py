"""GPU parity tests for the 2-D (duplex) beam search through the C ABI vs the CPU oracle.

The kernels map inputs to log space with glibc's logf algorithm and, in exp_mode 0 (exact exp, the
north-star parity target), evaluate LogSpace '+' with glibc's expf and log1pf algorithms (phred.cuh,
checked against libm over every bit pattern), so every score has the bits the oracle computes:
results must be identical in exp_mode 0 as in exp_mode 1 (the reference's default build: fastexp()
== 0, '+' is max).  Only exp_mode 2 (CUDA's own expf/log1pf, the opt-in fast mode) is compared with
a tolerance on the fraction of identical pairs.  Beams up to 8 run duplex_packed.cu (four pairs per
warp), wider ones duplex.cu; the cases below cover both."""
import numpy as np
import pytest

from helpers import band_envelope, duplex_pairs, load_kats, softmax_posteriors

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def seqs(res):
    return [res["labels"][o:o + n].tolist() for o, n in zip(res["offsets"], res["lens"])]


@pytest.mark.parametrize("exp_mode", [0, 1])
@pytest.mark.parametrize("kat", [pytest.param(k, id=k["name"]) for k in load_kats() if k["fn"] == "beam_search_duplex"])
def test_reference_kats(dec, oracle, kat, exp_mode):
    a = np.asarray(kat["post"], np.float32)
    b = np.asarray(kat["post2"], np.float32)
    st, labels = dec.beam_search_duplex(a, b, None, kat["beam_size"], kat["beam_cut_threshold"], kat["collapse_repeats"],
                                        exp_mode)
    assert st == 0
    assert oracle.to_text(kat["alphabet"], labels) == kat["expect_seq"]


CASES = [
    # n_pairs, T1, T2, C, beam, thr, collapse, band half width (None = default full envelope)
    (12, 200, 200, 5, 5, 0.0, True, 8), (12, 200, 230, 5, 5, 0.1, True, 12), (8, 300, 300, 5, 5, 0.0, False, 6),
    (6, 120, 120, 5, 5, 0.0, True, None), (6, 150, 140, 3, 5, 0.0, True, 10), (4, 200, 200, 5, 16, 0.0, True, 8),
    (4, 160, 160, 5, 32, 0.0, True, 5), (6, 100, 100, 5, 1, 0.0, True, 4), (3, 1, 7, 5, 5, 0.0, True, None),
    (4, 400, 400, 5, 5, 0.1, True, 32),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "N{}_T{}x{}_C{}_b{}_thr{}_c{}_w{}".format(*c))
def test_max_plus_mode_is_exact(dec, oracle, case):
    N, T1, T2, Cn, beam, thr, collapse, w = case
    rng = np.random.default_rng(1009 * (N + T1 + T2 + beam))
    p1, p2 = duplex_pairs(rng, N, T1, T2, Cn)
    env = None if w is None else band_envelope(N, T1, T2, w)
    res = dec.beam_search_duplex_batch(p1, p2, env, beam, thr, collapse, exp_mode=1)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, None if env is None else env.astype(np.uint64), beam, thr,
                                                    collapse, exp_mode=1)
    assert res["status"].tolist() == ost.tolist()
    got = seqs(res)
    want = [ol[n, :olen[n]].tolist() for n in range(N)]
    assert got == want


@pytest.mark.parametrize("case", CASES, ids=lambda c: "N{}_T{}x{}_C{}_b{}_thr{}_c{}_w{}".format(*c))
def test_exact_exp_mode_is_exact(dec, oracle, case):
    """exp_mode 0, the --no-default-features build the north star names as the parity target: identical."""
    N, T1, T2, Cn, beam, thr, collapse, w = case
    rng = np.random.default_rng(2003 * (N + T1 + T2 + beam))
    p1, p2 = duplex_pairs(rng, N, T1, T2, Cn)
    env = None if w is None else band_envelope(N, T1, T2, w)
    res = dec.beam_search_duplex_batch(p1, p2, env, beam, thr, collapse, exp_mode=0)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, None if env is None else env.astype(np.uint64), beam, thr,
                                                    collapse, exp_mode=0)
    assert res["status"].tolist() == ost.tolist()
    assert seqs(res) == [ol[n, :olen[n]].tolist() for n in range(N)]


def test_exact_exp_mode_soft_rows(dec, oracle):
    """Softmax rows (every label plausible on every row) keep both operands of most log-adds within a few units of
    each other, the regime where expf / log1pf rounding decides bits: still identical.  Beams 5 and 8 take the
    four-pairs-per-warp kernel, 32 the one-pair-per-warp kernel; 19 pairs leave the last warp partly empty."""
    rng = np.random.default_rng(31337)
    N, T1, T2 = 19, 250, 260
    p1, p2 = softmax_posteriors(rng, (N, T1, 5), scale=1.5), softmax_posteriors(rng, (N, T2, 5), scale=1.5)
    env = band_envelope(N, T1, T2, 10)
    for beam, thr in [(5, 0.0), (8, 0.05), (32, 0.0), (3, 0.0)]:
        for mode in (0, 1):
            res = dec.beam_search_duplex_batch(p1, p2, env, beam, thr, True, exp_mode=mode)
            ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, env.astype(np.uint64), beam, thr, True, exp_mode=mode)
            assert res["status"].tolist() == ost.tolist()
            assert seqs(res) == [ol[n, :olen[n]].tolist() for n in range(N)]


def test_packed_kernel_matches_one_pair_per_warp_kernel(dec, oracle):
    """Ragged windows (every pair its own band width and drift) through the device entry point: the pairs of one warp
    then run windows of different lengths side by side.  Checked against the oracle, both modes."""
    rng = np.random.default_rng(99)
    N, T1, T2 = 23, 180, 200
    p1, p2 = duplex_pairs(rng, N, T1, T2, 5)
    env = np.empty((N, T1, 2), np.int64)
    for n in range(N):
        w = 3 + (n * 5) % 17
        c = (np.arange(T1) * T2) // T1 + (n % 3) - 1
        hi = np.maximum.accumulate(np.clip(c + w + 1, 1, T2))
        lo = np.clip(c - w, 0, None)
        lo = np.minimum(lo, np.concatenate([[0], hi[:-1]]))
        lo = np.minimum(lo, hi - 1)
        env[n, :, 0], env[n, :, 1] = lo, hi
    for mode in (0, 1):
        res = dec.beam_search_duplex_batch(p1, p2, env, 5, 0.0, True, exp_mode=mode)
        ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, env.astype(np.uint64), 5, 0.0, True, exp_mode=mode)
        assert res["status"].tolist() == ost.tolist() and (ost == 0).all()
        assert seqs(res) == [ol[n, :olen[n]].tolist() for n in range(N)]


RING_SCRIPT = r"""
import json, sys, numpy as np
import fctc_b200, oracle, torch
from helpers import band_envelope, duplex_pairs
rng = np.random.default_rng(4242)
N, T = 9, 1500
p1, p2 = duplex_pairs(rng, N, T, T, 5)
env = band_envelope(N, T, T, 12)
dec = fctc_b200.Decoder(0)
out = {}
for mode in (0, 1):
    res = dec.beam_search_duplex_batch(p1, p2, env, 5, 0.0, True, exp_mode=mode)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, env.astype(np.uint64), 5, 0.0, True, exp_mode=mode)
    got = [res["labels"][o:o + n].tolist() for o, n in zip(res["offsets"], res["lens"])]
    out[str(mode)] = bool(res["status"].tolist() == ost.tolist() and got == [ol[n, :olen[n]].tolist() for n in range(N)])
# the device entry point reports a refused hand-over instead of retrying
dev = torch.device("cuda:0")
d1, d2, de = torch.from_numpy(p1).to(dev), torch.from_numpy(p2).to(dev), torch.from_numpy(env).to(dev)
cap = N * ((T + 31) // 32 * 32)
labels = torch.empty(cap, dtype=torch.uint8, device=dev); offsets = torch.empty(N, dtype=torch.int64, device=dev)
lens = torch.empty(N, dtype=torch.int32, device=dev); status = torch.empty(N, dtype=torch.int32, device=dev)
used = torch.zeros(1, dtype=torch.int64, device=dev)
nb = dec.duplex_scratch_bytes(N, T, T, 5, 5, 25, 0)
scratch = torch.empty(nb, dtype=torch.uint8, device=dev)
dec.beam_search_duplex_batch_device(d1.data_ptr(), d2.data_ptr(), N, T, T, 5, de.data_ptr(), 25, 5, 0.0, True, 1,
                                    labels.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                    used.data_ptr(), scratch.data_ptr(), nb, 0, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
out["device_status"] = status.cpu().tolist()
print(json.dumps(out))
"""


def test_ring_pool_reuse_and_overflow_retry():
    """The four-pairs-per-warp kernel keeps the nodes' SecondaryProbs cells in a pool of rings that dead nodes hand
    on to new ones.  With the default pool (1024 rings) a 1500-row pair reuses every ring several times; with a pool
    of 16 (FCTC_DUPLEX_RINGS, read once per process, hence the subprocess) the hand-over check refuses rings that
    are still alive, the pair is flagged, and the host entry point decodes it again with one ring per node.  Both
    ways the host result is the oracle's."""
    import json
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for rings, expect_refusals in ((0, False), (16, True)):
        envv = dict(os.environ)
        envv["PYTHONPATH"] = os.pathsep.join([root, os.path.join(root, "fast-ctc-decode_b200"), here])
        if rings:
            envv["FCTC_DUPLEX_RINGS"] = str(rings)
        r = subprocess.run([sys.executable, "-c", RING_SCRIPT], env=envv, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stderr[-2000:]
        out = json.loads(r.stdout.strip().splitlines()[-1])
        assert out["0"] and out["1"], f"rings={rings}: host results differ from the oracle"
        assert set(out["device_status"]) <= {0, 4}
        assert (4 in out["device_status"]) == expect_refusals, f"rings={rings}: device statuses {out['device_status']}"


CHUNK_SCRIPT = r"""
import json, numpy as np
import fctc_b200, oracle
from helpers import duplex_pairs, band_envelope
oracle.build()
dec = fctc_b200.Decoder(0)
rng = np.random.default_rng(99)
N, T = 37, 120
p1, p2 = duplex_pairs(rng, N, T, T + 7)
env = band_envelope(N, T, T + 7, 12)
ok = {}
for mode in (0, 1):
    for f16 in (False, True):
        a1, a2 = (p1.astype(np.float16), p2.astype(np.float16)) if f16 else (p1, p2)
        res = dec.beam_search_duplex_batch(a1, a2, env, 5, 0.0, True, exp_mode=mode)
        ol, olen, ost = oracle.beam_search_duplex_batch(a1.astype(np.float32), a2.astype(np.float32), env.astype(np.uint64), 5, 0.0,
                                                        True, exp_mode=mode)
        same = res["status"].tolist() == ost.tolist()
        for n in range(N):
            o, ln = int(res["offsets"][n]), int(res["lens"][n])
            same = same and res["labels"][o:o + ln].tolist() == ol[n, :olen[n]].tolist()
        ok[f"{mode}_{int(f16)}"] = bool(same)
print(json.dumps(ok))
"""


def test_host_batch_in_several_chunks():
    """The host entry point copies chunk k+1 into a second set of input buffers while chunk k's kernel runs.  With
    FCTC_DUPLEX_CHUNK_PAIRS=8 (read once per process, hence the subprocess) 37 pairs go through five chunks, the
    last a short one; float32 and binary16, exact-exp and max-plus: all equal to the oracle."""
    import json
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    envv = dict(os.environ)
    envv["PYTHONPATH"] = os.pathsep.join([root, os.path.join(root, "fast-ctc-decode_b200"), here])
    envv["FCTC_DUPLEX_CHUNK_PAIRS"] = "8"
    r = subprocess.run([sys.executable, "-c", CHUNK_SCRIPT], env=envv, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out and all(out.values()), out


@pytest.mark.parametrize("case", CASES[:5], ids=lambda c: "N{}_T{}x{}_C{}_b{}_thr{}_c{}_w{}".format(*c))
def test_cuda_exp_mode_tolerance(dec, oracle, case):
    """exp_mode 2 (CUDA's expf / log1pf): last-bit differences may flip a near-tie in a few pairs."""
    N, T1, T2, Cn, beam, thr, collapse, w = case
    rng = np.random.default_rng(2003 * (N + T1 + T2 + beam))
    p1, p2 = duplex_pairs(rng, N, T1, T2, Cn)
    env = None if w is None else band_envelope(N, T1, T2, w)
    res = dec.beam_search_duplex_batch(p1, p2, env, beam, thr, collapse, exp_mode=2)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, None if env is None else env.astype(np.uint64), beam, thr,
                                                    collapse, exp_mode=0)
    assert res["status"].tolist() == ost.tolist()
    got = seqs(res)
    want = [ol[n, :olen[n]].tolist() for n in range(N)]
    same = sum(g == w_ for g, w_ in zip(got, want))
    assert same >= int(np.ceil(0.75 * N)), f"only {same}/{N} pairs identical"
    for g, w_ in zip(got, want):
        assert abs(len(g) - len(w_)) <= max(2, len(w_) // 20)


def test_error_statuses(dec, oracle):
    import fctc_b200
    rng = np.random.default_rng(5)
    p1, p2 = duplex_pairs(rng, 4, 60, 60)
    p1[1] = np.nan                                     # tests/test_decode.py:370-374
    env = band_envelope(4, 60, 60, 6)
    env[2, 30:, 0] += 40                               # lower bound jumps past the last upper bound
    env[3, 10, :] = (20, 20)                           # empty window
    res = dec.beam_search_duplex_batch(p1, p2, env, 5, 0.0, True, exp_mode=0)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, env.astype(np.uint64), 5, 0.0, True, 0)
    assert res["status"].tolist() == ost.tolist() == [0, fctc_b200.INCOMPARABLE_VALUES, fctc_b200.INVALID_ENVELOPE,
                                                      fctc_b200.INVALID_ENVELOPE]
    st, labels = dec.beam_search_duplex(p1[1], p2[1])
    assert st == fctc_b200.INCOMPARABLE_VALUES
    assert fctc_b200.lib().fctc_status_message(3) == b"Invalid envelope values"
    with pytest.raises(fctc_b200.FctcError):
        dec.beam_search_duplex(p1[0], p2[0], beam_size=0)
    with pytest.raises(ValueError):
        dec.beam_search_duplex(p1[0], p2[0][:, :4])
    with pytest.raises(ValueError):
        dec.beam_search_duplex(p1[0], p2[0], np.zeros((59, 2), np.uint64))


def test_single_pair_with_envelope_and_strides(dec, oracle):
    rng = np.random.default_rng(6)
    p1, p2 = duplex_pairs(rng, 1, 90, 100)
    big1 = np.zeros((180, 10), np.float32)
    big1[::2, ::2] = p1[0]
    env = band_envelope(1, 90, 100, 9)[0].astype(np.uint64)
    st, labels = dec.beam_search_duplex(big1[::2, ::2], p2[0], env, 5, 0.0, True, exp_mode=1)
    ost, ol = oracle.beam_search_duplex(p1[0], p2[0], env, 5, 0.0, True, exp_mode=1)
    assert st == ost == 0 and labels.tolist() == ol.tolist()


def test_f16_posteriors_host_batch(dec, oracle):
    """fctc_beam_search_duplex_batch_f16: identical to the oracle on the exactly widened arrays, both exp modes."""
    rng = np.random.default_rng(556)
    p1, p2 = duplex_pairs(rng, 10, 260, 250, 5)
    h1, h2 = p1.astype(np.float16), p2.astype(np.float16)
    env = band_envelope(10, 260, 250, 9)
    for mode in (0, 1):
        res = dec.beam_search_duplex_batch(h1, h2, env, 5, 0.0, True, exp_mode=mode)
        ol, olen, ost = oracle.beam_search_duplex_batch(h1.astype(np.float32), h2.astype(np.float32), env.astype(np.uint64),
                                                        5, 0.0, True, exp_mode=mode)
        assert res["status"].tolist() == ost.tolist()
        assert seqs(res) == [ol[n, :olen[n]].tolist() for n in range(10)]


def test_f16_posteriors_device_batch(dec, oracle):
    """Binary16 posteriors through the device entry point: identical to the oracle on the exactly widened
    arrays (max-plus mode, which is bit-exact)."""
    import torch
    N, T1, T2, Cn, beam, w = 24, 300, 320, 5, 5, 12
    rng = np.random.default_rng(555)
    p1, p2 = duplex_pairs(rng, N, T1, T2, Cn)
    h1, h2 = p1.astype(np.float16), p2.astype(np.float16)
    env = band_envelope(N, T1, T2, w)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    d1, d2, de = torch.from_numpy(h1).to(dev), torch.from_numpy(h2).to(dev), torch.from_numpy(env).to(dev)
    stride = (T1 + 31) // 32 * 32
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    mw = 2 * w + 1
    nbytes = dec.duplex_scratch_bytes(N, T1, T2, Cn, beam, mw, 0)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dec.beam_search_duplex_batch_device(d1.data_ptr(), d2.data_ptr(), N, T1, T2, Cn, de.data_ptr(), mw, beam, 0.0, True, 1,
                                        labels.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                        used.data_ptr(), scratch.data_ptr(), nbytes, 0, st, f16=True)
    torch.cuda.synchronize()
    ol, olen, ost = oracle.beam_search_duplex_batch(h1.astype(np.float32), h2.astype(np.float32), env.astype(np.uint64),
                                                    beam, 0.0, True, exp_mode=1)
    lab, off, ln, stt = labels.cpu().numpy(), offsets.cpu().numpy(), lens.cpu().numpy(), status.cpu().numpy()
    assert stt.tolist() == ost.tolist()
    for n in range(N):
        assert lab[off[n]:off[n] + ln[n]].tolist() == ol[n, :olen[n]].tolist()
