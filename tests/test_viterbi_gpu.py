"""GPU parity tests for the Viterbi path: the CUDA kernels, called through the C ABI
(include/fctc.h), against the CPU oracle on identical inputs.  Bit-exact: labels, path and quality
characters (north star: "Viterbi sequence, path and quality string must be bit-exact")."""
import numpy as np
import pytest

from helpers import kat_post, load_kats, peaked_posteriors, softmax_posteriors

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def check_batch(oracle, post, res, qstring, qscale, qbias, collapse, reads=None):
    N = post.shape[0]
    reads = range(N) if reads is None else reads
    ol, oq, op, olen, ost = oracle.viterbi_batch(np.ascontiguousarray(post[list(reads)]), qstring, qscale, qbias,
                                                 collapse)
    slots = []
    for j, n in enumerate(reads):
        off, ln = int(res["offsets"][n]), int(res["lens"][n])
        assert ost[j] == oracle.OK and res["status"][n] == 0
        assert ln == olen[j], f"read {n}: length {ln} != oracle {olen[j]}"
        assert np.array_equal(res["labels"][off:off + ln], ol[j, :ln]), f"read {n}: labels differ"
        assert np.array_equal(res["path"][off:off + ln], op[j, :ln]), f"read {n}: path differs"
        if qstring:
            assert np.array_equal(res["quals"][off:off + ln].astype(np.uint32), oq[j, :ln]), f"read {n}: quality differs"
        assert off % 32 == 0
        slots.append((off, ln))
    # pool slots never overlap
    slots.sort()
    for (a, la), (b, _) in zip(slots, slots[1:]):
        assert a + la <= b


@pytest.mark.parametrize("kat", [pytest.param(k, id=k["name"]) for k in load_kats() if k["fn"] == "viterbi_search"])
def test_reference_kats_single_read(dec, oracle, kat):
    post = kat_post(kat["post"])
    st, labels, quals, path = dec.viterbi_search(post, kat["qstring"], kat["qscale"], kat["qbias"],
                                                 kat["collapse_repeats"])
    assert st == 0
    seq = oracle.to_text(kat["alphabet"], labels, quals)
    if kat["expect_seq"] is not None:
        assert seq == kat["expect_seq"]
    if kat["expect_path"] is not None:
        assert path.tolist() == kat["expect_path"]


def test_strided_single_read(dec, oracle):
    rng = np.random.default_rng(0)
    big = softmax_posteriors(rng, (300, 10))
    view = big[::3, ::2]  # non-contiguous in both axes, as numpy views reach the reference (src/lib.rs:198)
    st, labels, quals, path = dec.viterbi_search(view, True)
    ost, ol, oq, op = oracle.viterbi(view, True)
    assert st == ost == 0
    assert np.array_equal(labels, ol) and np.array_equal(quals, oq) and np.array_equal(path, op)
    # Fortran-ordered input
    f = np.asfortranarray(softmax_posteriors(rng, (257, 5)))
    st, labels, quals, path = dec.viterbi_search(f, True, 0.7, 2.0, False)
    ost, ol, oq, op = oracle.viterbi(f, True, 0.7, 2.0, False)
    assert np.array_equal(labels, ol) and np.array_equal(quals, oq) and np.array_equal(path, op)


SHAPES = [
    # N, T, C  — tile = 1024 rows: below, exact multiples, ragged tails, odd alphabets / alignments
    (7, 100, 5), (5, 1024, 5), (3, 4000, 5), (4, 4099, 5), (9, 1025, 5), (6, 777, 3), (5, 2049, 4), (3, 1500, 12),
    (4, 1333, 7), (2, 10000, 5), (3, 25, 5), (2, 1, 5), (3, 2, 2), (2, 3000, 1), (11, 4000, 5),
]


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("kind", ["softmax", "peaked"])
def test_batch_host_parity(dec, oracle, shape, kind):
    rng = np.random.default_rng(1000 * sum(shape) + len(kind))
    post = softmax_posteriors(rng, shape) if kind == "softmax" else peaked_posteriors(rng, shape)
    for qstring, qscale, qbias, collapse in [(True, 1.0, 0.0, True), (True, 0.7, 2.0, False), (False, 1.0, 0.0, True)]:
        res = dec.viterbi_batch(post, qstring, qscale, qbias, collapse, qual_bytes=1)
        check_batch(oracle, post, res, qstring, qscale, qbias, collapse)


def test_long_runs_across_tiles(dec, oracle):
    """Runs of one label that span several 1024-row tiles: the sequential f32 mean must carry over."""
    rng = np.random.default_rng(5)
    T = 5000
    post = np.zeros((4, T, 5), np.float32)
    post[0, :, 2] = rng.random(T, dtype=np.float32) * 0.5 + 0.5   # one run over the whole read
    post[1, :, 0] = 0.9
    post[1, 1000:3100, 3] = rng.random(2100, dtype=np.float32) * 0.05 + 0.95  # run crossing two tile edges
    post[2, :, 0] = 0.4
    post[2, 1020:1030, 1] = 0.8      # run straddling one edge
    post[2, 2047, 4] = 0.9           # emission on the last row of a tile
    post[2, 2048, 4] = 0.7           # continues on the first row of the next
    post[3, :, 0] = 1.0              # all blank
    for collapse in (True, False):
        res = dec.viterbi_batch(post, True, 1.0, 0.0, collapse)
        check_batch(oracle, post, res, True, 1.0, 0.0, collapse)
    assert res["lens"][3] == 0


def test_nan_ties_and_garbage(dec, oracle):
    """find_max semantics (src/search.rs:303-318): ties -> lowest index; NaN never wins unless first."""
    rng = np.random.default_rng(9)
    post = softmax_posteriors(rng, (3, 600, 5))
    post[0, ::7, :] = 0.2                      # exact ties -> blank
    post[0, 3::11, 1:] = 0.25                  # ties among labels -> label 1
    post[1, ::5, 2] = np.nan                   # NaN in a label column: ignored
    post[1, 1::9, 0] = np.nan                  # NaN in the blank column: blank "wins"
    post[2] = rng.normal(size=(600, 5)).astype(np.float32) * 3  # negative / >1 "probabilities"
    res = dec.viterbi_batch(post, True, 1.0, 0.0, True)
    check_batch(oracle, post, res, True, 1.0, 0.0, True)


def test_wide_quality_codes(dec, oracle):
    """qscale/qbias that push the phred code past 255 need 32-bit quality cells, like Rust's char."""
    import fctc_b200
    rng = np.random.default_rng(11)
    post = peaked_posteriors(rng, (3, 900, 5))
    with pytest.raises(fctc_b200.FctcError):
        dec.viterbi_batch(post, True, 30.0, 100.0, True, qual_bytes=1)
    res = dec.viterbi_batch(post, True, 30.0, 100.0, True, qual_bytes=4)
    check_batch(oracle, post, res, True, 30.0, 100.0, True)
    assert max(res["quals"][o:o + ln].max() for o, ln in zip(res["offsets"], res["lens"]) if ln) > 255
    # codes in the surrogate range are not chars: the reference would abort, we flag the read
    res = dec.viterbi_batch(post, True, 0.0, 56000.0, True, qual_bytes=4)
    assert (res["status"] == fctc_b200.INVALID_QUALITY_CHAR).all()


def test_generic_kernel_paths(dec, oracle):
    """Shapes the streaming kernel does not take: very wide alphabets and very long reads."""
    rng = np.random.default_rng(13)
    for shape in [(2, 300, 200), (1, 70000, 5)]:
        post = softmax_posteriors(rng, shape)
        res = dec.viterbi_batch(post, True, 1.0, 0.0, True)
        check_batch(oracle, post, res, True, 1.0, 0.0, True)


def test_capacity_and_argument_errors(dec):
    import fctc_b200
    rng = np.random.default_rng(17)
    post = softmax_posteriors(rng, (4, 500, 5))
    small = dict(labels=np.empty(64, np.uint8), path=np.empty(64, np.int32))
    with pytest.raises(fctc_b200.FctcError) as e:
        dec.viterbi_batch(post, False, out=small)
    assert e.value.code == fctc_b200.E_CAPACITY
    with pytest.raises(fctc_b200.FctcError):
        dec.viterbi_search(np.zeros((0, 5), np.float32))
    with pytest.raises(TypeError):
        dec.viterbi_search(np.zeros((5, 5), np.float64))


def test_device_batch_synthetic_large(dec, oracle):
    """Device-resident API on generator output (the benchmark's input), checked on a sample of
    reads copied back, plus size-independent invariants over the whole batch."""
    import torch
    N, T, Cn = 4096, 4000, 5
    dev = torch.device("cuda:0")
    for kind in (0, 1):
        post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream().cuda_stream
        dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, kind, seed=0xC7C00001, first_read=0, stream=stream)
        cap = N * ((T + 31) // 32 * 32)
        labels = torch.empty(cap, dtype=torch.uint8, device=dev)
        quals = torch.empty(cap, dtype=torch.uint8, device=dev)
        path = torch.empty(cap, dtype=torch.int32, device=dev)
        offsets = torch.empty(N, dtype=torch.int64, device=dev)
        lens = torch.empty(N, dtype=torch.int32, device=dev)
        status = torch.empty(N, dtype=torch.int32, device=dev)
        dec.viterbi_batch_device(post.data_ptr(), N, T, Cn, True, 0.7, 2.0, True, labels.data_ptr(), quals.data_ptr(),
                                 1, path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                 stream)
        torch.cuda.synchronize()
        res = dict(labels=labels.cpu().numpy(), quals=quals.cpu().numpy(), path=path.cpu().numpy(),
                   offsets=offsets.cpu().numpy(), lens=lens.cpu().numpy(), status=status.cpu().numpy())
        rows = post.sum(dim=2)
        assert torch.allclose(rows, torch.ones_like(rows), atol=1e-5)  # generator emits normalised rows
        sample = list(range(0, N, 97)) + [N - 1]
        check_batch(oracle, post.cpu().numpy(), res, True, 0.7, 2.0, True, reads=sample)
        # invariants over every read: status ok, labels non-blank, paths strictly increasing, counts add up
        assert (res["status"] == 0).all()
        lens_np = res["lens"].astype(np.int64)
        assert np.array_equal(res["offsets"], np.arange(N, dtype=np.int64) * ((T + 31) // 32 * 32))  # fixed slots
        # device-side recount of emissions: argmax, blank drop, repeat collapse
        mx = post.max(dim=2, keepdim=True).values  # first maximum, as find_max (torch.argmax leaves ties open)
        idx = torch.arange(Cn, device=dev).expand_as(post)
        arg = torch.where(post == mx, idx, torch.full_like(idx, Cn)).min(dim=2).values
        prev = torch.cat([torch.full((N, 1), -1, device=dev, dtype=arg.dtype), arg[:, :-1]], dim=1)
        emitted = ((arg != 0) & (arg != prev)).sum(dim=1).cpu().numpy()
        assert np.array_equal(emitted, lens_np)
        del post


# ---- binary16 posteriors: decoded as the exactly widened float32 values ----
F16_CASES = [
    # N, T, C, qstring, collapse
    (40, 4000, 5, True, True), (33, 1024, 5, True, False), (17, 8, 5, True, True), (9, 4008, 5, False, True),
    (12, 999, 5, True, True),   # T % 8 != 0: generic kernel
    (6, 640, 4, True, True), (5, 300, 12, True, True), (3, 1, 5, True, True),
]


@pytest.mark.parametrize("case", F16_CASES, ids=lambda c: "N{}_T{}_C{}_q{}_c{}".format(*c))
@pytest.mark.parametrize("kind", ["softmax", "peaked"])
def test_f16_posteriors_host_batch(dec, oracle, case, kind):
    N, T, Cn, qstring, collapse = case
    rng = np.random.default_rng(1000 + N + T + Cn)
    post = softmax_posteriors(rng, (N, T, Cn)) if kind == "softmax" else peaked_posteriors(rng, (N, T, Cn))
    half = post.astype(np.float16)
    half[0, T // 2] = np.float16(np.nan)      # non-finite rows travel through the widening unchanged
    half[-1, 0] = np.float16(np.inf)
    res = dec.viterbi_batch(half, qstring, 0.7, 2.0, collapse)
    check_batch(oracle, half.astype(np.float32), res, qstring, 0.7, 2.0, collapse)


def test_f16_posteriors_device_batch(dec, oracle):
    import torch
    N, T, Cn = 3000, 4000, 5
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, 0, 77, 0, st)
    half = post.half()
    stride = (T + 31) // 32 * 32
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    quals = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    dec.viterbi_batch_device(half.data_ptr(), N, T, Cn, True, 1.0, 0.0, True, labels.data_ptr(), quals.data_ptr(), 1,
                             path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), st, f16=True)
    torch.cuda.synchronize()
    res = dict(labels=labels.cpu().numpy(), quals=quals.cpu().numpy(), path=path.cpu().numpy(),
               offsets=offsets.cpu().numpy(), lens=lens.cpu().numpy(), status=status.cpu().numpy())
    reads = list(range(0, N, 37))
    check_batch(oracle, half.float().cpu().numpy(), res, True, 1.0, 0.0, True, reads=reads)


def check_packed_layout(res):
    lens = res["lens"].astype(np.int64)
    want = np.concatenate([[0], np.cumsum((lens + 31) & ~31)])
    assert np.array_equal(res["offsets"], want[:-1]), "packed offsets are the running sum of round_up(len, 32)"
    assert res["used"] == want[-1]
    for off, ln in zip(res["offsets"], lens):  # the gap up to the next read's start is cleared
        pad = ((ln + 31) & ~31) - ln
        assert not res["labels"][off + ln:off + ln + pad].any() and not res["path"][off + ln:off + ln + pad].any()


@pytest.mark.parametrize("shape", [(7, 100, 5), (4, 4099, 5), (9, 1025, 5), (6, 777, 3), (3, 1500, 12), (2, 1, 5),
                                   (3, 2, 2), (2100, 4000, 5)], ids=lambda s: "x".join(map(str, s)))
def test_packed_host_batch(dec, oracle, shape):
    """fctc_viterbi_batch_packed: the same per-read results as the slot form, squeezed together on the device
    before the read-back ((2100, 4000, 5) spans three pipeline chunks)."""
    rng = np.random.default_rng(77 + sum(shape))
    post = softmax_posteriors(rng, shape) if shape[0] % 2 else peaked_posteriors(rng, shape)
    for qstring, qscale, qbias, collapse, qb in [(True, 0.7, 2.0, True, 1), (False, 1.0, 0.0, False, 1), (True, 30.0, 100.0, True, 4)]:
        res = dec.viterbi_batch(post, qstring, qscale, qbias, collapse, qual_bytes=qb, packed=True)
        check_packed_layout(res)
        sample = None if shape[0] < 100 else list(range(0, shape[0], 37)) + [shape[0] - 1]
        check_batch(oracle, post, res, qstring, qscale, qbias, collapse, reads=sample)
        if shape[0] >= 100:  # every read against the slot form
            ref = dec.viterbi_batch(post, qstring, qscale, qbias, collapse, qual_bytes=qb)
            assert np.array_equal(ref["lens"], res["lens"]) and np.array_equal(ref["status"], res["status"])
            for n in range(shape[0]):
                a, b, ln = int(ref["offsets"][n]), int(res["offsets"][n]), int(ref["lens"][n])
                assert np.array_equal(ref["labels"][a:a + ln], res["labels"][b:b + ln])
                assert np.array_equal(ref["path"][a:a + ln], res["path"][b:b + ln])
                if qstring:
                    assert np.array_equal(ref["quals"][a:a + ln], res["quals"][b:b + ln])
            assert res["d2h_bytes"] < ref["d2h_bytes"]


def test_packed_f16_and_capacity(dec, oracle):
    import fctc_b200
    rng = np.random.default_rng(5)
    half = softmax_posteriors(rng, (6, 1200, 5)).astype(np.float16)
    res = dec.viterbi_batch(half, True, 1.0, 0.0, True, packed=True)
    check_packed_layout(res)
    check_batch(oracle, half.astype(np.float32), res, True, 1.0, 0.0, True)
    # a pool of exactly the used size is enough; one entry less is refused
    used = res["used"]
    out = dict(labels=np.empty(used, np.uint8), quals=np.empty(used, np.uint8), path=np.empty(used, np.int32))
    again = dec.viterbi_batch(half, True, 1.0, 0.0, True, packed=True, out=out)
    assert again["used"] == used and np.array_equal(again["labels"], res["labels"][:used])
    out = dict(labels=np.empty(used - 32, np.uint8), quals=np.empty(used - 32, np.uint8), path=np.empty(used - 32, np.int32))
    with pytest.raises(fctc_b200.FctcError) as err:
        dec.viterbi_batch(half, True, 1.0, 0.0, True, packed=True, out=out)
    assert err.value.code == fctc_b200.E_CAPACITY
