"""GPU parity tests for the 1-D beam search: CUDA kernel through the C ABI vs the CPU oracle on
identical inputs.  The kernel repeats the reference's f32 operations in the same order and the
oracle's tie rule (equal scores -> lower node id), so sequences and paths are expected to be
identical on every read (the north star only asks for >= 99.9 %)."""
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


def check_batch(oracle, post, res, beam, thr, collapse, reads=None):
    reads = list(range(post.shape[0])) if reads is None else list(reads)
    ol, op, olen, ost, _ = oracle.beam_search_batch(np.ascontiguousarray(post[reads]), beam, thr, collapse)
    bad = []
    for j, n in enumerate(reads):
        off, ln = int(res["offsets"][n]), int(res["lens"][n])
        assert res["status"][n] == ost[j], f"read {n}: status {res['status'][n]} != oracle {ost[j]}"
        if ost[j] != 0:
            assert ln == 0
            continue
        same = ln == olen[j] and np.array_equal(res["labels"][off:off + ln], ol[j, :ln]) and \
            np.array_equal(res["path"][off:off + ln], op[j, :ln])
        if not same:
            bad.append(n)
    assert not bad, f"{len(bad)} of {len(reads)} reads differ from the oracle: {bad[:10]}"


@pytest.mark.parametrize("kat", [pytest.param(k, id=k["name"]) for k in load_kats() if k["fn"] == "beam_search"])
def test_reference_kats_single_read(dec, oracle, kat):
    post = kat_post(kat["post"])
    st, labels, path = dec.beam_search(post, kat["beam_size"], kat["beam_cut_threshold"], kat["collapse_repeats"])
    assert st == 0
    if kat["expect_seq"] is not None:
        assert oracle.to_text(kat["alphabet"], labels) == kat["expect_seq"]
    if kat["expect_path"] is not None:
        assert path.tolist() == kat["expect_path"]
    assert len(labels) == len(path)


CASES = [
    # N, T, C, beam, threshold, collapse
    (24, 400, 5, 5, 0.1, True), (24, 400, 5, 5, 0.0, True), (16, 400, 5, 5, 0.1, False), (12, 600, 5, 32, 0.0, True),
    (12, 300, 5, 32, 0.0, False), (8, 500, 5, 1, 0.0, True), (8, 500, 5, 2, 0.05, True), (8, 300, 5, 8, 0.0, True),
    (6, 300, 5, 40, 0.0, True), (4, 200, 5, 64, 0.0, True), (6, 500, 3, 5, 0.1, True), (4, 400, 12, 5, 0.0, True),
    (4, 300, 12, 16, 0.01, True), (6, 700, 4, 7, 0.0, True), (5, 33, 5, 5, 0.1, True), (3, 1, 5, 5, 0.0, True),
    (3, 250, 2, 5, 0.0, True), (2, 4000, 5, 5, 0.1, True), (2, 2000, 5, 32, 0.0, True),
    # the narrow-beam kernel's corners (beam * C <= 32) and the first shapes past it
    (6, 300, 2, 16, 0.0, True), (6, 300, 3, 10, 0.02, True), (6, 300, 4, 8, 0.0, False), (6, 300, 5, 6, 0.0, True),
    (6, 300, 5, 7, 0.0, True), (6, 300, 3, 11, 0.0, True),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "N{}_T{}_C{}_b{}_thr{}_c{}".format(*c))
@pytest.mark.parametrize("kind", ["softmax", "peaked"])
def test_batch_host_parity(dec, oracle, case, kind):
    N, T, Cn, beam, thr, collapse = case
    rng = np.random.default_rng(7919 * (N + T + Cn + beam) + len(kind))
    post = softmax_posteriors(rng, (N, T, Cn), scale=(1.0 if beam > 8 else 3.0)) if kind == "softmax" \
        else peaked_posteriors(rng, (N, T, Cn))
    res = dec.beam_search_batch(post, beam, thr, collapse)
    check_batch(oracle, post, res, beam, thr, collapse)


@pytest.mark.parametrize("beam", [3, 5, 6, 16, 32])
@pytest.mark.parametrize("collapse", [True, False])
def test_score_ties_follow_the_oracle_rule(dec, oracle, beam, collapse):
    """Rows of equal or few distinct probabilities make many candidates tie exactly: the ranking's
    fast rounds must notice and the exact rounds must apply "equal score -> lower node id"."""
    rng = np.random.default_rng(100 + beam)
    T, Cn = 60, 5
    flat = np.full((T, Cn), 0.2, np.float32)
    coarse = rng.choice(np.array([0.125, 0.25, 0.5], np.float32), size=(T, Cn))
    half = flat.copy()
    half[::2] = coarse[::2]
    for post in (flat, coarse, half):
        st, labels, path = dec.beam_search(post, beam, 0.0, collapse)
        ost, ol, op = oracle.beam_search(post, beam, 0.0, collapse)
        assert st == ost == 0
        assert np.array_equal(labels, ol) and np.array_equal(path, op)


def test_unnormalised_readme_input(dec, oracle):
    """README.md:18-27: np.random.rand(100, 5) without normalisation, beam 5, threshold 0.1."""
    rng = np.random.default_rng(2)
    post = rng.random((6, 100, 5), dtype=np.float32)
    res = dec.beam_search_batch(post, 5, 0.1, True)
    check_batch(oracle, post, res, 5, 0.1, True)
    x = post[0]
    st, labels, path = dec.beam_search(x, beam_size=5, beam_cut_threshold=0.1)
    ost, ol, op = oracle.beam_search(x, 5, 0.1)
    assert st == ost == 0 and np.array_equal(labels, ol) and np.array_equal(path, op)


def test_error_statuses(dec, oracle):
    """SearchError paths (src/lib.rs:36-58): NaN input, beam emptied by the threshold."""
    import fctc_b200
    rng = np.random.default_rng(3)
    post = softmax_posteriors(rng, (5, 200, 5))
    post[1] = np.nan                      # tests/test_decode.py:100-104
    post[3, 50:, :] = 0.01                # every label below the threshold from row 50 on
    post[4, 120, 2] = np.nan              # a single NaN in the middle
    res = dec.beam_search_batch(post, 5, 0.1, True)
    assert res["status"].tolist() == [0, fctc_b200.INCOMPARABLE_VALUES, 0, fctc_b200.RAN_OUT_OF_BEAM,
                                      fctc_b200.INCOMPARABLE_VALUES]
    check_batch(oracle, post, res, 5, 0.1, True)
    st, labels, path = dec.beam_search(post[1])
    assert st == fctc_b200.INCOMPARABLE_VALUES and len(labels) == 0
    assert fctc_b200.lib().fctc_status_message(st) == b"Failed to compare values (NaNs in input?)"
    # argument checks of src/lib.rs:339-349
    for kwargs in [dict(beam_size=0), dict(beam_cut_threshold=-0.1), dict(beam_cut_threshold=0.2),
                   dict(beam_cut_threshold=1.1)]:
        with pytest.raises(fctc_b200.FctcError) as e:
            dec.beam_search(post[0], **kwargs)
        assert e.value.code == fctc_b200.E_ARGUMENT
    # empty input decodes to the empty sequence (src/search.rs:178: the loop does not run)
    st, labels, path = dec.beam_search(np.zeros((0, 5), np.float32))
    assert st == 0 and len(labels) == 0 and len(path) == 0


def test_strided_single_read(dec, oracle):
    rng = np.random.default_rng(4)
    big = softmax_posteriors(rng, (400, 10))
    view = big[::2, ::2]
    st, labels, path = dec.beam_search(view, 5, 0.0)
    ost, ol, op = oracle.beam_search(view, 5, 0.0)
    assert st == ost == 0 and np.array_equal(labels, ol) and np.array_equal(path, op)


def test_node_pool_overflow_is_retried(dec, oracle):
    """Flat posteriors with threshold 0 create (C-1)*beam proposals per row — more than the default
    pool; the host entry point must retry those reads with the hard bound and still agree."""
    rng = np.random.default_rng(5)
    post = np.full((3, 300, 5), 0.2, np.float32) + rng.random((3, 300, 5), dtype=np.float32) * 1e-3
    res = dec.beam_search_batch(post, 32, 0.0, False)
    assert (res["status"] == 0).all()
    check_batch(oracle, post, res, 32, 0.0, False)


def test_wide_beam_chunks_share_the_trie_pools(dec, oracle):
    """More reads than one chunk of the host pipeline at beam 32: the chunks' kernels are in flight together and take
    their trie pools from the ring of free slot numbers (csrc/kernels.h, BeamParams::slot_ring).  Every read must
    equal the oracle, and the result pool must hold the reads without overlap."""
    rng = np.random.default_rng(4242)
    N, T, Cn, beam = 60_000, 64, 5, 32
    post = peaked_posteriors(rng, (N, T, Cn))
    res = dec.beam_search_batch(post, beam, 0.0, True)
    assert (res["status"] == 0).all()
    order = np.argsort(res["offsets"], kind="stable")
    off, ln = res["offsets"][order], res["lens"][order].astype(np.int64)
    assert (off[1:] >= off[:-1] + (ln[:-1] + 31) // 32 * 32).all()
    check_batch(oracle, post, res, beam, 0.0, True)
    # and the same call again on the same handle: the ring is set up anew for every call
    res2 = dec.beam_search_batch(post[:55_000], beam, 0.0, True)
    check_batch(oracle, post, res2, beam, 0.0, True, reads=range(0, 55_000, 11))


def test_device_batch_synthetic(dec, oracle):
    """Device-resident API on generator output; sample of reads checked against the oracle."""
    import torch
    dev = torch.device("cuda:0")
    for (N, T, beam, thr, kind) in [(2048, 1000, 5, 0.1, 0), (1024, 1000, 32, 0.0, 0), (1024, 1000, 5, 0.1, 1)]:
        Cn = 5
        post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream().cuda_stream
        dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, kind, seed=0xC7C00002, first_read=0, stream=stream)
        cap = N * ((T + 31) // 32 * 32)
        labels = torch.empty(cap, dtype=torch.uint8, device=dev)
        path = torch.empty(cap, dtype=torch.int32, device=dev)
        offsets = torch.empty(N, dtype=torch.int64, device=dev)
        lens = torch.empty(N, dtype=torch.int32, device=dev)
        status = torch.empty(N, dtype=torch.int32, device=dev)
        used = torch.zeros(1, dtype=torch.int64, device=dev)
        nbytes = dec.beam_scratch_bytes(N, T, Cn, beam, 0)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        assert scratch.data_ptr() % 256 == 0
        dec.beam_search_batch_device(post.data_ptr(), N, T, Cn, beam, thr, True, labels.data_ptr(), path.data_ptr(),
                                     cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                     scratch.data_ptr(), nbytes, 0, stream)
        torch.cuda.synchronize()
        res = dict(labels=labels.cpu().numpy(), path=path.cpu().numpy(), offsets=offsets.cpu().numpy(),
                   lens=lens.cpu().numpy(), status=status.cpu().numpy())
        assert (res["status"] == 0).all()
        check_batch(oracle, post.cpu().numpy(), res, beam, thr, True, reads=range(0, N, 37))
        # invariants over all reads: paths strictly increasing and inside the read, labels non-blank
        lens_np = res["lens"].astype(np.int64)
        for n in range(0, N, 5):
            o, ln = int(res["offsets"][n]), int(lens_np[n])
            pth = res["path"][o:o + ln]
            assert (np.diff(pth) > 0).all() and (ln == 0 or (pth[0] >= 0 and pth[-1] < T))
            assert (res["labels"][o:o + ln] >= 1).all() and (res["labels"][o:o + ln] < Cn).all()


# ---- binary16 posteriors: decoded as the exactly widened float32 values --------------------------------
@pytest.mark.parametrize("case", [(12, 500, 5, 5, 0.1, True), (6, 400, 5, 32, 0.0, True), (6, 300, 3, 10, 0.02, False),
                                  (4, 300, 12, 5, 0.0, True)],
                         ids=lambda c: "N{}_T{}_C{}_b{}_thr{}_c{}".format(*c))
def test_f16_posteriors_host_batch(dec, oracle, case):
    N, T, Cn, beam, thr, collapse = case
    rng = np.random.default_rng(4242 + T + Cn + beam)
    half = softmax_posteriors(rng, (N, T, Cn)).astype(np.float16)
    res = dec.beam_search_batch(half, beam, thr, collapse)
    ol, op, olen, ost, _ = oracle.beam_search_batch(half.astype(np.float32), beam, thr, collapse)
    for n in range(N):
        off, ln = int(res["offsets"][n]), int(res["lens"][n])
        assert res["status"][n] == ost[n]
        if ost[n] == 0:
            assert ln == olen[n]
            assert np.array_equal(res["labels"][off:off + ln], ol[n, :ln]) and np.array_equal(res["path"][off:off + ln], op[n, :ln])
