"""GPU parity tests of the state-conditioned (crf) decoders through the C ABI and the drop-in module
vs the CPU oracle: the reference's only crf vector (src/search.rs:431-510) plus random k-mer
transition models, bit-exact (labels, path, quality characters)."""
import numpy as np
import pytest

from helpers import load_kats

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def crf_posteriors(rng, T, S, C, peak=0.0):
    x = rng.random((T, S, C), dtype=np.float32)
    if peak:
        x[..., 0] += peak
    x /= x.sum(axis=2, keepdims=True, dtype=np.float32)
    return x.astype(np.float32)


@pytest.mark.parametrize("kat", [pytest.param(k, id=k["name"]) for k in load_kats() if k["fn"].startswith("crf_")])
def test_reference_kats(dec, oracle, kat):
    import fast_ctc_decode as fcd
    post = np.asarray(kat["post"], np.float32)
    init = np.asarray(kat["init_state"], np.float32)
    if kat["fn"] == "crf_greedy_search":
        st, labels, quals, path = dec.crf_greedy_search(post, init, kat["qstring"], kat["qscale"], kat["qbias"])
        assert st == 0
        assert oracle.to_text(kat["alphabet"], labels, quals) == kat["expect_seq"]
        assert path.tolist() == kat["expect_path"]
        seq, p = fcd.crf_greedy_search(post, init, kat["alphabet"], kat["qstring"], kat["qscale"], kat["qbias"])
    else:
        st, labels, path = dec.crf_beam_search(post, init, kat["beam_size"], kat["beam_cut_threshold"])
        assert st == 0
        assert oracle.to_text(kat["alphabet"], labels) == kat["expect_seq"]
        assert path.tolist() == kat["expect_path"]
        seq, p = fcd.crf_beam_search(post, init, kat["alphabet"], kat["beam_size"], kat["beam_cut_threshold"])
    assert seq == kat["expect_seq"] and p == kat["expect_path"]


@pytest.mark.parametrize("shape", [(300, 16, 5), (500, 64, 5), (200, 4, 3), (150, 25, 6), (64, 1024, 5)],
                         ids=lambda s: "x".join(map(str, s)))
def test_random_models_match_oracle(dec, oracle, shape):
    T, S, C = shape
    rng = np.random.default_rng(31 * T + S)
    for peak in (0.0, 1.5):
        post = crf_posteriors(rng, T, S, C, peak)
        init = rng.random(S, dtype=np.float32)
        st, labels, quals, path = dec.crf_greedy_search(post, init, True, 0.7, 2.0)
        ost, ol, oq, op = oracle.crf_greedy(post, init, True, 0.7, 2.0)
        assert st == ost == 0
        assert np.array_equal(labels, ol) and np.array_equal(quals, oq) and np.array_equal(path, op)
        for beam, thr in ((5, 0.0), (5, 0.05), (32, 0.0), (1, 0.0)):
            st, labels, path = dec.crf_beam_search(post, init, beam, thr)
            ost, ol, op = oracle.crf_beam_search(post, init, beam, thr)
            assert st == ost, (beam, thr)
            if st == 0:
                assert np.array_equal(labels, ol) and np.array_equal(path, op), (beam, thr)


def test_reference_abort_conditions_are_reported(dec, oracle):
    import fctc_b200
    rng = np.random.default_rng(9)
    post = crf_posteriors(rng, 100, 16, 5)
    init = rng.random(16, dtype=np.float32)
    bad = post.copy()
    bad[40] = np.nan
    st, labels, quals, path = dec.crf_greedy_search(bad, init)
    assert st == fctc_b200.REFERENCE_PANIC and oracle.crf_greedy(bad, init)[0] == oracle.PANIC
    # 6 states, 4 bases: (state*4) % 6 + label can leave [0, 6)
    post6 = crf_posteriors(rng, 100, 6, 5)
    init6 = np.array([0, 0, 0, 0, 0, 1], np.float32)
    st, *_ = dec.crf_greedy_search(post6, init6)
    assert (st == fctc_b200.REFERENCE_PANIC) == (oracle.crf_greedy(post6, init6)[0] == oracle.PANIC)
    st, labels, path = dec.crf_beam_search(bad, init, 5, 0.0)
    assert st in (fctc_b200.INCOMPARABLE_VALUES, fctc_b200.REFERENCE_PANIC)
    assert oracle.crf_beam_search(bad, init, 5, 0.0)[0] in (oracle.INCOMPARABLE, oracle.PANIC)


def test_device_batch_greedy(dec, oracle):
    import torch
    rng = np.random.default_rng(12)
    N, T, S, C = 200, 120, 16, 5
    post = np.stack([crf_posteriors(rng, T, S, C, 1.0) for _ in range(N)])
    init = rng.random((N, S), dtype=np.float32)
    dev = torch.device("cuda:0")
    dpost, dinit = torch.from_numpy(post).to(dev), torch.from_numpy(init).to(dev)
    stride = (T + 31) // 32 * 32
    labels = torch.empty(N * stride, dtype=torch.uint8, device=dev)
    quals = torch.empty(N * stride, dtype=torch.uint8, device=dev)
    path = torch.empty(N * stride, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    rc = dec._L.fctc_crf_greedy_batch_device(dec._h, dpost.data_ptr(), dinit.data_ptr(), N, T, S, C, S, 1, 1.0, 0.0,
                                             labels.data_ptr(), quals.data_ptr(), 1, path.data_ptr(), N * stride,
                                             offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                             torch.cuda.current_stream().cuda_stream or None)
    assert rc == 0
    torch.cuda.synchronize()
    lab, q, pth, ln = labels.cpu().numpy(), quals.cpu().numpy(), path.cpu().numpy(), lens.cpu().numpy()
    assert (status.cpu().numpy() == 0).all()
    for n in range(N):
        ost, ol, oq, op = oracle.crf_greedy(post[n], init[n], True)
        o = n * stride
        assert ln[n] == len(ol) and np.array_equal(lab[o:o + ln[n]], ol) and np.array_equal(q[o:o + ln[n]], oq)
        assert np.array_equal(pth[o:o + ln[n]], op)


def test_crf_duplex_matches_oracle(dec, oracle):
    """duplex::crf_beam_search (src/duplex.rs:652-834).  The reference has no vector for it (parity
    unpinned); compared with this repository's restatement: exactly in the max-plus mode (no
    exp/log1p), by tolerance in the exact-exp mode."""
    import fast_ctc_decode as fcd
    rng = np.random.default_rng(21)
    for (T1, T2, S, C, beam, w) in [(80, 80, 16, 5, 5, 8), (60, 70, 4, 3, 5, None), (100, 100, 64, 5, 8, 6)]:
        hid = np.where(rng.random(T1) < 0.6, 0, rng.integers(1, C, T1))

        def render(T):
            idx = np.minimum((np.arange(T) * T1) // T, T1 - 1)
            x = (0.1 * rng.random((T, S, C), dtype=np.float32)).astype(np.float32)
            x[np.arange(T), :, hid[idx]] += 1.0
            x /= x.sum(axis=2, keepdims=True, dtype=np.float32)
            return x.astype(np.float32)
        p1, p2 = render(T1), render(T2)
        i1, i2 = rng.random(S, dtype=np.float32), rng.random(S, dtype=np.float32)
        env = None
        if w is not None:
            c = (np.arange(T1) * T2) // T1
            env = np.stack([np.maximum(0, c - w), np.minimum(T2, c + w + 1)], axis=1).astype(np.uint64)
        st, labels = dec.crf_beam_search_duplex(p1, i1, p2, i2, env, beam, 0.0, exp_mode=1)
        ost, ol = oracle.crf_beam_search_duplex(p1, i1, p2, i2, env, beam, 0.0, exp_mode=1)
        assert st == ost == 0 and labels.tolist() == ol.tolist()
        st, labels = dec.crf_beam_search_duplex(p1, i1, p2, i2, env, beam, 0.0, exp_mode=0)
        ost, ol = oracle.crf_beam_search_duplex(p1, i1, p2, i2, env, beam, 0.0, exp_mode=0)
        assert st == ost == 0 and labels.tolist() == ol.tolist()  # glibc's expf / log1pf on the device: identical
        seq = fcd.crf_beam_search_duplex(p1, i1, p2, i2, "NACGT"[:C], envelope=env, beam_size=beam)
        assert isinstance(seq, str) and abs(len(seq) - len(ol)) <= max(2, len(ol) // 10)


def _encode16(x, kind):
    """The device array for an input kind and the float32 values it stands for."""
    import fctc_b200
    if kind == fctc_b200.IN_F32:
        return x, x
    if kind == fctc_b200.IN_F16:
        h = x.astype(np.float16)
        return h.view(np.uint16), h.astype(np.float32)
    b = fctc_b200.to_bf16_bits(x)
    return b, fctc_b200.from_bf16_bits(b)


@pytest.mark.parametrize("kind", [0, 1, 2], ids=["f32", "f16", "bf16"])
def test_crf_searches_over_16_bit_posteriors(dec, oracle, kind):
    """fctc_crf_greedy_batch_device_in / fctc_crf_beam_search_batch_device_in: binary16 and bfloat16 posteriors are
    widened exactly where the searches read them, so labels, paths and quality characters equal the oracle's on the
    widened float32 array (narrow, wide and beyond-32 beams)."""
    import torch
    rng = np.random.default_rng(77 + kind)
    N, T, S, C = 24, 150, 16, 5
    post = np.stack([crf_posteriors(rng, T, S, C, 1.0 if n % 2 else 0.0) for n in range(N)])
    init = rng.random((N, S), dtype=np.float32)
    arr, vals = _encode16(post, kind)
    dev = torch.device("cuda:0")
    dpost, dinit = torch.from_numpy(np.ascontiguousarray(arr).view(np.int16 if kind else np.float32)).to(dev), torch.from_numpy(init).to(dev)
    stride = (T + 31) // 32 * 32
    labels = torch.empty(N * stride, dtype=torch.uint8, device=dev)
    quals = torch.empty(N * stride, dtype=torch.uint8, device=dev)
    path = torch.empty(N * stride, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream or None
    rc = dec._L.fctc_crf_greedy_batch_device_in(dec._h, dpost.data_ptr(), kind, dinit.data_ptr(), N, T, S, C, S, 1, 0.9, 1.0,
                                                labels.data_ptr(), quals.data_ptr(), 1, path.data_ptr(), N * stride,
                                                offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), stream)
    assert rc == 0, dec.last_error()
    torch.cuda.synchronize()
    lab, q, pth, ln = labels.cpu().numpy(), quals.cpu().numpy(), path.cpu().numpy(), lens.cpu().numpy()
    assert (status.cpu().numpy() == 0).all()
    for n in range(N):
        ost, ol, oq, op = oracle.crf_greedy(vals[n], init[n], True, 0.9, 1.0)
        o = n * stride
        assert ln[n] == len(ol) and np.array_equal(lab[o:o + ln[n]], ol) and np.array_equal(q[o:o + ln[n]], oq)
        assert np.array_equal(pth[o:o + ln[n]], op)

    used = torch.zeros(32, dtype=torch.int64, device=dev)
    for beam, thr in ((5, 0.0), (5, 0.02), (32, 0.0), (40, 0.0)):
        nbytes = dec.beam_scratch_bytes(N, T, C, beam, (C - 1) * beam)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        rc = dec._L.fctc_crf_beam_search_batch_device_in(dec._h, dpost.data_ptr(), kind, dinit.data_ptr(), N, T, S, C, S, beam,
                                                         thr, labels.data_ptr(), path.data_ptr(), N * stride,
                                                         offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                                         used.data_ptr(), scratch.data_ptr(), nbytes, (C - 1) * beam, stream)
        assert rc == 0, dec.last_error()
        torch.cuda.synchronize()
        lab, pth, ln, off, st = (labels.cpu().numpy(), path.cpu().numpy(), lens.cpu().numpy(), offsets.cpu().numpy(),
                                 status.cpu().numpy())
        for n in range(N):
            ost, ol, op = oracle.crf_beam_search(vals[n], init[n], beam, thr)
            assert st[n] == ost, (beam, thr, n)
            if ost == 0:
                o = off[n]
                assert lab[o:o + ln[n]].tolist() == list(ol) and pth[o:o + ln[n]].tolist() == list(op), (beam, thr, n)
    # logits are not defined for state-conditioned posteriors
    rc = dec._L.fctc_crf_greedy_batch_device_in(dec._h, dpost.data_ptr(), 0x10, dinit.data_ptr(), N, T, S, C, S, 0, 1.0, 0.0,
                                                labels.data_ptr(), None, 1, path.data_ptr(), N * stride,
                                                offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), stream)
    assert rc < 0 and "input_kind" in dec.last_error()


def test_crf_duplex_over_binary16_posteriors(dec, oracle):
    """fctc_crf_beam_search_duplex_batch_device_in with FCTC_IN_F16: equal to the oracle on the widened arrays
    (exact-exp and max-plus modes)."""
    import torch
    rng = np.random.default_rng(91)
    N, T1, T2, S, C, beam, w = 6, 70, 64, 16, 5, 5, 8
    hid = np.where(rng.random((N, T1)) < 0.6, 0, rng.integers(1, C, (N, T1)))

    def render(T):
        idx = np.minimum((np.arange(T) * T1) // T, T1 - 1)
        x = (0.1 * rng.random((N, T, S, C), dtype=np.float32)).astype(np.float32)
        for n in range(N):
            x[n, np.arange(T), :, hid[n, idx]] += 1.0
        x /= x.sum(axis=3, keepdims=True, dtype=np.float32)
        return x.astype(np.float16)
    h1, h2 = render(T1), render(T2)
    v1, v2 = h1.astype(np.float32), h2.astype(np.float32)
    i1, i2 = rng.random((N, S), dtype=np.float32), rng.random((N, S), dtype=np.float32)
    c = (np.arange(T1) * T2) // T1
    env1 = np.stack([np.maximum(0, c - w), np.minimum(T2, c + w + 1)], axis=1).astype(np.int64)
    env = np.ascontiguousarray(np.broadcast_to(env1, (N, T1, 2)))
    dev = torch.device("cuda:0")
    d1, d2 = torch.from_numpy(h1.view(np.int16)).to(dev), torch.from_numpy(h2.view(np.int16)).to(dev)
    di1, di2, denv = torch.from_numpy(i1).to(dev), torch.from_numpy(i2).to(dev), torch.from_numpy(env).to(dev)
    stride = max((T1 + 31) // 32 * 32, 32)
    labels = torch.empty(N * stride, dtype=torch.uint8, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(32, dtype=torch.int64, device=dev)
    npr = (C - 1) * beam
    nbytes = dec.crf_duplex_scratch_bytes(N, T1, T2, C, beam, 2 * w + 1, npr)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    for mode in (0, 1):
        rc = dec._L.fctc_crf_beam_search_duplex_batch_device_in(
            dec._h, d1.data_ptr(), di1.data_ptr(), S, d2.data_ptr(), di2.data_ptr(), S, 1, N, T1, T2, S, C, denv.data_ptr(),
            2 * w + 1, beam, 0.0, mode, labels.data_ptr(), N * stride, offsets.data_ptr(), lens.data_ptr(),
            status.data_ptr(), used.data_ptr(), scratch.data_ptr(), nbytes, npr,
            torch.cuda.current_stream().cuda_stream or None)
        assert rc == 0, dec.last_error()
        torch.cuda.synchronize()
        lab, ln, off, st = labels.cpu().numpy(), lens.cpu().numpy(), offsets.cpu().numpy(), status.cpu().numpy()
        for n in range(N):
            ost, ol = oracle.crf_beam_search_duplex(v1[n], i1[n], v2[n], i2[n], env1.astype(np.uint64), beam, 0.0, exp_mode=mode)
            assert st[n] == ost == 0 and lab[off[n]:off[n] + ln[n]].tolist() == ol.tolist(), (mode, n)
    rc = dec._L.fctc_crf_beam_search_duplex_batch_device_in(
        dec._h, d1.data_ptr(), di1.data_ptr(), S, d2.data_ptr(), di2.data_ptr(), S, 2, N, T1, T2, S, C, denv.data_ptr(),
        2 * w + 1, beam, 0.0, 0, labels.data_ptr(), N * stride, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
        used.data_ptr(), scratch.data_ptr(), nbytes, npr, None)
    assert rc < 0
