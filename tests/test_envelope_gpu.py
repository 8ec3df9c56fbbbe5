"""Duplex envelope built on the device from the two reads' Viterbi paths:
equal to a host restatement of the same rule, valid by the reference's envelope check, and usable —
the duplex search run inside it matches the oracle run inside the same envelope."""
import numpy as np
import pytest

from helpers import duplex_pairs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def host_envelope(path1, off1, len1, path2, off2, len2, T1, T2, pad):
    N = len(len1)
    env = np.zeros((N, T1, 2), np.int64)
    for n in range(N):
        L1, L2 = int(len1[n]), int(len2[n])
        a = path1[off1[n]:off1[n] + L1]
        b = path2[off2[n]:off2[n] + L2]
        t = np.arange(T1, dtype=np.int64)
        prop = (t * T2) // T1
        if L1 <= 0 or L2 <= 0:
            c = np.minimum(prop, T2 - 1)
        else:
            k = np.searchsorted(a, t, side="right")
            anchor = b[np.maximum(k - 1, 0).astype(np.int64) * L2 // L1]
            c = np.where(k == 0, np.minimum(prop, b[0]), anchor).astype(np.int64)
        hi = np.minimum(T2, c + pad + 1)
        lo = np.maximum(0, c - pad)
        lo[1:] = np.minimum(lo[1:], hi[:-1])
        env[n, :, 0], env[n, :, 1] = lo, hi
    return env


def viterbi_on_device(dec, post_np):
    import torch
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    N, T, Cn = post_np.shape
    post = torch.from_numpy(post_np).to(dev)
    stride = (T + 31) // 32 * 32
    cap = N * stride
    out = dict(labels=torch.empty(cap, dtype=torch.uint8, device=dev), path=torch.empty(cap, dtype=torch.int32, device=dev),
               offsets=torch.empty(N, dtype=torch.int64, device=dev), lens=torch.empty(N, dtype=torch.int32, device=dev),
               status=torch.empty(N, dtype=torch.int32, device=dev))
    dec.viterbi_batch_device(post.data_ptr(), N, T, Cn, False, 1.0, 0.0, True, out["labels"].data_ptr(), 0, 1,
                             out["path"].data_ptr(), cap, out["offsets"].data_ptr(), out["lens"].data_ptr(),
                             out["status"].data_ptr(), st)
    return post, out


@pytest.mark.parametrize("N,T1,T2,pad", [(16, 300, 300, 12), (9, 400, 333, 20), (5, 64, 200, 6), (4, 257, 100, 0)])
def test_envelope_matches_host_rule_and_decodes(dec, oracle, N, T1, T2, pad):
    import torch
    rng = np.random.default_rng(31 * N + T1 + T2)
    p1, p2 = duplex_pairs(rng, N, T1, T2, 5)
    p1[0, :, :] = 0.0
    p1[0, :, 0] = 1.0   # an all-blank read: no bases, diagonal band
    d1, v1 = viterbi_on_device(dec, p1)
    d2, v2 = viterbi_on_device(dec, p2)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    env = torch.empty((N, T1, 2), dtype=torch.int64, device=dev)
    widest = torch.full((1,), -7, dtype=torch.int32, device=dev)
    dec.duplex_envelope_device(v1["path"].data_ptr(), v1["offsets"].data_ptr(), v1["lens"].data_ptr(), v2["path"].data_ptr(),
                               v2["offsets"].data_ptr(), v2["lens"].data_ptr(), N, T1, T2, pad, env.data_ptr(),
                               widest.data_ptr(), st)
    torch.cuda.synchronize()
    got = env.cpu().numpy()
    want = host_envelope(v1["path"].cpu().numpy(), v1["offsets"].cpu().numpy(), v1["lens"].cpu().numpy(),
                         v2["path"].cpu().numpy(), v2["offsets"].cpu().numpy(), v2["lens"].cpu().numpy(), T1, T2, pad)
    assert np.array_equal(got, want)
    # the reference's validity rule (src/duplex.rs:484-488): lo < hi and lo <= previous hi
    assert (got[:, :, 0] < got[:, :, 1]).all() and (got[:, 1:, 0] <= got[:, :-1, 1]).all()
    assert (got[:, :, 0] >= 0).all() and (got[:, :, 1] <= T2).all()
    assert int(widest.item()) == int((got[:, :, 1] - got[:, :, 0]).max())
    # and it is usable: the device search inside it equals the oracle inside it (max-plus mode is bit-exact)
    res = dec.beam_search_duplex_batch(p1, p2, got, 5, 0.0, True, exp_mode=1)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, got.astype(np.uint64), 5, 0.0, True, exp_mode=1)
    assert res["status"].tolist() == ost.tolist()
    for n in range(N):
        o, ln = int(res["offsets"][n]), int(res["lens"][n])
        assert res["labels"][o:o + ln].tolist() == ol[n, :olen[n]].tolist()
