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


# ---- envelope from a banded alignment of the label strings (fctc_duplex_envelope_aligned_device) ---------------------
def host_align(a, b, band):
    """Restatement of the alignment rule in align.cu: unit-cost edit distance inside the band j in [i*L2//L1 - band,
    i*L2//L1 + band], ties diagonal > up > left; partner[k] for every base k of a."""
    L1, L2 = len(a), len(b)
    INF = 1 << 29
    D = [dict() for _ in range(L1 + 1)]
    mv = [dict() for _ in range(L1 + 1)]
    for j in range(0, min(L2, band) + 1):
        D[0][j] = j
    for i in range(1, L1 + 1):
        c = i * L2 // L1
        for j in range(max(0, c - band), min(L2, c + band) + 1):
            dg = D[i - 1].get(j - 1, INF) + (a[i - 1] != b[j - 1]) if j >= 1 else INF
            up = D[i - 1].get(j, INF) + 1
            best, m = (dg, 0) if dg <= up else (up, 1)
            left = D[i].get(j - 1, INF) + 1
            if left < best:
                best, m = left, 2
            if best < INF:
                D[i][j], mv[i][j] = best, m
    partner = [0] * L1
    if L2 not in D[L1]:
        return [k * L2 // L1 for k in range(L1)]
    i, j = L1, L2
    while i > 0:
        m = mv[i][j]
        if m == 0:
            partner[i - 1] = j - 1
            i, j = i - 1, j - 1
        elif m == 1:
            partner[i - 1] = max(j - 1, 0)
            i -= 1
        else:
            j -= 1
    return partner


def host_envelope_aligned(path1, off1, len1, path2, off2, len2, partners, T1, T2, pad):
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
            pj = np.clip(np.asarray(partners[n], np.int64), 0, L2 - 1)
            anchor = b[pj[np.maximum(k - 1, 0)]]
            c = np.where(k == 0, np.minimum(prop, b[0]), anchor).astype(np.int64)
        hi = np.minimum(T2, c + pad + 1)
        lo = np.maximum(0, c - pad)
        lo[1:] = np.minimum(lo[1:], hi[:-1])
        env[n, :, 0], env[n, :, 1] = lo, hi
    return env


def indel_pairs(rng, N, n_bases, T, C=5, p_indel=0.08):
    """Two renderings of one base sequence; the second has lost and gained bases in stretches (indels), both with
    random dwells.  Rows: dominant label 1.0, the rest small noise, normalised."""
    def render(seq):
        dw = rng.integers(1, 4, len(seq))
        rows = []
        for s, d in zip(seq, dw):
            rows += [s] + [0] * (d - 1)       # emission row, then blanks
        rows = (rows + [0] * T)[:T]
        x = (0.03 * rng.random((T, C), dtype=np.float32)).astype(np.float32)
        x[np.arange(T), rows] = 1.0
        return (x / x.sum(axis=1, keepdims=True, dtype=np.float32)).astype(np.float32)
    p1, p2 = np.zeros((N, T, C), np.float32), np.zeros((N, T, C), np.float32)
    for n in range(N):
        seq = rng.integers(1, C, n_bases).tolist()
        seq = [s for i, s in enumerate(seq) if i == 0 or s != seq[i - 1]]  # no repeats: Viterbi would collapse them
        other = []
        for s in seq:
            r = rng.random()
            if r < p_indel / 2:
                continue                       # deletion
            other.append(s)
            if r > 1 - p_indel / 2:
                other.append(int(rng.integers(1, C)))  # insertion
        other = [s for i, s in enumerate(other) if i == 0 or s != other[i - 1]]
        p1[n], p2[n] = render(seq), render(other)
    return p1, p2


@pytest.mark.parametrize("N,n_bases,T,band,pad", [(12, 120, 320, 24, 10), (6, 90, 256, 8, 6), (5, 60, 200, 127, 4)])
def test_aligned_envelope_matches_host_rule_and_decodes(dec, oracle, N, n_bases, T, band, pad):
    import torch
    rng = np.random.default_rng(7 * N + n_bases + band)
    p1, p2 = indel_pairs(rng, N, n_bases, T)
    p2[1, :, :] = 0.0
    p2[1, :, 0] = 1.0   # a pair whose second read has no bases: diagonal band
    d1, v1 = viterbi_on_device(dec, p1)
    d2, v2 = viterbi_on_device(dec, p2)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    stride = (T + 31) // 32 * 32
    env = torch.empty((N, T, 2), dtype=torch.int64, device=dev)
    widest = torch.zeros(1, dtype=torch.int32, device=dev)
    partner = torch.full((N * stride,), -1, dtype=torch.int32, device=dev)
    nb = dec.align_scratch_bytes(N, T, band)
    scratch = torch.empty(max(nb, 256), dtype=torch.uint8, device=dev)
    dec.duplex_envelope_aligned_device(v1["labels"].data_ptr(), v1["path"].data_ptr(), v1["offsets"].data_ptr(),
                                       v1["lens"].data_ptr(), v2["labels"].data_ptr(), v2["path"].data_ptr(),
                                       v2["offsets"].data_ptr(), v2["lens"].data_ptr(), N, T, T, band, pad, env.data_ptr(),
                                       widest.data_ptr(), partner.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    torch.cuda.synchronize()
    h = {k: (v1[k].cpu().numpy(), v2[k].cpu().numpy()) for k in ("labels", "path", "offsets", "lens")}
    got_partner = partner.cpu().numpy()
    partners = []
    for n in range(N):
        o1, l1, o2, l2 = int(h["offsets"][0][n]), int(h["lens"][0][n]), int(h["offsets"][1][n]), int(h["lens"][1][n])
        a, b = h["labels"][0][o1:o1 + l1].tolist(), h["labels"][1][o2:o2 + l2].tolist()
        want = host_align(a, b, band) if l1 and l2 else []
        if l1 and l2:
            assert got_partner[o1:o1 + l1].tolist() == want, f"pair {n}: alignment differs"
        partners.append(want)
    got = env.cpu().numpy()
    want_env = host_envelope_aligned(h["path"][0], h["offsets"][0], h["lens"][0], h["path"][1], h["offsets"][1], h["lens"][1],
                                     partners, T, T, pad)
    assert np.array_equal(got, want_env)
    # the reference's validity rule (src/duplex.rs:484-488)
    assert (got[:, :, 0] < got[:, :, 1]).all() and (got[:, 1:, 0] <= got[:, :-1, 1]).all()
    assert (got[:, :, 0] >= 0).all() and (got[:, :, 1] <= T).all()
    assert int(widest.item()) == int((got[:, :, 1] - got[:, :, 0]).max())
    # usable: the device search inside it equals the oracle inside it
    res = dec.beam_search_duplex_batch(p1, p2, got, 5, 0.0, True, exp_mode=0)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, got.astype(np.uint64), 5, 0.0, True, exp_mode=0)
    assert res["status"].tolist() == ost.tolist()
    for n in range(N):
        o, ln = int(res["offsets"][n]), int(res["lens"][n])
        assert res["labels"][o:o + ln].tolist() == ol[n, :olen[n]].tolist()


def test_aligned_envelope_is_tighter_than_the_proportional_one(dec):
    """On pairs with indels the proportional anchor drifts; windows widen where the anchors jump.  The aligned anchor
    follows the pair: its windows stay close to 2*pad+1 rows plus the dwell jitter."""
    import torch
    rng = np.random.default_rng(2024)
    N, T, pad, band = 16, 1600, 8, 48
    p1, p2 = indel_pairs(rng, N, 700, T, p_indel=0.12)
    d1, v1 = viterbi_on_device(dec, p1)
    d2, v2 = viterbi_on_device(dec, p2)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    stride = (T + 31) // 32 * 32
    env_a = torch.empty((N, T, 2), dtype=torch.int64, device=dev)
    env_p = torch.empty((N, T, 2), dtype=torch.int64, device=dev)
    partner = torch.empty((N * stride,), dtype=torch.int32, device=dev)
    scratch = torch.empty(dec.align_scratch_bytes(N, T, band) // 3, dtype=torch.uint8, device=dev)  # a third: pairs in turns
    dec.duplex_envelope_aligned_device(v1["labels"].data_ptr(), v1["path"].data_ptr(), v1["offsets"].data_ptr(),
                                       v1["lens"].data_ptr(), v2["labels"].data_ptr(), v2["path"].data_ptr(),
                                       v2["offsets"].data_ptr(), v2["lens"].data_ptr(), N, T, T, band, pad, env_a.data_ptr(), 0,
                                       partner.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    dec.duplex_envelope_device(v1["path"].data_ptr(), v1["offsets"].data_ptr(), v1["lens"].data_ptr(), v2["path"].data_ptr(),
                               v2["offsets"].data_ptr(), v2["lens"].data_ptr(), N, T, T, pad, env_p.data_ptr(), 0, st)
    torch.cuda.synchronize()
    a, p = env_a.cpu().numpy(), env_p.cpu().numpy()
    # how far the window centre is from the true partner row: compare the two anchors through the paths themselves
    path1, path2 = v1["path"].cpu().numpy(), v2["path"].cpu().numpy()
    off1, len1, off2, len2 = (v1["offsets"].cpu().numpy(), v1["lens"].cpu().numpy(), v2["offsets"].cpu().numpy(),
                              v2["lens"].cpu().numpy())
    drift_a = drift_p = 0.0
    part = partner.cpu().numpy()
    for n in range(N):
        L1, L2 = int(len1[n]), int(len2[n])
        k = np.arange(L1)
        ja = np.clip(part[off1[n]:off1[n] + L1], 0, L2 - 1)
        jp = k * L2 // L1
        drift_a += np.abs(ja - jp).mean()      # bases the proportional anchor is off by, on average
    assert drift_a / N > 2.0, "the test pairs should drift by several bases"
    assert (a[:, :, 1] - a[:, :, 0]).max() <= (p[:, :, 1] - p[:, :, 0]).max() + 2 * 3


def test_move_table_and_dwells(dec):
    import torch
    rng = np.random.default_rng(12)
    N, T = 9, 333
    p1, _ = duplex_pairs(rng, N, T, T, 5)
    p1[2, :, :] = 0.0
    p1[2, :, 0] = 1.0
    d1, v1 = viterbi_on_device(dec, p1)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    moves = torch.full((N, T), 7, dtype=torch.uint8, device=dev)
    dwell = torch.full_like(v1["path"], -1)
    dec.move_table_device(v1["path"].data_ptr(), v1["offsets"].data_ptr(), v1["lens"].data_ptr(), v1["status"].data_ptr(),
                          N, T, moves.data_ptr(), dwell.data_ptr(), st)
    torch.cuda.synchronize()
    path, off, ln = v1["path"].cpu().numpy(), v1["offsets"].cpu().numpy(), v1["lens"].cpu().numpy()
    mv, dw = moves.cpu().numpy(), dwell.cpu().numpy()
    for n in range(N):
        rows = path[off[n]:off[n] + ln[n]]
        want = np.zeros(T, np.uint8)
        want[rows] = 1
        assert np.array_equal(mv[n], want)
        assert dw[off[n]:off[n] + ln[n]].tolist() == np.diff(np.append(rows, T)).tolist()
    assert ln[2] == 0 and not mv[2].any()
