"""GPU tests at BASELINE.json's full sizes.  The oracle cannot decode 4x10^9 rows in a test, so these
check size-independent properties over the WHOLE batch on the device (emission recount from an
independent argmax, ordering and range invariants, status words) plus bit-exact comparison with the
oracle on reads sampled across the batch.  Skipped when the GPU has too little free memory."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEED = 0xC7C00001


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def need_free(gib):
    import torch
    torch.cuda.empty_cache()
    free_b, _ = torch.cuda.mem_get_info(0)
    if free_b < gib * 2 ** 30:
        pytest.skip(f"needs {gib} GiB of free device memory, {free_b / 2 ** 30:.0f} available")


def first_argmax(post):
    """torch restatement of find_max (first maximum)."""
    import torch
    mx = post.max(dim=2, keepdim=True).values
    idx = torch.arange(post.shape[2], device=post.device).expand_as(post)
    return torch.where(post == mx, idx, torch.full_like(idx, post.shape[2])).min(dim=2).values


def test_viterbi_1m_reads(dec, oracle):
    """configs[1]: 1M reads x T=4000 x C=5, qstring with qscale/qbias, collapse_repeats."""
    import torch
    need_free(125)
    N, T, Cn = 1_000_000, 4000, 5
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, 0, SEED, 0, st)
    stride = (T + 31) // 32 * 32
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    quals = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    dec.viterbi_batch_device(post.data_ptr(), N, T, Cn, True, 0.7, 2.0, True, labels.data_ptr(), quals.data_ptr(), 1,
                             path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), st)
    torch.cuda.synchronize()
    assert int((status != 0).sum()) == 0
    assert torch.equal(offsets, torch.arange(N, device=dev, dtype=torch.int64) * stride)
    # emission recount over every read, in chunks (independent of the kernel's argmax / scan)
    chunk = 32768
    for lo in range(0, N, chunk):
        arg = first_argmax(post[lo:lo + chunk])
        prev = torch.cat([torch.full((arg.shape[0], 1), -1, device=dev, dtype=arg.dtype), arg[:, :-1]], dim=1)
        want = ((arg != 0) & (arg != prev)).sum(dim=1).to(torch.int32)
        assert torch.equal(want, lens[lo:lo + chunk]), f"emission counts differ in reads {lo}..{lo + chunk}"
        del arg, prev, want
    # slot contents: labels in range, paths strictly increasing, qualities within phred's range
    lab2 = labels.view(N, stride)
    pth2 = path.view(N, stride)
    q2 = quals.view(N, stride)
    pos = torch.arange(stride, device=dev).unsqueeze(0)
    for lo in range(0, N, 65536):
        sl = slice(lo, lo + 65536)
        valid = pos < lens[sl].unsqueeze(1)
        assert bool(((lab2[sl] >= 1) & (lab2[sl] < Cn))[valid].all())
        inc = (pth2[sl][:, 1:] > pth2[sl][:, :-1]) | ~valid[:, 1:]
        assert bool(inc.all())
        qv = q2[sl][valid]
        assert int(qv.min()) >= 33 and int(qv.max()) <= 33 + 30 + 1   # 0.7 * 40 + 2 -> code <= 63
    # bit-exact against the oracle on reads sampled across the batch
    sample = np.linspace(0, N - 1, 2048).astype(np.int64)
    sub = post[torch.from_numpy(sample).to(dev)].cpu().numpy()
    ol, oq, op, olen, ost = oracle.viterbi_batch(sub, True, 0.7, 2.0, True)
    sidx = torch.from_numpy(sample).to(dev)
    lab_s, pth_s, q_s, len_s = lab2[sidx].cpu().numpy(), pth2[sidx].cpu().numpy(), q2[sidx].cpu().numpy(), lens[sidx].cpu().numpy()
    for j in range(len(sample)):
        ln = int(len_s[j])
        assert ln == olen[j]
        assert np.array_equal(lab_s[j, :ln], ol[j, :ln])
        assert np.array_equal(pth_s[j, :ln], op[j, :ln])
        assert np.array_equal(q_s[j, :ln].astype(np.uint32), oq[j, :ln])


@pytest.mark.parametrize("N,T,beam,thr,gib", [(262_144, 4000, 5, 0.1, 45), (65_536, 10_000, 32, 0.0, 125)],
                         ids=["beam5_262144x4000", "beam32_65536x10000"])
def test_beam_search_full_batches(dec, oracle, N, T, beam, thr, gib):
    """configs[2] and configs[3]."""
    import torch
    need_free(gib)
    Cn = 5
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, 0, SEED + 1, 0, st)
    stride = (T + 31) // 32 * 32
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    free_b, _ = torch.cuda.mem_get_info(0)
    nbytes = min(dec.beam_scratch_bytes(N, T, Cn, beam, 0), int(free_b * 0.9) // 256 * 256)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dec.beam_search_batch_device(post.data_ptr(), N, T, Cn, beam, thr, True, labels.data_ptr(), path.data_ptr(), cap,
                                 offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                 scratch.data_ptr(), nbytes, 0, st)
    torch.cuda.synchronize()
    del scratch
    st_np = status.cpu().numpy()
    overflow = int((st_np == 4).sum())
    assert set(np.unique(st_np)) <= {0, 4} and overflow <= N // 1000   # trie-pool overflows are flagged, rare
    off_np, len_np = offsets.cpu().numpy(), lens.cpu().numpy().astype(np.int64)
    # pool slots are disjoint and inside the used prefix
    order = np.argsort(off_np, kind="stable")
    ends = off_np[order] + (len_np[order] + 31) // 32 * 32
    assert (off_np[order][1:] >= ends[:-1]).all() and ends.max() <= int(used.item())
    # labels valid and paths strictly increasing over the whole pool (checked per read on a stride)
    lab_np, pth_np = labels.cpu().numpy(), path.cpu().numpy()
    for n in range(0, N, 997):
        o, ln = off_np[n], len_np[n]
        assert (lab_np[o:o + ln] >= 1).all() and (lab_np[o:o + ln] < Cn).all()
        assert (np.diff(pth_np[o:o + ln]) > 0).all() and (ln == 0 or (0 <= pth_np[o] and pth_np[o + ln - 1] < T))
    # identical to the oracle on sampled reads
    sample = np.linspace(0, N - 1, 1024 if beam > 8 else 4096).astype(np.int64)
    sub = post[torch.from_numpy(sample).to(dev)].cpu().numpy()
    ol, op, olen, ost, _ = oracle.beam_search_batch(sub, beam, thr, True)
    for j, n in enumerate(sample):
        if st_np[n] != 0:
            continue
        o, ln = off_np[n], len_np[n]
        assert ln == olen[j] and np.array_equal(lab_np[o:o + ln], ol[j, :ln]) and np.array_equal(pth_np[o:o + ln], op[j, :ln])


def test_duplex_full_batch(dec, oracle):
    """configs[4]: 16,384 pairs x T~4000 x C=5, beam 5, band +-32, exact exp."""
    import torch
    need_free(110)
    N, T, Cn, band = 16_384, 4000, 5, 32
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    a = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    b = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    dec.synth_posteriors_device(a.data_ptr(), N, T, Cn, 1, SEED, 0, st)
    dec.synth_posteriors_device(b.data_ptr(), N, T, Cn, 2, SEED, 0, st)
    c = np.arange(T)
    env1 = np.stack([np.maximum(0, c - band), np.minimum(T, c + band + 1)], axis=1).astype(np.int64)
    env = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(env1, (N, T, 2)))).to(dev)
    stride = (T + 31) // 32 * 32
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    mw = 2 * band + 1
    free_b, _ = torch.cuda.mem_get_info(0)
    nbytes = min(dec.duplex_scratch_bytes(N, T, T, Cn, 5, mw, 0), int(free_b * 0.9) // 256 * 256)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dec.beam_search_duplex_batch_device(a.data_ptr(), b.data_ptr(), N, T, T, Cn, env.data_ptr(), mw, 5, 0.0, True, 0,
                                        labels.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                        used.data_ptr(), scratch.data_ptr(), nbytes, 0, st)
    torch.cuda.synchronize()
    del scratch
    st_np = status.cpu().numpy()
    assert set(np.unique(st_np)) <= {0, 4} and int((st_np == 4).sum()) <= N // 100
    off_np, len_np, lab_np = offsets.cpu().numpy(), lens.cpu().numpy(), labels.cpu().numpy()
    assert (len_np[st_np == 0] > 0).all()
    sample = np.linspace(0, N - 1, 256).astype(np.int64)
    idx = torch.from_numpy(sample).to(dev)
    ol, olen, ost = oracle.beam_search_duplex_batch(a[idx].cpu().numpy(), b[idx].cpu().numpy(),
                                                    np.ascontiguousarray(np.broadcast_to(env1, (len(sample), T, 2))).astype(np.uint64),
                                                    5, 0.0, True, 0)
    same = 0
    for j, n in enumerate(sample):
        if st_np[n] != 0:
            same += 1
            continue
        o, ln = off_np[n], len_np[n]
        same += int(ln == olen[j] and np.array_equal(lab_np[o:o + ln], ol[j, :ln]))
    assert same == len(sample)   # exact-exp mode evaluates glibc's expf / log1pf: identical, not merely close
    print(f"duplex full batch: {same}/{len(sample)} sampled pairs identical to the oracle")
