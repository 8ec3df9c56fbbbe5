"""GPU test of fctc_beam_search_batch_device_static, the beam-search entry point for callers that
cannot allocate device memory at call time (HBM already handed to their own arena, or allocator
calls not allowed on the calling thread).  It keeps its trie scratch in a fixed 64 MiB array inside
the library instead of taking caller scratch.

Checked here, through the C ABI: two calls in a row (different shapes, so the second one runs on
top of whatever the first left in the shared array) each give, read by read, exactly the labels,
paths, lengths and statuses of the regular entry point fctc_beam_search_batch_device — and of the
CPU oracle on a sample of the reads; the amount of free device memory is the same before and after
each call (once the first call on the device has looked the array up); reads whose trie outgrows a
deliberately small pool are flagged the same way by both entry points; and a shape whose trie pools
cannot fit the array is refused with a capacity error instead of running.
This file sorts last on purpose: a run that stops at the first failure has then run everything else."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _decode(dec, torch, post, beam, thr, static, nodes_per_row=0, check_free=False):
    dev = post.device
    N, T, Cn = post.shape
    cap = N * ((T + 31) // 32 * 32)
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]  # the driver's figure: sees every allocation on the device
        dec.beam_search_batch_device_static(post.data_ptr(), N, T, Cn, beam, thr, True, labels.data_ptr(),
                                            path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(),
                                            status.data_ptr(), used.data_ptr(), nodes_per_row, stream)
        torch.cuda.synchronize()
        if check_free:
            assert torch.cuda.mem_get_info(dev)[0] == free_before, "the call changed the amount of free device memory"
    else:
        nbytes = dec.beam_scratch_bytes(N, T, Cn, beam, nodes_per_row)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        dec.beam_search_batch_device(post.data_ptr(), N, T, Cn, beam, thr, True, labels.data_ptr(), path.data_ptr(),
                                     cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                     scratch.data_ptr(), nbytes, nodes_per_row, stream)
    torch.cuda.synchronize()
    lab, pth = labels.cpu().numpy(), path.cpu().numpy()
    off, ln, st = offsets.cpu().numpy(), lens.cpu().numpy(), status.cpu().numpy()
    return [(int(st[n]), lab[off[n]:off[n] + ln[n]].tolist(), pth[off[n]:off[n] + ln[n]].tolist()) for n in range(N)]


def test_static_scratch_entry_point_matches_regular_one(oracle):
    import torch

    import fctc_b200
    dec = fctc_b200.Decoder(0)
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    try:
        # the first call on a device looks the array's address up (and makes the driver load the module that
        # holds it): from the second call on, free device memory must not move
        warm = torch.empty((8, 64, 5), dtype=torch.float32, device=dev)
        dec.synth_posteriors_device(warm.data_ptr(), 8, 64, 5, 0, 1, 0, stream)
        _decode(dec, torch, warm, 5, 0.1, static=True)
        # calls in a row on the shared array: a wide beam first, then narrower shapes on top of its leftovers
        for seed, (N, T, beam, thr) in enumerate([(96, 600, 32, 0.0), (300, 1000, 5, 0.1), (64, 500, 8, 0.0)]):
            post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
            dec.synth_posteriors_device(post.data_ptr(), N, T, 5, 0, 0xC7C00010 + seed, 0, stream)
            want = _decode(dec, torch, post, beam, thr, static=False)
            got = _decode(dec, torch, post, beam, thr, static=True, check_free=True)
            assert all(s == 0 for s, _, _ in want)
            assert got == want  # integer outputs: exact
            assert sum(len(l) for _, l, _ in got) > 0.3 * N * T
            # ... and the CPU oracle itself on a sample of the reads
            pick = list(range(0, N, max(1, N // 12)))
            hp = post[pick].cpu().numpy()
            ol, op, olen, ost, _ = oracle.beam_search_batch(hp, beam, thr, True)
            for j, n in enumerate(pick):
                assert got[n] == (int(ost[j]), ol[j, :olen[j]].tolist(), op[j, :olen[j]].tolist())
        # a pool of one proposal per row is too small for beam 5: both entry points flag the same reads
        want = _decode(dec, torch, post[:, :400].contiguous(), 5, 0.0, static=False, nodes_per_row=1)
        got = _decode(dec, torch, post[:, :400].contiguous(), 5, 0.0, static=True, nodes_per_row=1, check_free=True)
        assert got == want and any(s == fctc_b200.NODE_POOL_OVERFLOW for s, _, _ in got)
        # beyond the built-in array: 4 pools of 96 proposals x 200,000 rows do not fit 64 MiB
        big = torch.zeros((1, 200_000, 5), dtype=torch.float32, device=dev)
        with pytest.raises(fctc_b200.FctcError) as e:
            _decode(dec, torch, big, 32, 0.0, static=True)
        assert e.value.code == fctc_b200.E_CAPACITY and "static scratch too small" in str(e.value)
    finally:
        dec.close()
