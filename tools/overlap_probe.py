#!/usr/bin/env python3
"""Does an H2D copy overlap a running beam-search kernel?  (development aid)"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import torch  # noqa: E402

import fctc_b200  # noqa: E402

N, T, beam, thr = 14208, 4000, 5, 0.1
dev = torch.device("cuda:0")
dec = fctc_b200.Decoder(0)
post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
post2 = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
dec.synth_posteriors_device(post.data_ptr(), N, T, 5, 0, 0xC7C00001, 0, torch.cuda.current_stream().cuda_stream)
h = torch.empty((N, T, 5), dtype=torch.float32).pin_memory()
h.copy_(post)
stride = (T + 31) // 32 * 32
cap = N * stride
labels = torch.empty(cap, dtype=torch.uint8, device=dev)
path = torch.empty(cap, dtype=torch.int32, device=dev)
offsets = torch.empty(N, dtype=torch.int64, device=dev)
lens = torch.empty(N, dtype=torch.int32, device=dev)
status = torch.empty(N, dtype=torch.int32, device=dev)
used = torch.zeros(1, dtype=torch.int64, device=dev)
nbytes = dec.beam_scratch_bytes(N, T, 5, beam, 0)
scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize()


def kernel(stream):
    used.zero_()
    dec.beam_search_batch_device(post.data_ptr(), N, T, 5, beam, thr, True, labels.data_ptr(), path.data_ptr(), cap,
                                 offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                 scratch.data_ptr(), nbytes, 0, stream.cuda_stream)


def timed(fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


def copy_only():
    with torch.cuda.stream(sb):
        post2.copy_(h, non_blocking=True)


def both():
    kernel(sa)
    with torch.cuda.stream(sb):
        post2.copy_(h, non_blocking=True)


for _ in range(2):
    timed(lambda: kernel(sa)); timed(copy_only); timed(both)
print(f"kernel {timed(lambda: kernel(sa)):.1f} ms, copy {timed(copy_only):.1f} ms, both {timed(both):.1f} ms")
