#!/usr/bin/env python3
"""Device-resident timing of the Viterbi kernel on binary16 posteriors (development aid)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import torch  # noqa: E402

import fctc_b200  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
T = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
dev = torch.device("cuda:0")
dec = fctc_b200.Decoder(0)
st = torch.cuda.current_stream().cuda_stream
post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
dec.synth_posteriors_device(post.data_ptr(), N, T, 5, kind, 0xC7C00001, 0, st)
half = post.half()
del post
cap = N * ((T + 31) // 32 * 32)
labels = torch.empty(cap, dtype=torch.uint8, device=dev)
quals = torch.empty(cap, dtype=torch.uint8, device=dev)
path = torch.empty(cap, dtype=torch.int32, device=dev)
offsets = torch.empty(N, dtype=torch.int64, device=dev)
lens = torch.empty(N, dtype=torch.int32, device=dev)
status = torch.empty(N, dtype=torch.int32, device=dev)
for q in (1, 0):
    ts = []
    for _ in range(8):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dec.viterbi_batch_device(half.data_ptr(), N, T, 5, q, 1.0, 0.0, True, labels.data_ptr(), quals.data_ptr(), 1,
                                 path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), st, f16=True)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts[3:])[2]
    em = int(lens.sum().item())
    byt = N * T * 10 + em * (5 + q) + N * 16
    print(f"f16 N={N} T={T} kind={kind} qstring={q}: {ms:.3f} ms rows/s={N*T/ms*1e3:.3e} in+out={byt/ms/1e6:.0f} GB/s")
