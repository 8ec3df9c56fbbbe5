#!/usr/bin/env python3
"""Timing of the device FASTQ assembly after a Viterbi decode (development aid)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import torch  # noqa: E402

import fctc_b200  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
T = 4000
dev = torch.device("cuda:0")
dec = fctc_b200.Decoder(0)
st = torch.cuda.current_stream().cuda_stream
post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
dec.synth_posteriors_device(post.data_ptr(), N, T, 5, 0, 1, 0, st)
stride = (T + 31) // 32 * 32
cap = N * stride
labels = torch.empty(cap, dtype=torch.uint8, device=dev)
quals = torch.empty(cap, dtype=torch.uint8, device=dev)
path = torch.empty(cap, dtype=torch.int32, device=dev)
offsets = torch.empty(N, dtype=torch.int64, device=dev)
lens = torch.empty(N, dtype=torch.int32, device=dev)
status = torch.empty(N, dtype=torch.int32, device=dev)
dec.viterbi_batch_device(post.data_ptr(), N, T, 5, True, 1.0, 0.0, True, labels.data_ptr(), quals.data_ptr(), 1,
                         path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), st)
out_cap = 2 * cap + N * 40
out = torch.empty(out_cap, dtype=torch.uint8, device=dev)
rec = torch.empty(N + 1, dtype=torch.int64, device=dev)
nb = dec.fastq_scratch_bytes(N)
scratch = torch.empty(nb, dtype=torch.uint8, device=dev)
ts = []
for _ in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dec.fastq_records_device(labels.data_ptr(), quals.data_ptr(), offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), N,
                             "NACGT", "read_", 0, out.data_ptr(), out_cap, rec.data_ptr(), scratch.data_ptr(), nb, st)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[2]
total = int(rec[-1])
print(f"fastq N={N}: {ms:.3f} ms, {total/1e6:.1f} MB of text, {2*total/ms/1e6:.0f} GB/s read+written, {int(lens.sum())/ms/1e6:.1f} G labels/s")
