#!/usr/bin/env python3
"""Host-to-host timing of fctc_beam_search_batch on pinned memory (development aid)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import torch  # noqa: E402

import fctc_b200  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
T = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
beam = int(sys.argv[3]) if len(sys.argv) > 3 else 5
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.1
f16 = len(sys.argv) > 5 and sys.argv[5] == "f16"
dev = torch.device("cuda:0")
dec = fctc_b200.Decoder(0)
post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
dec.synth_posteriors_device(post.data_ptr(), N, T, 5, 0, 0xC7C00001, 0, torch.cuda.current_stream().cuda_stream)
h = torch.empty((N, T, 5), dtype=torch.float16 if f16 else torch.float32).pin_memory()
h.copy_(post)
del post
stride = (T + 31) // 32 * 32
out = dict(labels=torch.empty(N * stride, dtype=torch.uint8).pin_memory().numpy(),
           path=torch.empty(N * stride, dtype=torch.int32).pin_memory().numpy(),
           offsets=torch.empty(N, dtype=torch.int64).pin_memory().numpy(),
           lens=torch.empty(N, dtype=torch.int32).pin_memory().numpy(),
           status=torch.empty(N, dtype=torch.int32).pin_memory().numpy())
hn = h.numpy()
for _ in range(2):
    dec.beam_search_batch(hn, beam, thr, True, out=out)
ts = []
for _ in range(4):
    t0 = time.perf_counter()
    dec.beam_search_batch(hn, beam, thr, True, out=out)
    ts.append(time.perf_counter() - t0)
dt = min(ts)
print(f"e2e {'f16 ' if f16 else ''}beam={beam} N={N} T={T} chunk={os.environ.get('FCTC_BEAM_CHUNK_READS', 'default')}: {dt*1e3:.1f} ms  rows/s={N*T/dt:.3e}  "
      f"h2d {N*T*(10 if f16 else 20)/dt/1e9:.1f} GB/s")
