#!/usr/bin/env python3
"""What the PCIe link gives for the byte counts of one end-to-end Viterbi step (2.62 GB in, 0.79 GB out,
pinned memory, two streams): the ceiling of bench.py's e2e number.  Development aid."""
import torch, time
h = torch.empty(2_621_440_000, dtype=torch.uint8).pin_memory()
d = torch.empty_like(h, device="cuda")
o = torch.empty(786_956_288, dtype=torch.uint8, device="cuda")
ho = torch.empty(786_956_288, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for both in (False, True):
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): ho.copy_(o, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("both" if both else "h2d only", f"{dt*1e3:.1f} ms  h2d {h.numel()/dt/1e9:.1f} GB/s")
