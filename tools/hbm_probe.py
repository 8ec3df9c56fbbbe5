#!/usr/bin/env python3
"""What this box's HBM gives to plain library kernels, for comparison with the Viterbi kernel's figure in bench.py: a device
copy (read + write bytes, as MEASURED_PEAKS.json counts it) and a read-only pass (torch.sum), each as the best single launch
and as 40 launches back to back (sustained, under the box's power cap), with the SM clock and power sampled during the run."""
import json
import subprocess
import sys
import torch

def smi():
    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-i", "0"],
                         capture_output=True, text=True).stdout.strip()
    return out

def timed(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

def main():
    n = 5 * 2**30  # 20 GiB of float32, the size of the Viterbi input
    a = torch.empty(n, dtype=torch.float32, device="cuda").normal_()
    b = torch.empty(n // 2, dtype=torch.float32, device="cuda")
    res = {}
    for name, fn, nbytes in (("copy_10GiB", lambda: b.copy_(a[: n // 2]), 2 * 4 * (n // 2)), ("sum_20GiB", lambda: a.sum(), 4 * n)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        best = min(timed(fn, 1) for _ in range(10))
        sustained = timed(fn, 40)
        res[name] = {"best_gbs": nbytes / best / 1e6, "sustained_gbs": nbytes / sustained / 1e6, "smi_after": smi()}
    print(json.dumps(res))

if __name__ == "__main__":
    main()
