#!/usr/bin/env python3
"""Per-read start time, duration and SM of the wide beam kernel (development probe).  Needs a library built with
-DFCTC_BEAM_TIMING:  FCTC_BUILD_OUT=/tmp/lib_timing.so FCTC_NVCC_EXTRA=-DFCTC_BEAM_TIMING python fast-ctc-decode_b200/build.py --force;
FCTC_LIB=/tmp/lib_timing.so python tools/beam_timing_probe.py 4736 10000 [beam [nodes_per_row]].  Round 2 findings: read
durations spread from 0.75x to 1.7x the median, so a launch of a single round of reads takes about twice a steady-state
round; and at T = 10,000 a flat 96 GiB cap on the recommended scratch held 4,292 of the 4,736 pools the resident warps want
(444 reads of the first round started late) — the cap now follows the device's memory."""
import os, sys
ROOT = os.getcwd()
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import numpy as np, torch
import fctc_b200
N, T = int(sys.argv[1]), int(sys.argv[2])
BEAM = int(sys.argv[3]) if len(sys.argv) > 3 else 32
NPR = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dev = torch.device("cuda:0"); dec = fctc_b200.Decoder(0); st = torch.cuda.current_stream().cuda_stream
post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
dec.synth_posteriors_device(post.data_ptr(), N, T, 5, 0, 0xC7C00001, 0, st)
cap = N * ((T + 31) // 32 * 32)
labels = torch.empty(cap, dtype=torch.uint8, device=dev); path = torch.empty(cap, dtype=torch.int32, device=dev)
offsets = torch.empty(N, dtype=torch.int64, device=dev); lens = torch.empty(N, dtype=torch.int32, device=dev)
status = torch.empty(N, dtype=torch.int32, device=dev); used = torch.zeros(1, dtype=torch.int64, device=dev)
nb = dec.beam_scratch_bytes(N, T, 5, BEAM, NPR); scratch = torch.empty(nb, dtype=torch.uint8, device=dev)
for rep in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dec.beam_search_batch_device(post.data_ptr(), N, T, 5, BEAM, 0.0, True, labels.data_ptr(), path.data_ptr(), cap, offsets.data_ptr(),
                                 lens.data_ptr(), status.data_ptr(), used.data_ptr(), scratch.data_ptr(), nb, NPR, st)
    e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
o = offsets.cpu().numpy().astype(np.uint64); ln = lens.cpu().numpy().astype(np.int64)
start = (o & np.uint64(0xffffffff)).astype(np.int64); sm = (o >> np.uint64(48)).astype(np.int64)
d = ((o >> np.uint64(32)) & np.uint64(0xffff)).astype(np.int64) * 8
start -= start.min()
print(f"scratch {nb/2**30:.1f} GiB status!=0: {int((status != 0).sum())}"); print(f"N={N} T={T} kernel {ms:.1f} ms; read duration us: min {d.min()} p10 {np.percentile(d,10):.0f} median {np.median(d):.0f} p90 {np.percentile(d,90):.0f} max {d.max()}")
print("start us: p0,p50,p90,p99,max", start.min(), np.median(start), np.percentile(start,90), np.percentile(start,99), start.max())
late = start > 1000
print("reads starting later than 1 ms:", int(late.sum()), "on SMs:", len(np.unique(sm[late])), "SMs used:", len(np.unique(sm)))
print("reads per SM min/max", np.bincount(sm).min(), np.bincount(sm).max())
print("corr(duration, emitted labels) %.3f; labels min/median/max %d %d %d" % (np.corrcoef(d, ln)[0, 1], ln.min(), np.median(ln), ln.max()))
per_sm = np.array([d[sm == k].mean() for k in np.unique(sm)])
within = np.array([d[sm == k].std() for k in np.unique(sm)])
print("per-SM mean duration us: min %.0f median %.0f max %.0f; std of a read's duration inside one SM: median %.0f" %
      (per_sm.min(), np.median(per_sm), per_sm.max(), np.median(within)))
edges = np.percentile(start, [0, 20, 40, 60, 80, 100])
for a, b in zip(edges[:-1], edges[1:]):
    m = (start >= a) & (start <= b)
    print("  reads starting in [%.0f, %.0f] us: median duration %.0f us, p90 %.0f" % (a, b, np.median(d[m]), np.percentile(d[m], 90)))
first = start < 1000
if first.any():
    k = np.argsort(d[first])
    print("first-round reads: fastest 5 durations", d[first][k[:5]], "on SMs", sm[first][k[:5]], "; slowest 5", d[first][k[-5:]], "on SMs", sm[first][k[-5:]])
