#!/usr/bin/env python3
"""Quick device-resident timing of the beam-search kernel (development aid)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import torch  # noqa: E402

import fctc_b200  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
    beam = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.1
    reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
    Cn = 5
    dev = torch.device("cuda:0")
    dec = fctc_b200.Decoder(0)
    post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, 0, 0xC7C00001, 0, st)
    cap = N * ((T + 31) // 32 * 32)
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    nbytes = dec.beam_scratch_bytes(N, T, Cn, beam, 0)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    times = []
    for _ in range(reps):
        used.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dec.beam_search_batch_device(post.data_ptr(), N, T, Cn, beam, thr, True, labels.data_ptr(), path.data_ptr(), cap,
                                     offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), used.data_ptr(),
                                     scratch.data_ptr(), nbytes, 0, st)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = min(times)
    bad = int((status != 0).sum().item())
    print(f"beam N={N} T={T} beam={beam} thr={thr}: {ms:.2f} ms  rows/s={N*T/ms*1e3:.3e}  not-ok reads={bad}  "
          f"scratch={nbytes/2**30:.2f} GiB  all={[round(t, 2) for t in times]}")


if __name__ == "__main__":
    main()
