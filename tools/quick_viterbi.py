#!/usr/bin/env python3
"""Quick device-resident timing of the Viterbi kernel (development aid; bench.py is the contract)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import torch  # noqa: E402

import fctc_b200  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
    Cn = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    dev = torch.device("cuda:0")
    dec = fctc_b200.Decoder(0)
    post = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    dec.synth_posteriors_device(post.data_ptr(), N, T, Cn, kind, 0xC7C00001, 0, st)
    cap = N * ((T + 31) // 32 * 32)
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    quals = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    for qstring in (1, 0):
        times = []
        for it in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dec.viterbi_batch_device(post.data_ptr(), N, T, Cn, qstring, 1.0, 0.0, True, labels.data_ptr(),
                                     quals.data_ptr(), 1, path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(),
                                     status.data_ptr(), st)
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        ms = sorted(times[3:])[len(times[3:]) // 2]
        emitted = int(lens.sum().item())
        in_b = N * T * Cn * 4
        out_b = emitted * (1 + 4 + (1 if qstring else 0)) + N * 16
        print(f"N={N} T={T} kind={kind} qstring={qstring}: {ms:.3f} ms  rows/s={N*T/ms*1e3:.3e}  "
              f"in={in_b/ms/1e6:.1f} GB/s  in+out={(in_b+out_b)/ms/1e6:.1f} GB/s  emitted/row={emitted/(N*T):.3f}  all={times}")


if __name__ == "__main__":
    main()
