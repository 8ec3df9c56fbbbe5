#!/usr/bin/env python3
"""Quick device-resident timing of the duplex kernel (development aid)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import fctc_b200  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 2368
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    band = int(sys.argv[3]) if len(sys.argv) > 3 else 32
    mode = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
    npr = int(sys.argv[6]) if len(sys.argv) > 6 else 0   # nodes_per_row (0 = default)
    Cn, beam = 5, 5
    dev = torch.device("cuda:0")
    dec = fctc_b200.Decoder(0)
    st = torch.cuda.current_stream().cuda_stream
    a = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    b = torch.empty((N, T, Cn), dtype=torch.float32, device=dev)
    dec.synth_posteriors_device(a.data_ptr(), N, T, Cn, 1, 0xC7C00001, 0, st)
    dec.synth_posteriors_device(b.data_ptr(), N, T, Cn, 2, 0xC7C00001, 0, st)
    c = np.arange(T)
    env = np.stack([np.maximum(0, c - band), np.minimum(T, c + band + 1)], axis=1).astype(np.int64)
    env = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(env, (N, T, 2)))).to(dev)
    cap = N * ((T + 31) // 32 * 32)
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    used = torch.zeros(1, dtype=torch.int64, device=dev)
    mw = 2 * band + 1
    nbytes = dec.duplex_scratch_bytes(N, T, T, Cn, beam, mw, npr)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    times = []
    for _ in range(reps):
        used.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dec.beam_search_duplex_batch_device(a.data_ptr(), b.data_ptr(), N, T, T, Cn, env.data_ptr(), mw, beam, 0.0, True, mode,
                                            labels.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(),
                                            used.data_ptr(), scratch.data_ptr(), nbytes, npr, st)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = min(times)
    print(f"duplex N={N} T={T} band={band} mode={mode}: {ms:.2f} ms rows/s={N*T/ms*1e3:.3e} not-ok={int((status != 0).sum())} "
          f"scratch={nbytes/2**30:.2f} GiB mean_len={float(lens.float().mean()):.1f}")


if __name__ == "__main__":
    main()
