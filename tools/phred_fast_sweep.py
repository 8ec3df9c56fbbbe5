#!/usr/bin/env python3
"""How far the Viterbi kernel's fast quality value (hardware lg2, one fused multiply-add) lies from the reference's
(glibc log10f, three rounded float32 operations), measured on the device: every float32 miss probability from 1e-4 to 1
(all that 1 - mean can be for a probability), plus a stride through (1, 2^40] for sums of invalid "probabilities".
Prints, per binade and (qscale, qbias), the largest |fast - exact| and its ratio to the eps the kernel allows there
(include/fctc.h, fctc_test_libm_device which = 5 / 6).  The eps constants in csrc/phred.cuh were set from this table."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import fctc_b200  # noqa: E402


def main():
    dec = fctc_b200.Decoder(0)
    lo = int(np.float32(1e-4).view(np.uint32))
    one = 0x3F800000
    bits = np.concatenate([np.arange(lo, one + 1, dtype=np.uint32), np.arange(one, 0x53800000, 4099, dtype=np.uint32)])
    x = torch.from_numpy(bits.view(np.float32).copy()).cuda()
    out = torch.empty_like(x)
    eps = torch.empty_like(x)
    st = torch.cuda.current_stream().cuda_stream
    expo = ((torch.from_numpy(bits.astype(np.int64)).cuda() >> 23) & 0xFF) - 127
    worst_ratio = 0.0
    for qscale, qbias in [(1.0, 0.0), (0.7, 2.0), (2.5, -3.0), (1.0, 33.0), (0.1, 0.0), (40.0, 100.0), (-1.0, 50.0)]:
        y = torch.tensor([qscale, qbias], dtype=torch.float32, device="cuda")
        dec.libm_device(x.data_ptr(), y.data_ptr(), out.data_ptr(), x.numel(), 5, st)
        dec.libm_device(x.data_ptr(), y.data_ptr(), eps.data_ptr(), x.numel(), 6, st)
        torch.cuda.synchronize()
        d = out.abs()
        ratio = d / eps
        print(f"qscale {qscale} qbias {qbias}: max |fast - exact| {float(d.max()):.3e}, max / eps {float(ratio.max()):.4f}")
        for e in sorted(set(expo.unique().tolist())):
            m = expo == e
            if e <= 0 or e % 8 == 0:
                print(f"   binade 2^{e:>3}: max diff {float(d[m].max()):.3e}  eps there {float(eps[m].min()):.3e}  ratio {float(ratio[m].max()):.4f}")
        worst_ratio = max(worst_ratio, float(ratio.max()))
    print("worst ratio", worst_ratio)


if __name__ == "__main__":
    main()
