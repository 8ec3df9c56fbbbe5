#!/usr/bin/env python3
"""Fraction of duplex pairs whose GPU sequence equals the CPU oracle's (exact-exp mode)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200"), os.path.join(ROOT, "tests")]
import numpy as np  # noqa: E402

import fctc_b200  # noqa: E402
import oracle  # noqa: E402
from helpers import band_envelope, duplex_pairs  # noqa: E402

N, T, band = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(123)
p1, p2 = duplex_pairs(rng, N, T, T, 5, noise=float(sys.argv[4]) if len(sys.argv) > 4 else 0.08)
env = band_envelope(N, T, T, band)
dec = fctc_b200.Decoder(0)
for mode in (0, 1):
    res = dec.beam_search_duplex_batch(p1, p2, env, 5, 0.0, True, exp_mode=mode)
    ol, olen, ost = oracle.beam_search_duplex_batch(p1, p2, env.astype(np.uint64), 5, 0.0, True, exp_mode=mode)
    same = sum(res["labels"][o:o + n].tolist() == ol[i, :olen[i]].tolist()
               for i, (o, n) in enumerate(zip(res["offsets"], res["lens"])))
    print(f"mode {mode}: {same}/{N} pairs identical ({100.0 * same / N:.2f} %), statuses equal: {(res['status'] == ost).all()}")
