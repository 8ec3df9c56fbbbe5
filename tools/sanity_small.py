#!/usr/bin/env python3
"""Small run through every kernel (for compute-sanitizer: keep it short)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200"), os.path.join(ROOT, "tests")]
import numpy as np  # noqa: E402

import fctc_b200  # noqa: E402
from helpers import band_envelope, duplex_pairs, peaked_posteriors, softmax_posteriors  # noqa: E402

rng = np.random.default_rng(0)
dec = fctc_b200.Decoder(0)
for shape in [(9, 700, 5), (3, 130, 12), (2, 300, 100), (4, 129, 3)]:
    post = softmax_posteriors(rng, shape)
    for q, c in [(True, True), (True, False), (False, True)]:
        r = dec.viterbi_batch(post, q, 0.7, 2.0, c)
        assert (r["status"] == 0).all()
post = peaked_posteriors(rng, (6, 400, 5))
for beam, thr in [(5, 0.1), (32, 0.0), (40, 0.0)]:
    r = dec.beam_search_batch(post, beam, thr, True)
    assert (r["status"] == 0).all()
p1, p2 = duplex_pairs(rng, 4, 120, 130)
for mode in (0, 1):
    r = dec.beam_search_duplex_batch(p1, p2, band_envelope(4, 120, 130, 8), 5, 0.0, True, mode)
    assert (r["status"] == 0).all()
crf = rng.random((90, 16, 5), dtype=np.float32)
crf /= crf.sum(axis=2, keepdims=True)
init = rng.random(16, dtype=np.float32)
assert dec.crf_greedy_search(crf, init, True)[0] == 0
assert dec.crf_beam_search(crf, init, 5, 0.0)[0] == 0
assert dec.crf_beam_search_duplex(crf, init, crf, init, None, 5, 0.0)[0] == 0
dec.close()
print("sanity ok")
