#!/usr/bin/env python3
"""Latency of the single-read drop-in calls (one [T, 5] read per call), next to the CPU oracle on one core.
Development aid for INTEGRATION.md's guidance on when to batch."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200")]
import numpy as np  # noqa: E402

import fast_ctc_decode as fcd  # noqa: E402
import oracle  # noqa: E402

rng = np.random.default_rng(0)
for T in (400, 4000):
    z = (rng.random((T, 5), dtype=np.float32) * 2 - 1) * 3
    e = np.exp(z - z.max(axis=-1, keepdims=True))
    post = (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)

    def timeit(fn, reps=20):
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return (time.perf_counter() - t0) / reps * 1e3
    print(f"T={T}: viterbi_search {timeit(lambda: fcd.viterbi_search(post, 'NACGT', True)):.3f} ms "
          f"(CPU oracle {timeit(lambda: oracle.viterbi(post, True)):.3f}), "
          f"beam_search(5, 0.1) {timeit(lambda: fcd.beam_search(post, 'NACGT', 5, 0.1)):.3f} ms "
          f"(CPU oracle {timeit(lambda: oracle.beam_search(post, 5, 0.1)):.3f}), "
          f"beam_search(32) {timeit(lambda: fcd.beam_search(post, 'NACGT', 32, 0.0), 5):.3f} ms "
          f"(CPU oracle {timeit(lambda: oracle.beam_search(post, 32, 0.0), 5):.3f})")
