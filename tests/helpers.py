"""Shared helpers for the test-suite (CPU and GPU)."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_kats():
    with open(os.path.join(GOLDEN, "reference_kats.json")) as f:
        return json.load(f)["kats"]


def kat_post(spec):
    """Materialise a posterior matrix stored either literally or as a recipe."""
    if isinstance(spec, dict):
        assert spec["recipe"] == "spaced_emissions"  # tests/test_decode.py:122-135
        x = np.zeros((spec["T"], spec["C"]), np.float32)
        x[:, 0] = spec["stay"]
        for base, pos in enumerate(range(0, spec["T"], spec["step"])):
            x[pos, base % (spec["C"] - 1) + 1] = 1.0
        return x
    return np.asarray(spec, dtype=np.float32)


def softmax_posteriors(rng, shape, scale=3.0):
    """Softmax-normalised synthetic posteriors, float32 (SURVEY.md §8d primary distribution)."""
    z = (rng.random(shape, dtype=np.float32) * 2.0 - 1.0) * np.float32(scale)
    e = np.exp(z - z.max(axis=-1, keepdims=True))
    return (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)


def peaked_posteriors(rng, shape, p_blank=0.65):
    """Nanopore-like peaked rows: one dominant label per row, blank with probability p_blank."""
    *lead, C = shape
    n = int(np.prod(lead))
    dom = np.where(rng.random(n) < p_blank, 0, rng.integers(1, max(C, 2), n) % C)
    x = (0.05 * rng.random((n, C), dtype=np.float32) ** 2).astype(np.float32)
    x[np.arange(n), dom] = 1.0
    x /= x.sum(axis=1, keepdims=True, dtype=np.float32)
    return x.reshape(shape).astype(np.float32)
