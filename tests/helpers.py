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


def duplex_pairs(rng, n_pairs, T1, T2, C=5, p_blank=0.6, noise=0.08):
    """Template/complement-like posterior pairs: one hidden label track rendered twice with
    independent noise; the second rendering is resampled to T2 rows."""
    hid = np.where(rng.random((n_pairs, T1)) < p_blank, 0, rng.integers(1, C, (n_pairs, T1)))

    def render(track, T):
        idx = np.minimum((np.arange(T) * track.shape[1]) // max(T, 1), track.shape[1] - 1)
        tr = track[:, idx]
        x = (noise * rng.random((n_pairs, T, C), dtype=np.float32)).astype(np.float32)
        n, t = np.meshgrid(np.arange(n_pairs), np.arange(T), indexing="ij")
        x[n, t, tr] = 1.0
        flip = rng.random((n_pairs, T)) < 0.05  # a few disagreeing rows
        x[flip] = rng.random((int(flip.sum()), C), dtype=np.float32)
        x /= x.sum(axis=2, keepdims=True, dtype=np.float32)
        return x.astype(np.float32)
    return render(hid, T1), render(hid, T2)


def band_envelope(n_pairs, T1, T2, half_width):
    """Monotone band [max(0, c-w), min(T2, c+w+1)) around the diagonal."""
    c = (np.arange(T1) * T2) // max(T1, 1)
    lo = np.maximum(0, c - half_width)
    hi = np.minimum(T2, c + half_width + 1)
    env = np.stack([lo, hi], axis=1).astype(np.int64)
    return np.ascontiguousarray(np.broadcast_to(env, (n_pairs, T1, 2)))
