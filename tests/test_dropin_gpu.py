"""Acceptance suite of the drop-in module `fast_ctc_decode` (C++ extension over the C ABI): the
behaviours the reference's Python tests pin (tests/test_decode.py in the reference checkout) —
argument forms, defaults, alphabets as str/list/tuple, multi-character labels, exception classes
and messages, and the exact outputs of its hand-made matrices — plus the new batched calls."""
import numpy as np
import pytest

from helpers import kat_post, load_kats

pytestmark = pytest.mark.gpu

ALPHABET = "NACGT"


@pytest.fixture(scope="module")
def fcd():
    import fast_ctc_decode
    return fast_ctc_decode


def unit_rows(rng, samples=100, n=5):
    """tests/test_decode.py:14-17: rand rows scaled to unit L2 norm (not a probability simplex)."""
    x = rng.random((samples, n)).astype(np.float32)
    return x / np.linalg.norm(x, ord=2, axis=1, keepdims=True)


def test_module_surface(fcd):
    for name in ("viterbi_search", "beam_search", "beam_search_duplex", "crf_greedy_search", "crf_beam_search",
                 "crf_beam_search_duplex"):
        assert callable(getattr(fcd, name))                       # src/lib.rs:619-625
    assert fcd.__version__ == "0.3.7"                             # Cargo.toml:3


@pytest.mark.parametrize("alphabet", [ALPHABET, list(ALPHABET), tuple(ALPHABET), "NRUST"])
def test_beam_search_alphabet_forms(fcd, alphabet):
    """tests/test_decode.py:19-37, 56-60"""
    probs = unit_rows(np.random.default_rng(1))
    seq, path = fcd.beam_search(probs, alphabet, 5, 0.1)
    assert isinstance(seq, str) and isinstance(path, list)
    assert len(seq) == len(path) and len(set(seq)) == len(alphabet) - 1


def test_beam_search_argument_forms(fcd):
    """tests/test_decode.py:39-54, 67-71"""
    probs = unit_rows(np.random.default_rng(2))
    a = fcd.beam_search(network_output=probs, alphabet=ALPHABET, beam_size=5, beam_cut_threshold=0.1)
    b = fcd.beam_search(probs, ALPHABET, 5, 0.1)
    assert a == b
    seq, path = fcd.beam_search(probs, ALPHABET)                 # defaults: beam 5, threshold 0.0, collapse
    assert (seq, path) == fcd.beam_search(probs, ALPHABET, 5, 0.0, True)
    with pytest.raises(TypeError):
        fcd.beam_search(probs)
    seq, path = fcd.beam_search(probs, ALPHABET, 5, 0.0)
    assert len(seq) == len(path) and len(set(seq)) == 4


def test_beam_search_value_errors(fcd):
    """tests/test_decode.py:62-98 and the messages of src/lib.rs:333-349"""
    probs = unit_rows(np.random.default_rng(3))
    with pytest.raises(ValueError, match="beam_size cannot be 0"):
        fcd.beam_search(probs, ALPHABET, 0, 0.1)
    with pytest.raises(ValueError, match="beam_cut_threshold must be at least 0.0"):
        fcd.beam_search(probs, ALPHABET, 5, -0.1)
    with pytest.raises(ValueError, match="beam_cut_threshold cannot be more than 0.2$"):
        fcd.beam_search(probs, ALPHABET, 5, 1.0 / len(ALPHABET))
    with pytest.raises(ValueError):
        fcd.beam_search(probs, ALPHABET, 5, 1.1)
    with pytest.raises(ValueError, match="alphabet size 4 does not match probability matrix inner dimension 5"):
        fcd.beam_search(probs, "NAGC", 5, 0.1)
    with pytest.raises(ValueError):
        fcd.beam_search(probs, "NAGCTX", 5, 0.1)
    with pytest.raises(TypeError):
        fcd.beam_search(probs.astype(np.float64), ALPHABET)


def test_nan_input_raises_runtime_error(fcd):
    """tests/test_decode.py:100-104, 370-374"""
    nan = np.full((100, 5), np.nan, np.float32)
    with pytest.raises(RuntimeError, match="Failed to compare values"):
        fcd.beam_search(nan, ALPHABET)
    ok = unit_rows(np.random.default_rng(4))
    with pytest.raises(RuntimeError, match="Failed to compare values"):
        fcd.beam_search_duplex(nan, ok, ALPHABET)


@pytest.mark.parametrize("alphabet,samples,thr", [("NAG", 100, 0.1), ("NABCDEFGHIJK", 10000, 0.0)])
def test_beam_search_other_alphabet_sizes(fcd, alphabet, samples, thr):
    """tests/test_decode.py:106-120"""
    probs = unit_rows(np.random.default_rng(5), samples, len(alphabet))
    seq, path = fcd.beam_search(probs, alphabet, 5, beam_cut_threshold=thr)
    assert len(seq) == len(path) and len(set(seq)) == len(alphabet) - 1


def test_viterbi_random_data(fcd):
    """tests/test_decode.py:201-225"""
    probs = unit_rows(np.random.default_rng(6))
    seq, path = fcd.viterbi_search(probs, ALPHABET)
    assert len(seq) == len(path) and len(set(seq)) == 4
    seq, path = fcd.viterbi_search(probs, ALPHABET, qstring=True)
    assert len(seq) == 2 * len(path)
    with pytest.raises(TypeError):
        fcd.viterbi_search(probs)
    for bad in ("NACG", "NACGTR"):
        with pytest.raises(ValueError, match="alphabet size does not match probability matrix dimensions"):
            fcd.viterbi_search(probs, bad)
    with pytest.raises(ValueError, match="Empty alphabet given"):
        fcd.viterbi_search(probs, "")


@pytest.mark.parametrize("kat", [pytest.param(k, id=k["name"]) for k in load_kats()
                                 if k["fn"] in ("viterbi_search", "beam_search", "beam_search_duplex")])
def test_reference_known_answers(fcd, kat):
    post = kat_post(kat["post"])
    if kat["fn"] == "viterbi_search":
        seq, path = fcd.viterbi_search(post, kat["alphabet"], kat["qstring"], kat["qscale"], kat["qbias"],
                                       kat["collapse_repeats"])
    elif kat["fn"] == "beam_search":
        seq, path = fcd.beam_search(post, kat["alphabet"], kat["beam_size"], kat["beam_cut_threshold"],
                                    kat["collapse_repeats"])
    else:
        seq = fcd.beam_search_duplex(post, np.asarray(kat["post2"], np.float32), kat["alphabet"])
        path = None
        assert isinstance(seq, str)
    if kat["expect_seq"] is not None:
        assert seq == kat["expect_seq"]
    if kat.get("expect_path") is not None:
        assert path == kat["expect_path"]


def test_duplex_envelope_argument(fcd, oracle):
    rng = np.random.default_rng(8)
    a, b = unit_rows(rng, 60), unit_rows(rng, 70)
    env = np.stack([np.maximum(0, np.arange(60) - 8), np.minimum(70, np.arange(60) + 12)], axis=1).astype(np.uint64)
    seq = fcd.beam_search_duplex(a, b, ALPHABET, envelope=env, beam_size=5, beam_cut_threshold=0.0)
    st, labels = oracle.beam_search_duplex(a, b, env, 5, 0.0, True, 0)
    assert st == 0 and seq == oracle.to_text(ALPHABET, labels)  # exact-exp mode is bit-identical to a glibc host
    with pytest.raises(ValueError, match="the lengths of network_output_1 and envelope do not match"):
        fcd.beam_search_duplex(a, b, ALPHABET, envelope=env[:-1])
    with pytest.raises(ValueError, match="the inner axis of envelope must have size 2"):
        fcd.beam_search_duplex(a, b, ALPHABET, envelope=np.zeros((60, 3), np.uint64))
    with pytest.raises(ValueError, match="inner axes of the network outputs do not match"):
        fcd.beam_search_duplex(a, b[:, :4], ALPHABET)
    fcd.set_exp_mode("fastexp")
    try:
        seq2 = fcd.beam_search_duplex(a, b, ALPHABET, envelope=env)
        st, labels = oracle.beam_search_duplex(a, b, env, 5, 0.0, True, 1)
        assert seq2 == oracle.to_text(ALPHABET, labels)
    finally:
        fcd.set_exp_mode("exact")


def test_batched_calls_match_single_calls(fcd):
    rng = np.random.default_rng(9)
    batch = np.stack([unit_rows(rng, 300) for _ in range(6)])
    multi = ["N", "AAA", "CCC", "GGG", "TTTT"]
    got = fcd.viterbi_search_batch(batch, multi, qstring=True, qscale=0.7, qbias=2.0)
    assert got == [fcd.viterbi_search(x, multi, True, 0.7, 2.0) for x in batch]
    got = fcd.beam_search_batch(batch, ALPHABET, beam_size=5, beam_cut_threshold=0.1)
    assert got == [fcd.beam_search(x, ALPHABET, 5, 0.1) for x in batch]
    got = fcd.beam_search_duplex_batch(batch[:3], batch[3:], ALPHABET)
    assert got == [fcd.beam_search_duplex(x, y, ALPHABET) for x, y in zip(batch[:3], batch[3:])]


def test_short_lived_threads_reuse_handles(fcd):
    """The module keeps one library handle per live thread; a thread that ends returns its handle for the next one.
    Forty threads in a row must not grow device memory use beyond the first few."""
    import threading

    import torch
    rng = np.random.default_rng(10)
    post = unit_rows(rng, 400)
    want = fcd.viterbi_search(post, ALPHABET)
    results = []

    def work():
        results.append(fcd.viterbi_search(post, ALPHABET) == want and fcd.beam_search(post, ALPHABET) is not None)

    def run(n):
        for _ in range(n):
            t = threading.Thread(target=work)
            t.start()
            t.join()
    run(4)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    run(40)
    torch.cuda.synchronize()
    assert all(results) and len(results) == 44
    assert free_before - torch.cuda.mem_get_info(0)[0] < (8 << 20), "threads that ended left their device buffers behind"
