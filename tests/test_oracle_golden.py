"""Pins the CPU oracle (oracle/ctc_oracle.cpp) against every known-answer vector the reference's
own tests hold for the decode path.  CPU only."""
import numpy as np
import pytest

from helpers import kat_post, load_kats

KATS = load_kats()


def _by_fn(fn):
    return [pytest.param(k, id=k["name"]) for k in KATS if k["fn"] == fn]


def test_phred_scores(oracle):
    (kat,) = [k for k in KATS if k["fn"] == "phred"]
    got = "".join(chr(oracle.phred(np.float32(p), kat["qscale"], kat["qbias"])[0]) for p in kat["probs_f32"])
    assert got == kat["expect"]


@pytest.mark.parametrize("kat", _by_fn("viterbi_search"))
def test_viterbi_kats(oracle, kat):
    post = kat_post(kat["post"])
    st, labels, quals, path = oracle.viterbi(post, kat["qstring"], kat["qscale"], kat["qbias"], kat["collapse_repeats"])
    assert st == oracle.OK
    seq = oracle.to_text(kat["alphabet"], labels, quals)
    if kat["expect_seq"] is not None:
        assert seq == kat["expect_seq"]
    if kat["expect_path"] is not None:
        assert path.tolist() == kat["expect_path"]
    assert len(labels) == len(path)


@pytest.mark.parametrize("kat", _by_fn("beam_search"))
def test_beam_search_kats(oracle, kat):
    post = kat_post(kat["post"])
    st, labels, path = oracle.beam_search(post, kat["beam_size"], kat["beam_cut_threshold"], kat["collapse_repeats"])
    assert st == oracle.OK
    if kat["expect_seq"] is not None:
        assert oracle.to_text(kat["alphabet"], labels) == kat["expect_seq"]
    if kat["expect_path"] is not None:
        assert path.tolist() == kat["expect_path"]
    assert len(labels) == len(path)


@pytest.mark.parametrize("kat", _by_fn("crf_greedy_search"))
def test_crf_greedy_kats(oracle, kat):
    post = np.asarray(kat["post"], np.float32)
    init = np.asarray(kat["init_state"], np.float32)
    st, labels, quals, path = oracle.crf_greedy(post, init, kat["qstring"], kat["qscale"], kat["qbias"])
    assert st == oracle.OK
    assert oracle.to_text(kat["alphabet"], labels, quals) == kat["expect_seq"]
    assert path.tolist() == kat["expect_path"]


@pytest.mark.parametrize("kat", _by_fn("crf_beam_search"))
def test_crf_beam_kats(oracle, kat):
    post = np.asarray(kat["post"], np.float32)
    init = np.asarray(kat["init_state"], np.float32)
    st, labels, path = oracle.crf_beam_search(post, init, kat["beam_size"], kat["beam_cut_threshold"])
    assert st == oracle.OK
    assert oracle.to_text(kat["alphabet"], labels) == kat["expect_seq"]
    assert path.tolist() == kat["expect_path"]


@pytest.mark.parametrize("exp_mode", [0, 1], ids=["exact_exp", "fastexp_as_shipped"])
@pytest.mark.parametrize("kat", _by_fn("beam_search_duplex"))
def test_duplex_kats(oracle, kat, exp_mode):
    # The reference's CI runs these with the default feature set (fastexp, which evaluates to 0);
    # the north_star parity target is the exact-exp build.  Both must give the asserted strings.
    a = np.asarray(kat["post"], np.float32)
    b = np.asarray(kat["post2"], np.float32)
    st, labels = oracle.beam_search_duplex(a, b, kat["envelope"], kat["beam_size"], kat["beam_cut_threshold"],
                                           kat["collapse_repeats"], exp_mode)
    assert st == oracle.OK
    assert oracle.to_text(kat["alphabet"], labels) == kat["expect_seq"]


def test_tree_assembly(oracle):
    """src/tree.rs:201-269"""
    import ctypes as C
    L = oracle.lib()
    t = L.oracle_trie_new(2)
    try:
        def data(node):
            d = C.c_int64(0)
            return d.value if L.oracle_trie_data(t, node, C.byref(d)) else None

        def label(node):
            v = L.oracle_trie_label(t, node)
            return None if v < 0 else v

        def child(node, lab):
            v = L.oracle_trie_child(t, node, lab)
            return None if v < 0 else v

        assert child(-1, 0) is None and child(-1, 1) is None and data(-1) is None and label(-1) is None
        steps = [(-1, 0, 100, 0), (-1, 1, 101, 1), (0, 0, 102, 2), (0, 1, 103, 3), (3, 1, 104, 4), (1, 0, 105, 5)]
        for parent, lab, d, want in steps:
            assert L.oracle_trie_add(t, parent, lab, d) == want
            assert child(parent, lab) == want and data(want) == d and label(want) == lab
        assert label(-1) is None
        assert [child(-1, 0), child(-1, 1), child(0, 0), child(0, 1), child(1, 0)] == [0, 1, 2, 3, 5]
        assert [child(1, 1), child(2, 0), child(2, 1), child(3, 0)] == [None] * 4
        assert child(3, 1) == 4 and data(4) == 104 and label(4) == 1
        labs = (C.c_int64 * 8)()
        dat = (C.c_int64 * 8)()
        n = L.oracle_trie_ancestors(t, 4, labs, dat, 8)
        assert list(labs[:n]) == [1, 1, 0]
        assert list(zip(labs[:n], dat[:n])) == [(1, 104), (1, 103), (0, 100)]
    finally:
        L.oracle_trie_free(t)


def _sec(oracle, offset, pairs):
    lab = np.asarray([p[0] for p in pairs], np.float32)
    gap = np.asarray([p[1] for p in pairs], np.float32)
    L = oracle.lib()

    def get_gap(at):
        return L.oracle_secondary_get_gap(offset, lab.ctypes.data, gap.ctypes.data, len(pairs), at)

    def upd(lo, hi):
        return L.oracle_secondary_update_max(offset, lab.ctypes.data, gap.ctypes.data, len(pairs), lo, hi, 0)
    return get_gap, upd


NEG_INF = float("-inf")


def _ln(x):
    return float(np.log(np.float32(x)))


def test_secondary_probs_get(oracle):
    """src/duplex.rs:842-893"""
    pairs = [(0, 0.1), (0, 0.2), (0, 0.3)]
    g = [pytest.approx(_ln(v), rel=1e-6) for v in (0.1, 0.2, 0.3)]
    get, _ = _sec(oracle, 0, pairs)
    assert [get(i) for i in (-1, 0, 1, 2, 3)] == [NEG_INF, g[0], g[1], g[2], NEG_INF]
    get, _ = _sec(oracle, 3, pairs)
    assert [get(i) for i in (-1, 0, 2, 3, 4, 5, 6)] == [NEG_INF] * 3 + g + [NEG_INF]
    get, _ = _sec(oracle, -1, pairs)
    assert [get(i) for i in (-2, -1, 0, 1, 2)] == [NEG_INF] + g + [NEG_INF]
    for off, ats in [(0, (-1, 0, 1)), (-1, (-2, -1, 0)), (4, (3, 4, 5))]:
        get, _ = _sec(oracle, off, [])
        assert [get(i) for i in ats] == [NEG_INF] * 3


def test_secondary_probs_update_max(oracle):
    """src/duplex.rs:895-993"""
    _, upd = _sec(oracle, 0, [])
    big = 2 ** 63 - 1
    for lo, hi in [(0, 0), (-1, 0), (0, 1), (-1, 1), (-big - 1, big)]:
        assert upd(lo, hi) == NEG_INF
    pairs = [(0, 0.1), (0.3, 0), (0.2, 0), (0.4, 0), (0, 0.5)]
    _, upd = _sec(oracle, 2, pairs)
    want = {(0, 0): None, (0, 2): None, (0, 3): 0.1, (2, 2): None, (2, 3): 0.1, (2, 4): 0.3, (2, 5): 0.3,
            (2, 6): 0.4, (2, 7): 0.5, (6, 7): 0.5, (7, 7): None, (2, 10): 0.5, (3, 10): 0.5, (8, 10): None}
    for (lo, hi), v in want.items():
        got = upd(lo, hi)
        if v is None:
            assert got == NEG_INF, (lo, hi)
        else:
            assert got == pytest.approx(_ln(v), rel=1e-6), (lo, hi)


def test_fastexp_as_shipped_is_zero(oracle):
    """the default-feature exp approximation (src/fastexp.rs:29-62) evaluates to 0 on
    little-endian hosts, so LogSpace '+' degenerates to max in the shipped wheel."""
    L = oracle.lib()
    for x in (0.0, -0.1, -1.0, -5.0, -20.0, -499.0, -600.0):
        assert L.oracle_fastexp_as_shipped(x) == 0.0
    assert L.oracle_logadd(-1.0, -2.0, 1) == -1.0
    assert L.oracle_logadd(-1.0, -2.0, 0) == pytest.approx(np.log(np.exp(-1.0) + np.exp(-2.0)), rel=1e-6)
    assert L.oracle_logadd(NEG_INF, -2.0, 0) == -2.0


def test_error_statuses(oracle):
    """tests/test_decode.py:100-104, 370-374 and src/search.rs:274-277"""
    nan = np.full((100, 5), np.nan, np.float32)
    assert oracle.beam_search(nan)[0] == oracle.INCOMPARABLE
    rng = np.random.default_rng(1)
    ok = rng.random((100, 5), dtype=np.float32)
    assert oracle.beam_search_duplex(nan, ok)[0] == oracle.INCOMPARABLE
    low = np.full((10, 5), 0.01, np.float32)
    assert oracle.beam_search(low, 5, 0.1)[0] == oracle.RAN_OUT_OF_BEAM
    env = np.array([[0, 5]] * 5 + [[7, 9]] * 5, np.uint64)  # lower bound jumps past the last upper bound
    assert oracle.beam_search_duplex(ok[:10], ok[:10], env)[0] == oracle.INVALID_ENVELOPE
    assert oracle.beam_search(np.zeros((0, 5), np.float32))[0:2][0] == oracle.OK  # empty input: ("", [])
    assert oracle.viterbi(np.zeros((0, 5), np.float32))[0] == oracle.PANIC       # src/search.rs:329 assert


def test_random_shapes_match_reference_suite_properties(oracle):
    """tests/test_decode.py:15-23,201-210: len(seq) == len(path), every base occurs."""
    rng = np.random.default_rng(7)
    for C in (3, 5, 12):
        x = rng.random((100 if C < 12 else 10000, C), dtype=np.float32)
        x = (x / np.linalg.norm(x, ord=2, axis=1, keepdims=True)).astype(np.float32)
        st, labels, path = oracle.beam_search(x, 5, 0.1 if C < 12 else 0.0)
        assert st == oracle.OK and len(labels) == len(path) and len(set(labels.tolist())) == C - 1
        st, labels, quals, path = oracle.viterbi(x, True)
        assert st == oracle.OK and len(labels) == len(path) == len(quals)
