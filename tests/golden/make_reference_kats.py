#!/usr/bin/env python3
"""Writes tests/golden/reference_kats.json: the known-answer vectors that the reference's own
tests hold for the decode path, transcribed as data (inputs + expected outputs).

The reference cannot be built or imported in this image (Rust, no cargo), so these vectors are
not *generated* by running it; they are the literal matrices and expected results asserted in:
  /root/reference/src/search.rs:431-601      (crf_test_greedy, test_phred_scores, test_viterbi,
                                              test_viterbi_blank_bounds)
  /root/reference/tests/test_decode.py:122-189, 227-355, 376-412
  /root/reference/tests/fast_ctc_wasm.test.js:12,29-46
Run:  python tests/golden/make_reference_kats.py
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def mat(x):
    return np.asarray(x, dtype=np.float32).tolist()


def main():
    kats = []

    # ---- src/search.rs:513-525 test_phred_scores ---------------------------------------------
    kats.append({
        "name": "rust_phred_scores", "fn": "phred", "source": "src/search.rs:513-525",
        "probs_f32": [float(np.float32(v)) for v in
                      [0.0, 0.5, np.float32(1.0) - np.float32(1e-1), np.float32(1.0) - np.float32(1e-2),
                       np.float32(1.0) - np.float32(1e-3), np.float32(1.0) - np.float32(1e-4),
                       np.float32(1.0) - np.float32(1e-5), np.float32(1.0) - np.float32(1e-6), 1.0]],
        "qscale": 1.0, "qbias": 0.0, "expect": "!$+5?IIII",
    })

    # ---- src/search.rs:528-554 test_viterbi; same matrix in tests/fast_ctc_wasm.test.js:12 ----
    m10 = [[0.0, 0.4, 0.6], [0.0, 0.3, 0.7], [0.3, 0.3, 0.4], [0.4, 0.3, 0.3], [0.4, 0.3, 0.3],
           [0.3, 0.3, 0.4], [0.1, 0.4, 0.5], [0.1, 0.5, 0.4], [0.8, 0.1, 0.1], [0.1, 0.1, 0.8]]
    nag = ["N", "A", "G"]
    kats.append({"name": "rust_viterbi_10x3", "fn": "viterbi_search", "source": "src/search.rs:528-548",
                 "post": mat(m10), "alphabet": nag, "qstring": False, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "GGAG", "expect_path": [0, 5, 7, 9]})
    kats.append({"name": "rust_viterbi_10x3_q", "fn": "viterbi_search", "source": "src/search.rs:550-553",
                 "post": mat(m10), "alphabet": nag, "qstring": True, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "GGAG%$$(", "expect_path": [0, 5, 7, 9]})
    kats.append({"name": "js_viterbi_10x3", "fn": "viterbi_search", "source": "tests/fast_ctc_wasm.test.js:39-46",
                 "post": mat(m10), "alphabet": nag, "qstring": False, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "GGAG", "expect_path": [0, 5, 7, 9]})
    kats.append({"name": "js_beam_10x3", "fn": "beam_search", "source": "tests/fast_ctc_wasm.test.js:29-36",
                 "post": mat(m10), "alphabet": nag, "beam_size": 5, "beam_cut_threshold": 0.0,
                 "collapse_repeats": True, "expect_seq": "GAGAG", "expect_path": [0, 1, 2, 4, 6]})

    # ---- src/search.rs:557-601 test_viterbi_blank_bounds -----------------------------------------
    m13 = [[0.6, 0.2, 0.2], [0.6, 0.2, 0.2]] + m10 + [[0.4, 0.3, 0.3]]
    for q, col, seq, path, src in [
        (False, True, "GGAG", [2, 7, 9, 11], "src/search.rs:576-579"),
        (True, True, "GGAG%$$(", [2, 7, 9, 11], "src/search.rs:581-584"),
        (False, False, "GGGGGAG", [2, 3, 4, 7, 8, 9, 11], "src/search.rs:586-589"),
        (True, False, "GGGGGAG%&##$$(", [2, 3, 4, 7, 8, 9, 11], "src/search.rs:591-594"),
    ]:
        kats.append({"name": f"rust_viterbi_13x3_q{int(q)}_c{int(col)}", "fn": "viterbi_search", "source": src,
                     "post": mat(m13), "alphabet": nag, "qstring": q, "qscale": 1.0, "qbias": 0.0,
                     "collapse_repeats": col, "expect_seq": seq, "expect_path": path})
    kats.append({"name": "rust_beam_13x3_collapse", "fn": "beam_search", "source": "src/search.rs:596-597",
                 "post": mat(m13), "alphabet": nag, "beam_size": 5, "beam_cut_threshold": 0.0,
                 "collapse_repeats": True, "expect_seq": "GAGAG", "expect_path": None})
    kats.append({"name": "rust_beam_13x3_nocollapse", "fn": "beam_search", "source": "src/search.rs:599-600",
                 "post": mat(m13), "alphabet": nag, "beam_size": 5, "beam_cut_threshold": 0.0,
                 "collapse_repeats": False, "expect_seq": "GGGAGAG", "expect_path": None})

    # ---- src/search.rs:431-510 crf_test_greedy ---------------------------------------------------
    crf = np.zeros((7, 4, 5), np.float32)
    crf[0, 2, 0] = 1.0
    crf[1, 2, 2] = 0.9
    crf[2, 1, 4] = 0.7
    crf[3, 3, 0] = 1.0
    crf[4, 3, 1] = 0.99
    crf[5, 0, 1] = 0.9
    crf[6, 0, 3] = 0.999
    init = [0.0, 0.0, 1.0, 0.0, 0.0]
    nacgt = ["N", "A", "C", "G", "T"]
    kats.append({"name": "rust_crf_greedy", "fn": "crf_greedy_search", "source": "src/search.rs:485-489",
                 "post": crf.tolist(), "init_state": init, "alphabet": nacgt, "qstring": False, "qscale": 1.0,
                 "qbias": 0.0, "expect_seq": "CTAAG", "expect_path": [1, 2, 4, 5, 6]})
    kats.append({"name": "rust_crf_greedy_q", "fn": "crf_greedy_search", "source": "src/search.rs:491-495",
                 "post": crf.tolist(), "init_state": init, "alphabet": nacgt, "qstring": True, "qscale": 1.0,
                 "qbias": 0.0, "expect_seq": "CTAAG+&5+?", "expect_path": [1, 2, 4, 5, 6]})
    kats.append({"name": "rust_crf_beam", "fn": "crf_beam_search", "source": "src/search.rs:497-509",
                 "post": crf.tolist(), "init_state": init, "alphabet": nacgt, "beam_size": 5,
                 "beam_cut_threshold": 0.01, "expect_seq": "CTAAG", "expect_path": [1, 2, 4, 5, 6]})

    # ---- tests/test_decode.py (alphabet NACGT, beam 5, threshold 0.1) -----------------------------
    def stay(w, n=5):
        x = np.zeros((w, n), np.float32)
        x[:, 0] = 0.5
        return x

    x = stay(5000)
    emit = np.arange(0, 5000, 4)
    for base, pos in enumerate(emit):
        x[pos, base % 4 + 1] = 1.0
    # stored as a recipe, not 25,000 floats
    spaced = {"recipe": "spaced_emissions", "T": 5000, "C": 5, "stay": 0.5, "step": 4}
    kats.append({"name": "py_beam_path_T5000", "fn": "beam_search", "source": "tests/test_decode.py:122-135",
                 "post": spaced, "alphabet": nacgt, "beam_size": 5, "beam_cut_threshold": 0.1,
                 "collapse_repeats": True, "expect_seq": None, "expect_path": emit.tolist()})
    kats.append({"name": "py_viterbi_path_T5000", "fn": "viterbi_search", "source": "tests/test_decode.py:227-240",
                 "post": spaced, "alphabet": nacgt, "qstring": False, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": None, "expect_path": emit.tolist()})

    x = stay(20)
    for idx in [6, 13, 18]:
        x[idx, 0] = 0.0
        x[idx, 1] = 1.0
    kats.append({"name": "py_beam_AAA", "fn": "beam_search", "source": "tests/test_decode.py:137-152",
                 "post": mat(x), "alphabet": nacgt, "beam_size": 5, "beam_cut_threshold": 0.1,
                 "collapse_repeats": True, "expect_seq": "AAA", "expect_path": [6, 13, 18]})
    kats.append({"name": "py_viterbi_AAA", "fn": "viterbi_search", "source": "tests/test_decode.py:242-257",
                 "post": mat(x), "alphabet": nacgt, "qstring": False, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "AAA", "expect_path": [6, 13, 18]})
    kats.append({"name": "py_viterbi_AAA_q", "fn": "viterbi_search", "source": "tests/test_decode.py:259-277",
                 "post": mat(x), "alphabet": nacgt, "qstring": True, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "AAAIII", "expect_path": [6, 13, 18]})

    multi = ["N", "AAA", "CCC", "GGG", "TTTT"]
    x = stay(20)
    for k, idx in enumerate([6, 13, 18]):
        x[idx, 0] = 0.0
        x[idx, k + 1] = 1.0
    kats.append({"name": "py_beam_multichar", "fn": "beam_search", "source": "tests/test_decode.py:154-171",
                 "post": mat(x), "alphabet": multi, "beam_size": 5, "beam_cut_threshold": 0.1,
                 "collapse_repeats": True, "expect_seq": "AAACCCGGG", "expect_path": [6, 13, 18]})
    kats.append({"name": "py_viterbi_multichar", "fn": "viterbi_search", "source": "tests/test_decode.py:321-338",
                 "post": mat(x), "alphabet": multi, "qstring": False, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "AAACCCGGG", "expect_path": [6, 13, 18]})

    x = stay(20)
    for idx in [6, 13, 18]:
        x[idx:idx + 3, 0] = 0.0
        x[idx:idx + 3, 1] = 1.0
    kats.append({"name": "py_beam_spread", "fn": "beam_search", "source": "tests/test_decode.py:173-189",
                 "post": mat(x), "alphabet": nacgt, "beam_size": 5, "beam_cut_threshold": 0.1,
                 "collapse_repeats": True, "expect_seq": "AAA", "expect_path": [6, 13, 18]})

    x = stay(20)
    for r, c, v in [(3, 1, 0.99), (4, 1, 0.99), (6, 2, 0.999), (7, 2, 0.999), (9, 4, 0.6), (10, 4, 0.7),
                    (11, 4, 0.8), (13, 4, 0.4), (14, 4, 0.5), (15, 4, 0.6)]:
        x[r, 0] = 0.0
        x[r, c] = v
    kats.append({"name": "py_viterbi_mean_qscores", "fn": "viterbi_search", "source": "tests/test_decode.py:279-319",
                 "post": mat(x), "alphabet": nacgt, "qstring": True, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "ACTT5?&$", "expect_path": None})

    off = [[0.7, 0.1, 0.2], [0.7, 0.1, 0.2], [0.2, 0.3, 0.5], [0.2, 0.2, 0.6], [0.3, 0.3, 0.4],
           [0.2, 0.2, 0.6], [0.2, 0.3, 0.5], [0.7, 0.1, 0.2], [0.7, 0.1, 0.2]]
    kats.append({"name": "py_viterbi_off_path", "fn": "viterbi_search", "source": "tests/test_decode.py:340-355",
                 "post": mat(off), "alphabet": ["N", "A", "B"], "qstring": False, "qscale": 1.0, "qbias": 0.0,
                 "collapse_repeats": True, "expect_seq": "B", "expect_path": None})

    # ---- duplex: tests/test_decode.py:376-412 -----------------------------------------------------
    nab = ["N", "A", "B"]
    same = [[0.01, 0.98, 0.01]] * 4 + [[0.9, 0.05, 0.05], [0.7, 0.05, 0.35], [0.9, 0.05, 0.05]] + \
           [[0.01, 0.98, 0.01]] * 3 + [[0.01, 0.01, 0.98]] * 4
    kats.append({"name": "py_duplex_identical", "fn": "beam_search_duplex", "source": "tests/test_decode.py:376-395",
                 "post": mat(same), "post2": mat(same), "alphabet": nab, "envelope": None, "beam_size": 5,
                 "beam_cut_threshold": 0.0, "collapse_repeats": True, "expect_seq": "AAB"})
    dx = [[0.01, 0.98, 0.01], [0.01, 0.34, 0.65], [0.01, 0.98, 0.01], [0.01, 0.01, 0.98]]
    dy = [[0.0, 1.0, 0.0], [0.0, 1.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]
    kats.append({"name": "py_duplex_disagree_1d", "fn": "beam_search", "source": "tests/test_decode.py:399-405",
                 "post": mat(dx), "alphabet": nab, "beam_size": 5, "beam_cut_threshold": 0.0,
                 "collapse_repeats": True, "expect_seq": "ABAB", "expect_path": None})
    kats.append({"name": "py_duplex_disagree", "fn": "beam_search_duplex", "source": "tests/test_decode.py:406-412",
                 "post": mat(dx), "post2": mat(dy), "alphabet": nab, "envelope": None, "beam_size": 5,
                 "beam_cut_threshold": 0.0, "collapse_repeats": True, "expect_seq": "AB"})

    with open(os.path.join(HERE, "reference_kats.json"), "w") as f:
        json.dump({"reference": "nanoporetech/fast-ctc-decode v0.3.7", "kats": kats}, f, indent=0)
    print(f"wrote {len(kats)} known-answer vectors")


if __name__ == "__main__":
    main()
