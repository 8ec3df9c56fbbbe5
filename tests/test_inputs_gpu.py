"""Input kinds beyond float32 posteriors (include/fctc.h: FCTC_IN_*): binary16 / bfloat16 storage and logits with the
softmax fused into the kernels' load stage.

EXACT softmax: the device evaluates the fixed recipe (first maximum, glibc expf, index-order sum, IEEE division), so
decoding logits must equal, bit for bit, the CPU oracle's decode of the host-side softmax of the same (widened)
logits — labels, paths and quality characters.  FAST softmax (hardware 2^x, one reciprocal) is a few ulp off:
Viterbi labels/paths must still agree (random rows have no near-ties at that scale), quality characters may differ
by one code on a small fraction of labels, beam sequences must agree on nearly all reads."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def logits(rng, shape, scale=4.0):
    return ((rng.random(shape, dtype=np.float32) * 2 - 1) * np.float32(scale) + np.float32(1.5)).astype(np.float32)


def encode(z, elem):
    """The array handed to the library and the float32 values it stands for."""
    import fctc_b200
    if elem == fctc_b200.IN_F32:
        return z, z
    if elem == fctc_b200.IN_F16:
        h = z.astype(np.float16)
        return h, h.astype(np.float32)
    b = fctc_b200.to_bf16_bits(z)
    return b, fctc_b200.from_bf16_bits(b)


def reads(res, qual):
    out = []
    for n, (o, ln) in enumerate(zip(res["offsets"], res["lens"])):
        rec = [int(res["status"][n]), res["labels"][o:o + ln].tolist(), res["path"][o:o + ln].tolist()]
        if qual:
            rec.append(res["quals"][o:o + ln].tolist())
        out.append(rec)
    return out


def oracle_viterbi(oracle, post, qstring, qscale, qbias, collapse):
    ol, oq, op, olen, ost = oracle.viterbi_batch(post, qstring, qscale, qbias, collapse)
    out = []
    for n in range(post.shape[0]):
        rec = [int(ost[n]), ol[n, :olen[n]].tolist(), op[n, :olen[n]].tolist()]
        if qstring:
            rec.append(oq[n, :olen[n]].tolist())
        out.append(rec)
    return out


SHAPES = [(6, 1024, 5), (5, 1000, 5), (3, 777, 5), (4, 640, 4), (3, 300, 12), (2, 9, 5)]


@pytest.mark.parametrize("elem", [0, 1, 2], ids=["f32", "f16", "bf16"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_viterbi_posteriors_in_16_bit_storage(dec, oracle, shape, elem):
    """Posteriors stored as binary16 / bfloat16: identical to the oracle on the widened values."""
    rng = np.random.default_rng(17 * sum(shape) + elem)
    z = logits(rng, shape)
    post = oracle.softmax(z)
    arr, vals = encode(post, elem)
    for qstring, collapse in [(True, True), (False, False)]:
        res = dec.viterbi_batch_in(arr, elem, qstring, 0.7, 2.0, collapse)
        assert reads(res, qstring) == oracle_viterbi(oracle, vals, qstring, 0.7, 2.0, collapse)


@pytest.mark.parametrize("elem", [0, 1, 2], ids=["f32", "f16", "bf16"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_viterbi_logits_exact_softmax(dec, oracle, shape, elem):
    import fctc_b200
    rng = np.random.default_rng(23 * sum(shape) + elem)
    arr, vals = encode(logits(rng, shape), elem)
    post = oracle.softmax(vals)
    for qstring, collapse in [(True, True), (True, False), (False, True)]:
        res = dec.viterbi_batch_in(arr, elem | fctc_b200.IN_LOGITS_EXACT, qstring, 1.0, 0.0, collapse)
        assert reads(res, qstring) == oracle_viterbi(oracle, post, qstring, 1.0, 0.0, collapse)


def test_viterbi_logits_exact_special_rows(dec, oracle):
    """Rows with equal logits (ties go to the lowest label, blank first), huge spreads (exp underflows to 0), infinities
    and NaNs: the device follows the recipe through the same special values as the host."""
    import fctc_b200
    rng = np.random.default_rng(3)
    z = logits(rng, (4, 512, 5))
    z[0, ::3, :] = 0.25                        # all equal
    z[0, 1::7, 1:] = 2.0                       # labels tie above the blank
    z[1, ::2, 2] += 200.0                      # every other exp underflows
    z[1, 1::4, 0] = -np.inf                    # a label that cannot win
    z[2, ::5, 3] = np.inf                      # inf - inf: NaN row
    z[2, 2::5, 1] = np.nan
    z[3] *= 40.0
    post = oracle.softmax(z)
    res = dec.viterbi_batch_in(z, fctc_b200.IN_LOGITS_EXACT, True, 1.0, 0.0, True)
    assert reads(res, True) == oracle_viterbi(oracle, post, True, 1.0, 0.0, True)


@pytest.mark.parametrize("elem", [0, 2], ids=["f32", "bf16"])
def test_viterbi_logits_fast_softmax(dec, oracle, elem):
    import fctc_b200
    rng = np.random.default_rng(31 + elem)
    arr, vals = encode(logits(rng, (8, 2048, 5)), elem)
    post = oracle.softmax(vals)
    want = oracle_viterbi(oracle, post, True, 1.0, 0.0, True)
    got = reads(dec.viterbi_batch_in(arr, elem | fctc_b200.IN_LOGITS_FAST, True, 1.0, 0.0, True), True)
    n_lab = n_off = 0
    for g, w in zip(got, want):
        assert g[:3] == w[:3]                  # status, labels, emission rows
        d = np.abs(np.asarray(g[3], np.int64) - np.asarray(w[3], np.int64))
        assert d.max(initial=0) <= 1
        n_lab += d.size
        n_off += int((d != 0).sum())
    assert n_off <= max(2, n_lab // 200), f"{n_off} of {n_lab} quality characters differ"
    # without quality strings no probability is used: the logits are decoded as they are, exactly
    got = reads(dec.viterbi_batch_in(arr, elem | fctc_b200.IN_LOGITS_FAST, False, 1.0, 0.0, True), False)
    assert got == [w[:3] for w in want]


@pytest.mark.parametrize("elem", [0, 1, 2], ids=["f32", "f16", "bf16"])
@pytest.mark.parametrize("beam,thr", [(5, 0.1), (32, 0.0), (8, 0.0)])
def test_beam_logits_exact_softmax(dec, oracle, beam, thr, elem):
    import fctc_b200
    rng = np.random.default_rng(41 * beam + elem)
    for shape in [(7, 600, 5), (3, 333, 4)]:
        arr, vals = encode(logits(rng, shape, 3.0), elem)
        post = oracle.softmax(vals)
        ol, op, olen, ost, _ = oracle.beam_search_batch(post, beam, thr if shape[2] == 5 else 0.0, True)
        want = [[int(ost[n]), ol[n, :olen[n]].tolist(), op[n, :olen[n]].tolist()] for n in range(shape[0])]
        res = dec.beam_search_batch_in(arr, elem | fctc_b200.IN_LOGITS_EXACT, beam, thr if shape[2] == 5 else 0.0, True)
        assert reads(res, False) == want
        # posteriors in the same storage type
        parr, pvals = encode(post, elem)
        ol, op, olen, ost, _ = oracle.beam_search_batch(pvals, beam, 0.0, True)
        want = [[int(ost[n]), ol[n, :olen[n]].tolist(), op[n, :olen[n]].tolist()] for n in range(shape[0])]
        assert reads(dec.beam_search_batch_in(parr, elem, beam, 0.0, True), False) == want


def test_beam_logits_fast_softmax(dec, oracle):
    import fctc_b200
    rng = np.random.default_rng(77)
    z = logits(rng, (64, 500, 5), 3.0)
    post = oracle.softmax(z)
    ol, op, olen, ost, _ = oracle.beam_search_batch(post, 5, 0.1, True)
    res = dec.beam_search_batch_in(z, fctc_b200.IN_LOGITS_FAST, 5, 0.1, True)
    same = sum(res["labels"][o:o + ln].tolist() == ol[n, :olen[n]].tolist()
               for n, (o, ln) in enumerate(zip(res["offsets"], res["lens"])))
    assert same >= 60, f"only {same}/64 reads identical with the fast softmax"


def test_input_kind_errors(dec):
    import fctc_b200
    z = np.zeros((2, 16, 5), np.float32)
    with pytest.raises(fctc_b200.FctcError):
        dec.viterbi_batch_in(z, 3)
    with pytest.raises(fctc_b200.FctcError):
        dec.viterbi_batch_in(z, 0x30)
    with pytest.raises(TypeError):
        dec.viterbi_batch_in(z, fctc_b200.IN_F16)
