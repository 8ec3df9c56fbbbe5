"""The several-GPUs-in-one-call entry points (fctc_viterbi_batch_multi / fctc_beam_search_batch_multi): contiguous
read ranges, one host thread and one handle per device, results gathered into one set of host arrays in read order.
On a box with one GPU the device list names it twice, which still exercises the range arithmetic, the threads and
the rebasing of offsets; with two or more GPUs the ranges really run on different devices.  Checked against the
single-device call and the CPU oracle."""
import numpy as np
import pytest

from helpers import peaked_posteriors, softmax_posteriors

pytestmark = pytest.mark.gpu


def device_lists():
    import fctc_b200
    n = fctc_b200.lib().fctc_device_count()
    lists = [[0], [0, 0], [0, 0, 0]]
    if n >= 2:
        lists += [[0, 1], [1, 0, 1]]
    if n >= 4:
        lists += [[0, 1, 2, 3]]
    return lists


def reads_of(res, qual=False):
    out = []
    for n, (o, ln) in enumerate(zip(res["offsets"], res["lens"])):
        rec = (int(res["status"][n]), res["labels"][o:o + ln].tolist(), res["path"][o:o + ln].tolist())
        if qual:
            rec += (res["quals"][o:o + ln].tolist(),)
        out.append(rec)
    return out


@pytest.mark.parametrize("shape", [(37, 1000, 5), (5, 257, 5), (2, 64, 4), (1, 33, 5)], ids=lambda s: "x".join(map(str, s)))
def test_viterbi_multi_matches_single_device_and_oracle(oracle, shape):
    import fctc_b200
    rng = np.random.default_rng(sum(shape))
    post = softmax_posteriors(rng, shape)
    dec = fctc_b200.Decoder(0)
    want = reads_of(dec.viterbi_batch(post, True, 0.7, 2.0, True), qual=True)
    dec.close()
    ol, oq, op, olen, ost = oracle.viterbi_batch(post, True, 0.7, 2.0, True)
    for n in range(shape[0]):
        assert want[n] == (0, ol[n, :olen[n]].tolist(), op[n, :olen[n]].tolist(), oq[n, :olen[n]].tolist())
    stride = (shape[1] + 31) // 32 * 32
    for devs in device_lists():
        res = fctc_b200.viterbi_batch_multi(devs, post, True, 0.7, 2.0, True)
        assert res["offsets"].tolist() == [n * stride for n in range(shape[0])]  # slot form, read order
        assert reads_of(res, qual=True) == want, f"devices {devs}"


@pytest.mark.parametrize("beam,thr", [(5, 0.1), (32, 0.0)])
def test_beam_multi_matches_single_device_and_oracle(oracle, beam, thr):
    import fctc_b200
    rng = np.random.default_rng(beam)
    post = peaked_posteriors(rng, (23, 700, 5))
    ol, op, olen, ost, _ = oracle.beam_search_batch(post, beam, thr, True)
    want = [(int(ost[n]), ol[n, :olen[n]].tolist(), op[n, :olen[n]].tolist()) for n in range(23)]
    stride = (700 + 31) // 32 * 32
    for devs in device_lists():
        res = fctc_b200.beam_search_batch_multi(devs, post, beam, thr, True)
        assert reads_of(res) == want, f"devices {devs}"
        # range g keeps to its own part of the pool: entries [lo_g * stride, hi_g * stride)
        G = len(devs)
        for g in range(G):
            lo, hi = 23 * g // G, 23 * (g + 1) // G
            o, ln = res["offsets"][lo:hi], res["lens"][lo:hi]
            assert (o >= lo * stride).all() and (o + ln <= hi * stride).all()


def test_multi_errors():
    import fctc_b200
    post = np.zeros((4, 10, 5), np.float32)
    with pytest.raises(fctc_b200.FctcError):
        fctc_b200.viterbi_batch_multi([], post)
    with pytest.raises(fctc_b200.FctcError) as e:
        fctc_b200.viterbi_batch_multi([0, 9999], post)
    assert "9999" in str(e.value)
    with pytest.raises(fctc_b200.FctcError):
        fctc_b200.beam_search_batch_multi([0], post, beam_size=0)
    fctc_b200.lib().fctc_multi_release()
    res = fctc_b200.viterbi_batch_multi([0, 0], post)  # handles come back after a release
    assert (res["lens"] == 0).all() and (res["status"] == 0).all()
