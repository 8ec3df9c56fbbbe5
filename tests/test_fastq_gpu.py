"""FASTQ record assembly on the device: the text produced from the
Viterbi kernel's slot-form result equals the records a host loop builds from the same arrays."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ALPHABET = "NACGT"


@pytest.fixture(scope="module")
def dec():
    import fctc_b200
    d = fctc_b200.Decoder(0)
    yield d
    d.close()


def host_records(labels, quals, offsets, lens, status, prefix, first):
    """CPU restatement: what a caller of the reference does with (seq + qual) per read."""
    out, offs = bytearray(), []
    for n in range(len(lens)):
        offs.append(len(out))
        if status[n] != 0:
            continue
        o, ln = int(offsets[n]), int(lens[n])
        seq = "".join(ALPHABET[l] for l in labels[o:o + ln])
        qual = bytes(quals[o:o + ln]).decode("latin-1")
        out += f"@{prefix}{first + n}\n{seq}\n+\n{qual}\n".encode("latin-1")
    offs.append(len(out))
    return bytes(out), np.asarray(offs, np.int64)


@pytest.mark.parametrize("N,T,first,prefix", [(1, 40, 0, "r"), (77, 333, 9, "read_"), (5000, 120, 999_999_990, ""),
                                               (3000, 1000, 7, "run42/ch7/")])
def test_records_match_host_assembly(dec, N, T, first, prefix):
    import torch
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    post = torch.empty((N, T, 5), dtype=torch.float32, device=dev)
    dec.synth_posteriors_device(post.data_ptr(), N, T, 5, 0, 11 + N, 0, st)
    stride = (T + 31) // 32 * 32
    cap = N * stride
    labels = torch.empty(cap, dtype=torch.uint8, device=dev)
    quals = torch.empty(cap, dtype=torch.uint8, device=dev)
    path = torch.empty(cap, dtype=torch.int32, device=dev)
    offsets = torch.empty(N, dtype=torch.int64, device=dev)
    lens = torch.empty(N, dtype=torch.int32, device=dev)
    status = torch.empty(N, dtype=torch.int32, device=dev)
    dec.viterbi_batch_device(post.data_ptr(), N, T, 5, True, 1.0, 0.0, True, labels.data_ptr(), quals.data_ptr(), 1,
                             path.data_ptr(), cap, offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), st)
    if N > 3:
        status[3] = 2  # a failed read leaves no record
        status[N - 1] = 1
    out_cap = int(2 * cap + N * (32 + len(prefix)))
    out = torch.zeros(out_cap, dtype=torch.uint8, device=dev)
    rec = torch.empty(N + 1, dtype=torch.int64, device=dev)
    nbytes = dec.fastq_scratch_bytes(N)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dec.fastq_records_device(labels.data_ptr(), quals.data_ptr(), offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), N,
                             ALPHABET, prefix, first, out.data_ptr(), out_cap, rec.data_ptr(), scratch.data_ptr(), nbytes, st)
    torch.cuda.synchronize()
    want, want_offs = host_records(labels.cpu().numpy(), quals.cpu().numpy(), offsets.cpu().numpy(), lens.cpu().numpy(),
                                   status.cpu().numpy(), prefix, first)
    got_offs = rec.cpu().numpy()
    assert np.array_equal(got_offs, want_offs)
    assert bytes(out[:int(got_offs[-1])].cpu().numpy()) == want

    # too small a buffer: nothing is written, the needed size is still reported
    small = torch.zeros(16, dtype=torch.uint8, device=dev)
    dec.fastq_records_device(labels.data_ptr(), quals.data_ptr(), offsets.data_ptr(), lens.data_ptr(), status.data_ptr(), N,
                             ALPHABET, prefix, first, small.data_ptr(), 16, rec.data_ptr(), scratch.data_ptr(), nbytes, st)
    torch.cuda.synchronize()
    assert int(rec[-1]) == len(want) and (len(want) <= 16 or int(small.sum()) == 0)
