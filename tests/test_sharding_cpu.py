"""Multi-process (world size 2, gloo, CPU) test of the N>1 host logic: index sharding with no
data-path collective, per-rank decoding of the shard, and the max-over-ranks time reduction that
bench.py performs.  The decode itself runs on the CPU oracle here (no GPU in this container)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_partition_the_reads():
    sys.path.insert(0, os.path.join(ROOT, "fast-ctc-decode_b200"))
    from fctc_b200.sharding import shard_range, weak_shard_first_read
    for n in (0, 1, 7, 8, 1000, 1_000_000, 262_144, 65_536):
        for world in (1, 2, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert [weak_shard_first_read(10, r) for r in range(3)] == [0, 10, 20]
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, tmpdir):
    import torch
    import torch.distributed as dist
    sys.path[:0] = [ROOT, os.path.join(ROOT, "fast-ctc-decode_b200"), os.path.join(ROOT, "tests")]
    import oracle
    from fctc_b200.sharding import shard_range
    from helpers import softmax_posteriors
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        n_reads, T, Cn = 13, 300, 5
        post = softmax_posteriors(np.random.default_rng(42), (n_reads, T, Cn))  # same bytes on every rank
        lo, hi = shard_range(n_reads, rank, world)
        _, _, _, lens, status = oracle.viterbi_batch(np.ascontiguousarray(post[lo:hi]), True)
        # results stay per rank; only test bookkeeping travels (bench.py: barrier + MAX of device time)
        mine = torch.zeros(n_reads, dtype=torch.int64)
        mine[lo:hi] = torch.from_numpy(lens)
        dist.all_reduce(mine, op=dist.ReduceOp.SUM)
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        if rank == 0:
            _, _, _, full_lens, _ = oracle.viterbi_batch(post, True)
            ok = bool((mine.numpy() == full_lens).all()) and t.item() == float(world) and (status == 0).all()
            with open(os.path.join(tmpdir, "ok"), "w") as f:
                f.write("1" if ok else "0")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_decode(tmp_path, oracle):
    import socket

    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").read_text() == "1"
