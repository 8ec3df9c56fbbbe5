#!/usr/bin/env python3
"""What the host side gives for the byte counts of one end-to-end Viterbi step: the ceiling of bench.py's e2e number.

Every rank copies 2.62 GB host -> device (32,768 reads x 4000 x 5 float32) and 0.50 GB device -> host (the packed
results of the same reads) between pinned host buffers and its own GPU, on two streams, all ranks at once — bare
cudaMemcpyAsync, no kernel.  Run alone for one GPU, or under torchrun for N ranks:

  python tools/pcie_probe.py
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py

Prints one JSON line (rank 0): per-rank and aggregate GB/s for H2D alone and for H2D + D2H together, the NUMA node
of every rank's GPU and the cores it ran on.  With --bind (default) a rank is pinned to its GPU's NUMA node before
it allocates the pinned buffers, as bench.py does; --no-bind shows the unbound figure."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-bind", action="store_true")
    ap.add_argument("--h2d-bytes", type=int, default=32768 * 4000 * 5 * 4)
    ap.add_argument("--d2h-bytes", type=int, default=32768 * 2560 * 6)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    torch.cuda.set_device(local_rank)
    binding = {"numa_node": None, "cores": len(os.sched_getaffinity(0)), "note": "unbound (--no-bind)"}
    if not args.no_bind:
        import bench
        binding = bench.bind_to_gpu_numa_node(torch, local_rank, local_world)
    if world > 1:
        dist.init_process_group("gloo")
    h = torch.empty(args.h2d_bytes, dtype=torch.uint8).pin_memory()
    h.fill_(1)  # first touch on this rank's cores
    ho = torch.empty(args.d2h_bytes, dtype=torch.uint8).pin_memory()
    ho.fill_(1)
    d = torch.empty(args.h2d_bytes, dtype=torch.uint8, device="cuda")
    o = torch.empty(args.d2h_bytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    out = {}
    for both in (False, True):
        best = 1e9
        for _ in range(4):
            barrier()
            t0 = time.perf_counter()
            with torch.cuda.stream(s1):
                d.copy_(h, non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    ho.copy_(o, non_blocking=True)
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        out["both" if both else "h2d_only"] = best
    mine = {"rank": rank, "h2d_only_gbs": args.h2d_bytes / out["h2d_only"] / 1e9,
            "both_h2d_gbs": args.h2d_bytes / out["both"] / 1e9, "both_d2h_gbs": args.d2h_bytes / out["both"] / 1e9,
            "both_ms": out["both"] * 1e3, "binding": binding}
    allr = [mine]
    if world > 1:
        allr = [None] * world
        dist.all_gather_object(allr, mine)
    if rank == 0:
        slow = max(r["both_ms"] for r in allr)
        print(json.dumps({"ranks": world, "h2d_bytes_per_rank": args.h2d_bytes, "d2h_bytes_per_rank": args.d2h_bytes,
                          "aggregate_h2d_gbs_h2d_only": sum(r["h2d_only_gbs"] for r in allr),
                          "aggregate_h2d_gbs_both": world * args.h2d_bytes / (slow * 1e-3) / 1e9,
                          "aggregate_d2h_gbs_both": world * args.d2h_bytes / (slow * 1e-3) / 1e9,
                          "slowest_rank_ms": slow,
                          "rows_per_s_ceiling": world * 32768 * 4000 / (slow * 1e-3), "per_rank": allr}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
