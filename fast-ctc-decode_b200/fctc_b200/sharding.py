"""Static read-index sharding: reads are independent, so N GPUs take contiguous
index ranges and never exchange data.  Pure index arithmetic, shared by bench.py and the tests."""


def shard_range(n_reads, rank, world):
    """Contiguous range [lo, hi) of rank `rank` when n_reads are split over `world` ranks
    ([g*N/G, (g+1)*N/G) — sizes differ by at most one read)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    lo = (n_reads * rank) // world
    hi = (n_reads * (rank + 1)) // world
    return lo, hi


def weak_shard_first_read(reads_per_gpu, rank):
    """First global read index of rank `rank` in the weak-scaling benchmark (every rank decodes
    reads_per_gpu reads of one global job)."""
    return reads_per_gpu * rank
