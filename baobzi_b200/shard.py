"""Point sharding for the multi-GPU path: the tree is replicated, the points of a batch are split into
contiguous ranges, one per rank; there is no data-path collective.  Only timing and
throughput aggregation cross ranks (max of the per-rank device times, sum of the per-rank point counts)."""


def shard_range(n, world, rank):
    """Contiguous range [lo, hi) of rank `rank` out of `world` over n points: [k*n/N, (k+1)*n/N)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return (rank * n) // world, ((rank + 1) * n) // world


def aggregate(local_ms, local_points, dist=None):
    """(max over ranks of the elapsed ms, total points) -- via torch.distributed when initialised."""
    if dist is None or not dist.is_initialized():
        return float(local_ms), int(local_points)
    import torch
    t = torch.tensor([float(local_ms)], dtype=torch.float64)
    p = torch.tensor([int(local_points)], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(p, op=dist.ReduceOp.SUM)
    return float(t.item()), int(p.item())
