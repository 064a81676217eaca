"""N>1 host logic on CPU: world_size-2 gloo processes shard a batch by contiguous range, each rank evaluates its
shard (with the oracle standing in as the checker -- there is no GPU here), and the concatenation equals the
unsharded evaluation; the timing/throughput aggregation takes the max time and the total point count."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, load_golden


def test_shard_ranges_partition_everything():
    from baobzi_b200.shard import shard_range
    for n in (0, 1, 7, 1000, 10**9 + 7):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, world, k) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, path, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from baobzi_b200.shard import aggregate, shard_range
    from oracle import ref
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(123)  # same batch on every rank; each evaluates only its range
    x = a["lower_left"] + rng.random((10001, t.dim)) * (a["upper_right"] - a["lower_left"])
    lo, hi = shard_range(len(x), world, rank)
    y = t.eval_multi(x[lo:hi])
    np.save(os.path.join(out_dir, "y%d.npy" % rank), y)
    ms, pts = aggregate(10.0 * (rank + 1), hi - lo, dist)
    np.save(os.path.join(out_dir, "agg%d.npy" % rank), np.array([ms, pts]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_evaluation_equals_unsharded(ref, tmp_path):
    import torch.multiprocessing as mp
    path, _ = load_golden("t2d_o8")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = 2
    mp.start_processes(_worker, args=(world, port, path, str(tmp_path)), nprocs=world, join=True, start_method="spawn")
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(123)
    x = a["lower_left"] + rng.random((10001, t.dim)) * (a["upper_right"] - a["lower_left"])
    whole = t.eval_multi(x)
    parts = np.concatenate([np.load(tmp_path / ("y%d.npy" % k)) for k in range(world)])
    assert np.array_equal(parts.view(np.int64), whole.view(np.int64))
    for k in range(world):
        ms, pts = np.load(tmp_path / ("agg%d.npy" % k))
        assert ms == 20.0 and pts == 10001  # max over ranks, sum over ranks
