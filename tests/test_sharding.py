"""Multi-GPU host logic on CPU: stream partitioning, and the world_size-2 gloo path bench.py
uses (barrier, max-over-ranks of the elapsed time, summed sample counts)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lpcnet_b200 import sharding


@pytest.mark.parametrize("n,world", [(1024, 1), (1024, 2), (1000, 8), (7, 4), (3, 8), (65536, 8)])
def test_shard_range_partitions_all_streams(n, world):
    seen = []
    for r in range(world):
        lo, hi = sharding.shard_range(n, r, world)
        assert 0 <= lo <= hi <= n
        seen += list(range(lo, hi))
        for s in (lo, hi - 1):
            if lo < hi:
                assert sharding.owner_of(s, n, world) == r
    assert seen == list(range(n))
    sizes = [sharding.shard_range(n, r, world)[1] - sharding.shard_range(n, r, world)[0] for r in range(world)]
    assert max(sizes) - min(sizes) <= 1


def test_aggregate_rate_uses_slowest_rank():
    assert sharding.aggregate_rate([100, 100], [1.0, 2.0]) == 100.0


def _worker(rank, world, port, n_streams, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(n_streams, rank, world)
    dist.barrier()
    elapsed = torch.tensor([1.0 + rank], dtype=torch.float64)      # pretend rank 1 is slower
    samples = torch.tensor([(hi - lo) * 160], dtype=torch.float64)
    dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    dist.all_reduce(samples, op=dist.ReduceOp.SUM)
    if rank == 0:
        q.put((float(elapsed), float(samples)))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 1001, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    elapsed, samples = q.get()
    assert elapsed == 2.0 and samples == 1001 * 160
