"""N>1 host logic on CPU: two gloo ranks shard a stream set, agree on the slowest rank's time and
report the whole-job throughput (the path bench.py takes under torchrun)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from concentus_b200.sharding import aggregate_throughput, shard_streams


def test_shards_partition_the_streams():
    for n in (1, 7, 4096, 1025):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                seen.extend(shard_streams(n, r, world))
            assert seen == list(range(n))
            sizes = [len(shard_streams(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard_streams(1001, rank, world)
    seconds = 2.0 if rank == 0 else 3.0            # rank 1 is the slow one
    value, t = aggregate_throughput(float(len(mine)), seconds, dist)
    counts = torch.tensor([len(mine)], dtype=torch.int64)
    dist.all_reduce(counts)
    out[rank] = (value, t, int(counts.item()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_aggregate():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, 29533, out), nprocs=world, join=True)
    for r in range(world):
        value, t, total = out[r]
        assert total == 1001
        assert t == 3.0
        assert abs(value - 1001 / 3.0) < 1e-9
