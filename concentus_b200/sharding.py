"""Multi-GPU plumbing for the batch decoder: streams are independent, so a job is sharded by
stream index with no data-path collective.  One process per GPU; torch.distributed is
used only to agree on timing (barrier + max over ranks), as bench.py's contract requires."""


def shard_streams(n_streams, rank, world):
    """Contiguous block of stream indices owned by `rank` (blocks differ by at most one stream).
    A multistream (surround) decoder counts as one stream, so its elementary streams stay together."""
    base, extra = divmod(n_streams, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def aggregate_throughput(units_this_rank, seconds_this_rank, dist=None):
    """Whole-job throughput: all ranks' units divided by the slowest rank's time."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_this_rank / seconds_this_rank, seconds_this_rank
    import torch
    t = torch.tensor([seconds_this_rank], dtype=torch.float64)
    u = torch.tensor([units_this_rank], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t, u = t.cuda(), u.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(u.item()) / float(t.item()), float(t.item())
