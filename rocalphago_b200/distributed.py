"""Multi-GPU plumbing: one process per GPU, positions sharded by rank, no collective on the data path.

Positions are independent, so a batch is split into contiguous per-rank slices, every rank
evaluates its slice with its own replica of the weights, and only the results (S*S floats per
position for a policy, one for a value) are gathered -- with NCCL over NVLink on GPUs, with gloo
in the CPU tests of this host logic.
"""
import torch
import torch.distributed as dist


def shard_bounds(n, rank, world):
    """[begin, end) of rank's contiguous slice of n items; sizes differ by at most one and the
    larger slices come first."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("rank %d of world %d" % (rank, world))
    base, extra = divmod(int(n), world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_sizes(n, world):
    return [shard_bounds(n, r, world)[1] - shard_bounds(n, r, world)[0] for r in range(world)]


def gather_results(local, n_total, group=None):
    """All ranks contribute their slice of the results (rows in shard_bounds order); every rank
    gets the full (n_total, ...) tensor back.  Uneven slices are padded for the collective."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    rank = dist.get_rank(group)
    sizes = shard_sizes(n_total, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError("rank %d holds %d rows, its shard has %d" % (rank, local.shape[0], sizes[rank]))
    width = max(sizes)
    padded = local
    if local.shape[0] < width:
        pad = torch.zeros((width - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        padded = torch.cat([local, pad], dim=0)
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded.contiguous(), group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], dim=0)


def evaluate_sharded(eval_fn, packed_all, group=None):
    """eval_fn(packed_slice) -> results for that slice; returns the gathered results on every rank.
    `packed_all` is the full batch (every rank holds or can index it); each rank touches only its slice."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    b, e = shard_bounds(packed_all.shape[0], rank, world)
    return gather_results(eval_fn(packed_all[b:e]), packed_all.shape[0], group=group)
