"""Sharding of independent units (edges, start/goal queries) over the ranks of one box, and the gather
of the fixed-size result records -- the only collective of the path.

The path has no exchange step: rank r owns the contiguous block [r*n/G, (r+1)*n/G) of the units, maps and
tables are replicated, and results come back with one all_gather of the per-query records
(torch.distributed: NCCL on the GPUs, gloo in the CPU tests)."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous block of rank `rank` out of `world`: sizes differ by at most one, blocks tile [0, n)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return rank * n // world, (rank + 1) * n // world


def shard_sizes(n, world):
    return [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]


def gather_records(local, n_total, device=None, group=None):
    """all_gather of a 1-D structured numpy array of records (the rank's shard, in shard_range order).
    Returns the n_total records in global order on every rank.  Shards may differ in size by one: every
    rank pads to the largest shard so that one fixed-size collective carries them."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    sizes = shard_sizes(n_total, world)
    if len(local) != sizes[rank]:
        raise ValueError(f"rank {rank} holds {len(local)} records, its shard has {sizes[rank]}")
    if world == 1:
        return local.copy()
    item = local.dtype.itemsize
    cap = max(sizes)
    buf = np.zeros(cap * item, np.uint8)
    buf[:len(local) * item] = np.ascontiguousarray(local).view(np.uint8).reshape(-1)
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t, group=group)
    out = np.empty(n_total, local.dtype)
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        out[lo:hi] = parts[r].cpu().numpy()[:(hi - lo) * item].view(local.dtype)
    return out
