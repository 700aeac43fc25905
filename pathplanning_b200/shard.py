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


def shard_indices(n, rank, world, layout="block", cost=None):
    """Unit indices of rank `rank`.  layout "block": the contiguous block of shard_range; "strided": rank, rank + world,
    ... (independent random units: every rank gets the same mix of short and long ones); with `cost` (expected cost
    per unit, any positive scale): greedy longest-first assignment to the least loaded rank, ties to the lower rank --
    deterministic, every rank computes the same partition."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    if cost is not None:
        cost = np.asarray(cost, np.float64)
        if cost.shape != (n,):
            raise ValueError("cost must have one entry per unit")
        load = np.zeros(world)
        count = np.zeros(world, np.int64)
        owner = np.empty(n, np.int64)
        cap = -(-n // world)  # equal shard sizes (+-1) so that one fixed-size gather carries the records
        for i in np.argsort(-cost, kind="stable"):
            open_ = np.flatnonzero(count < cap)
            r = open_[np.argmin(load[open_])]
            owner[i] = r
            load[r] += cost[i]
            count[r] += 1
        return np.flatnonzero(owner == rank)
    if layout == "block":
        lo, hi = shard_range(n, rank, world)
        return np.arange(lo, hi)
    if layout == "strided":
        return np.arange(rank, n, world)
    raise ValueError(f"unknown layout {layout}")


def gather_records(local, n_total, device=None, group=None, indices=None):
    """all_gather of a 1-D structured numpy array of records (the rank's shard, in shard_range order, or in the order of
    `indices` = this rank's shard_indices(...)).  Returns the n_total records in global order on every rank.  Shards may
    differ in size: every rank pads to the largest shard so that one fixed-size collective carries them; with `indices`
    the index lists travel the same way."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if indices is not None:
        indices = np.ascontiguousarray(indices, np.int64)
        if len(indices) != len(local):
            raise ValueError(f"rank {rank} holds {len(local)} records for {len(indices)} indices")
        if world == 1:
            out = np.empty(n_total, local.dtype)
            out[indices] = local
            return out
        cap = -(-n_total // world) + 1
        item = local.dtype.itemsize
        buf = np.zeros(8 + cap * (item + 8), np.uint8)
        buf[:8] = np.array([len(local)], np.int64).view(np.uint8)
        buf[8:8 + len(local) * 8] = indices.view(np.uint8)
        buf[8 + cap * 8:8 + cap * 8 + len(local) * item] = np.ascontiguousarray(local).view(np.uint8).reshape(-1)
        t = torch.from_numpy(buf)
        if device is not None:
            t = t.to(device)
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t, group=group)
        out = np.empty(n_total, local.dtype)
        seen = 0
        for r in range(world):
            b = parts[r].cpu().numpy()
            m = int(b[:8].view(np.int64)[0])
            idx = b[8:8 + m * 8].view(np.int64)
            out[idx] = b[8 + cap * 8:8 + cap * 8 + m * item].view(local.dtype)
            seen += m
        if seen != n_total:
            raise ValueError(f"gathered {seen} records, expected {n_total}")
        return out
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
