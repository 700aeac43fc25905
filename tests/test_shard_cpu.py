"""The multi-GPU path on CPU: sharding arithmetic and the gather of the result records with
world_size-2 (and 3, uneven shards) gloo process groups.  No compute calls: records are synthetic."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pathplanning_b200.shard import shard_range, shard_sizes  # noqa: E402

RECORD = np.dtype([("status", "<i4"), ("n_expanded", "<i4"), ("cost", "<f8"), ("hash", "<u8")], align=True)


def test_shard_ranges_tile_the_batch():
    for n in (0, 1, 7, 4096, 1_000_003):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            sizes = shard_sizes(n, world)
            assert sum(sizes) == n and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _records(n):
    rec = np.zeros(n, RECORD)
    rec["status"] = np.arange(n) % 3 - 1
    rec["n_expanded"] = 1000 + np.arange(n)
    rec["cost"] = np.arange(n) * 0.25
    rec["hash"] = (np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
    return rec


def _worker(rank, world, port, n, q):
    import torch.distributed as dist
    from pathplanning_b200.shard import gather_records
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = _records(n)
        lo, hi = shard_range(n, rank, world)
        got = gather_records(full[lo:hi], n)
        ok = got.tobytes() == full.tobytes()
        # a rank that passes the wrong shard is rejected before the collective
        bad = False
        try:
            gather_records(full[lo:hi + 1] if hi < n else full[lo:hi - 1], n)
        except ValueError:
            bad = True
        dist.barrier()
        q.put((rank, ok, bad))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("world,n", [(2, 4096), (2, 4097), (3, 1000)])
def test_gather_records_gloo(world, n):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in results) == list(range(world))
    assert all(ok and bad for _, ok, bad in results)
