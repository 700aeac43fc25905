"""ppx_comm_*  on two GPUs of one box: the world model broadcast from rank 0, uneven shards of fixed-size
records gathered on every rank, variable-length pose arrays gathered by counts -> offsets.  One process per GPU, the
128-byte NCCL id handed over by the parent.  Skipped on a single-GPU box (the round's scaling run covers N > 1)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rank_main(rank, world, uid, q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import torch
        import pathplanning_b200 as pp
        from pathplanning_b200.lib import ppx
        from common import load_npz, c2_world_desc
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        ctx = pp.Context(rank)
        buf = (C.c_char * 128).from_buffer_copy(uid)
        ctx._check(ppx.ppx_comm_init(ctx.handle, buf, rank, world))
        g = load_npz("edges_c2.npz")
        # 1. the map: only rank 0 uploads it
        if rank == 0:
            desc, occ, sq, pc = c2_world_desc(g)
            ctx.upload_map(desc, sq, occ, pc)
        ctx._check(ppx.ppx_comm_broadcast_map(ctx.handle, 0))
        ctx._check(ppx.ppx_ctx_synchronize(ctx.handle))
        ctx.map = c2_world_desc(g)[0]
        got = ctx.is_state_valid(g["sv_pts"], pp.ValidatorParams(1.0, 0.1))
        ok_map = bool(np.array_equal(got, g["sv_valid"]))
        edges = pp.make_rs_edges(g["starts"], g["steer"], g["dir"], g["len"])
        v, l = ctx.check_edges_rs(edges, 2.0, pp.ValidatorParams(1.0, 0.1))
        ok_map = ok_map and bool(np.array_equal(v, g["r10_valid"]))
        # 2. uneven shards of 24-byte records
        rec = np.dtype([("a", "<i8"), ("b", "<f8"), ("c", "<u8")])
        sizes = [1000 + 7 * r for r in range(world)]

        def records(r):
            x = np.zeros(sizes[r], rec)
            x["a"] = np.arange(sizes[r]) + 100000 * r
            x["b"] = np.sqrt(np.arange(sizes[r]) + r)
            x["c"] = np.uint64(0x9E3779B97F4A7C15) * (np.arange(sizes[r], dtype=np.uint64) + np.uint64(r))
            return x
        local = torch.from_numpy(records(rank).view(np.uint8).copy()).to(dev)
        total = sum(sizes)
        out = torch.zeros(total * rec.itemsize, dtype=torch.uint8, device=dev)
        counts = (C.c_int64 * world)()
        ctx._check(ppx.ppx_comm_gather_results(ctx.handle, local.data_ptr(), sizes[rank], rec.itemsize, out.data_ptr(), total, counts))
        ctx._check(ppx.ppx_ctx_synchronize(ctx.handle))
        want = np.concatenate([records(r) for r in range(world)])
        ok_rec = list(counts) == sizes and out.cpu().numpy().view(rec).tobytes() == want.tobytes()
        # 3. variable-length pose arrays
        maxp = 37

        def paths(r):
            rng = np.random.default_rng(5 + r)
            n = 50 + 11 * r
            cnt = rng.integers(-1, maxp + 1, n).astype(np.int32)  # -1: no path
            poses = rng.normal(size=(n, maxp, 3))
            return cnt, poses
        cnt, poses = paths(rank)
        d_cnt = torch.from_numpy(cnt).to(dev)
        d_poses = torch.from_numpy(poses).to(dev)
        allp = [paths(r) for r in range(world)]
        tq = sum(len(c) for c, _ in allp)
        tp = sum(int(np.clip(c, 0, maxp).sum()) for c, _ in allp)
        d_off = torch.zeros(tq + 1, dtype=torch.int64, device=dev)
        d_all = torch.zeros(tp * 3, dtype=torch.float64, device=dev)
        qc = (C.c_int64 * world)()
        total_poses = C.c_int64()
        ctx._check(ppx.ppx_comm_gather_paths(ctx.handle, len(cnt), d_cnt.data_ptr(), d_poses.data_ptr(), maxp, qc, d_off.data_ptr(), tq + 1,
                                             d_all.data_ptr(), tp, C.byref(total_poses)))
        ctx._check(ppx.ppx_ctx_synchronize(ctx.handle))
        want_off = np.concatenate([[0], np.cumsum(np.concatenate([np.clip(c, 0, maxp) for c, _ in allp]))])
        want_poses = np.concatenate([p[i, :max(0, c[i])] for c, p in allp for i in range(len(c))])
        ok_paths = (total_poses.value == tp and np.array_equal(d_off.cpu().numpy(), want_off)
                    and np.array_equal(d_all.cpu().numpy().reshape(-1, 3), want_poses))
        ctx._check(ppx.ppx_comm_destroy(ctx.handle))
        ctx.close()
        q.put((rank, ok_map, bool(ok_rec), bool(ok_paths), ""))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, False, False, False, traceback.format_exc()))


def test_comm_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from pathplanning_b200.lib import ppx
    uid = (C.c_char * 128)()
    assert ppx.ppx_comm_unique_id(uid) == 0
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, world, bytes(uid), q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, ok_map, ok_rec, ok_paths, err in sorted(results):
        assert not err, err
        assert ok_map and ok_rec and ok_paths, (rank, ok_map, ok_rec, ok_paths)
