"""BASELINE.json's full sizes, checked through size-independent properties (the oracle finishes
fractions of these batches in seconds, the full batches in minutes):

* C3, 1 M Reeds-Shepp edges: per-edge results do not depend on the batch they are in (permutation,
  prefix), the FP32 cell filter changes no bit against the all-exact evaluation, and ALL 1 M edges are
  solved and checked by the compiled reference, forked over the host cores (word, verdict, sample count,
  last ratio).
* C4, 4096 Hybrid A* queries: a query's record does not depend on the shard it is searched in
  (what makes sharding by query over 1/2/4/8 GPUs result-neutral), and the first 1024 queries are searched by
  the reference on all host cores: status, expansion count and the CRC of the expansion sequence.
"""
import multiprocessing as mp
import os
import zlib
import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_npz
from pathplanning_b200 import synth

pytestmark = pytest.mark.gpu
N = 1_000_000


def _cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def _fork_map(fn, chunks):
    """fn(chunk) in one forked process per chunk (the reference keeps process-global state: RNG, logger)."""
    ctxmp = mp.get_context("fork")
    procs = []
    for c in chunks:
        a, b = ctxmp.Pipe(duplex=False)

        def run(conn=b, chunk=c):
            try:
                conn.send(fn(chunk))
            except Exception as e:  # pragma: no cover
                conn.send(e)
            conn.close()
        p = ctxmp.Process(target=run)
        p.start()
        procs.append((p, a))
    out = []
    for p, a in procs:
        r = a.recv()
        p.join()
        if isinstance(r, Exception):
            raise r
        out.append(r)
    return out


@pytest.fixture(scope="module")
def c3_full(ctx):
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    ctx.build_map(desc, occ)
    s, g = synth.random_rs_queries(N, seed=1000)  # batch 0 of bench.py
    sol = ctx.rs_solve(s, g, pp.RSCostParams(2.0))
    return s, g, sol["edges"]


def test_c3_full_batch_properties(ctx, c3_full, ref):
    s, g, edges = c3_full
    vp = pp.ValidatorParams(1.0, 0.1)
    v, l, ns = ctx.check_edges_rs(edges, 2.0, vp, want_samples=True)
    assert len(v) == N and 0.05 < v.mean() < 0.2 and 4.0 < ns.mean() < 6.0
    assert np.all(l[v] == 1.0) and np.all((l[~v] >= 0.0) & (l[~v] < 1.0))
    # idempotence + permutation invariance
    perm = np.random.default_rng(0).permutation(N)
    v2, l2, ns2 = ctx.check_edges_rs(edges[perm], 2.0, vp, want_samples=True)
    assert np.array_equal(v2, v[perm]) and np.array_equal(ns2, ns[perm])
    assert np.array_equal(l2.view(np.uint32), l[perm].view(np.uint32))
    # a prefix checked alone gives the prefix of the results
    v3, l3, ns3 = ctx.check_edges_rs(edges[:123457], 2.0, vp, want_samples=True)
    assert np.array_equal(v3, v[:123457]) and np.array_equal(ns3, ns[:123457])
    # the filter changes no bit against the all-exact evaluation of the full batch
    ctx.set_option("rs_filter", 0)
    try:
        v4, l4, ns4 = ctx.check_edges_rs(edges, 2.0, vp, want_samples=True)
    finally:
        ctx.set_option("rs_filter", 1)
    assert np.array_equal(v4, v) and np.array_equal(ns4, ns) and np.array_equal(l4.view(np.uint32), l.view(np.uint32))
    # fused solve + check agrees with check-only on pre-solved edges
    r = ctx.connect_check_rs(s, g, pp.RSCostParams(2.0), vp)
    assert np.array_equal(r["valid"], v) and np.array_equal(r["last"].view(np.uint32), l.view(np.uint32))
    # ALL edges against the compiled reference (Connect + IsPathValid), one process per host core
    cfg = synth.C3
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for rect in synth.config_rects(cfg):
        w.add_rectangle(*rect)
    w.update()
    word = ctx.rs_solve(s, g, pp.RSCostParams(2.0))["word"]
    k = _cores()
    bounds = [(i * N // k, (i + 1) * N // k) for i in range(k)]

    def part(b):
        rr = w.check_rs_edges(s[b[0]:b[1]], g[b[0]:b[1]], 2.0)
        return rr["valid"], rr["samples"].astype(np.int64), rr["last"], rr["word"]
    parts = _fork_map(part, bounds)
    w.close()
    rv = np.concatenate([p[0] for p in parts])
    rs_ = np.concatenate([p[1] for p in parts])
    rl = np.concatenate([p[2] for p in parts])
    rw = np.concatenate([p[3] for p in parts])
    assert len(rv) == N
    assert np.array_equal(rw, word)
    assert np.array_equal(rv, v)
    assert np.array_equal(rs_, ns.astype(np.int64))
    assert np.allclose(rl, l, rtol=1e-6, atol=0)


def test_c4_full_batch_is_shard_independent(ctx, ref):
    cfg = synth.C2
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for rect in synth.config_rects(cfg):
        w.add_rectangle(*rect)
    w.update()
    occ, sq = w.grids()
    pc, _ = w.gvd_update()
    ctx.upload_map(w.map_desc(), sq, occ, pc)
    vp = pp.ValidatorParams(1.0, 0.1)
    rng = np.random.default_rng(12345)
    nq = 4096
    pts = np.empty((0, 3))
    while len(pts) < 2 * nq:
        p = np.stack([rng.uniform(-18.5, 18.5, 4 * nq), rng.uniform(-18.5, 18.5, 4 * nq), rng.uniform(-np.pi, np.pi, 4 * nq)], 1)
        pts = np.concatenate([pts, p[ctx.is_state_valid(p, vp)]])
    starts, goals = pts[:nq].copy(), pts[nq:2 * nq].copy()
    seeds = (1000 + np.arange(nq)).astype(np.uint64)
    planner = pp.HybridAStar(ctx, pp.HybridParams(), vp)
    full = planner.search(starts, goals, seeds)["results"]
    assert 0.8 < (full["status"] == 0).mean() < 1.0
    # the first 1024 queries on the reference (same field: the reference's own grids were uploaded), all host cores
    nref = 1024
    logged = planner.search(starts[:nref], goals[:nref], seeds[:nref], max_log=12000)
    assert logged["results"].tobytes() == full[:nref].tobytes()
    k = _cores()

    def part(c):
        H = ref.Hybrid(w)
        out = []
        for i in range(c, nref, k):
            r = H.search(starts[i], goals[i], int(seeds[i]))
            out.append((i, r["status"], r["n_expanded"], zlib.crc32(np.ascontiguousarray(r["exp_keys"][:, :3], np.int32).tobytes())))
        H.close()
        return out
    recs = sorted(x for p in _fork_map(part, list(range(k))) for x in p)
    assert len(recs) == nref
    bad = []
    for i, st, nexp, crc in recs:
        res = full[i]
        good = int(res["status"]) == st and int(res["n_expanded"]) == nexp and nexp <= 12000
        good = good and zlib.crc32(np.ascontiguousarray(logged["log"][i]["key"][:nexp], np.int32).tobytes()) == crc
        if not good:
            bad.append(i)
    assert not bad, f"{len(bad)} of {nref} queries expand differently from the reference: {bad[:10]}"
    # shards of 8 "GPUs", searched one after the other on this one: identical records
    for r in (0, 3, 7):
        lo, hi = r * nq // 8, (r + 1) * nq // 8
        part = planner.search(starts[lo:hi], goals[lo:hi], seeds[lo:hi])["results"]
        assert part.tobytes() == full[lo:hi].tobytes()
    # N3 / N2 at the same size, through properties that do not need the reference: the pose array of every solved
    # query starts at the start pose, ends within one step of the goal, has no gap above 1.5 steps, its first
    # sample is a cusp index; the smoother (contracting parameter set) keeps the end points, moves interior points
    # by less than a metre and returns NaN only with status Failure; shards give the same arrays as the full batch.
    out = planner.search(starts, goals, seeds, max_path=192)
    assert int(out["results"]["path_len"].max()) <= 192
    rs = planner.resample(goals, out["results"], out["paths"], 0.1, max_points=1024)
    cnt = rs["count"]
    solved = out["results"]["status"] == 0
    assert ((cnt > 0) == (solved & (out["results"]["path_len"] >= 2))).all() and (cnt >= 0).all()
    for q in np.flatnonzero(cnt > 0):
        p = rs["poses"][q, :cnt[q]]
        assert np.allclose(p[0], starts[q], atol=1e-12)
        assert np.hypot(*(p[-1, :2] - goals[q, :2])) <= 0.1 + 1e-9
        gaps = np.hypot(*np.diff(p[:, :2], axis=0).T)
        c = rs["cusp"][q, :cnt[q]]
        regular = ~(c[:-1] | c[1:])  # a snapped cusp sample may sit anywhere (upstream quirk): skip its two gaps
        assert gaps[regular].max(initial=0.0) <= 0.2 + 1e-9
        assert rs["cusp"][q, 0] and not rs["cusp"][q, cnt[q]:].any()
    part = planner.resample(goals[512:1024], out["results"][512:1024], out["paths"][512:1024], 0.1, max_points=1024)
    assert np.array_equal(part["poses"], rs["poses"][512:1024]) and np.array_equal(part["count"], cnt[512:1024])
    gold = load_npz("smooth.npz")
    sm = pp.Smoother(ctx, pp.SmootherParams(max_curvature=0.5, curvature_weight=0.0, path_weight=0.1, max_iterations=200))
    sm.set_gvd(gold["c2_near_obstacle"].astype(np.int32), gold["c2_near_edge"].astype(np.int32))
    so = sm.smooth(rs["poses"], rs["cusp"], cnt)
    assert ((so["status"] == -1) == (cnt <= 0)).mean() > 0.85  # a few smoothed paths graze an obstacle: Failure as well
    finite = 0
    for q in np.flatnonzero(cnt > 4):
        a, b = rs["poses"][q, :cnt[q]], so["poses"][q, :cnt[q]]
        if not np.isfinite(b).all():  # 0 * inf of the inactive curvature term at a mis-snapped cusp (upstream): Failure
            assert so["status"][q] == -1
            continue
        finite += 1
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[-1], b[-1]) and np.array_equal(a[:, 2], b[:, 2])
        assert np.hypot(*(a[:, :2] - b[:, :2]).T).max() < 1.0
    assert finite > 0.9 * (cnt > 4).sum()
    sm.close()
    planner.close()
    w.close()
