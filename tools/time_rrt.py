#!/usr/bin/env python
"""BASELINE config C5: RRT* in SE(2) on the C3 map (Reeds-Shepp and straight connections, occupancy validator,
maxConnectionDistance 1.0, goalBias 0.05), trees grown to a node budget; plus brute-force kNN throughput at
100 k points.  Run on the GPU box:  python tools/time_rrt.py [max_nodes ...]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [10000, 100000]
    ctx = pp.Context(0)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ)
    vp = pp.ValidatorParams(1.0, 0.1)
    rng = np.random.default_rng(5)
    pts = np.stack([rng.uniform(-95, 95, 4000), rng.uniform(-95, 95, 4000), rng.uniform(-np.pi, np.pi, 4000)], 1)
    pts = pts[ctx.is_state_valid(pts, vp)]
    out = []
    for max_nodes in sizes:
        for conn, name in ((1, "reeds_shepp"), (0, "straight")):
            for ntrees in (1, 148):
                starts, goals = pts[:ntrees], pts[ntrees:2 * ntrees]
                seeds = (100 + np.arange(ntrees)).astype(np.uint64)
                t0 = time.perf_counter()
                r = ctx.rrt_run(starts, goals, seeds, cfg["lb"], cfg["ub"], dim=3, connection=conn, star=1, validator=1,
                                max_iteration=50 * max_nodes, max_nodes=max_nodes, max_connection_distance=1.0, goal_bias=0.05,
                                min_turning_radius=2.0, vp=vp, want_trees=False)["results"]
                dt = time.perf_counter() - t0
                rec = dict(config="C5", connection=name, trees=ntrees, max_nodes=max_nodes, seconds=dt,
                           nodes=int(r["n_nodes"].sum()), iterations=int(r["iterations"].sum()),
                           edges_checked=int(r["n_edges_checked"].sum()), nodes_per_s=float(r["n_nodes"].sum() / dt),
                           iterations_per_s=float(r["iterations"].sum() / dt), edges_per_s=float(r["n_edges_checked"].sum() / dt))
                print(json.dumps(rec), flush=True)
                out.append(rec)
    # kNN: 100 k points in SE(2) coordinates, k = 11 (RRT*'s neighbourhood at 100 k nodes), batches of queries
    P = np.stack([rng.uniform(-100, 100, 100000), rng.uniform(-100, 100, 100000), rng.uniform(-np.pi, np.pi, 100000)], 1)
    for nq in (1, 1024, 65536):
        Q = np.stack([rng.uniform(-100, 100, nq), rng.uniform(-100, 100, nq), rng.uniform(-np.pi, np.pi, nq)], 1)
        ctx.knn(P, Q, 11)
        t0 = time.perf_counter()
        ctx.knn(P, Q, 11)
        dt = time.perf_counter() - t0
        rec = dict(config="kNN", points=100000, queries=nq, k=11, seconds=dt, queries_per_s=nq / dt, pairs_per_s=nq * 1e5 / dt)
        print(json.dumps(rec), flush=True)
        out.append(rec)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "rrt_c5.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
