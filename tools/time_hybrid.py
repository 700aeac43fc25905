#!/usr/bin/env python
"""Times the batched Hybrid A* graph search (wavefront + A*) on the C2 map.  Run on the GPU box."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402


def c2_queries(ctx, n, seed=12345):
    rng = np.random.default_rng(seed)
    vp = pp.ValidatorParams(1.0, 0.1)
    starts, goals = np.empty((0, 3)), np.empty((0, 3))
    while len(starts) < n:
        p = np.stack([rng.uniform(-18.5, 18.5, 4 * n), rng.uniform(-18.5, 18.5, 4 * n), rng.uniform(-np.pi, np.pi, 4 * n)], 1)
        p = p[ctx.is_state_valid(p, vp)]
        m = len(p) // 2
        starts = np.concatenate([starts, p[:m]])
        goals = np.concatenate([goals, p[m:2 * m]])
    return starts[:n].copy(), goals[:n].copy()


def main():
    ctx = pp.Context(0)
    cfg = synth.C2
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    sq = synth.exact_sqdist(occ)
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.upload_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]),
                   sq, occ, None)
    t0 = time.perf_counter()
    planner = pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(1.0, 0.1))
    print(f"ppx_hybrid_create (table build): {(time.perf_counter() - t0) * 1e3:.1f} ms")
    sizes = [int(a) for a in sys.argv[1:]] or [1, 64, 512, 4096]
    for n in sizes:
        starts, goals = c2_queries(ctx, n)
        seeds = (np.arange(n) + 1000).astype(np.uint64)
        for rep in range(2):
            t0 = time.perf_counter()
            out = planner.search(starts, goals, seeds, max_path=128)
            dt = time.perf_counter() - t0
        r = out["results"]
        print(f"{n} queries: {dt * 1e3:.1f} ms -> {n / dt:.0f} plans/s; success {int((r['status'] == 0).sum())}, "
              f"mean expansions {r['n_expanded'].mean():.0f}, max {r['n_expanded'].max()}, total children {int(r['n_children'].sum())}", flush=True)
    planner.close()
    ctx.close()


if __name__ == "__main__":
    main()
