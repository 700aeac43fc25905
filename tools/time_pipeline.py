#!/usr/bin/env python
"""Times search -> resample (N3) -> smooth (N2) for a batch of C2 queries through the host-buffer C ABI.
Run on the GPU box.  The smoother's GVD cell grids come from tests/golden/smooth.npz (reference-generated data)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402
from time_hybrid import c2_queries  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    ctx = pp.Context(0)
    cfg = synth.C2
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    sq = synth.exact_sqdist(occ)
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.upload_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]),
                   sq, occ, None)
    gold = np.load(os.path.join(ROOT, "tests", "golden", "smooth.npz"))
    planner = pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(1.0, 0.1))
    starts, goals = c2_queries(ctx, n)
    seeds = (np.arange(n) + 1000).astype(np.uint64)
    max_points = 1024
    for rep in range(2):
        t0 = time.perf_counter()
        out = planner.search(starts, goals, seeds, max_path=128)
        t1 = time.perf_counter()
        rs = planner.resample(goals, out["results"], out["paths"], 0.1, max_points=max_points)
        t2 = time.perf_counter()
    cnt = rs["count"]
    print(f"{n} queries: search {1e3 * (t1 - t0):.1f} ms, resample {1e3 * (t2 - t1):.1f} ms "
          f"(solved {int((cnt > 0).sum())}, mean {cnt[cnt > 0].mean():.0f} / max {cnt.max()} points, truncated {int((cnt < 0).sum())})")
    mc = float(np.float32(0.5))
    for name, sp in (("default (diverges, 2000 it)", pp.SmootherParams(max_curvature=mc)),
                     ("curvature_weight=0 (2000 it)", pp.SmootherParams(max_curvature=mc, curvature_weight=0.0))):
        sm = pp.Smoother(ctx, sp)
        sm.set_gvd(gold["c2_near_obstacle"].astype(np.int32), gold["c2_near_edge"].astype(np.int32))
        for rep in range(2):
            t0 = time.perf_counter()
            o = sm.smooth(rs["poses"], rs["cusp"], rs["count"])
            dt = time.perf_counter() - t0
        st = o["status"]
        print(f"smooth [{name}]: {dt * 1e3:.1f} ms -> {int((cnt > 0).sum()) / dt:.0f} paths/s; statuses "
              f"{dict(zip(*np.unique(st, return_counts=True)))}; mean iterations {o['iterations'][cnt > 0].mean():.0f}", flush=True)
        sm.close()
    planner.close()
    ctx.close()


if __name__ == "__main__":
    main()
