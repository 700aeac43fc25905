#!/usr/bin/env python
"""A/B of the two Hybrid A* search engines on the C4 workload (device-built GVD with Voronoi edges): wavefront and
search device time per batch size, results compared.  Run on the GPU box:  python tools/time_search.py [sizes...]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402


ENGINES = [int(e) for e in os.environ.get("PPX_ENGINES", "0,1").split(",")]
REPS = int(os.environ.get("PPX_REPS", "2"))


def main():
    ctx = pp.Context(0)
    cfg = synth.C2
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    ctx.build_gvd(desc, occ, voronoi=True)
    planner = pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(1.0, 0.1))
    rng = np.random.default_rng(12345)
    vp = pp.ValidatorParams(1.0, 0.1)
    nmax = 4096
    pts = np.empty((0, 3))
    while len(pts) < 2 * nmax:
        p = np.stack([rng.uniform(-18.5, 18.5, 4 * nmax), rng.uniform(-18.5, 18.5, 4 * nmax), rng.uniform(-np.pi, np.pi, 4 * nmax)], 1)
        pts = np.concatenate([pts, p[ctx.is_state_valid(p, vp)]])
    S, G = pts[:nmax].copy(), pts[nmax:2 * nmax].copy()
    seeds = (np.arange(nmax) + 1000).astype(np.uint64)
    sizes = [int(a) for a in sys.argv[1:]] or [1, 64, 512, 4096]
    ref = None
    for n in sizes:
        outs = {}
        for engine in ENGINES:
            ctx.set_option("search_engine", engine)
            for rep in range(REPS):
                t0 = time.perf_counter()
                out = planner.search(S[:n], G[:n], seeds[:n], max_path=128)
                dt = time.perf_counter() - t0
            wf, se = planner.timing()
            r = out["results"]
            outs[engine] = out
            print(f"engine {engine} n={n}: host {dt * 1e3:.1f} ms, wavefront {wf:.2f} ms, search {se:.2f} ms -> "
                  f"{n / ((wf + se) * 1e-3):.0f} plans/s; solved {int((r['status'] == 0).sum())}, expansions {int(r['n_expanded'].sum())}, "
                  f"max {int(r['n_expanded'].max())}", flush=True)
        if len(ENGINES) == 2:
            same = outs[0]["results"].tobytes() == outs[1]["results"].tobytes() and outs[0]["paths"].tobytes() == outs[1]["paths"].tobytes()
            print(f"  engines identical: {same}", flush=True)
    if os.environ.get("PPX_NO_LONGEST"):
        return
    # the longest single query alone (critical path)
    r = outs[ENGINES[-1]]["results"]
    worst = int(np.argmax(r["n_expanded"]))
    for engine in ENGINES:
        ctx.set_option("search_engine", engine)
        for rep in range(2):
            planner.search(S[worst:worst + 1], G[worst:worst + 1], seeds[worst:worst + 1])
        wf, se = planner.timing()
        print(f"engine {engine} longest query ({int(r['n_expanded'][worst])} expansions) alone: wavefront {wf:.2f} ms, search {se:.2f} ms "
              f"= {se * 1e3 / max(1, int(r['n_expanded'][worst])):.2f} us/expansion", flush=True)
    planner.close()
    ctx.close()


if __name__ == "__main__":
    main()
