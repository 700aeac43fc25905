#!/usr/bin/env python
"""Small launches of the kernels that have no bench line of their own, for ncu:
k_edt_columns / k_edt_rows / k_voronoi_edges / k_pathcost (map build on the C3 grid), k_rrt (one RRT* tree, Reeds-Shepp
and straight connections, to PPX_PROF_NODES nodes), k_knn_cta_per_query / k_knn_thread_per_query (100 k points).

    ncu --set full -k regex:'k_edt|k_rrt|k_knn|k_voronoi|k_pathcost' ... python tools/prof_misc.py
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402


def main():
    nodes = int(os.environ.get("PPX_PROF_NODES", "5000"))
    ctx = pp.Context(0)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    t0 = time.perf_counter()
    ctx.build_map(desc, occ)
    ctx.build_gvd(desc, occ)
    print("map + gvd build", time.perf_counter() - t0, flush=True)
    vp = pp.ValidatorParams(1.0, 0.1)
    rng = np.random.default_rng(5)
    pts = np.stack([rng.uniform(-95, 95, 400), rng.uniform(-95, 95, 400), rng.uniform(-np.pi, np.pi, 400)], 1)
    pts = pts[ctx.is_state_valid(pts, vp)]
    for conn in (0, 1):
        t0 = time.perf_counter()
        r = ctx.rrt_run(pts[:1], pts[1:2], np.array([100], np.uint64), cfg["lb"], cfg["ub"], dim=3, connection=conn, star=1,
                        validator=1, max_iteration=50 * nodes, max_nodes=nodes, max_connection_distance=1.0, goal_bias=0.05,
                        min_turning_radius=2.0, vp=vp, want_trees=False)["results"]
        print("rrt connection", conn, "nodes", int(r["n_nodes"][0]), "iterations", int(r["iterations"][0]),
              time.perf_counter() - t0, flush=True)
    P = np.stack([rng.uniform(-100, 100, 100000), rng.uniform(-100, 100, 100000), rng.uniform(-np.pi, np.pi, 100000)], 1)
    for nq in (256, 65536):
        Q = np.stack([rng.uniform(-100, 100, nq), rng.uniform(-100, 100, nq), rng.uniform(-np.pi, np.pi, nq)], 1)
        t0 = time.perf_counter()
        ctx.knn(P, Q, 11)
        print("knn", nq, time.perf_counter() - t0, flush=True)


if __name__ == "__main__":
    main()
