#!/usr/bin/env python
"""BASELINE config C5, one tree: RRT* over SE(2) on the C3 map from (-90,-90,0) towards (90,90,1), seed 7, goalBias 0.05,
maxConnectionDistance 1.0, to a node budget (default 100 k), straight and Reeds-Shepp connections -- the workload of the
`c5` object of bench.py.  PPX_RRT_CLUSTER=1|2|4|8|16 fixes the cluster size (default: chosen from the number of trees).

    python tools/time_rrt_single.py [max_nodes ...]
"""
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
    sizes = [int(a) for a in sys.argv[1:]] or [100000]
    ctx = pp.Context(0)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ)
    vp = pp.ValidatorParams(1.0, 0.1)
    for nodes in sizes:
        for conn, name in ((0, "straight"), (1, "reeds_shepp")):
            t0 = time.perf_counter()
            r = ctx.rrt_run(np.array([(-90.0, -90.0, 0.0)]), np.array([(90.0, 90.0, 1.0)]), np.array([7], np.uint64), cfg["lb"], cfg["ub"],
                            dim=3, connection=conn, star=1, validator=1, max_iteration=10**7, max_nodes=nodes,
                            max_connection_distance=1.0, goal_bias=0.05, min_turning_radius=2.0, vp=vp, want_trees=True)
            dt = time.perf_counter() - t0
            res = r["results"][0]
            n = int(res["n_nodes"])
            digest = int(np.frombuffer(np.ascontiguousarray(r["nodes"][0][:n]).tobytes(), np.uint64).sum(dtype=np.uint64)
                         ^ np.uint64(int(r["parents"][0][:n].astype(np.int64).sum())))
            print(json.dumps(dict(cluster=os.environ.get("PPX_RRT_CLUSTER", "auto"), connection=name, max_nodes=nodes, nodes=n,
                                  iterations=int(res["iterations"]), edges_checked=int(res["n_edges_checked"]), seconds=dt,
                                  us_per_iteration=dt / max(1, int(res["iterations"])) * 1e6, tree_digest=digest)), flush=True)


if __name__ == "__main__":
    main()
