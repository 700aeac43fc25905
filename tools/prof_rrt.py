import os, sys, time, numpy as np
sys.path.insert(0, ".")
import pathplanning_b200 as pp
from pathplanning_b200 import synth
ctx = pp.Context(0)
cfg = synth.C3
occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ)
vp = pp.ValidatorParams(1.0, 0.1)
for conn, nodes in ((0, int(os.environ.get("NODES_ST", "20000"))), (1, int(os.environ.get("NODES_RS", "300")))):
    t0 = time.perf_counter()
    r = ctx.rrt_run(np.array([(-90.0, -90.0, 0.0)]), np.array([(90.0, 90.0, 1.0)]), np.array([7], np.uint64), cfg["lb"], cfg["ub"], dim=3, connection=conn, star=1, validator=1, max_iteration=10**7, max_nodes=nodes, max_connection_distance=1.0, goal_bias=0.05, min_turning_radius=2.0, vp=vp, want_trees=False)["results"][0]
    print(conn, int(r["n_nodes"]), int(r["iterations"]), time.perf_counter() - t0, flush=True)
