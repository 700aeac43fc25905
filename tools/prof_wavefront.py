import sys, numpy as np, torch
sys.path.insert(0, ".")
import pathplanning_b200 as pp
from pathplanning_b200 import synth
from pathplanning_b200.lib import ppx
ctx = pp.Context(0)
cfg = synth.C2
occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ)
rng = np.random.default_rng(3)
n = 4096
goals = np.stack([rng.uniform(-18, 18, n), rng.uniform(-18, 18, n), np.zeros(n)], 1)
d_goals = torch.from_numpy(goals).cuda()
d_cost = torch.empty((n, rows, cols), dtype=torch.float32, device="cuda")
for _ in range(2):
    assert ppx.ppx_wavefront_dev(ctx.handle, n, d_goals.data_ptr(), d_cost.data_ptr(), None) == 0
    ctx.synchronize()
print("done")
