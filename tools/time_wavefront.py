#!/usr/bin/env python
"""Times ppx_wavefront_dev on the C2 (400x400) and C3 (2000x2000) maps: single goal latency and
batched throughput.  Run on the GPU box."""
import ctypes
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402
from pathplanning_b200.lib import ppx  # noqa: E402


def run(cfg, batches):
    ctx = pp.Context(0)
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    sq = synth.exact_sqdist(occ)
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.upload_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]),
                   sq, occ, None)
    rng = np.random.default_rng(3)
    ext = cfg["ub"][0] * 0.9
    for n in batches:
        goals = np.stack([rng.uniform(-ext, ext, n), rng.uniform(-ext, ext, n), np.zeros(n)], 1)
        d_goals = torch.from_numpy(goals).cuda()
        d_cost = torch.empty((n, rows, cols), dtype=torch.float32, device="cuda")
        for rep in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rc = ppx.ppx_wavefront_dev(ctx.handle, n, d_goals.data_ptr(), d_cost.data_ptr(), None)
            assert rc == 0
            ctx.synchronize()
            dt = time.perf_counter() - t0
        reached = int(torch.isfinite(d_cost).sum().item())
        print(f"{cfg['name']} {rows}x{cols}: {n} goals in {dt * 1e3:.2f} ms -> {n / dt:.1f} wavefronts/s, "
              f"{reached / dt / 1e6:.1f} Mcells/s", flush=True)
        del d_cost
    ctx.close()


if __name__ == "__main__":
    run(synth.C2, [1, 148, 1024, 4096])
    run(synth.C3, [1, 16])
