#!/usr/bin/env python
"""The host-buffer edge check (ppx_check_edges_rs, BASELINE config C3) from pinned memory: wall time per 1M-edge call,
next to the bare 72 MB H2D copy of the same process (the copy rate differs from process to process, 38-52 GB/s on the
box used: the call follows it at +0.08 ... +0.15 ms).  Run on the GPU box."""
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


def main():
    ctx = pp.Context(0)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ)
    n = 1_000_000
    vp = pp.ValidatorParams(1.0, 0.1).to_c()
    s, g = synth.random_rs_queries(n, seed=1000)
    edges = ctx.rs_solve(s, g, pp.RSCostParams(2.0))["edges"]
    h_in = torch.from_numpy(np.frombuffer(edges.tobytes(), np.uint8).copy()).pin_memory()
    h_valid = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_last = torch.empty(n, dtype=torch.float32).pin_memory()
    d = torch.empty(n * 72, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        d.copy_(h_in, non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        d.copy_(h_in, non_blocking=True)
    torch.cuda.synchronize()
    copy_ms = (time.perf_counter() - t0) * 100
    for _ in range(3):
        assert ppx.ppx_check_edges_rs(ctx.handle, n, h_in.data_ptr(), 2.0, ctypes.byref(vp), h_valid.data_ptr(), h_last.data_ptr(), None) == 0
    t0 = time.perf_counter()
    for _ in range(20):
        ppx.ppx_check_edges_rs(ctx.handle, n, h_in.data_ptr(), 2.0, ctypes.byref(vp), h_valid.data_ptr(), h_last.data_ptr(), None)
    ms = (time.perf_counter() - t0) * 50
    print(f"{ms:.3f} ms per 1M-edge call = {n / ms / 1e6:.3f} G edges/s; bare 72 MB H2D {copy_ms:.3f} ms "
          f"= {72e-3 / copy_ms * 1e3:.1f} GB/s; valid {int(h_valid.sum())}", flush=True)


if __name__ == "__main__":
    main()
