#!/usr/bin/env python
"""A/B of the two Reeds-Shepp edge-check engines on the C3 bench workload: serial launches, device time per 1M edges.
Run on the GPU box:  python tools/time_check.py"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathplanning_b200 as pp  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402
from pathplanning_b200.lib import ppx  # noqa: E402


def main():
    ctx = pp.Context(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    ctx.build_map(desc, occ)
    n = 1_000_000
    vp = pp.ValidatorParams(1.0, 0.1).to_c()
    cp = pp.RSCostParams(2.0).to_c()
    dev = torch.device("cuda", 0)
    d_edges = []
    for b in range(4):
        s, g = synth.random_rs_queries(n, seed=1000 + 17 * b)
        ds, dg = torch.from_numpy(s).to(dev), torch.from_numpy(g).to(dev)
        e = torch.empty(n * 72, dtype=torch.uint8, device=dev)
        assert ppx.ppx_rs_solve_dev(ctx.handle, n, ds.data_ptr(), dg.data_ptr(), ctypes.byref(cp), e.data_ptr(), None, None, None) == 0
        d_edges.append(e)
    outs = [(torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.float32, device=dev)) for _ in range(4)]
    d_samples = torch.empty(n, dtype=torch.int32, device=dev)
    res = {}
    for engine in (0, 1):
        ctx.set_option("rs_engine", engine)
        for overlap in (0, 1):
            ctx.set_option("rs_overlap", overlap)
            for i in range(8):
                ppx.ppx_check_edges_rs_dev(ctx.handle, n, d_edges[i % 4].data_ptr(), 2.0, ctypes.byref(vp), outs[i % 4][0].data_ptr(),
                                           outs[i % 4][1].data_ptr(), None)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for i in range(40):
                ppx.ppx_check_edges_rs_dev(ctx.handle, n, d_edges[i % 4].data_ptr(), 2.0, ctypes.byref(vp), outs[i % 4][0].data_ptr(),
                                           outs[i % 4][1].data_ptr(), None)
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 40
            print(f"engine {engine} overlap {overlap}: {ms * 1e3:.1f} us per 1M edges = {n / ms / 1e6:.2f} G edges/s", flush=True)
        ctx.set_option("rs_overlap", 0)
        ppx.ppx_check_edges_rs_dev(ctx.handle, n, d_edges[0].data_ptr(), 2.0, ctypes.byref(vp), outs[0][0].data_ptr(), outs[0][1].data_ptr(),
                                   d_samples.data_ptr())
        torch.cuda.synchronize()
        res[engine] = (outs[0][0].clone(), outs[0][1].clone(), d_samples.clone())
    same = all(torch.equal(a, b) for a, b in zip(res[0], res[1]))
    print("engines identical:", same, "mean samples", float(res[1][2].float().mean()))
    # solve (Connect) and solve + check from (start, goal) pairs
    s, g = synth.random_rs_queries(n, seed=1000)
    ds, dg = torch.from_numpy(s).to(dev), torch.from_numpy(g).to(dev)
    d_word = torch.empty(n, dtype=torch.int32, device=dev)
    d_len = torch.empty(n, dtype=torch.float64, device=dev)
    e = torch.empty(n * 72, dtype=torch.uint8, device=dev)
    sres = {}
    for engine in (0, 1):
        ctx.set_option("rs_engine", engine)
        for what in ("solve", "solve+check"):
            def call():
                if what == "solve":
                    rc = ppx.ppx_rs_solve_dev(ctx.handle, n, ds.data_ptr(), dg.data_ptr(), ctypes.byref(cp), e.data_ptr(), d_word.data_ptr(), d_len.data_ptr(), None)
                else:
                    rc = ppx.ppx_connect_check_rs_dev(ctx.handle, n, ds.data_ptr(), dg.data_ptr(), ctypes.byref(cp), ctypes.byref(vp),
                                                      outs[0][0].data_ptr(), outs[0][1].data_ptr(), d_word.data_ptr(), d_len.data_ptr(), None)
                assert rc == 0, ppx.ppx_last_error(ctx.handle)
            for i in range(3):
                call()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for i in range(10):
                call()
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            print(f"engine {engine} {what}: {ms:.3f} ms per 1M pairs = {n / ms / 1e6:.3f} G/s", flush=True)
            sres[(engine, what)] = (d_word.clone(), d_len.clone(), e.clone() if what == "solve" else outs[0][0].clone(), outs[0][1].clone())
    for what in ("solve", "solve+check"):
        a, b = sres[(0, what)], sres[(1, what)]
        print(what, "engines identical:", [bool(torch.equal(x, y)) for x, y in zip(a, b)])
    ctx.close()


if __name__ == "__main__":
    main()
