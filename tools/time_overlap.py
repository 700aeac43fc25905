#!/usr/bin/env python
"""k_check_rs with the drain tail of one batch overlapped by the next batch: two contexts (two streams, each with its
own copy of the C3 map and its own output buffers) take the 1M-edge batches alternately.  Compared with the same
launches on one context.  Evidence for DESIGN.md 4.1 (how much of a launch is tail); run on the GPU box."""
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

N, NB, STEPS = 1_000_000, 4, 40


def main():
    dev = torch.device("cuda", 0)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    ctxs, streams = [], []
    for _ in range(2):
        c = pp.Context(0)
        st = torch.cuda.Stream(device=dev)
        c.set_stream(st.cuda_stream)
        c.build_map(desc, occ)
        ctxs.append(c)
        streams.append(st)
    vp = pp.ValidatorParams(1.0, 0.1).to_c()
    cp = pp.RSCostParams(2.0).to_c()
    edges = []
    for b in range(NB):
        s, g = synth.random_rs_queries(N, seed=1000 + 17 * b)
        ds, dg = torch.from_numpy(s).to(dev), torch.from_numpy(g).to(dev)
        e = torch.empty(N * 72, dtype=torch.uint8, device=dev)
        assert ppx.ppx_rs_solve_dev(ctxs[0].handle, N, ds.data_ptr(), dg.data_ptr(), ctypes.byref(cp), e.data_ptr(), None, None, None) == 0
        edges.append(e)
    torch.cuda.synchronize()
    outs = [(torch.empty(N, dtype=torch.uint8, device=dev), torch.empty(N, dtype=torch.float32, device=dev)) for _ in range(NB)]

    def launch(c, b):
        v, l = outs[b]
        assert ppx.ppx_check_edges_rs_dev(c.handle, N, edges[b].data_ptr(), 2.0, ctypes.byref(vp), v.data_ptr(), l.data_ptr(), None) == 0

    # reference results: every batch on one context
    for b in range(NB):
        launch(ctxs[0], b)
    torch.cuda.synchronize()
    want = [(v.clone(), l.clone()) for v, l in outs]

    def timed(two):
        for v, l in outs:
            v.zero_()
            l.zero_()
        torch.cuda.synchronize()
        for i in range(8):
            launch(ctxs[i % 2 if two else 0], i % NB)
        torch.cuda.synchronize()
        e0, e1, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event()
        e0.record(streams[0])
        streams[1].wait_event(e0)
        for i in range(STEPS):
            launch(ctxs[i % 2 if two else 0], i % NB)
        eb.record(streams[1])
        streams[0].wait_event(eb)
        e1.record(streams[0])
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / STEPS
        for b in range(NB):
            assert torch.equal(outs[b][0], want[b][0]) and torch.equal(outs[b][1], want[b][1]), "overlapped results differ"
        return ms

    for rep in range(2):
        a, b = timed(False), timed(True)
        ctxs[0].set_option("rs_overlap", 1)
        c = timed(False)
        ctxs[0].set_option("rs_overlap", 0)
        print(f"one context: {a * 1e3:.1f} us per 1M-edge launch = {N / a / 1e6:.2f} G edges/s; two contexts alternating: "
              f"{b * 1e3:.1f} us per launch = {N / b / 1e6:.2f} G edges/s; one context with rs_overlap (programmatic dependent "
              f"launch): {c * 1e3:.1f} us = {N / c / 1e6:.2f} G edges/s (results identical)", flush=True)


if __name__ == "__main__":
    main()
