#!/usr/bin/env python
"""bench.py -- BASELINE config C3: batched collision check of 1M Reeds-Shepp edges (disc
footprint r = 1.0 m) on a 2000x2000 occupancy grid at 0.1 m.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode check|solve_check]

One "step" = one pass of StateValidatorOccupancyMap::IsPathValid over one batch of 1M pre-solved
Reeds-Shepp edges (mode check, default) or PathConnectionReedsShepp::Connect + IsPathValid over 1M
start/goal pairs (mode solve_check).  Rank r of N owns its own batch (weak scaling, no data-path
collective; NCCL only reduces the timing).  Prints ONE JSON line on rank 0.

  value     edges/s with the edge records already resident in HBM (CUDA events on the launch stream)
  e2e       edges/s through the host-buffer C-ABI call (pinned host memory, H2D + kernel + D2H)
  roofline  algorithmic bytes (80 + 4*S per edge, S = samples visited) / kernel time vs measured HBM peak
  cpu_baseline / --impl reference: the UNMODIFIED reference (oracle/_ref/libpp_ref.so) timed on all
            host cores, one process per core, on a bounded sample of the same edges.
"""
import argparse
import ctypes
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_EDGES = 1_000_000
N_BATCHES = 4  # rotating batches: 4 x 72 MB of edge records > 126 MB L2
RMIN = 2.0
RADIUS, MIN_STEP = 1.0, 0.1
WORKLOAD = "C3: 1M Reeds-Shepp edges x disc footprint r=1.0m on 2000x2000 grid @0.1m (parking lot, p=0.5)"
L2_NOTE = "GPU arm: 4 rotating edge batches per rank (288 MB of records in mode check) > 126 MB L2, no flush needed"


def bench_config(mode):
    """The `config` object, identical on both arms."""
    return {"workload": WORKLOAD, "mode": mode, "l2": L2_NOTE}


_REAL_STDOUT = None


def _stdout_to_stderr():
    """stdout must carry exactly ONE JSON line.  Libraries print there too (NCCL's version banner comes out of
    ncclCommInitRank whatever NCCL_DEBUG_FILE says), so fd 1 points at stderr while the bench runs and the
    JSON line is written to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def _synth():
    """pathplanning_b200/synth.py loaded by path: importing the package would dlopen libppx.so, which the reference arm
    must not do (its record has to show no product code loaded)."""
    import importlib.util
    if "ppx_synth" in sys.modules:
        return sys.modules["ppx_synth"]
    spec = importlib.util.spec_from_file_location("ppx_synth", os.path.join(ROOT, "pathplanning_b200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ppx_synth"] = mod
    spec.loader.exec_module(mod)
    return mod


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# --------------------------------------------------------------------------------------------------
# Reference arm: the compiled reference on the host cores (test infrastructure, see oracle/)
def _ref_world():
    import refapi
    synth = _synth()
    cfg = synth.C3
    w = refapi.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in synth.config_rects(cfg):
        w.add_rectangle(*r)
    w.update()
    w.set_validator_params(RADIUS, MIN_STEP)
    return w


def _ref_worker(w, starts, goals, mode, reps, conn):
    """Child process: solve (untimed for mode check) then time the reference loop."""
    import refapi
    try:
        if mode == "check":
            r = w.check_rs_edges(starts, goals, RMIN)
            times = [w.check_rs_segments(starts, r["steer"], r["dir"], r["len"], RMIN)["seconds"] for _ in range(reps)]
        else:
            times = [w.check_rs_edges(starts, goals, RMIN, full=False)["seconds"] for _ in range(reps)]
        conn.send(times)
    except Exception as e:  # pragma: no cover
        conn.send(e)
    conn.close()


def run_reference(args, per_core):
    """Times the reference's own CPU implementation with one process per host core.
    Returns (edges_per_s, ms_per_step, cores, sample description)."""
    import multiprocessing as mp
    synth = _synth()
    cores = host_cores()
    w = _ref_world()  # built once, inherited by fork (process-global Random<> statics stay per process)
    n = per_core * cores
    starts, goals = synth.random_rs_queries(n, seed=1000)
    reps = args.warmup + args.steps
    ctxmp = mp.get_context("fork")
    procs, pipes = [], []
    for c in range(cores):
        a, b = ctxmp.Pipe(duplex=False)
        sl = slice(c * per_core, (c + 1) * per_core)
        p = ctxmp.Process(target=_ref_worker, args=(w, starts[sl], goals[sl], args.mode, reps, b))
        p.start()
        procs.append(p)
        pipes.append(a)
    results = [a.recv() for a in pipes]
    for p in procs:
        p.join()
    for r in results:
        if isinstance(r, Exception):
            raise r
    # all cores run concurrently; a step takes as long as the slowest core
    per_step = np.max(np.array(results)[:, args.warmup:], axis=0)
    sec = float(np.mean(per_step))
    sample = f"{n} edges of the workload ({per_core}/core), {args.steps} timed passes, one process per core"
    return n / sec, sec * 1e3, cores, sample


def _ref_hybrid_worker(queries, field, conn):
    """Child process: the reference's wavefront + GVD update + graph search (hybrid_a_star.cpp:248-257) per query.
    The GVD is attached before the obstacle brushfire runs (example.py:47-48), so the potential is the Voronoi field;
    `field` = (sqdist, voro_sqdist) built by the device (ppx_gvd_build) replaces the reference's two distance grids, so
    that both arms search the same potential (the reference recomputes it from those grids at every GVD::Update)."""
    import zlib
    import refapi
    synth = _synth()
    try:
        cfg = synth.C2
        w = refapi.World(cfg["lb"], cfg["ub"], cfg["resolution"])
        for r in synth.config_rects(cfg):
            w.add_rectangle(*r)
        w.gvd_update()
        diff = w.set_gvd_distances(*field) if field is not None else (0, 0)
        w.set_validator_params(RADIUS, MIN_STEP)
        H = refapi.Hybrid(w)
        t0 = time.perf_counter()
        recs = []
        for qi, s, g, seed in queries:
            r = H.search(s, g, seed)
            recs.append((qi, int(r["status"]), int(r["n_expanded"]), zlib.crc32(np.ascontiguousarray(r["exp_keys"][:, :3], np.int32).tobytes()),
                         float(r["cost"])))
        conn.send((time.perf_counter() - t0, recs, diff))
    except Exception as e:  # pragma: no cover
        conn.send(e)
    conn.close()


def run_reference_hybrid(starts, goals, seeds, per_core, field=None):
    """plans/s of the reference on all host cores (one process per core) over the first per_core*cores queries, and the
    per-query records (status, expansions, CRC of the expansion sequence, cost) the GPU run is checked against."""
    import multiprocessing as mp
    cores = host_cores()
    n = min(len(starts), per_core * cores)
    ctxmp = mp.get_context("fork")
    procs, pipes = [], []
    for c in range(cores):
        idx = range(c, n, cores)
        qs = [(i, tuple(starts[i]), tuple(goals[i]), int(seeds[i])) for i in idx]
        a, b = ctxmp.Pipe(duplex=False)
        p = ctxmp.Process(target=_ref_hybrid_worker, args=(qs, field, b))
        p.start()
        procs.append(p)
        pipes.append(a)
    results = [a.recv() for a in pipes]
    for p in procs:
        p.join()
    for r in results:
        if isinstance(r, Exception):
            raise r
    sec = max(r[0] for r in results)
    recs = sorted(rec for r in results for rec in r[1])
    base = {"value": n / sec, "unit": "plans/s", "cores": cores, "kind": "reference",
            "sample": f"first {n} of the queries, wavefront + GVD update + graph search each, one process per core"}
    return base, recs, results[0][2]


def reference_main(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import refapi
    if not refapi.available():
        _emit({"impl": "reference", "unavailable": "oracle/_ref/libpp_ref.so not built"})
        return
    value, ms, cores, sample = run_reference(args, per_core=args.cpu_edges_per_core)
    per_step_edges = args.cpu_edges_per_core * cores
    line = {
        "impl": "reference", "metric": "collision-checked RS edges/s", "value": value, "unit": "edges/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": bench_config(args.mode),
        "config_detail": {"edges_per_step": per_step_edges, "note": "bounded sample of the workload per step, all host cores"},
        "cpu_baseline": {"value": value, "unit": "edges/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# --------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed regions run."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.active = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._halt.is_set():
            if self.active.is_set():
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in names.items():
                        if mask & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(0.002)

    def stop(self):
        self._halt.set()

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def gpu_numa_affinity(index, apply=True):
    """Binds this process to the CPUs local to GPU `index` (sysfs local_cpulist of its PCI function) so that pinned host
    buffers are first-touched on the GPU's NUMA node.  Returns the node number, or None when the box does not say."""
    try:
        import torch
        p = torch.cuda.get_device_properties(index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        node = int(open(base + "/numa_node").read().strip())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if apply and cpus and node >= 0:
            os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def c4_queries(ctx, pp, n, seed=12345):
    """BASELINE config C4: n (start, goal) pairs uniform in +-18.5 m with uniform heading, rejection-sampled with
    IsStateValid on the C2 map; query q uses RNG seed 1000 + q."""
    rng = np.random.default_rng(seed)
    vp = pp.ValidatorParams(RADIUS, MIN_STEP)
    pts = np.empty((0, 3))
    while len(pts) < 2 * n:
        p = np.stack([rng.uniform(-18.5, 18.5, 4 * n), rng.uniform(-18.5, 18.5, 4 * n), rng.uniform(-np.pi, np.pi, 4 * n)], 1)
        pts = np.concatenate([pts, p[ctx.is_state_valid(p, vp)]])
    return pts[:n].copy(), pts[n:2 * n].copy(), (np.arange(n) + 1000).astype(np.uint64)


def hybrid_section(args, rank, world, local_rank, stream):
    """Config C4: Hybrid A* queries on the C2 map, sharded by query over the ranks (strong scaling, interleaved shards,
    no data-path collective), results gathered with one NCCL all_gather of the fixed-size result records.

    The potential both arms search is the Voronoi field the DEVICE builds (ppx_gvd_build); the CPU arm gets the
    device-built distance grids injected into the reference's GVD, and a sample of the timed batch is compared with the
    reference inside the run: status, expansion count, CRC of the expansion sequence, cost."""
    import zlib
    import torch
    import torch.distributed as dist
    import pathplanning_b200 as pp
    from pathplanning_b200 import synth
    from pathplanning_b200.lib import ppx, HybridLimitsC
    from pathplanning_b200.api import HYBRID_RESULT
    from pathplanning_b200.shard import shard_indices
    dev = torch.device("cuda", local_rank)
    ctx = pp.Context(local_rank)
    ctx.set_stream(stream.cuda_stream)
    cfg = synth.C2
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    t0 = time.perf_counter()
    field = ctx.build_gvd(desc, occ, voronoi=True)
    gvd_ms = (time.perf_counter() - t0) * 1e3
    planner = pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(RADIUS, MIN_STEP))
    nq = args.hybrid_queries
    starts, goals, seeds = c4_queries(ctx, pp, nq)
    mine = shard_indices(nq, rank, world, layout="strided")
    m = len(mine)
    d_s = torch.from_numpy(starts[mine].copy()).to(dev)
    d_g = torch.from_numpy(goals[mine].copy()).to(dev)
    d_seed = torch.from_numpy(seeds[mine].astype(np.int64)).to(dev)
    d_res = torch.zeros(m * HYBRID_RESULT.itemsize, dtype=torch.uint8, device=dev)
    lim = HybridLimitsC(0, 0, 0, 0)
    launches0 = ctx.launch_count

    def run_dev():
        rc = ppx.ppx_hybrid_search_dev(planner.handle, m, d_s.data_ptr(), d_g.data_ptr(), d_seed.data_ptr(), ctypes.byref(lim),
                                       d_res.data_ptr(), None, None)
        assert rc == 0, ppx.ppx_last_error(ctx.handle)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run_dev()  # warm-up (allocates the workspaces)
    run_dev()
    run_dev()
    barrier()
    reps = 3
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(reps):
        run_dev()
    ev1.record(stream)
    barrier()
    t = torch.tensor([ev0.elapsed_time(ev1) / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev = float(t.item())
    wf_ms, search_ms = planner.timing()  # this rank's last repetition
    per_rank = None
    if world > 1:  # the step is the slowest rank's: show every rank's phases (last repetition) and its shard's work
        mine_ms = torch.tensor([wf_ms, search_ms, ev0.elapsed_time(ev1) / reps], dtype=torch.float64, device=dev)
        all_ms = [torch.zeros_like(mine_ms) for _ in range(world)]
        dist.all_gather(all_ms, mine_ms)
        per_rank = [{"rank": r, "wavefront_ms": float(a[0]), "search_ms": float(a[1]), "step_ms": float(a[2])} for r, a in enumerate(all_ms)]
    launches = (ctx.launch_count - launches0) // (reps + 3)
    local = d_res.cpu().numpy().view(HYBRID_RESULT).copy()  # the device-timed run's records of this shard
    # through the host-buffer call (goal de-duplication, H2D, wavefronts, search, D2H of results + paths)
    planner.search(starts[mine], goals[mine], seeds[mine], max_path=128)
    barrier()
    t0 = time.perf_counter()
    out = planner.search(starts[mine], goals[mine], seeds[mine], max_path=128)
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    same_as_timed = bool(out["results"].tobytes() == local.tobytes())
    # gather of the fixed-size records: the only collective of the path.  Through the C ABI (ppx_comm_gather_results: NCCL
    # on the context's stream, device buffers in, device buffer out, uneven shards); the NCCL id travels over
    # torch.distributed's store.  One D2H copy of the gathered records afterwards.
    gather_ms = 0.0
    allres = local if world == 1 else None
    if world > 1:
        uid = [None]
        if rank == 0:
            buf = (ctypes.c_char * 128)()
            assert ppx.ppx_comm_unique_id(buf) == 0
            uid[0] = bytes(buf)
        dist.broadcast_object_list(uid, src=0)
        rc = ppx.ppx_comm_init(ctx.handle, (ctypes.c_char * 128).from_buffer_copy(uid[0]), rank, world)
        assert rc == 0, ppx.ppx_last_error(ctx.handle)
        d_all = torch.zeros(nq * HYBRID_RESULT.itemsize, dtype=torch.uint8, device=dev)
        counts = (ctypes.c_int64 * world)()

        def gather():
            rc = ppx.ppx_comm_gather_results(ctx.handle, d_res.data_ptr(), m, HYBRID_RESULT.itemsize, d_all.data_ptr(), nq, counts)
            assert rc == 0, ppx.ppx_last_error(ctx.handle)
        gather()  # warm-up of the collective
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        gather()
        g1.record(stream)
        torch.cuda.synchronize()
        gather_ms = g0.elapsed_time(g1)
        flat = d_all.cpu().numpy().view(HYBRID_RESULT)
        # rank order, interleaved shards: rank r's k-th record is query r + k * world
        allres = np.empty(nq, HYBRID_RESULT)
        off = 0
        for r in range(world):
            allres[r::world] = flat[off:off + counts[r]]
            off += counts[r]
        assert off == nq
        ppx.ppx_comm_destroy(ctx.handle)
    solved, expanded = int((allres["status"] == 0).sum()), int(allres["n_expanded"].sum())
    section = {
        "metric": "batched Hybrid A* plans/s", "workload": f"C4: {nq} independent start/goal queries on the C2 parking lot (400x400 @0.1m), "
        "wavefront + graph search per query, sharded by query (interleaved)", "scaling": "strong", "queries_per_gpu": m,
        "value": nq / (ms_dev * 1e-3), "unit": "plans/s", "ms": ms_dev, "wavefront_ms": wf_ms, "search_ms": search_ms,
        "e2e": {"value": nq / float(te.item()), "unit": "plans/s", "h2d_bytes": int(m * 56), "d2h_bytes": int(m * (40 + 128 * 40)),
                "records_equal_device_timed_run": same_as_timed},
        "solved": solved, "expansions": expanded, "longest_query_expansions": int(allres["n_expanded"].max()),
        "gather_ms": gather_ms, "gpu_launches": int(launches), "per_rank": per_rank,
        "field": {"what": "Voronoi potential built on the device (ppx_gvd_build: distance map, Voronoi edges, Voronoi distance "
                          "map, PathCostMap); the CPU arm searches the same field", "build_ms_host_call": gvd_ms,
                  "voronoi_edge_cells": int((field["voro_sqdist"] == 0).sum())},
        "note": "smoothing (SURVEY N2) is not part of this figure, neither here nor in the CPU figure",
    }
    # roofline of k_hybrid_search_cta: it is bound by instruction issue on dependent chains, not by HBM.  Warp-instructions
    # per expansion come from the committed ncu capture of this workload; the peak is 4 schedulers x SMs x clock.
    ipath = os.path.join(ROOT, "profiles", "hybrid_issue.json")
    if os.path.exists(ipath):
        ij = json.load(open(ipath))
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        clock_ghz = ij.get("sm_clock_ghz", 1.965)
        achieved = expanded / (search_ms * 1e-3) * ij["warp_instructions_per_expansion"] / 1e9 / max(1, world) if world == 1 else None
        if achieved is not None:
            peak = 4 * sms * clock_ghz
            section["roofline"] = {"bound": "issue", "kernel": "k_hybrid_search_cta", "achieved": achieved, "peak": peak,
                                   "unit": "G warp-instructions/s", "frac": achieved / peak,
                                   "warp_instructions_per_expansion": ij["warp_instructions_per_expansion"],
                                   "source": ij.get("source"), "note": "search phase only (search_ms); HBM is not its bound"}
    if world > 1:
        # weak-scaling companion: every rank searches all nq queries; a strong-scaled shard of nq/N queries is bounded
        # below by its longest single query (a chain of ~7000 dependent expansions)
        d_sw = torch.from_numpy(starts.copy()).to(dev)
        d_gw = torch.from_numpy(goals.copy()).to(dev)
        d_seedw = torch.from_numpy(seeds.astype(np.int64)).to(dev)
        d_resw = torch.zeros(nq * HYBRID_RESULT.itemsize, dtype=torch.uint8, device=dev)

        def run_weak():
            rc = ppx.ppx_hybrid_search_dev(planner.handle, nq, d_sw.data_ptr(), d_gw.data_ptr(), d_seedw.data_ptr(),
                                           ctypes.byref(lim), d_resw.data_ptr(), None, None)
            assert rc == 0, ppx.ppx_last_error(ctx.handle)

        run_weak()
        barrier()
        w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0.record(stream)
        run_weak()
        w1.record(stream)
        barrier()
        tw = torch.tensor([w0.elapsed_time(w1)], dtype=torch.float64, device=dev)
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        section["weak"] = {"queries_per_gpu": nq, "value": world * nq / (float(tw.item()) * 1e-3), "unit": "plans/s",
                           "ms": float(tw.item())}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            base, recs, diff = run_reference_hybrid(starts, goals, seeds, args.cpu_plans_per_core,
                                                    field=(field["sqdist"], field["voro_sqdist"]))
            section["cpu_baseline"] = base
            # the same queries on the device with expansion logs, tied to the timed batch through the result records
            k = len(recs)
            chk = planner.search(starts[:k], goals[:k], seeds[:k], max_log=12000)
            mismatched, untied = [], 0
            for qi, st, nexp, crc, cost in recs:
                r = chk["results"][qi]
                tr = allres[qi]
                if not (int(tr["status"]) == int(r["status"]) and int(tr["n_expanded"]) == int(r["n_expanded"])
                        and int(tr["expanded_hash"]) == int(r["expanded_hash"])):
                    untied += 1
                good = int(r["status"]) == st and int(r["n_expanded"]) == nexp and nexp <= 12000
                if good:
                    good = zlib.crc32(np.ascontiguousarray(chk["log"][qi]["key"][:nexp], np.int32).tobytes()) == crc
                if good and st == 0:
                    good = abs(float(r["cost"]) - cost) <= 1e-6 * max(1.0, abs(cost))
                if not good:
                    mismatched.append(int(qi))
            section["checked_vs_reference"] = {
                "queries": k, "matching": k - len(mismatched), "mismatched": mismatched[:16], "not_tied_to_timed_batch": untied,
                "compared": "status, expansion count, CRC32 of the expansion sequence (discrete keys), cost within 1e-6; the logged "
                            "run equals the timed batch in status / expansion count / expansion hash",
                "reference_field_cells_replaced": {"obstacle_distance": diff[0], "voronoi_distance": diff[1],
                                                   "of": int(rows * cols)}}
        except Exception as e:  # pragma: no cover
            section["cpu_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"}
    planner.close()
    ctx.close()
    return section


def _ref_extras_worker(conn, c5_nodes):
    """Child process (one core): the reference's side of the extra configurations -- C1 grid A*, C2 single Hybrid A* query,
    C5 restated RRT* (bounded), C3 solve + check and the rectangular-footprint port on a bounded sample."""
    import refapi
    synth = _synth()
    out = {}
    try:
        # C1: AStarN2 on the 200 x 200 grid with 60 rectangle outlines, Euclidean functors
        w = refapi.World((-50.0, -50.0, -math.pi), (50.0, 50.0, math.pi), 0.5)
        rng = np.random.default_rng(3)
        for _ in range(60):
            w.add_rectangle(2.0 + 10.0 * rng.uniform(), 1.0, rng.uniform(-45, 45), rng.uniform(-45, 45), rng.uniform(-math.pi, math.pi))
        w.update()
        r = w.astar_n2((1, 1), (198, 198))
        out["c1"] = dict(ms=r["seconds"] * 1e3, explored=r["explored"], cost=r["cost"], status=r["status"])
        w.close()
        # C2: one Hybrid A* query on the 400 x 400 parking lot (wavefront + GVD update + graph search)
        cfg = synth.C2
        w = refapi.World(cfg["lb"], cfg["ub"], cfg["resolution"])
        for rect in synth.config_rects(cfg):
            w.add_rectangle(*rect)
        w.gvd_update()
        w.set_validator_params(RADIUS, MIN_STEP)
        H = refapi.Hybrid(w)
        t0 = time.perf_counter()
        r = H.search((-18.0, -18.0, 0.0), (17.0, 17.5, math.pi / 2), 42)
        out["c2"] = dict(ms=(time.perf_counter() - t0) * 1e3, status=r["status"], cost=r["cost"], n_expanded=r["n_expanded"],
                         phases_ms=[t * 1e3 for t in r["times"]])
        H.close()
        w.close()
        # C3 map: solve + check and the rectangle port on 100k pairs; C5 restated RRT* to c5_nodes nodes
        cfg = synth.C3
        w = refapi.World(cfg["lb"], cfg["ub"], cfg["resolution"])
        for rect in synth.config_rects(cfg):
            w.add_rectangle(*rect)
        w.update()
        w.set_validator_params(RADIUS, MIN_STEP)
        s, g = synth.random_rs_queries(100000, seed=1000)
        r = w.check_rs_edges(s, g, RMIN, full=False)
        out["solve_check"] = dict(edges=100000, seconds=r["seconds"])
        r2 = w.check_rs_edges(s[:20000], g[:20000], RMIN)
        rr = w.check_rs_segments_rect(s[:20000], r2["steer"], r2["dir"], r2["len"], RMIN, RECT_FOOTPRINT)
        out["rect"] = dict(edges=20000, seconds=rr["seconds"], valid=int(rr["valid"].sum()), samples=int(rr["samples"].sum()))
        for conn_kind, name in ((0, "straight"), (1, "reeds_shepp")):
            r = refapi.rrtstar_se2_restated(w, cfg["lb"], cfg["ub"], C5_START, C5_GOAL, 7, 10**7, c5_nodes, 1.0, 0.05, conn_kind, RMIN)
            out["c5_" + name] = dict(nodes=r["n_nodes"], iterations=r["iterations"], edges=r["edges_checked"], seconds=r["seconds"])
        w.close()
        conn.send(out)
    except Exception as e:  # pragma: no cover
        conn.send(e)
    conn.close()


RECT_FOOTPRINT = (4.5, 2.0, 1.3, 3, 0.1)  # length, width, rear axle -> centre, covering discs, min step
C5_START, C5_GOAL = (-90.0, -90.0, 0.0), (90.0, 90.0, 1.0)


def extras_section(args, local_rank, stream):
    """The other BASELINE configurations in the driver-run line (N = 1, rank 0): C1 grid A*, C2 single Hybrid A* query, C3
    solve + check and the rectangular-footprint variant, C5 one RRT* tree to 100 k nodes -- each beside the reference (or
    its restated loop, kind "port") timed in this run on one host core, results compared where both exist."""
    import multiprocessing as mp
    import torch
    import pathplanning_b200 as pp
    from pathplanning_b200 import synth
    from pathplanning_b200.lib import ppx
    out = {}
    ref = None
    if not args.no_cpu_baseline:
        ctxmp = mp.get_context("fork")
        a, b = ctxmp.Pipe(duplex=False)
        proc = ctxmp.Process(target=_ref_extras_worker, args=(b, args.cpu_c5_nodes))
        proc.start()
    ctx = pp.Context(local_rank)
    ctx.set_stream(stream.cuda_stream)
    vp = pp.ValidatorParams(RADIUS, MIN_STEP)

    def timed(fn, reps=3):
        fn()
        t = []
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn()
            torch.cuda.synchronize()
            t.append(time.perf_counter() - t0)
        return min(t), r
    # C1
    rng = np.random.default_rng(3)
    rects = [(2.0 + 10.0 * rng.uniform(), 1.0, rng.uniform(-45, 45), rng.uniform(-45, 45), rng.uniform(-math.pi, math.pi)) for _ in range(60)]
    lb, ub = (-50.0, -50.0, -math.pi), (50.0, 50.0, math.pi)
    occ = synth.rasterize_rectangles(rects, lb, ub, 0.5)
    rows, cols, origin = synth.map_geometry(lb, ub, 0.5)
    ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=0.5, world_origin=origin, lb=lb, ub=ub), occ)
    sec, r = timed(lambda: ctx.astar_n2([(1, 1)], [(198, 198)], max_path=1024))
    nb = 4096
    free = np.argwhere(occ < 0)
    pick = np.random.default_rng(4).integers(len(free), size=(2, nb))
    secb, rb = timed(lambda: ctx.astar_n2(free[pick[0]], free[pick[1]]))
    out["c1"] = {"workload": "C1: grid A* (AStarN2, Euclidean functors) on a 200x200 grid, 60 rectangle outlines, (1,1)->(198,198)",
                 "ms_single_query": sec * 1e3, "explored": int(r["results"][0]["n_explored"]), "cost": float(r["results"][0]["cost"]),
                 "batch": {"queries": nb, "value": nb / secb, "unit": "plans/s", "solved": int((rb["results"]["status"] == 0).sum())},
                 "note": "host call: H2D + kernel + D2H; a single query is one chain of dependent expansions on one warp"}
    # C2
    cfg = synth.C2
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.build_gvd(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ, voronoi=True)
    planner = pp.HybridAStar(ctx, pp.HybridParams(), vp)
    sec, r = timed(lambda: planner.search(np.array([[-18.0, -18.0, 0.0]]), np.array([[17.0, 17.5, math.pi / 2]]), np.array([42], np.uint64), max_path=256))
    wf_ms, se_ms = planner.timing()
    out["c2"] = {"workload": "C2: one Hybrid A* query on the 400x400 parking lot, (-18,-18,0) -> (17,17.5,pi/2), seed 42, Voronoi field built on the device",
                 "ms": sec * 1e3, "wavefront_ms": wf_ms, "search_ms": se_ms, "status": int(r["results"][0]["status"]),
                 "cost": float(r["results"][0]["cost"]), "n_expanded": int(r["results"][0]["n_expanded"])}
    planner.close()
    # C3 map: solve + check, rectangle variant, C5
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    ctx.build_map(pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"]), occ)
    n = N_EDGES
    s, g = synth.random_rs_queries(n, seed=1000)
    dev = torch.device("cuda", local_rank)
    d_s, d_g = torch.from_numpy(s).to(dev), torch.from_numpy(g).to(dev)
    d_valid = torch.empty(n, dtype=torch.uint8, device=dev)
    d_last = torch.empty(n, dtype=torch.float32, device=dev)
    cp, vpc = pp.RSCostParams(RMIN).to_c(), vp.to_c()

    def dev_ms(fn, reps=10):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    ms = dev_ms(lambda: ppx.ppx_connect_check_rs_dev(ctx.handle, n, d_s.data_ptr(), d_g.data_ptr(), ctypes.byref(cp), ctypes.byref(vpc),
                                                      d_valid.data_ptr(), d_last.data_ptr(), None, None, None))
    hs, hg = torch.from_numpy(s).pin_memory(), torch.from_numpy(g).pin_memory()
    hv, hl = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory()
    esec, _ = timed(lambda: ppx.ppx_connect_check_rs(ctx.handle, n, hs.data_ptr(), hg.data_ptr(), ctypes.byref(cp), ctypes.byref(vpc),
                                                      hv.data_ptr(), hl.data_ptr(), None, None, None))
    out["solve_check"] = {"workload": "C3 solve + check: PathConnectionReedsShepp::Connect + IsPathValid from 1M (start, goal) pairs",
                          "value": n / (ms * 1e-3), "unit": "edges/s", "ms_per_step": ms,
                          "e2e": {"value": n / esec, "unit": "edges/s", "h2d_bytes_per_step": n * 48, "d2h_bytes_per_step": n * 5},
                          "valid_fraction": float(d_valid.float().mean().item())}
    sol = ctx.rs_solve(s[:200000], g[:200000], pp.RSCostParams(RMIN))
    rsec, rr = timed(lambda: ctx.check_edges_rs_rect(sol["edges"], RMIN, RECT_FOOTPRINT))
    out["rect_footprint"] = {"workload": "C3 check-only with the rectangular footprint variant (4.5 x 2.0 m, 3 covering discs; NOT a reference "
                             "behaviour, defined in include/ppx.h)", "value": 200000 / rsec, "unit": "edges/s (host call, 200k edges)",
                             "valid_fraction": float(rr[0].mean()), "mean_samples": float(rr[2].mean())}
    c5 = {}
    for conn_kind, name in ((0, "straight"), (1, "reeds_shepp")):
        t0 = time.perf_counter()
        r = ctx.rrt_run(np.array([C5_START]), np.array([C5_GOAL]), np.array([7], np.uint64), cfg["lb"], cfg["ub"], dim=3, connection=conn_kind,
                        star=1, validator=1, max_iteration=10**7, max_nodes=100000, max_connection_distance=1.0, goal_bias=0.05,
                        min_turning_radius=RMIN, vp=vp, want_trees=False)["results"][0]
        dt = time.perf_counter() - t0
        c5[name] = {"nodes": int(r["n_nodes"]), "iterations": int(r["iterations"]), "edges_checked": int(r["n_edges_checked"]), "seconds": dt,
                    "nodes_per_s": float(r["n_nodes"]) / dt, "nn_queries_per_s": 2.0 * float(r["iterations"]) / dt,
                    "edges_per_s": float(r["n_edges_checked"]) / dt}
    out["c5"] = {"workload": "C5: ONE RRT* tree over SE(2) on the C3 map to 100k nodes, maxConnectionDistance 1.0, goalBias 0.05, seed 7, "
                 "occupancy validator, brute-force kNN in the loop", **c5}
    ctx.close()
    if not args.no_cpu_baseline:
        ref = a.recv()
        proc.join()
        if isinstance(ref, Exception):
            out["cpu_baseline_error"] = f"{type(ref).__name__}: {ref}"
        else:
            out["c1"]["cpu_baseline"] = {"ms": ref["c1"]["ms"], "cores": 1, "kind": "reference", "explored": ref["c1"]["explored"], "cost": ref["c1"]["cost"],
                                         "matches": bool(ref["c1"]["explored"] == out["c1"]["explored"] and ref["c1"]["cost"] == out["c1"]["cost"])}
            out["c2"]["cpu_baseline"] = {"ms": ref["c2"]["ms"], "cores": 1, "kind": "reference", "n_expanded": ref["c2"]["n_expanded"], "cost": ref["c2"]["cost"],
                                         "matches": bool(ref["c2"]["n_expanded"] == out["c2"]["n_expanded"]
                                                         and abs(ref["c2"]["cost"] - out["c2"]["cost"]) <= 1e-6 * abs(ref["c2"]["cost"])),
                                         "note": "the reference searches its own brushfire field here (12 of 160000 Voronoi-distance cells differ from the device's)"}
            out["solve_check"]["cpu_baseline"] = {"value": ref["solve_check"]["edges"] / ref["solve_check"]["seconds"], "unit": "edges/s", "cores": 1,
                                                  "kind": "reference", "sample": "first 100000 pairs of the workload"}
            out["rect_footprint"]["cpu_baseline"] = {"value": ref["rect"]["edges"] / ref["rect"]["seconds"], "unit": "edges/s", "cores": 1, "kind": "port",
                                                     "sample": "first 20000 edges"}
            for name in ("straight", "reeds_shepp"):
                rc = ref["c5_" + name]
                out["c5"][name]["cpu_baseline"] = {"nodes_per_s": rc["nodes"] / rc["seconds"], "iterations_per_s": rc["iterations"] / rc["seconds"],
                                                   "cores": 1, "kind": "port", "sample": f"the same tree to {rc['nodes']} nodes (restated loop over the "
                                                   "reference's sampler / connection / validator; O(n) brute-force neighbours like the device)"}
    return out


def ours_main(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    # CPU baseline first (rank 0, N=1 only), in a subprocess, before this process touches CUDA
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--mode", args.mode,
                                  "--steps", "3", "--warmup", "1", "--cpu-edges-per-core", str(args.cpu_edges_per_core)],
                                 capture_output=True, text=True, timeout=900)
            ref_line = json.loads(out.stdout.strip().splitlines()[-1])
            cpu_baseline = ref_line.get("cpu_baseline") or {"unavailable": ref_line.get("unavailable")}
        except Exception as e:  # pragma: no cover
            cpu_baseline = {"unavailable": f"{type(e).__name__}: {e}"}

    import torch
    import torch.distributed as dist
    import pathplanning_b200 as pp
    from pathplanning_b200 import synth
    from pathplanning_b200.lib import ppx

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = pp.Context(local_rank)
    # A dedicated (non-null) stream shared by torch and the context: ppx_ctx_set_stream(NULL) means
    # "the context's own stream", so torch's legacy default stream (handle 0) must not be passed --
    # the events below would then bracket nothing (that was the bug of the first r1 run).
    stream = torch.cuda.Stream(device=torch.device("cuda", local_rank))
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx.set_stream(stream.cuda_stream)

    # world model: replicated on every rank (setup, untimed)
    cfg = synth.C3
    occ = synth.rasterize_rectangles(synth.config_rects(cfg), cfg["lb"], cfg["ub"], cfg["resolution"])
    sq = synth.exact_sqdist(occ)
    rows, cols, origin = synth.map_geometry(cfg["lb"], cfg["ub"], cfg["resolution"])
    desc = pp.MapDesc(rows=rows, cols=cols, resolution=cfg["resolution"], world_origin=origin, lb=cfg["lb"], ub=cfg["ub"])
    # the obstacle distance map is built on the device (ppx_map_build) and checked against the host EDT
    ctx.build_map(desc, occ)
    t0 = time.perf_counter()
    sq_dev = ctx.build_map(desc, occ)
    map_build_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(sq_dev, sq), "device distance map differs from the exact EDT"
    vp = pp.ValidatorParams(RADIUS, MIN_STEP).to_c()
    cp = pp.RSCostParams(RMIN).to_c()

    # this rank's edge batches, resident in HBM.  Batch 0 of rank 0 is the edge set the CPU arm samples.
    n = N_EDGES
    dev = torch.device("cuda", local_rank)
    d_starts, d_goals, d_edges = [], [], []
    for b in range(N_BATCHES):
        s, g = synth.random_rs_queries(n, seed=1000 + 17 * (rank * N_BATCHES + b))
        d_starts.append(torch.from_numpy(s).to(dev))
        d_goals.append(torch.from_numpy(g).to(dev))
        e = torch.empty(n * 72, dtype=torch.uint8, device=dev)
        rc = ppx.ppx_rs_solve_dev(ctx.handle, n, d_starts[b].data_ptr(), d_goals[b].data_ptr(), ctypes.byref(cp),
                                  e.data_ptr(), None, None, None)
        assert rc == 0, ppx.ppx_last_error(ctx.handle)
        d_edges.append(e)
    # every batch has its own output arrays (what the overlapped launches of "rs_overlap" require)
    outs = [(torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.float32, device=dev)) for _ in range(N_BATCHES)]
    d_valid, d_last = outs[0]
    d_samples = torch.empty(n, dtype=torch.int32, device=dev)
    d_word = torch.empty(n, dtype=torch.int32, device=dev)

    def step(b, samples=False):
        if args.mode == "check":
            rc = ppx.ppx_check_edges_rs_dev(ctx.handle, n, d_edges[b].data_ptr(), RMIN, ctypes.byref(vp), outs[b][0].data_ptr(),
                                            outs[b][1].data_ptr(), d_samples.data_ptr() if samples else None)
        else:
            rc = ppx.ppx_connect_check_rs_dev(ctx.handle, n, d_starts[b].data_ptr(), d_goals[b].data_ptr(), ctypes.byref(cp),
                                              ctypes.byref(vp), d_valid.data_ptr(), d_last.data_ptr(), d_word.data_ptr(), None, None)
        assert rc == 0, ppx.ppx_last_error(ctx.handle)

    # S: mean samples visited per edge (measured once, untimed) -> algorithmic bytes per edge
    if args.mode == "check":
        step(0, samples=True)
        torch.cuda.synchronize()
        mean_samples = float(d_samples.float().mean().item())
    else:
        rc = ppx.ppx_check_edges_rs_dev(ctx.handle, n, d_edges[0].data_ptr(), RMIN, ctypes.byref(vp), d_valid.data_ptr(),
                                        d_last.data_ptr(), d_samples.data_ptr())
        torch.cuda.synchronize()
        mean_samples = float(d_samples.float().mean().item())
    valid_frac = float(d_valid.float().mean().item())
    in_bytes = 72 if args.mode == "check" else 48
    bytes_per_edge = in_bytes + 8 + 4.0 * mean_samples

    # Launch mode of the timed loop.  "rs_overlap": consecutive launches on the stream overlap by programmatic dependent
    # launch, so that the drain of one batch (a few long sample chains) is filled by the next batch.
    # Used only after it has reproduced the serial results of every batch here; the serial figure is reported as well.
    overlap_ok, want = False, None
    if args.mode == "check" and not args.serial_only:
        for b in range(N_BATCHES):
            step(b)
        torch.cuda.synchronize()
        want = [(v.clone(), l.clone()) for v, l in outs]
        for v, l in outs:
            v.zero_()
            l.zero_()
        try:
            ctx.set_option("rs_overlap", 1)
            for i in range(3 * N_BATCHES):
                step(i % N_BATCHES)
            torch.cuda.synchronize()
            overlap_ok = all(torch.equal(outs[b][0], want[b][0]) and torch.equal(outs[b][1], want[b][1]) for b in range(N_BATCHES))
        except Exception:  # pragma: no cover
            overlap_ok = False
        if not overlap_ok:
            ctx.set_option("rs_overlap", 0)

    sampler = ClockSampler(local_rank)
    sampler.start()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(args.warmup, 3)):
        step(i % N_BATCHES)
    barrier()
    launches0 = ctx.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active.set()
    wall0 = time.perf_counter()
    ev0.record(stream)
    for i in range(args.steps):
        step(i % N_BATCHES)
    ev1.record(stream)
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    sampler.active.clear()
    launches = ctx.launch_count - launches0
    ms = ev0.elapsed_time(ev1)
    # the events must bracket the kernels: device time and host wall time of the same region agree
    if not (0.5 * wall_ms <= ms <= 1.05 * wall_ms + 0.5):
        raise RuntimeError(f"timing self-check failed: events {ms:.3f} ms vs wall {wall_ms:.3f} ms")
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * n / (ms_per_step * 1e-3)

    serial = None
    if overlap_ok:
        # the overlapped loop must have left every batch's serial results behind; then the same loop, strictly serial
        if not all(torch.equal(outs[b][0], want[b][0]) and torch.equal(outs[b][1], want[b][1]) for b in range(N_BATCHES)):
            raise RuntimeError("overlapped launches gave different results")
        ctx.set_option("rs_overlap", 0)
        for i in range(3):
            step(i % N_BATCHES)
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        for i in range(args.steps):
            step(i % N_BATCHES)
        s1.record(stream)
        barrier()
        ts = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        serial_ms = float(ts.item()) / args.steps
        serial = {"value": world * n / (serial_ms * 1e-3), "unit": "edges/s", "ms_per_step": serial_ms,
                  "what": "the same loop with strictly serial launches (option off): every launch waits for the last sample "
                          "chain of the previous one"}

    # Host side of the e2e figure: pin this rank to the CPUs local to its GPU before the pinned buffers are allocated
    # (first touch places them on that NUMA node), then measure what the box gives: H2D bandwidth of rank 0 alone and of all
    # ranks at once -- the ceiling of the host-buffer call, which moves 72 B per edge over PCIe.
    numa = gpu_numa_affinity(local_rank, apply=world > 1)  # at N = 1 the CPU baselines below use every host core
    pcie = None
    if True:
        hb = torch.empty(72_000_000, dtype=torch.uint8).pin_memory()
        db = torch.empty(72_000_000, dtype=torch.uint8, device=dev)

        def h2d_gbs(reps=5):
            db.copy_(hb, non_blocking=True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                db.copy_(hb, non_blocking=True)
            torch.cuda.synchronize()
            return reps * hb.numel() / (time.perf_counter() - t0) / 1e9
        barrier()
        alone = h2d_gbs() if rank == 0 else 0.0
        barrier()
        together = h2d_gbs()
        tt = torch.tensor([together, -together, together], dtype=torch.float64, device=dev)
        if world > 1:
            tmin = tt[1:2].clone()
            dist.all_reduce(tmin, op=dist.ReduceOp.MAX)
            tsum = tt[2:3].clone()
            dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
            tt = torch.stack([tt[0], tmin[0], tsum[0]])
        pcie = {"h2d_gbs_rank0_alone": alone, "h2d_gbs_all_ranks_min": float(-tt[1].item()), "h2d_gbs_all_ranks_sum": float(tt[2].item()),
                "numa_node_rank0": numa, "e2e_ceiling_edges_per_s": float(tt[2].item()) * 1e9 / 72.0,
                "note": "pinned 72 MB copies; the e2e call cannot exceed sum / 72 B per edge"}
        del hb, db
    # e2e: the host-buffer C-ABI call a reference-side caller makes (pinned host memory)
    if args.mode == "check":
        h_in = [d_edges[b].cpu().pin_memory() for b in range(2)]
        h2d = n * 72
    else:
        h_in = [(d_starts[b].cpu().pin_memory(), d_goals[b].cpu().pin_memory()) for b in range(2)]
        h2d = n * 48
    h_valid = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_last = torch.empty(n, dtype=torch.float32).pin_memory()
    d2h = n * 5

    def e2e_step(b):
        if args.mode == "check":
            rc = ppx.ppx_check_edges_rs(ctx.handle, n, h_in[b].data_ptr(), RMIN, ctypes.byref(vp), h_valid.data_ptr(),
                                        h_last.data_ptr(), None)
        else:
            rc = ppx.ppx_connect_check_rs(ctx.handle, n, h_in[b][0].data_ptr(), h_in[b][1].data_ptr(), ctypes.byref(cp),
                                          ctypes.byref(vp), h_valid.data_ptr(), h_last.data_ptr(), None, None, None)
        assert rc == 0, ppx.ppx_last_error(ctx.handle)

    for i in range(3):
        e2e_step(i % 2)
    barrier()
    e2e_steps = max(3, min(args.steps, 20))
    sampler.active.set()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(i % 2)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    sampler.active.clear()
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * n * e2e_steps / float(te.item())
    sampler.stop()

    hybrid = None
    if not args.no_hybrid:
        try:
            hybrid = hybrid_section(args, rank, world, local_rank, stream)
        except Exception as e:  # pragma: no cover
            hybrid = {"error": f"{type(e).__name__}: {e}"}

    extras = None
    if rank == 0 and world == 1 and not args.no_extras:
        try:
            extras = extras_section(args, local_rank, stream)
        except Exception as e:  # pragma: no cover
            extras = {"error": f"{type(e).__name__}: {e}"}

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        achieved = bytes_per_edge * n / (ms_per_step * 1e-3) / 1e9
        if serial:
            serial["roofline_frac"] = bytes_per_edge * n / (serial["ms_per_step"] * 1e-3) / 1e9 / peak
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get(args.mode)
        line = {
            "metric": "collision-checked RS edges/s", "value": value, "unit": "edges/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(args.mode),  # identical on the reference arm
            "config_detail": {"edges_per_gpu_per_step": n,
                       "launches": ("overlapped: ppx_ctx_set_option(rs_overlap) = programmatic dependent launch on one stream, own "
                                    "output arrays per batch, results verified against serial launches; see \"serial\"")
                       if overlap_ok else "serial",
                       "mean_samples_per_edge": mean_samples, "valid_fraction": valid_frac,
                       "l2": f"{N_BATCHES} rotating edge batches ({N_BATCHES * n * in_bytes / 1e6:.0f} MB) > 126 MB L2; the 16 MB clearance grid is meant to stay L2-resident"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": "profiles/traffic.json (ncu --set full capture of one launch, not measured in "
                         "this run)" if traffic is not None else None, "peak_source": peak_src, "bytes_per_edge": bytes_per_edge,
                         "note": "FP64-issue / dependent-L2-latency bound, not HBM bound (DESIGN.md)"},
            "cpu_baseline": cpu_baseline,
            "map_build": {"ms": map_build_ms, "cells": int(rows * cols),
                          "what": "ppx_map_build host call: H2D occupancy + device EDT + derived grids + D2H distance map"},
            "e2e": {"value": e2e_value, "unit": "edges/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "pcie": pcie},
            "gpu_launches": launches,
            "clocks": sampler.summary(),
            "serial": serial,
            "hybrid": hybrid,
            "extras": extras,
        }
        _emit(line)
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="check", choices=["check", "solve_check"])
    ap.add_argument("--cpu-edges-per-core", type=int, default=250_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--serial-only", action="store_true", help="time strictly serial launches (no rs_overlap)")
    ap.add_argument("--no-hybrid", action="store_true", help="skip the Hybrid A* (config C4) section of the JSON line")
    ap.add_argument("--hybrid-queries", type=int, default=4096)
    ap.add_argument("--cpu-plans-per-core", type=int, default=16)
    ap.add_argument("--no-extras", action="store_true", help="skip the C1 / C2 / C5 / solve+check / rectangle objects of the JSON line")
    ap.add_argument("--cpu-c5-nodes", type=int, default=20000, help="node budget of the CPU (restated) RRT* baseline")
    args = ap.parse_args()
    _stdout_to_stderr()
    if args.impl == "reference":
        reference_main(args)
    else:
        ours_main(args)


if __name__ == "__main__":
    main()
