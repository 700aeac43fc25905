#!/usr/bin/env python
"""Generates the golden fixtures in this directory from the reference itself.

Runs only in the build container: it parses the reference's own known-answer test
(/root/reference/planner/tests/test_reeds_shepp.cpp) and calls the compiled, unmodified
reference (oracle/_ref/libpp_ref.so, built by oracle/Makefile).  The fixtures are committed so
that the CPU and GPU test suites never need /root/reference at run time.

    python tests/golden/make_golden.py
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refapi  # noqa: E402
from pathplanning_b200 import synth  # noqa: E402

REF_TEST = "/root/reference/planner/tests/test_reeds_shepp.cpp"
REF_HEADER = "/root/reference/planner/src/geometry/reeds_shepp.h"


def path_word_names():
    src = open(REF_HEADER).read()
    body = src[src.index("enum class PathWords"):]
    body = body[body.index("{") + 1:body.index("NumPathWords")]
    body = re.sub(r"//.*", "", body)
    names = [t.strip().split("=")[0].strip() for t in body.split(",") if t.strip()]
    assert names[0] == "NoPath"
    return names[1:]


def parse_kats():
    names = path_word_names()
    assert len(names) == 48
    src = open(REF_TEST).read()
    src = src[src.index("int main()"):]
    cases = []
    start = goal = None
    for line in src.splitlines():
        m = re.search(r"start = \{([^}]*)\}", line)
        if m:
            start = [float(eval(x.replace("M_PI", "3.14159265358979323846"))) for x in m.group(1).split(",")]
        m = re.search(r"goal = \{([^}]*)\}", line)
        if m:
            goal = [float(eval(x.replace("M_PI", "3.14159265358979323846"))) for x in m.group(1).split(",")]
        m = re.search(r"Planner::Test\(start, goal, \{(.*)\}\);", line)
        if m:
            toks = [t.strip() for t in m.group(1).split(",") if t.strip()]
            words = []
            for t in toks:
                if t == "optimalWord":
                    words.append(current)
                else:
                    words.append(names.index(t.split("::")[-1]))
            cases.append(dict(start=start, goal=goal, words=words))
        m = re.search(r"optimalWord = Planner::ReedsShepp::PathWords::(\w+);", line)
        if m:
            current = names.index(m.group(1))
    return names, cases


def main():
    bad, wrap = refapi.selfcheck()
    assert bad == 0, "reference build shows trap H1"
    names, cases = parse_kats()
    assert len(cases) == 48, len(cases)
    # reference outputs for the KATs (unit radius, unit multipliers, like the reference test)
    for c in cases:
        sh = refapi.rs_shortest(c["start"], c["goal"], 1.0)
        assert sh["word"] in c["words"], (c, sh["word"])
        c["ref_shortest_word"] = int(sh["word"])
        c["ref_final"] = sh["final"].tolist()
        c["ref_length"] = sh["length"]
        op = refapi.rs_optimal_batch([c["start"]], [c["goal"]], 1.0)
        assert int(op["word"][0]) in c["words"]
        c["ref_optimal_word"] = int(op["word"][0])
        c["ref_cost"] = float(op["cost"][0])
    json.dump(dict(source="planner/tests/test_reeds_shepp.cpp:38-300", word_names=names, heading_wrap_h2=wrap, cases=cases),
              open(os.path.join(HERE, "rs_kat.json"), "w"), indent=1)

    # random Reeds-Shepp solves (r=2, default multipliers and a non-default set)
    rng = np.random.default_rng(20261104)
    n = 4000
    starts = np.stack([rng.uniform(-20, 20, n), rng.uniform(-20, 20, n), rng.uniform(-np.pi, np.pi, n)], 1)
    d = rng.uniform(0.05, 14, n)
    a = rng.uniform(-np.pi, np.pi, n)
    goals = np.stack([starts[:, 0] + d * np.cos(a), starts[:, 1] + d * np.sin(a), rng.uniform(-np.pi, np.pi, n)], 1)
    out = {}
    for tag, (rmin, rev, fwd, sw) in dict(default=(2.0, 1.0, 1.0, 0.0), weighted=(1.5, 2.0, 1.1, 0.7)).items():
        r = refapi.rs_optimal_batch(starts, goals, rmin, rev, fwd, sw)
        out.update({f"{tag}_params": np.array([rmin, rev, fwd, sw]), f"{tag}_word": r["word"].astype(np.int8),
                    f"{tag}_steer": r["steer"].astype(np.int8), f"{tag}_dir": r["dir"].astype(np.int8),
                    f"{tag}_len": r["len"], f"{tag}_length": r["length"], f"{tag}_cost": r["cost"]})
    np.savez_compressed(os.path.join(HERE, "rs_random.npz"), starts=starts, goals=goals, **out)

    # collision-checked edges on the C2 parking lot (400x400 @ 0.1 m)
    cfg = synth.C2
    rects = synth.config_rects(cfg)
    w = refapi.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in rects:
        w.add_rectangle(*r)
    w.update()
    occ, sq = w.grids()
    pathcost, vorosq = w.gvd_update()
    info = w.info()
    n = 20000
    rng = np.random.default_rng(42)
    starts = np.stack([rng.uniform(-21, 21, n), rng.uniform(-21, 21, n), rng.uniform(-np.pi, np.pi, n)], 1)
    d = rng.uniform(0.0, 12, n)
    d[:50] = 0.0  # zero-length paths (start == goal)
    a = rng.uniform(-np.pi, np.pi, n)
    goals = np.stack([starts[:, 0] + d * np.cos(a), starts[:, 1] + d * np.sin(a), rng.uniform(-np.pi, np.pi, n)], 1)
    goals[:50, 2] = starts[:50, 2]
    edges = {}
    for tag, (radius, step) in dict(r10=(1.0, 0.1), r03=(0.3, 0.05)).items():
        w.set_validator_params(radius, step)
        r = w.check_rs_edges(starts, goals, 2.0)
        edges.update({f"{tag}_params": np.array([radius, step], np.float32), f"{tag}_valid": r["valid"],
                      f"{tag}_last": r["last"], f"{tag}_samples": r["samples"].astype(np.int16)})
        if tag == "r10":
            edges.update(word=r["word"].astype(np.int8), steer=r["steer"].astype(np.int8), dir=r["dir"].astype(np.int8),
                         len=r["len"], length=r["length"], cost=r["cost"])
    # constant-steer primitives and straight segments on the same map
    m = 6000
    cs_start = np.stack([rng.uniform(-20, 20, m), rng.uniform(-20, 20, m), rng.uniform(-np.pi, np.pi, m)], 1)
    deltas = np.array([0.0, 0.45, -0.45, 0.9151, -0.9151])
    cs_steer = deltas[rng.integers(0, 5, m)]
    cs_len = np.where(rng.random(m) < 0.5, 1.5, rng.uniform(0.0, 6.0, m))
    cs_dir = rng.integers(0, 2, m).astype(np.int32)
    w.set_validator_params(1.0, 0.1)
    cs = w.check_csteer(cs_start, cs_steer, cs_len, cs_dir, 2.6)
    ln_a = np.stack([rng.uniform(-20, 20, m), rng.uniform(-20, 20, m), rng.uniform(-np.pi, np.pi, m)], 1)
    ln_b = np.stack([rng.uniform(-20, 20, m), rng.uniform(-20, 20, m), rng.uniform(-np.pi, np.pi, m)], 1)
    ln_b[:, :2] = ln_a[:, :2] + (ln_b[:, :2] - ln_a[:, :2]) * rng.uniform(0, 0.3, (m, 1))
    ln = w.check_line_edges(ln_a, ln_b)
    # IsStateValid on a pose cloud straddling the map border
    pts = np.stack([rng.uniform(-20.5, 20.5, m), rng.uniform(-20.5, 20.5, m), rng.uniform(-4, 4, m)], 1)
    pts[:8] = [[-20, -20, 0], [20, 20, 0], [19.999, 19.999, 0], [-20.05, 0, 0], [0, -20.05, 0], [20.0, 0, 0], [0, 20.0, 0], [0, 0, 3.2]]
    sv = w.is_state_valid(pts)
    np.savez_compressed(
        os.path.join(HERE, "edges_c2.npz"), rects=np.array(rects), lb=np.array(cfg["lb"]), ub=np.array(cfg["ub"]),
        resolution=np.float32(cfg["resolution"]), world_origin=np.array(info["world_origin"]),
        occupancy=occ.astype(np.int16), sqdist=sq, pathcost=pathcost, vorosq=vorosq,
        starts=starts, goals=goals, **edges,
        cs_start=cs_start, cs_steer=cs_steer, cs_len=cs_len, cs_dir=cs_dir, cs_valid=cs["valid"], cs_last=cs["last"],
        cs_final=cs["final"], ln_a=ln_a, ln_b=ln_b, ln_valid=ln["valid"], ln_last=ln["last"], sv_pts=pts, sv_valid=sv)
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


def c2_world():
    cfg = synth.C2
    w = refapi.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in synth.config_rects(cfg):
        w.add_rectangle(*r)
    w.update()
    return w


WAVEFRONT_GOALS = [
    (17.0, 17.5, np.pi / 2),  # the C2 query goal
    (-16.0, -14.0, 0.0),  # centre of the first parked car: enclosed by its outline
    (0.3, 0.2, 0.0),  # map centre
    (-19.95, 19.95, 0.0),  # corner cell
    (25.0, 0.0, 0.0),  # outside the map: nothing reached
    (-16.0, -11.75, 0.0),  # on a car outline (occupied goal cell)
]


def main_wavefront():
    """ObstaclesHeuristic::Update on the C2 parking lot: two full cost grids + digests of all."""
    import hashlib
    w = c2_world()
    costs, digests, reached = [], [], []
    for g in WAVEFRONT_GOALS:
        cost, explored, _ = w.wavefront(g)
        assert np.array_equal(explored, np.isfinite(cost))
        costs.append(cost)
        digests.append(hashlib.sha256(cost.tobytes()).hexdigest())
        reached.append(int(explored.sum()))
    np.savez_compressed(os.path.join(HERE, "wavefront_c2.npz"), goals=np.array(WAVEFRONT_GOALS), cost0=costs[0], cost1=costs[1],
                        digests=np.array(digests), reached=np.array(reached))
    print("wavefront reached:", reached)


HYBRID_QUERIES_FIXED = [
    ((-18.0, -18.0, 0.0), (17.0, 17.5, np.pi / 2), 42),  # BASELINE config C2
    ((5.0, 2.0, 1.0), (5.0, 2.0, 1.0), 1),  # start == goal
    ((-18.0, -18.0, 0.0), (-16.0, -14.0, 0.0), 2),  # goal enclosed by a car outline: Failure
    ((0.0, 2.0, 0.0), (6.0, 2.5, 0.3), 3),  # short hop, solved by an early Reeds-Shepp shot
]


def hybrid_queries(w, n_random=12):
    rng = np.random.default_rng(2026)
    qs = list(HYBRID_QUERIES_FIXED)
    while len(qs) < len(HYBRID_QUERIES_FIXED) + n_random:
        p = np.stack([rng.uniform(-18.5, 18.5, 2), rng.uniform(-18.5, 18.5, 2), rng.uniform(-np.pi, np.pi, 2)], 1)
        if w.is_state_valid(p).all():
            qs.append((tuple(p[0]), tuple(p[1]), 1000 + len(qs)))
    return qs


def main_hybrid():
    """Hybrid A* on the C2 parking lot: heuristic table, heuristic lookups, GetNeighborStates on random
    parents, and whole graph searches with their expansion sequences."""
    w = c2_world()
    w.set_validator_params(1.0, 0.1)
    H = refapi.Hybrid(w)
    table, off = H.table()
    assert np.array_equal(table, table.astype(np.float32).astype(np.float64))  # float-precision costs
    goal = HYBRID_QUERIES_FIXED[0][1]
    H.prepare_goal(goal)
    rng = np.random.default_rng(77)
    m = 3000
    hp = np.stack([rng.uniform(-21, 21, m), rng.uniform(-21, 21, m), rng.uniform(-4, 4, m)], 1)
    hv, hnh, hob = H.heuristic(hp)
    par = np.stack([rng.uniform(-19.5, 19.5, 1200), rng.uniform(-19.5, 19.5, 1200), rng.uniform(-3.5, 3.5, 1200)], 1)
    par = par[w.is_state_valid(par)][:400]
    # a cluster near the goal so that Reeds-Shepp shots are attempted and succeed
    near = np.stack([goal[0] + rng.uniform(-4, 2, 60), goal[1] + rng.uniform(-6, 1, 60), rng.uniform(-np.pi, np.pi, 60)], 1)
    par = np.concatenate([par, near[w.is_state_valid(near)]])
    seeds = (np.arange(len(par)) + 11).astype(np.uint64)
    ex_n = np.zeros(len(par), np.int32)
    ex = {k: np.zeros((len(par), 11) + sh, dt) for k, sh, dt in (("poses", (3,), np.float64), ("keys", (3,), np.int32),
          ("dirs", (), np.int32), ("costs", (), np.float64), ("kinds", (), np.int32), ("heuristics", (), np.float64))}
    ex_ph = np.zeros(len(par))
    for i in range(len(par)):
        e = H.expand(par[i], int(seeds[i]))
        ex_n[i] = e["n"]
        ex_ph[i] = e["parent_heuristic"]
        for k in ex:
            ex[k][i, :e["n"]] = e[k]
    out = dict(table=table.astype(np.float32), table_off=off, heur_goal=np.array(goal), heur_poses=hp, heur_combined=hv,
               heur_nonholo=hnh, heur_obstacles=hob, ex_parents=par, ex_seeds=seeds, ex_n=ex_n, ex_parent_heuristic=ex_ph,
               **{"ex_" + k: v for k, v in ex.items()})
    qs = hybrid_queries(w)
    out["q_start"] = np.array([q[0] for q in qs])
    out["q_goal"] = np.array([q[1] for q in qs])
    out["q_seed"] = np.array([q[2] for q in qs], np.uint64)
    st, cost, nexp, nchild = [], [], [], []
    for i, (s, g, seed) in enumerate(qs):
        r = H.search(s, g, seed)
        st.append(r["status"]); cost.append(r["cost"]); nexp.append(r["n_expanded"]); nchild.append(r["n_children"])
        out[f"q{i}_keys"] = r["exp_keys"].astype(np.int16)
        out[f"q{i}_poses"] = r["exp_poses"]
        out[f"q{i}_path"] = r["path"]
        out[f"q{i}_kinds"] = r["kinds"].astype(np.int8)
        out[f"q{i}_lengths"] = r["lengths"]
        print("query", i, "status", r["status"], "cost", r["cost"], "expanded", r["n_expanded"], "children", r["n_children"])
    out.update(q_status=np.array(st), q_cost=np.array(cost), q_expanded=np.array(nexp), q_children=np.array(nchild))
    np.savez_compressed(os.path.join(HERE, "hybrid_c2.npz"), **out)
    print("hybrid_c2.npz", os.path.getsize(os.path.join(HERE, "hybrid_c2.npz")))


def main_resample():
    """N3: the resampled pose arrays (hybrid_a_star.cpp:260-291) of the hybrid_c2 queries, at two step sizes."""
    w = c2_world()
    w.set_validator_params(1.0, 0.1)
    H = refapi.Hybrid(w)
    qs = hybrid_queries(w)
    out = {}
    for i, (s, g, seed) in enumerate(qs):
        r = H.search(s, g, seed)
        if r["status"] != 0:
            continue
        for tag, interp in (("a", 0.1), ("b", 0.25)):
            poses, cusp, length = H.resampled(interp)
            out[f"q{i}{tag}_poses"] = poses
            out[f"q{i}{tag}_cusp"] = np.flatnonzero(cusp).astype(np.int32)
            out[f"q{i}{tag}_length"] = np.array(length)
        print("query", i, "samples", len(poses), "cusps", int(cusp.sum()))
    out["interp"] = np.array([0.1, 0.25], np.float32)
    np.savez_compressed(os.path.join(HERE, "hybrid_resample_c2.npz"), **out)
    print("hybrid_resample_c2.npz", os.path.getsize(os.path.join(HERE, "hybrid_resample_c2.npz")))


OPEN_WORLD = dict(lb=(-10.0, -10.0, -np.pi), ub=(10.0, 10.0, np.pi), resolution=0.1,
                  rects=[(2.0, 2.0, 0.0, 5.0, 0.0), (3.0, 1.0, 4.0, -4.0, 0.5), (1.0, 4.0, -5.0, 0.0, 0.0)])
OPEN_QUERIES = [((-8.0, -8.0, 0.0), (8.0, 8.0, 0.78), 11), ((-8.0, 6.0, -1.0), (7.0, -7.0, 0.0), 12),
                ((0.0, 0.0, 0.0), (0.0, -8.0, 3.0), 13), ((8.0, 2.0, 3.1), (-8.0, 3.0, 3.1), 14),
                ((-2.0, -6.0, 1.57), (-2.0, 8.0, 1.57), 15), ((6.0, 6.0, -2.0), (-7.0, -7.0, 2.0), 16)]


def _smooth_configs():
    import pathplanning_b200 as pp
    mc = float(np.float32(1.0 / 2.0))
    return {"d": pp.SmootherParams(max_curvature=mc),  # the defaults HybridAStar::Initialize sets: the descent diverges
            "s": pp.SmootherParams(max_curvature=mc, max_iterations=1),  # its first step: every gradient term once
            "c": pp.SmootherParams(max_curvature=mc, curvature_weight=0.0),  # without the curvature term: converges
            "p": pp.SmootherParams(max_curvature=mc, curvature_weight=0.0, path_weight=0.1, max_iterations=200)}


SMOOTH_CONFIGS = _smooth_configs()


def open_world():
    w = refapi.World(OPEN_WORLD["lb"], OPEN_WORLD["ub"], OPEN_WORLD["resolution"])
    for r in OPEN_WORLD["rects"]:
        w.add_rectangle(*r)
    w.update()
    return w


def main_smooth():
    """N2: Smoother::Smooth (smoother.cpp:33-158) on the resampled solutions of (a) the hybrid_c2 queries and (b) six
    queries on an open 20 m x 20 m lot, with the GVD cell grids the smoother reads."""
    out = {}
    for tag, w, qs in (("c2", c2_world(), None), ("open", open_world(), OPEN_QUERIES)):
        w.set_validator_params(1.0, 0.1)
        H = refapi.Hybrid(w)
        if qs is None:
            qs = hybrid_queries(w)
        k = 0
        for i, (s, g, seed) in enumerate(qs):
            r = H.search(s, g, seed)
            if r["status"] != 0:
                continue
            poses, cusp, length = H.resampled(0.1)
            if len(poses) == 0:
                continue
            if k == 0:
                a, b = H.gvd_cells()
                out[f"{tag}_near_obstacle"] = a.astype(np.int16)
                out[f"{tag}_near_edge"] = b.astype(np.int16)
                if tag == "open":
                    occ, sq = w.grids()
                    pc, _ = w.gvd_update()
                    info = w.info()
                    out.update(open_occupancy=occ.astype(np.int8), open_sqdist=sq, open_pathcost=pc,
                               open_world_origin=np.array(info["world_origin"]), open_resolution=np.float32(info["resolution"]),
                               open_lb=np.array(OPEN_WORLD["lb"]), open_ub=np.array(OPEN_WORLD["ub"]))
            out[f"{tag}{k}_in"] = poses
            out[f"{tag}{k}_cusp"] = np.flatnonzero(cusp).astype(np.int32)
            for cfg, sp in SMOOTH_CONFIGS.items():
                H.set_smoother(sp)
                st, sm = H.smooth(poses, cusp)
                out[f"{tag}{k}{cfg}_out"] = sm[:, :2]
                out[f"{tag}{k}{cfg}_status"] = np.array(st)
                print(tag, "query", i, "cfg", cfg, "points", len(poses), "status", st, "nan", bool(np.isnan(sm).any()),
                      "moved", float(np.nanmax(np.abs(sm[:, :2] - poses[:, :2]))))
            k += 1
        out[f"{tag}_n"] = np.array(k)
        H.close()
        w.close()
    np.savez_compressed(os.path.join(HERE, "smooth.npz"), **out)
    print("smooth.npz", os.path.getsize(os.path.join(HERE, "smooth.npz")))


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "base"):
        main()
    if which in ("all", "wavefront"):
        main_wavefront()
    if which in ("all", "hybrid"):
        main_hybrid()
    if which in ("all", "resample"):
        main_resample()
    if which in ("all", "smooth"):
        main_smooth()
