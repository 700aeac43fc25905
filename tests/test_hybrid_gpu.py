"""GPU parity tests of Hybrid A* (A1-A6, A8 lookup, A9) through the C ABI against the fixture
generated from the compiled reference (tests/golden/hybrid_c2.npz) and the live reference.

Bars (BASELINE.json north_star): discrete results -- successor keys / kinds / directions, expansion
sequences, statuses -- bit-exact; continuous costs, heuristics and poses within 1e-6 relative (the
device libm is not the host's, so a last-ulp difference in a pose is expected and allowed).
"""
import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_npz, c2_world_desc, assert_close_rel

pytestmark = pytest.mark.gpu
REL = 1e-6


@pytest.fixture(scope="module")
def fx():
    return load_npz("hybrid_c2.npz")


@pytest.fixture(scope="module")
def planner(ctx):
    g = load_npz("edges_c2.npz")
    desc, occ, sq, pc = c2_world_desc(g)
    ctx.upload_map(desc, sq, occ, pc)
    h = pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(1.0, 0.1))
    yield h
    h.close()


def test_nonholo_table(planner, fx):
    """NonHolonomicHeuristic::Build: 41 x 41 x 73 Reeds-Shepp costs (float precision)."""
    table, off = planner.table()
    assert table.shape == fx["table"].shape and np.array_equal(off, fx["table_off"])
    ref = fx["table"].astype(np.float64)
    assert_close_rel(table, ref, REL, "table")
    assert (table != ref).mean() < 1e-3  # float-rounded costs: all but a handful are bit-identical


def test_heuristics(planner, fx):
    v, nh, ob = planner.heuristic(fx["heur_goal"], fx["heur_poses"])
    assert_close_rel(ob, fx["heur_obstacles"], REL, "obstacle heuristic")
    assert_close_rel(nh, fx["heur_nonholo"], REL, "non-holonomic heuristic")
    assert_close_rel(v, fx["heur_combined"], REL, "combined heuristic")


def _keys_match(keys, ref_keys, ref_poses, spatial=1.0, angular=0.0872, eps=1e-9):
    """Discrete poses must be identical, except for a component whose REFERENCE continuous value lies
    within eps of a discretisation boundary: there the reference's own key flips with the last ulp of
    its libm (e.g. a Reeds-Shepp shot ending on a goal with an integer coordinate), so either
    neighbouring cell is accepted."""
    keys, ref_keys = np.atleast_2d(keys), np.atleast_2d(ref_keys)
    diff = keys != ref_keys
    if not diff.any():
        return True
    q = np.atleast_2d(ref_poses) / np.array([spatial, spatial, angular])
    near = np.abs(q - np.round(q)) < eps
    return bool((near | ~diff).all() and (np.abs(keys - ref_keys)[:, :2] <= 1).all())


def test_expand(planner, fx):
    cnt, ch, ph = planner.expand(fx["heur_goal"], fx["ex_parents"], fx["ex_seeds"])
    assert np.array_equal(cnt, fx["ex_n"])
    assert_close_rel(ph, fx["ex_parent_heuristic"], REL, "parent heuristic")
    for i in range(len(cnt)):
        n = cnt[i]
        c = ch[i, :n]
        assert _keys_match(c["key"], fx["ex_keys"][i, :n], fx["ex_poses"][i, :n]), i
        assert np.array_equal(c["direction"], fx["ex_dirs"][i, :n]), i
        assert np.array_equal(c["kind"], fx["ex_kinds"][i, :n]), i
        assert np.abs(c["pose"] - fx["ex_poses"][i, :n]).max(initial=0) < 1e-9, i
        assert_close_rel(c["cost"], fx["ex_costs"][i, :n], REL, f"cost {i}")
        hr = fx["ex_heuristics"][i, :n]
        assert np.all(np.abs(c["heuristic"] - hr) <= 1e-9 + REL * np.abs(hr)), f"heuristic {i}"


def _check_query(fx, i, res, path, log):
    assert int(res["status"]) == fx["q_status"][i], (i, res)
    assert int(res["n_expanded"]) == fx["q_expanded"][i], (i, res)
    assert int(res["n_children"]) == fx["q_children"][i], (i, res)
    keys = fx[f"q{i}_keys"].astype(np.int32)
    n = len(keys)
    assert np.array_equal(log["key"][:n], keys[:, :3]), i
    assert np.array_equal(log["direction"][:n], keys[:, 3]), i
    assert np.abs(log["pose"][:n] - fx[f"q{i}_poses"]).max(initial=0) < 1e-9, i
    if fx["q_status"][i] == 0:
        assert_close_rel([res["cost"]], [fx["q_cost"][i]], REL, f"cost {i}")
        m = int(res["path_len"])
        assert m == len(fx[f"q{i}_path"])
        assert np.abs(path["pose"][:m] - fx[f"q{i}_path"]).max() < 1e-9
        assert np.array_equal(path["kind"][1:m], fx[f"q{i}_kinds"][1:])
        assert_close_rel(path["length"][1:m], fx[f"q{i}_lengths"][1:], REL, "action lengths")


def test_search_golden_batch(planner, fx):
    """All fixture queries in ONE batched call (one warp each), expansion logs compared entry by entry."""
    out = planner.search(fx["q_start"], fx["q_goal"], fx["q_seed"], max_path=256, max_log=8192)
    for i in range(len(fx["q_seed"])):
        _check_query(fx, i, out["results"][i], out["paths"][i], out["log"][i])


def test_search_single_equals_batch(planner, fx):
    a = planner.search(fx["q_start"][:1], fx["q_goal"][:1], fx["q_seed"][:1], max_path=256, max_log=2048)
    _check_query(fx, 0, a["results"][0], a["paths"][0], a["log"][0])


def test_search_limits(planner, fx):
    out = planner.search(fx["q_start"][:1], fx["q_goal"][:1], fx["q_seed"][:1], max_expansions=100)
    assert int(out["results"][0]["status"]) == -6 and int(out["results"][0]["n_expanded"]) == 100
    out = planner.search(fx["q_start"][:1], fx["q_goal"][:1], fx["q_seed"][:1], max_nodes=64)
    assert int(out["results"][0]["status"]) == -4


def test_search_live_reference(ctx, ref, planner):
    """200 fresh queries against the reference running here: statuses, expansion counts and hashes of
    the expansion sequences (compared through the logs), costs within 1e-6."""
    g = load_npz("edges_c2.npz")
    w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
    for r in g["rects"]:
        w.add_rectangle(*r)
    w.update()
    w.set_validator_params(1.0, 0.1)
    H = ref.Hybrid(w)
    rng = np.random.default_rng(99)
    starts, goals = [], []
    while len(starts) < 200:
        p = np.stack([rng.uniform(-18.5, 18.5, 2), rng.uniform(-18.5, 18.5, 2), rng.uniform(-np.pi, np.pi, 2)], 1)
        if w.is_state_valid(p).all():
            starts.append(p[0])
            goals.append(p[1])
    starts, goals = np.array(starts), np.array(goals)
    seeds = np.arange(200, dtype=np.uint64) + 5000
    out = planner.search(starts, goals, seeds, max_path=256, max_log=12000)
    differing = []
    for i in range(200):
        r = H.search(starts[i], goals[i], int(seeds[i]))
        res = out["results"][i]
        n = r["n_expanded"]
        same = (int(res["status"]) == r["status"] and int(res["n_expanded"]) == n and n <= 12000
                and np.array_equal(out["log"][i]["key"][:n], r["exp_keys"][:, :3]))
        if not same:
            differing.append(i)
            continue
        if r["status"] == 0:
            assert_close_rel([res["cost"]], [r["cost"]], REL, f"cost {i}")
    H.close()
    w.close()
    assert not differing, f"queries with a different expansion sequence: {differing}"


def test_resample_golden(planner, fx):
    """N3 (hybrid_a_star.cpp:260-291): solution chains resampled on the device vs the arrays the reference built
    (tests/golden/hybrid_resample_c2.npz): sample counts and cusp indices identical, poses within 1e-6."""
    gold = load_npz("hybrid_resample_c2.npz")
    out = planner.search(fx["q_start"], fx["q_goal"], fx["q_seed"], max_path=256)
    checked = 0
    for tag, interp in (("a", 0.1), ("b", 0.25)):
        rs = planner.resample(fx["q_goal"], out["results"], out["paths"], interp, max_points=1024)
        for i in range(len(fx["q_seed"])):
            n = int(rs["count"][i])
            if fx["q_status"][i] != 0:
                assert n == 0
                continue
            ref_poses = gold[f"q{i}{tag}_poses"]
            assert n == len(ref_poses), (i, tag, n, len(ref_poses))
            assert np.array_equal(np.flatnonzero(rs["cusp"][i, :n]), gold[f"q{i}{tag}_cusp"]), (i, tag)
            assert not rs["cusp"][i, n:].any()
            if n:
                assert_close_rel([rs["length"][i]], [float(gold[f"q{i}{tag}_length"])], REL, "composite length")
                d = np.abs(rs["poses"][i, :n] - ref_poses)
                d[:, 2] = np.minimum(d[:, 2], 2 * np.pi - d[:, 2])
                assert d.max() < 1e-6 * max(1.0, np.abs(ref_poses).max()), (i, tag, d.max())
                checked += 1
    assert checked >= 20


def test_resample_limits(planner, fx):
    out = planner.search(fx["q_start"][:1], fx["q_goal"][:1], fx["q_seed"][:1], max_path=256)
    rs = planner.resample(fx["q_goal"][:1], out["results"], out["paths"], 0.1, max_points=16)
    assert int(rs["count"][0]) == -1  # 232 samples do not fit
    out = planner.search(fx["q_start"][:1], fx["q_goal"][:1], fx["q_seed"][:1], max_path=4)
    rs = planner.resample(fx["q_goal"][:1], out["results"], out["paths"], 0.1, max_points=1024)
    assert int(rs["count"][0]) == -1  # the chain itself was cut by max_path


def test_resample_live_reference(ref, planner):
    """100 fresh queries: device search + device resampling vs the reference running here."""
    g = load_npz("edges_c2.npz")
    w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
    for r in g["rects"]:
        w.add_rectangle(*r)
    w.update()
    w.set_validator_params(1.0, 0.1)
    H = ref.Hybrid(w)
    rng = np.random.default_rng(314)
    starts, goals = [], []
    while len(starts) < 100:
        p = np.stack([rng.uniform(-18.5, 18.5, 2), rng.uniform(-18.5, 18.5, 2), rng.uniform(-np.pi, np.pi, 2)], 1)
        if w.is_state_valid(p).all():
            starts.append(p[0])
            goals.append(p[1])
    starts, goals = np.array(starts), np.array(goals)
    seeds = np.arange(100, dtype=np.uint64) + 9000
    out = planner.search(starts, goals, seeds, max_path=256, max_log=12000)
    rs = planner.resample(goals, out["results"], out["paths"], 0.1, max_points=2048)
    solved = 0
    for i in range(100):
        r = H.search(starts[i], goals[i], int(seeds[i]))
        n = r["n_expanded"]
        if not (int(out["results"][i]["status"]) == r["status"] and int(out["results"][i]["n_expanded"]) == n
                and np.array_equal(out["log"][i]["key"][:n], r["exp_keys"][:, :3])):
            continue  # expansion sequence differs (covered by test_search_live_reference)
        if r["status"] != 0:
            assert int(rs["count"][i]) == 0
            continue
        ref_poses, ref_cusp, ref_len = H.resampled(0.1)
        m = int(rs["count"][i])
        assert m == len(ref_poses), (i, m, len(ref_poses))
        assert np.array_equal(rs["cusp"][i, :m], ref_cusp), i
        d = np.abs(rs["poses"][i, :m] - ref_poses)
        d[:, 2] = np.minimum(d[:, 2], 2 * np.pi - d[:, 2])
        assert d.max() < 1e-6 * max(1.0, np.abs(ref_poses).max()), (i, d.max())
        solved += 1
    H.close()
    w.close()
    assert solved >= 60


def test_search_engines_identical(ctx, planner, fx):
    """The CTA-per-query engine (master warp + worker warps, the default) against the one-warp-per-query engine:
    result records (status, counts, node numbering, cost bits, expansion hash), solution chains and expansion logs
    are byte-identical, on the fixture queries and on 600 random ones -- more than there are resident CTAs on a
    small grid, so slots are reused."""
    rng = np.random.default_rng(2024)
    vp = pp.ValidatorParams(1.0, 0.1)
    pts = np.empty((0, 3))
    while len(pts) < 1200:
        p = np.stack([rng.uniform(-18.5, 18.5, 4000), rng.uniform(-18.5, 18.5, 4000), rng.uniform(-np.pi, np.pi, 4000)], 1)
        pts = np.concatenate([pts, p[ctx.is_state_valid(p, vp)]])
    starts = np.concatenate([fx["q_start"], pts[:600]])
    goals = np.concatenate([fx["q_goal"], pts[600:1200]])
    seeds = np.concatenate([fx["q_seed"].astype(np.uint64), np.arange(600, dtype=np.uint64) + 77])
    outs = []
    for engine in (0, 1):
        ctx.set_option("search_engine", engine)
        try:
            outs.append(planner.search(starts, goals, seeds, max_path=256, max_log=3000))
        finally:
            ctx.set_option("search_engine", 1)
    a, b = outs
    assert a["results"].tobytes() == b["results"].tobytes(), np.flatnonzero(a["results"] != b["results"])[:10]
    assert a["paths"].tobytes() == b["paths"].tobytes()
    assert a["log"].tobytes() == b["log"].tobytes()
    assert (a["results"]["status"] == 0).sum() > 300
