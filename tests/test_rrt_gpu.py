"""GPU parity tests of the nearest-neighbour kernel (A10) and the RRT / RRT* engine (A11).

Pins: the reference's two 4-point Tree cases (planner/tests/test_tree.cpp:91-130),
Tree<Point2d,2> of the reference over the brute-force flann stand-in on random points, and whole
RRTR2 / RRTStarR2 runs of the reference templates in free space (node sequences bit-exact: the
straight-line arithmetic has no transcendental function).  RRT* over Pose2d cannot be instantiated
from the reference as is (rrt_star.h:181); its oracle is the restated loop in oracle/ref_harness.cpp
(reference components, restated tree) -- parity unpinned beyond that.
"""
import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_npz, c2_world_desc

pytestmark = pytest.mark.gpu


def _numpy_knn(points, queries, k):
    d = np.zeros((len(queries), len(points)))
    for c in range(points.shape[1]):
        diff = queries[:, None, c] - points[None, :, c]
        d += diff * diff
    idx = np.lexsort((np.broadcast_to(np.arange(len(points)), d.shape), d), axis=1)[:, :k]
    return idx.astype(np.int32), np.take_along_axis(d, idx, 1)


def test_knn_reference_tree_cases(ctx):
    """planner/tests/test_tree.cpp:91-130: 4 collinear points A(0,0) B(1,0) C(2,0) D(3,0)."""
    pts = np.array([[0.0, 0.0], [1.0, 0.0], [2.0, 0.0], [3.0, 0.0]])
    idx, _ = ctx.knn(pts, np.array([[1.2, 0.0]]), 1)
    assert idx[0, 0] == 1
    idx, _ = ctx.knn(pts, np.array([[2.6, 0.0]]), 2)
    assert set(idx[0].tolist()) == {2, 3}


@pytest.mark.parametrize("dim,npts,nq,k", [(2, 5000, 3, 7), (2, 20000, 4000, 11), (3, 100000, 2000, 11), (3, 7, 600, 10),
                                           # every compile-time list length of the kernels (1, 4, 8, 16, 32), both mappings
                                           (3, 3000, 2, 1), (3, 3000, 700, 1), (2, 3000, 2, 4), (3, 3000, 700, 3), (3, 3000, 5, 8),
                                           (2, 3000, 700, 16), (3, 3000, 4, 17), (3, 3000, 700, 32), (2, 40, 3, 32)])
def test_knn_matches_bruteforce(ctx, dim, npts, nq, k):
    rng = np.random.default_rng(dim * 1000 + k)
    pts = rng.uniform(-50, 50, (npts, dim))
    pts[npts // 2] = pts[3]  # an exact duplicate: equal distances resolve to the lower index
    q = rng.uniform(-50, 50, (nq, dim))
    q[0] = pts[3]
    idx, d2 = ctx.knn(pts, q, k)
    ridx, rd2 = _numpy_knn(pts, q[:200], min(k, npts))
    kk = min(k, npts)
    assert np.array_equal(idx[:200, :kk], ridx)
    assert np.array_equal(d2[:200, :kk], rd2)
    if k > npts:
        assert (idx[:, npts:] == -1).all()


@pytest.mark.parametrize("offset,spread", [(1.0e6, 0.5), (-3.0e4, 40.0), (0.0, 1.0e-3), (1.0e9, 1.0e3)])
def test_knn_exact_far_from_origin(ctx, offset, spread):
    """Coordinates far from the origin (float spacing coarser than the point spacing) and tightly clustered points: the
    neighbour order is exact in FP64 (a guard for any reduced-precision pre-test of the distance tile)."""
    rng = np.random.default_rng(77)
    pts = offset + rng.uniform(-spread, spread, (20000, 3))
    pts[:, 2] = rng.uniform(-np.pi, np.pi, len(pts))
    q = offset + rng.uniform(-spread, spread, (700, 3))
    q[:, 2] = rng.uniform(-np.pi, np.pi, len(q))
    q[5] = pts[17]
    pts[9000] = pts[17]
    idx, d2 = ctx.knn(pts, q, 11)
    ridx, rd2 = _numpy_knn(pts, q[:150], 11)
    assert np.array_equal(idx[:150], ridx)
    assert np.array_equal(d2[:150], rd2)


def test_knn_reference_tree_random(ctx, ref):
    rng = np.random.default_rng(8)
    pts = rng.uniform(-5, 5, (3000, 2))
    q = rng.uniform(-5, 5, (64, 2))
    idx, _ = ctx.knn(pts, q, 5)
    assert np.array_equal(idx, ref.tree_knn_r2(pts, q, 5))


@pytest.mark.parametrize("star", [0, 1])
def test_rrt_r2_free_space_reference(ctx, ref, star):
    """Whole runs of the reference's RRTR2 / RRTStarR2 (planner/tests/test_rrt.cpp, test_rrt_star.cpp setting):
    identical trees, node by node."""
    lb, ub = (0.0, 0.0), (5.0, 5.0)
    seeds = np.arange(6, dtype=np.uint64) + 31
    starts = np.tile([0.0, 0.0, 0.0], (6, 1))
    goals = np.tile([2.0, 2.0, 0.0], (6, 1))
    out = ctx.rrt_run(starts, goals, seeds, lb, ub, dim=2, connection=0, star=star, validator=0, max_iteration=3000,
                      max_nodes=2000, max_connection_distance=0.1, goal_bias=0.05)
    for i in range(6):
        r = ref.rrt_r2_run(lb, ub, (0.0, 0.0), (2.0, 2.0), int(seeds[i]), 3000, 2000, 0.1, 0.05, star)
        res = out["results"][i]
        assert int(res["status"]) == r["status"], (i, res)
        n = r["n_nodes"]
        assert int(res["n_nodes"]) == n
        assert np.array_equal(out["nodes"][i, :n, :2], r["nodes"])
        assert np.array_equal(out["parents"][i, :n], r["parents"])
        if star:
            assert np.array_equal(out["costs"][i, :n], r["costs"])
        if r["status"] == 0:
            assert int(res["path_len"]) == len(r["path"])


@pytest.mark.parametrize("reeds_shepp", [0, 1])
def test_rrtstar_se2_restated(ctx, ref, reeds_shepp):
    """RRT* over Pose2d on the C2 parking lot with the occupancy-map validator against the restated
    loop (reference sampling / connection / validator components)."""
    g = load_npz("edges_c2.npz")
    desc, occ, sq, pc = c2_world_desc(g)
    ctx.upload_map(desc, sq, occ, pc)
    w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
    for r in g["rects"]:
        w.add_rectangle(*r)
    w.update()
    w.set_validator_params(1.0, 0.1)
    lb, ub = tuple(g["lb"]), tuple(g["ub"])
    start, goal = (-18.0, -18.0, 0.0), (17.0, 17.5, np.pi / 2)
    seeds = np.array([5, 6], np.uint64)
    # Reeds-Shepp: steering is rejected more often (more iterations), and without goal bias the goal state is never
    # sampled, so the tree grows to the node budget instead of stopping at the first connection to the goal
    max_it = 200000 if reeds_shepp else 4000
    bias = 0.0 if reeds_shepp else 0.05
    out = ctx.rrt_run(np.tile(start, (2, 1)), np.tile(goal, (2, 1)), seeds, lb, ub, dim=3, connection=reeds_shepp, star=1,
                      validator=1, max_iteration=max_it, max_nodes=1500, max_connection_distance=1.0, goal_bias=bias,
                      min_turning_radius=2.0, vp=pp.ValidatorParams(1.0, 0.1))
    for i in range(2):
        r = ref.rrtstar_se2_restated(w, lb, ub, start, goal, int(seeds[i]), max_it, 1500, 1.0, bias, reeds_shepp, 2.0)
        res = out["results"][i]
        n = r["n_nodes"]
        assert int(res["status"]) == r["status"], (res, r["n_nodes"], r["iterations"])
        if not reeds_shepp:
            assert int(res["iterations"]) == r["iterations"]
        assert n > 100
        if not reeds_shepp:
            # straight connections: no transcendental function anywhere -> the whole tree is bit-identical
            assert int(res["n_nodes"]) == n and int(res["n_edges_checked"]) == r["edges_checked"]
            assert np.array_equal(out["parents"][i, :n], r["parents"])
            assert np.array_equal(out["nodes"][i, :n], r["nodes"])
            assert np.array_equal(out["costs"][i, :n], r["costs"])
            continue
        # Reeds-Shepp connections.  Steering truncates every connection at max_connection_distance, so
        # new nodes keep landing ON arcs of earlier connections: "parent a" and "a's child b" give the same
        # cost in exact arithmetic, and a re-steered arc reproduces an existing node up to the last ulp.
        # Both the strict `cost < bestCost` (rrt_star.h:92) and the exact-equality duplicate rule
        # (utils/tree.h:127-129) are then decided by the last bit of sin/cos, which differs between the
        # device libm and glibc.  What must hold: identical trees up to the first such event, the event is
        # an ulp-level duplicate / cost tie, and the run statistics stay together.
        m = min(n, int(res["n_nodes"]))
        close = np.abs(out["nodes"][i, :m] - r["nodes"][:m]).max(1) < 1e-9
        prefix = m if close.all() else int(np.argmin(close))
        assert n >= 1000 and prefix >= 1
        same_parent = out["parents"][i, :prefix] == r["parents"][:prefix]
        cg, cr = out["costs"][i, :prefix][~same_parent], r["costs"][:prefix][~same_parent]
        fin = np.isfinite(cr)
        assert np.array_equal(np.isfinite(cg), fin) and np.all(np.abs(cg[fin] - cr[fin]) <= 1e-9 * np.abs(cr[fin])), \
            "a differing parent must be a cost tie"
        if prefix < m:
            def near_duplicate(nodes, j):
                d = np.abs(nodes[:j] - nodes[j])
                d[:, 2] = np.abs((d[:, 2] + np.pi) % (2 * np.pi) - np.pi)
                return bool((d.max(1) < 1e-9).any())
            assert near_duplicate(out["nodes"][i], prefix) or near_duplicate(r["nodes"], prefix), \
                f"trees diverge at node {prefix} without an ulp-level duplicate"
        assert abs(int(res["n_nodes"]) - n) <= 0.1 * n
    w.close()


def test_rrtstar_se2_straight_100k_nodes(ctx, ref):
    """BASELINE config C5 at its size: RRT* over SE(2) on the C3 map to 100 000 nodes, straight connections, occupancy-map
    validator.  No transcendental function is involved, so the WHOLE tree -- every node, parent and cost, the iteration and
    edge counts -- must be bit-identical to the restated loop over the reference's sampler / connection / validator
    (oracle/ref_harness.cpp, "port": RRTStar<Pose2d> does not compile in the reference, rrt_star.h:181)."""
    from pathplanning_b200 import synth
    cfg = synth.C3
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in synth.config_rects(cfg):
        w.add_rectangle(*r)
    w.update()
    w.set_validator_params(1.0, 0.1)
    occ, sq = w.grids()
    ctx.upload_map(w.map_desc(), sq, occ, None)
    lb, ub = tuple(cfg["lb"]), tuple(cfg["ub"])
    start, goal = (-90.0, -90.0, 0.0), (90.0, 90.0, 1.0)
    nodes = 100000
    out = ctx.rrt_run(np.array([start]), np.array([goal]), np.array([7], np.uint64), lb, ub, dim=3, connection=0, star=1, validator=1,
                      max_iteration=10**7, max_nodes=nodes, max_connection_distance=1.0, goal_bias=0.05, min_turning_radius=2.0,
                      vp=pp.ValidatorParams(1.0, 0.1))
    r = ref.rrtstar_se2_restated(w, lb, ub, start, goal, 7, 10**7, nodes, 1.0, 0.05, 0, 2.0)
    w.close()
    res = out["results"][0]
    n = r["n_nodes"]
    assert n == nodes + 1 and int(res["n_nodes"]) == n
    assert int(res["status"]) == r["status"] and int(res["iterations"]) == r["iterations"] and int(res["n_edges_checked"]) == r["edges_checked"]
    assert np.array_equal(out["parents"][0, :n], r["parents"])
    assert np.array_equal(out["nodes"][0, :n], r["nodes"])
    assert np.array_equal(out["costs"][0, :n], r["costs"])
