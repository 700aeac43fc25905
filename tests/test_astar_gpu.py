"""GPU parity of ppx_astar_n2 (BASELINE config C1: grid A* on a 200x200 occupancy grid) against the compiled
reference's AStarN2 with Euclidean step-cost / heuristic functors (oracle harness ref_astar_n2): status, optimal cost
(bit for bit: only +, * and sqrt are involved), explored-state count and the path cell by cell."""
import numpy as np
import pytest

import pathplanning_b200 as pp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c1(ref):
    """200 x 200 cells at 0.5 m with 60 rectangle outlines (2..12 m x 1 m) at uniform poses in +-45 m."""
    w = ref.World((-50.0, -50.0, -np.pi), (50.0, 50.0, np.pi), 0.5)
    rng = np.random.default_rng(3)
    for _ in range(60):
        w.add_rectangle(2.0 + 10.0 * rng.uniform(), 1.0, rng.uniform(-45, 45), rng.uniform(-45, 45), rng.uniform(-np.pi, np.pi))
    w.update()
    yield w
    w.close()


@pytest.fixture(scope="module")
def actx():  # own context: the other modules keep their maps in the shared one
    c = pp.Context(0)
    yield c
    c.close()


def _check(ref_out, res, path):
    assert int(res["status"]) == ref_out["status"]
    assert int(res["n_explored"]) == ref_out["explored"]
    if ref_out["status"] == 0:
        assert res["cost"] == ref_out["cost"]
        assert int(res["path_len"]) == ref_out["path_len"]
        n = int(res["path_len"])
        assert np.array_equal(path[:n], ref_out["path"][:n])
        assert (path[n:] == -1).all()
    else:
        assert np.isinf(res["cost"]) and int(res["path_len"]) == 0


def test_astar_n2_matches_reference(actx, c1):
    occ, sq = c1.grids()
    assert occ.shape == (200, 200)
    actx.upload_map(c1.map_desc(), sq, occ, None)
    rng = np.random.default_rng(17)
    free = np.argwhere(occ < 0)
    pairs = [((1, 1), (198, 198)), ((198, 1), (1, 198)), ((100, 100), (100, 100))]
    for _ in range(40):
        a, b = free[rng.integers(len(free))], free[rng.integers(len(free))]
        pairs.append((tuple(a), tuple(b)))
    # a goal on an occupied cell is never generated as a neighbour: Failure after the reachable area is exhausted
    wall = np.argwhere(occ >= 0)
    pairs.append(((1, 1), tuple(wall[len(wall) // 2])))
    starts = np.array([p[0] for p in pairs], np.int32)
    goals = np.array([p[1] for p in pairs], np.int32)
    out = actx.astar_n2(starts, goals, max_path=1024)
    solved = 0
    for i, (s, g) in enumerate(pairs):
        r = c1.astar_n2(s, g)
        _check(r, out["results"][i], out["paths"][i])
        solved += r["status"] == 0
    assert solved >= 35 and int(out["results"][-1]["status"]) == -1
    # the C1 query itself, alone: same record as in the batch
    one = actx.astar_n2(starts[:1], goals[:1], max_path=1024)
    assert one["results"].tobytes() == out["results"][:1].tobytes() and np.array_equal(one["paths"][0], out["paths"][0])
    assert int(one["results"][0]["n_explored"]) > 5000


def test_astar_n2_limits(actx, c1):
    occ, sq = c1.grids()
    actx.upload_map(c1.map_desc(), sq, occ, None)
    out = actx.astar_n2([(1, 1)], [(198, 198)], max_nodes=100)
    assert int(out["results"][0]["status"]) == -4  # PPX_ERR_CAPACITY
    out = actx.astar_n2([(1, 1), (-1, 5)], [(198, 198), (10, 10)], max_path=8)
    assert int(out["results"][1]["status"]) == -1 and int(out["results"][1]["n_explored"]) == 0  # start outside the grid: Failure
    n = int(out["results"][0]["path_len"])
    assert n > 8 and (out["paths"][0] >= 0).all()  # the first max_path cells of a longer path


@pytest.mark.parametrize("bidirectional", [False, True])
@pytest.mark.parametrize("cost_kind,heuristic_kind", [(0, 0), (1, 1), (2, 2), (0, 3), (3, 0)])
def test_astar_functors_and_bidirectional(actx, c1, cost_kind, heuristic_kind, bidirectional):
    """ppx_astar_n2_ex: the functor table (Euclidean / Manhattan / octile / zero heuristic / per-cell cost grid) and
    BidirectionalAStarN2 (algo/bidirectional_a_star.h:41-204) against the compiled reference running the same functors as
    C++ std::function objects: status, explored-state count (both searches), optimal cost bit for bit, path cell by cell
    (forward path + reversed reverse path, the meeting cell twice)."""
    occ, sq = c1.grids()
    actx.upload_map(c1.map_desc(), sq, occ, None)
    rng = np.random.default_rng(23 + cost_kind)
    cell_cost = (1.0 + rng.uniform(0, 3, occ.shape)).astype(np.float32) if cost_kind == 3 else None
    free = np.argwhere(occ < 0)
    pairs = [((1, 1), (198, 198)), ((100, 100), (100, 100))]
    for _ in range(10):
        a, b = free[rng.integers(len(free))], free[rng.integers(len(free))]
        pairs.append((tuple(a), tuple(b)))
    wall = np.argwhere(occ >= 0)
    pairs.append(((1, 1), tuple(wall[len(wall) // 3])))
    starts = np.array([p[0] for p in pairs], np.int32)
    goals = np.array([p[1] for p in pairs], np.int32)
    out = actx.astar_n2_ex(starts, goals, cost_kind, heuristic_kind, cell_cost, bidirectional, max_path=2048)
    solved = 0
    for i, (s, g) in enumerate(pairs):
        r = c1.astar_n2_ex(s, g, cost_kind, heuristic_kind, cell_cost, bidirectional)
        _check(r, out["results"][i], out["paths"][i])
        solved += r["status"] == 0
    assert solved >= 9
