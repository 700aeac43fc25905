"""GPU parity tests of the holonomic-with-obstacles wavefront (A8) through the C ABI.

Bar: cell costs bit-exact (float32 bit patterns) and explored sets identical to
ObstaclesHeuristic::Update (algo/heuristics.cpp:106-153), whose values depend on the global pop
order of the reference's Frontier (first reach wins, LIFO ties -- SURVEY.md H5).
"""
import hashlib

import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_npz, c2_world_desc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2(ctx):
    g = load_npz("edges_c2.npz")
    desc, occ, sq, pc = c2_world_desc(g)
    ctx.upload_map(desc, sq, occ, pc)
    return g


def test_wavefront_golden(ctx, c2):
    w = load_npz("wavefront_c2.npz")
    cost, explored = ctx.wavefront(w["goals"])
    assert np.array_equal(cost[0].view(np.uint32), w["cost0"].view(np.uint32))
    assert np.array_equal(cost[1].view(np.uint32), w["cost1"].view(np.uint32))
    for i in range(len(w["goals"])):
        assert hashlib.sha256(cost[i].tobytes()).hexdigest() == str(w["digests"][i]), f"goal {i}"
        assert int(explored[i].sum()) == int(w["reached"][i])
        assert np.array_equal(explored[i], np.isfinite(cost[i]))
    assert not np.isfinite(cost[4]).any()  # goal outside the map


def test_wavefront_batch_is_order_independent(ctx, c2):
    """Many goals in flight at once give the same grids as one at a time."""
    w = load_npz("wavefront_c2.npz")
    rng = np.random.default_rng(5)
    goals = np.stack([rng.uniform(-20, 20, 300), rng.uniform(-20, 20, 300), np.zeros(300)], 1)
    goals[:6] = w["goals"]
    cost = ctx.wavefront(goals, want_explored=False)
    assert np.array_equal(cost[0].view(np.uint32), w["cost0"].view(np.uint32))
    for i in (1, 7, 123, 299):
        single = ctx.wavefront(goals[i:i + 1], want_explored=False)
        assert np.array_equal(single[0].view(np.uint32), cost[i].view(np.uint32))


def _live(ctx, ref, world, goals):
    occ, sq = world.grids()
    ctx.upload_map(world.map_desc(), sq, occ, None)
    cost, explored = ctx.wavefront(goals)
    for i, g in enumerate(goals):
        rc, re, _ = world.wavefront(g)
        bad = int((cost[i].view(np.uint32) != rc.view(np.uint32)).sum())
        assert bad == 0, f"goal {i}: {bad} cells differ"
        assert np.array_equal(explored[i], re)


def test_wavefront_live_random_clutter(ctx, ref):
    """Random single-cell clutter: exercises the corner rule and dense cost ties."""
    rng = np.random.default_rng(11)
    w = ref.World((-6.4, -4.8, -np.pi), (6.4, 4.8, np.pi), 0.1)
    rows, cols = w.info()["rows"], w.info()["cols"]
    mask = rng.random((rows, cols)) < 0.28
    rr, cc = np.nonzero(mask)
    w.set_cells(rr, cc, 0)
    w.update()
    goals = np.stack([rng.uniform(-6.4, 6.4, 6), rng.uniform(-4.8, 4.8, 6), np.zeros(6)], 1)
    _live(ctx, ref, w, goals)
    w.close()


def test_wavefront_live_empty_map(ctx, ref):
    w = ref.World((-10.0, -10.0, -np.pi), (10.0, 10.0, np.pi), 0.1)
    w.update()
    _live(ctx, ref, w, np.array([[0.0, 0.0, 0.0], [-9.99, -9.99, 0.0], [3.3, -7.7, 0.0]]))
    w.close()


def test_wavefront_live_c3(ctx, ref):
    """BASELINE-size map (2000x2000): large frozen buckets (several thousand cells)."""
    from pathplanning_b200 import synth
    cfg = synth.C3
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in synth.config_rects(cfg):
        w.add_rectangle(*r)
    w.update()
    _live(ctx, ref, w, np.array([[12.0, -30.5, 0.0]]))
    w.close()
