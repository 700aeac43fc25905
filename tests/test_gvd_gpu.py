"""GPU parity of ppx_map_build: the obstacle squared-distance map computed on the device
against GVD::ObstacleDistanceMap of the compiled reference (brushfire, gvd.cpp:30-89), bit for bit."""
import numpy as np
import pytest

import pathplanning_b200 as pp
from pathplanning_b200 import synth

pytestmark = pytest.mark.gpu
INT_MAX = 2**31 - 1


def _ref_world(ref, cfg, rects=None):
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in (synth.config_rects(cfg) if rects is None else rects):
        w.add_rectangle(*r)
    w.update()
    return w


@pytest.mark.parametrize("name", ["C2", "C3"])
def test_distance_map_matches_reference(ctx, ref, name):
    cfg = getattr(synth, name)
    w = _ref_world(ref, cfg)
    occ, sq_ref = w.grids()
    sq = ctx.build_map(w.map_desc(), occ)
    assert np.array_equal(sq, sq_ref), f"{(sq != sq_ref).sum()} cells differ"
    # the validator reads the derived clearance grid: same verdicts as with the uploaded reference grid
    rng = np.random.default_rng(3)
    ext = cfg["ub"][0]
    p = np.stack([rng.uniform(-ext, ext, 20000), rng.uniform(-ext, ext, 20000), rng.uniform(-np.pi, np.pi, 20000)], 1)
    a = ctx.is_state_valid(p, pp.ValidatorParams(1.0, 0.1))
    ctx.upload_map(w.map_desc(), sq_ref, occ, None)
    b = ctx.is_state_valid(p, pp.ValidatorParams(1.0, 0.1))
    assert np.array_equal(a, b)
    w.close()


def test_distance_map_random_clutter(ctx, ref):
    """Isolated random cells (the case where 8-neighbour brushfire propagation is most stressed)."""
    cfg = dict(lb=(-15.0, -10.0, -np.pi), ub=(15.0, 10.0, np.pi), resolution=0.1)
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    info = w.info()
    rng = np.random.default_rng(11)
    n = 400
    w.set_cells(rng.integers(0, info["rows"], n), rng.integers(0, info["cols"], n), 0)
    w.update()
    occ, sq_ref = w.grids()
    sq = ctx.build_map(w.map_desc(), occ)
    assert sq.shape == sq_ref.shape
    diff = sq != sq_ref
    assert not diff.any(), f"{diff.sum()} cells differ; first {np.argwhere(diff)[:3]}"
    w.close()


def test_distance_map_empty_and_single(ctx):
    desc = pp.MapDesc(rows=37, cols=53, resolution=0.5, world_origin=(0.0, 0.0), lb=(0, 0, -np.pi), ub=(18.5, 26.5, np.pi))
    occ = np.full((37, 53), -1, np.int32)
    sq = ctx.build_map(desc, occ)
    assert (sq == INT_MAX).all()  # GVD::ObstacleDistanceMap's initial value, gvd.cpp:22
    occ[5, 50] = 0
    sq = ctx.build_map(desc, occ)
    rr, cc = np.indices(occ.shape)
    assert np.array_equal(sq, (rr - 5) ** 2 + (cc - 50) ** 2)
    assert np.array_equal(sq, synth.exact_sqdist(occ))
