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


# ---- N4: obstacle outlines rasterised on the device (ppx_map_rasterize) ---------------------------------------
@pytest.mark.parametrize("name", ["C2", "C3"])
def test_rasterize_parking_lot(ctx, ref, name):
    """The bench maps: every rectangle outline through Shape::RasterizeLine on the device, ids as AddObstacle assigns
    them -> occupancy identical to the reference's; then the distance map built from it is the reference's too."""
    cfg = getattr(synth, name)
    rects = synth.config_rects(cfg)
    w = _ref_world(ref, cfg, rects)
    occ_ref, sq_ref = w.grids()
    polys = [ref.shape_vertices(0, dx, dy, x, y, th) for (dx, dy, x, y, th) in rects]
    occ = ctx.rasterize(w.map_desc(), polys)
    assert np.array_equal(occ, occ_ref), f"{(occ != occ_ref).sum()} cells differ"
    assert np.array_equal(ctx.build_map(w.map_desc(), occ), sq_ref)
    w.close()


def test_rasterize_mixed_shapes(ctx, ref):
    """Rectangles, regular polygons and circles at random poses, overlapping each other and the map border: a cell
    shared by several outlines keeps the id of the obstacle added last; cells outside the map are dropped."""
    cfg = dict(lb=(-20.0, -15.0, -np.pi), ub=(20.0, 15.0, np.pi), resolution=0.1)
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    rng = np.random.default_rng(5)
    polys = []
    for i in range(300):
        kind = int(rng.integers(0, 3))
        a = float(rng.uniform(0.3, 6.0))
        b = float(rng.uniform(0.3, 6.0)) if kind == 0 else float(rng.integers(3, 25))
        x, y, th = float(rng.uniform(-22, 22)), float(rng.uniform(-17, 17)), float(rng.uniform(-np.pi, np.pi))
        assert w.add_shape(kind, a, b, x, y, th) == 1
        polys.append(ref.shape_vertices(kind, a, b, x, y, th))
    w.update()
    occ_ref, sq_ref = w.grids()
    occ = ctx.rasterize(w.map_desc(), polys)
    assert np.array_equal(occ, occ_ref), f"{(occ != occ_ref).sum()} cells differ"
    assert len(np.unique(occ)) > 200  # ids survive
    # incremental use: the second half added onto the grid of the first half
    half = ctx.rasterize(w.map_desc(), polys[:150])
    full = ctx.rasterize(w.map_desc(), polys[150:], occupancy=half, first_id=150)
    assert np.array_equal(full, occ_ref)
    assert np.array_equal(ctx.build_map(w.map_desc(), occ), sq_ref)
    # RemoveObstacle: the outline cells of the removed obstacles become free even where another outline crosses them
    gone = [7, 42, 150, 299]
    for i in gone:
        assert w.remove_obstacle(i) == 1
    w.update()
    occ_ref2, sq_ref2 = w.grids()
    occ2 = ctx.rasterize(w.map_desc(), [polys[i] for i in gone], occupancy=occ, first_id=-1)
    assert np.array_equal(occ2, occ_ref2) and (occ2 != occ).any()
    assert np.array_equal(ctx.build_map(w.map_desc(), occ2), sq_ref2)
    w.close()


def test_rasterize_degenerate(ctx):
    desc = pp.MapDesc(rows=40, cols=60, resolution=0.25, world_origin=(-5.0, -7.5), lb=(-5, -7.5, -np.pi), ub=(5, 7.5, np.pi))
    assert (ctx.rasterize(desc, []) == -1).all()
    one = ctx.rasterize(desc, [np.array([[0.1, 0.1]])])  # a single vertex: the edge v -> v is its own cell
    assert (one >= 0).sum() == 1 and one[int((0.1 + 5.0) / 0.25), int((0.1 + 7.5) / 0.25)] == 0
    out = ctx.rasterize(desc, [np.array([[100.0, 100.0], [120.0, 130.0], [90.0, 140.0]])])  # entirely outside
    assert (out == -1).all()
