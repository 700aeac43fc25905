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


# ---- N1, Voronoi half: ppx_gvd_build ------------------------------------------------------------------------------
def _oracle_rule():
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
    import gvd_rule
    return gvd_rule


def _gvd_world(ref, cfg, rects=None, shapes=None):
    """Reference GVD attached before the obstacle brushfire (example.py:47-48): CheckVoro active."""
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in (rects or []):
        w.add_rectangle(*r)
    for s in (shapes or []):
        w.add_shape(*s)
    pc, vd = w.gvd_update()
    occ, sq = w.grids()
    return w, occ, sq, vd, pc


def _check_against_port(g, occ, res):
    """Device grids vs the numpy restatement (exact transforms + the order-free Voronoi rule): everything identical;
    the nearest-cell grids name a seed cell at exactly the stored distance (their tie choice is the device's own)."""
    rule = _oracle_rule()
    sq, _ = rule.exact_edt(occ >= 0)
    assert np.array_equal(g["sqdist"], sq)
    no = g["nearest_obstacle"]
    edges = rule.voronoi_edges(occ, g["sqdist"], no)
    assert np.array_equal(g["voro_sqdist"] == 0, edges)
    vsq, _ = rule.exact_edt(edges)
    assert np.array_equal(g["voro_sqdist"], vsq)
    assert np.array_equal(g["pathcost"].view(np.uint32), rule.path_cost(sq, vsq, res).view(np.uint32))
    rr, cc = np.indices(occ.shape)
    for near, seed, dist in ((no, occ >= 0, g["sqdist"]), (g["nearest_edge"], edges, g["voro_sqdist"])):
        if not seed.any():
            assert (near == -1).all()
            continue
        assert seed[near[..., 0], near[..., 1]].all()
        assert np.array_equal((near[..., 0] - rr) ** 2 + (near[..., 1] - cc) ** 2, dist)


ALL = ("sqdist", "voro_sqdist", "pathcost", "nearest_obstacle", "nearest_edge")


@pytest.mark.parametrize("name", ["C2", "C3"])
def test_gvd_build_parking_lot(ctx, ref, name):
    """The whole GVD on the device vs the compiled reference on the bench maps: obstacle distance map and Voronoi-edge
    set identical on C2; the Voronoi distance map differs only where the reference's 8-neighbour brushfire
    overestimates (device = exact transform <= reference), in < 0.05 % of the cells, and the potential is
    bit-identical wherever the two distance maps are."""
    cfg = getattr(synth, name)
    w, occ, sq_ref, vd_ref, pc_ref = _gvd_world(ref, cfg, rects=synth.config_rects(cfg))
    g = ctx.build_gvd(w.map_desc(), occ, voronoi=True, want=ALL)
    assert np.array_equal(g["sqdist"], sq_ref)
    e_dev, e_ref = g["voro_sqdist"] == 0, vd_ref == 0
    if name == "C2":
        assert np.array_equal(e_dev, e_ref) and e_ref.sum() > 4000
    else:
        assert (e_dev != e_ref).sum() <= 2e-3 * e_ref.sum(), (e_dev != e_ref).sum()
    dv = g["voro_sqdist"] != vd_ref
    assert dv.mean() < 5e-4, dv.sum()
    if name == "C2":
        assert (g["voro_sqdist"][dv] < vd_ref[dv]).all()
    same = ~dv
    assert np.array_equal(g["pathcost"].view(np.uint32)[same], pc_ref.view(np.uint32)[same])
    if name == "C2":
        assert np.abs(g["pathcost"] - pc_ref).max() < 1e-2
    _check_against_port(g, occ, w.resolution)
    w.close()


def test_gvd_build_without_voronoi_is_the_fixture_potential(ctx):
    """flags = 0: the Voronoi map of a GVD created after OccupancyMap::Update() -- no edge, distance INT_MAX -- gives
    the potential of tests/golden/edges_c2.npz (generated through that call order) bit for bit."""
    from common import load_npz, c2_world_desc
    gold = load_npz("edges_c2.npz")
    desc, occ, sq, pc = c2_world_desc(gold)
    g = ctx.build_gvd(desc, occ, voronoi=False, want=ALL)
    assert np.array_equal(g["sqdist"], sq)
    assert (g["voro_sqdist"] == INT_MAX).all() and (g["nearest_edge"] == -1).all()
    assert np.array_equal(g["pathcost"].view(np.uint32), pc.view(np.uint32))


def test_gvd_build_random_clutter_and_updates(ctx, ref):
    """Arbitrarily rotated shapes: here the reference's brushfire is itself not an exact transform (its m_distance
    overestimates in a few cells per map), so the bar is the measured deviation: obstacle distances
    < 0.05 % of the cells (device smaller), Voronoi-edge cells < 4 % of the edge set.  The device is checked exactly
    against the restated rule.  Then RemoveObstacle + AddObstacle: the reference updates incrementally
    (gvd.cpp:74-89), the device rebuilds; same bars."""
    from test_oracle_cpu import random_shapes
    rng = np.random.default_rng(1)
    cfg = dict(lb=(-20.0, -20.0, -np.pi), ub=(20.0, 20.0, np.pi), resolution=0.1)
    shapes = random_shapes(rng, 58)
    w, occ, sq_ref, vd_ref, pc_ref = _gvd_world(ref, cfg, shapes=shapes)

    def check(occ, sq_ref, vd_ref):
        g = ctx.build_gvd(w.map_desc(), occ, voronoi=True, want=ALL)
        d = g["sqdist"] != sq_ref
        assert d.mean() < 5e-4 and (g["sqdist"][d] < sq_ref[d]).all()
        e_dev, e_ref = g["voro_sqdist"] == 0, vd_ref == 0
        assert (e_dev != e_ref).sum() < 0.04 * e_ref.sum(), ((e_dev != e_ref).sum(), e_ref.sum())
        _check_against_port(g, occ, w.resolution)

    check(occ, sq_ref, vd_ref)
    for i in (3, 17, 40):
        assert w.remove_obstacle(i) == 1
    extra = random_shapes(rng, 5)
    for s in extra:
        assert w.add_shape(*s) == 1
    pc2, vd2 = w.gvd_update()
    occ2, sq2 = w.grids()
    assert (occ2 != occ).any()
    check(occ2, sq2, vd2)
    w.close()


def test_gvd_build_empty_map(ctx):
    desc = pp.MapDesc(rows=37, cols=53, resolution=0.5, world_origin=(0.0, 0.0), lb=(0, 0, -np.pi), ub=(18.5, 26.5, np.pi))
    g = ctx.build_gvd(desc, np.full((37, 53), -1, np.int32), voronoi=True, want=ALL)
    assert (g["sqdist"] == INT_MAX).all() and (g["voro_sqdist"] == INT_MAX).all() and (g["pathcost"] == 0).all()
    assert (g["nearest_obstacle"] == -1).all() and (g["nearest_edge"] == -1).all()
