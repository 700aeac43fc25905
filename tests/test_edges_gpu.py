"""GPU parity tests of the collision-checked edge path (A3, A4, A7) through the C ABI.

Bars (BASELINE.json north_star): collision verdicts and chosen path words bit-exact; `last`
ratios, Reeds-Shepp lengths and costs within 1e-6 relative (double transcendental functions on
the device are not the host libm, SURVEY.md section 7 hard part 4).
"""
import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_kat, load_npz, c2_world_desc, assert_close_rel

pytestmark = pytest.mark.gpu
REL = 1e-6


def _edges_from(g, prefix=""):
    return pp.make_rs_edges(g[prefix + "starts"] if prefix else g["starts"], g["steer"], g["dir"], g["len"])


def test_rs_kat_words(ctx):
    """The reference's own 48 known-answer cases (planner/tests/test_reeds_shepp.cpp:38-300)."""
    kat = load_kat()
    starts = np.array([c["start"] for c in kat["cases"]])
    goals = np.array([c["goal"] for c in kat["cases"]])
    for shortest in (0, 1):
        r = ctx.rs_solve(starts, goals, pp.RSCostParams(1.0, 1.0, 1.0, 0.0, shortest))
        for i, c in enumerate(kat["cases"]):
            assert int(r["word"][i]) in c["words"], (i, kat["word_names"][int(r["word"][i])], c)
        key = "ref_shortest_word" if shortest else "ref_optimal_word"
        assert [int(w) for w in r["word"]] == [c[key] for c in kat["cases"]]
        assert_close_rel(r["length"], [c["ref_length"] for c in kat["cases"]], REL, "KAT length")
        # closure: the path built from the solution ends on the goal (test_reeds_shepp.cpp:17-22)
        end = ctx.rs_interpolate(r["edges"], 1.0, np.array([1.0]))[:, 0, :]
        assert np.abs(end[:, :2] - goals[:, :2]).max() < 1e-6
        assert np.abs(end[:, 2] - goals[:, 2]).max() < 1e-6 * np.pi / 180


@pytest.mark.parametrize("tag", ["default", "weighted"])
def test_rs_random_golden(ctx, tag):
    g = load_npz("rs_random.npz")
    rmin, rev, fwd, sw = g[f"{tag}_params"]
    r = ctx.rs_solve(g["starts"], g["goals"], pp.RSCostParams(rmin, rev, fwd, sw))
    word = g[f"{tag}_word"].astype(np.int32)
    assert np.array_equal(r["word"], word), f"{(r['word'] != word).sum()} words differ"
    e = r["edges"]
    code = np.zeros(len(word), np.uint64)
    for i in range(5):
        code |= ((g[f"{tag}_steer"][:, i].astype(np.uint64) & 3) | ((g[f"{tag}_dir"][:, i].astype(np.uint64) & 3) << 2)) << np.uint64(4 * i)
    valid_motion = g[f"{tag}_dir"] != 2
    mask = np.zeros(len(word), np.uint64)
    for i in range(5):
        mask |= np.where(valid_motion[:, i], np.uint64(0xF), np.uint64(0)) << np.uint64(4 * i)
    assert np.array_equal(e["code"] & mask, code & mask)
    assert_close_rel(np.where(valid_motion, e["len"], 0.0), np.where(valid_motion, g[f"{tag}_len"], 0.0), 1e-6, "motion lengths")
    assert_close_rel(r["length"], g[f"{tag}_length"], REL, "length")
    assert_close_rel(r["cost"], g[f"{tag}_cost"], REL, "cost")


@pytest.fixture(scope="module")
def c2(ctx):
    g = load_npz("edges_c2.npz")
    desc, occ, sq, pc = c2_world_desc(g)
    ctx.upload_map(desc, sq, occ, pc)
    return g


def test_is_state_valid_golden(ctx, c2):
    got = ctx.is_state_valid(c2["sv_pts"], pp.ValidatorParams(1.0, 0.1))
    assert np.array_equal(got, c2["sv_valid"])


@pytest.mark.parametrize("tag", ["r10", "r03"])
def test_check_rs_edges_golden(ctx, c2, tag):
    """Check-only: the reference's pre-solved segments go in, verdict / last / sample count out."""
    radius, step = c2[f"{tag}_params"]
    edges = pp.make_rs_edges(c2["starts"], c2["steer"], c2["dir"], c2["len"])
    valid, last, samples = ctx.check_edges_rs(edges, 2.0, pp.ValidatorParams(float(radius), float(step)), want_samples=True)
    assert np.array_equal(valid, c2[f"{tag}_valid"]), f"{(valid != c2[f'{tag}_valid']).sum()} verdicts differ"
    assert np.array_equal(samples.astype(np.int64), c2[f"{tag}_samples"].astype(np.int64))
    assert_close_rel(last, c2[f"{tag}_last"], REL, "last")
    assert 0.05 < valid.mean() < 0.95  # the fixture exercises both outcomes


def test_connect_check_rs_golden(ctx, c2):
    """Fused solve + check from raw start/goal pairs."""
    r = ctx.connect_check_rs(c2["starts"], c2["goals"], pp.RSCostParams(2.0), pp.ValidatorParams(1.0, 0.1))
    assert np.array_equal(r["word"], c2["word"].astype(np.int32))
    assert np.array_equal(r["valid"], c2["r10_valid"])
    assert_close_rel(r["last"], c2["r10_last"], REL, "last")
    assert_close_rel(r["length"], c2["length"], REL, "length")
    assert_close_rel(r["cost"], c2["cost"], REL, "cost")


def test_check_csteer_golden(ctx, c2):
    e = pp.make_csteer_edges(c2["cs_start"], c2["cs_steer"], c2["cs_len"], c2["cs_dir"], wheelbase=2.6)
    valid, last, final = ctx.check_edges_csteer(e, pp.ValidatorParams(1.0, 0.1), want_final=True)
    assert np.array_equal(valid, c2["cs_valid"])
    assert_close_rel(last, c2["cs_last"], REL, "last")
    assert np.abs(final - c2["cs_final"]).max() < 1e-9


def test_check_line_golden(ctx, c2):
    e = np.zeros(len(c2["ln_a"]), dtype=pp.LINE_EDGE)
    e["from"], e["to"] = c2["ln_a"], c2["ln_b"]
    valid, last = ctx.check_edges_line(e, pp.ValidatorParams(1.0, 0.1))
    assert np.array_equal(valid, c2["ln_valid"])
    assert_close_rel(last, c2["ln_last"], REL, "last")


def test_empty_and_degenerate(ctx, c2):
    vp = pp.ValidatorParams(1.0, 0.1)
    v, l = ctx.check_edges_rs(np.zeros(0, dtype=pp.RS_EDGE), 2.0, vp)
    assert len(v) == 0 and len(l) == 0
    # a segment with no motion: zero-length path => validity of the initial state, last = 1
    e = pp.make_rs_edges(np.array([[0.0, 0.0, 0.0], [-15.9, -12.1, 0.0]]), np.zeros((2, 5), int), np.full((2, 5), 2), np.full((2, 5), np.inf))
    v, l = ctx.check_edges_rs(e, 2.0, vp)
    assert np.array_equal(v, ctx.is_state_valid(e["start"], vp)) and np.all(l == 1.0)


def test_live_reference_large(ctx, ref):
    """100k fresh edges on the C2 map against the compiled reference running here."""
    g = load_npz("edges_c2.npz")
    w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
    for r in g["rects"]:
        w.add_rectangle(*r)
    w.update()
    occ, sq = w.grids()
    ctx.upload_map(w.map_desc(), sq, occ, None)
    rng = np.random.default_rng(99)
    n = 100000
    starts = np.stack([rng.uniform(-20, 20, n), rng.uniform(-20, 20, n), rng.uniform(-np.pi, np.pi, n)], 1)
    d, a = rng.uniform(0.1, 15, n), rng.uniform(-np.pi, np.pi, n)
    goals = np.stack([starts[:, 0] + d * np.cos(a), starts[:, 1] + d * np.sin(a), rng.uniform(-np.pi, np.pi, n)], 1)
    rr = w.check_rs_edges(starts, goals, 2.0)
    got = ctx.connect_check_rs(starts, goals, pp.RSCostParams(2.0), pp.ValidatorParams(1.0, 0.1))
    assert np.array_equal(got["word"], rr["word"])
    assert np.array_equal(got["valid"], rr["valid"])
    assert_close_rel(got["last"], rr["last"], REL, "last")
    edges = pp.make_rs_edges(starts, rr["steer"], rr["dir"], rr["len"])
    v, l, s = ctx.check_edges_rs(edges, 2.0, pp.ValidatorParams(1.0, 0.1), want_samples=True)
    assert np.array_equal(v, rr["valid"]) and np.array_equal(s.astype(np.int64), rr["samples"].astype(np.int64))
    w.close()


# ---- the FP32 cell filter of k_check_rs (ppx_fastlane.cuh) ------------------------------------
@pytest.fixture(scope="module")
def c3_edges(ctx, ref):
    """C3 map (2000x2000) + 200k pre-solved Reeds-Shepp edges of the bench workload, solved and
    checked by the compiled reference."""
    from pathplanning_b200 import synth
    cfg = synth.C3
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in synth.config_rects(cfg):
        w.add_rectangle(*r)
    w.update()
    occ, sq = w.grids()
    ctx.upload_map(w.map_desc(), sq, occ, None)
    s, g = synth.random_rs_queries(200000, seed=4242)
    rr = w.check_rs_edges(s, g, 2.0)
    rr["map"] = (w.map_desc(), sq, occ)  # tests that run after another map was uploaded to the shared context restore it
    w.close()
    return pp.make_rs_edges(s, rr["steer"], rr["dir"], rr["len"]), rr


def test_rs_filter_matches_reference_c3(ctx, c3_edges):
    edges, rr = c3_edges
    v, l, s = ctx.check_edges_rs(edges, 2.0, pp.ValidatorParams(1.0, 0.1), want_samples=True)
    assert np.array_equal(v, rr["valid"])
    assert np.array_equal(s.astype(np.int64), rr["samples"].astype(np.int64))
    assert_close_rel(l, rr["last"], REL, "last")


@pytest.mark.parametrize("radius,step", [(1.0, 0.1), (0.3, 0.05), (2.5, 0.4)])
def test_rs_filter_ab_exact(ctx, c3_edges, radius, step):
    """Filter on vs. off (every sample through the exact FP64 path): bit-identical outputs."""
    edges, _ = c3_edges
    vp = pp.ValidatorParams(radius, step)
    a = ctx.check_edges_rs(edges, 2.0, vp, want_samples=True)
    ctx.set_option("rs_filter", 0)
    try:
        b = ctx.check_edges_rs(edges, 2.0, vp, want_samples=True)
    finally:
        ctx.set_option("rs_filter", 1)
    assert np.array_equal(a[0], b[0])
    assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
    assert np.array_equal(a[2], b[2])


def test_rs_filter_audit(ctx, c3_edges):
    """Every certified sample re-evaluated exactly on the device: same cell, and the FP32 position
    error stays far below the certification margin kCellEps = 1e-3 cells."""
    edges, rr = c3_edges
    a = ctx.check_edges_rs_audit(edges, 2.0, pp.ValidatorParams(1.0, 0.1))
    assert np.array_equal(a["valid"], rr["valid"])
    assert a["cell_mismatch"] == 0
    assert a["certified"] + a["parked"] == int(a["samples"].sum())
    assert a["parked"] < 0.05 * a["certified"]
    assert a["max_dev_cells"] < 1e-3 / 4, a["max_dev_cells"]
    print("audit:", {k: a[k] for k in ("certified", "parked", "cell_mismatch", "max_dev_cells")})


def test_rs_filter_odd_edges(ctx, c3_edges):
    """Edges the filter must hand to the exact path: out-of-map starts, huge headings, near-border paths."""
    _, _ = c3_edges
    rng = np.random.default_rng(5)
    n = 4000
    starts = np.stack([rng.uniform(-104, 104, n), rng.uniform(-104, 104, n), rng.uniform(-50, 50, n)], 1)
    d, a = rng.uniform(0.0, 30, n), rng.uniform(-np.pi, np.pi, n)
    goals = np.stack([starts[:, 0] + d * np.cos(a), starts[:, 1] + d * np.sin(a), rng.uniform(-np.pi, np.pi, n)], 1)
    vp = pp.ValidatorParams(1.0, 0.1)
    sol = ctx.rs_solve(starts, goals, pp.RSCostParams(2.0))
    a = ctx.check_edges_rs(sol["edges"], 2.0, vp, want_samples=True)
    ctx.set_option("rs_filter", 0)
    try:
        b = ctx.check_edges_rs(sol["edges"], 2.0, vp, want_samples=True)
    finally:
        ctx.set_option("rs_filter", 1)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])
    assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))


def test_rs_overlap_option(ctx, c3_edges):
    """"rs_overlap": consecutive device-pointer launches may overlap (programmatic dependent launch).  Twelve launches
    over three different edge sets with own output arrays, then stream-ordered copies back: every result equals the
    serial one, also for the last launches (a launch must not complete before its predecessor)."""
    import ctypes as C
    from pathplanning_b200.lib import ppx
    edges, _ = c3_edges
    n = 60000
    vp = pp.ValidatorParams(1.0, 0.1)
    sets = [np.ascontiguousarray(edges[i * n:(i + 1) * n]) for i in range(3)]
    want = [ctx.check_edges_rs(e, 2.0, vp) for e in sets]
    vpc = vp.to_c()

    def dalloc(nbytes):
        p = C.c_void_p()
        ctx._check(ppx.ppx_dev_alloc(ctx.handle, nbytes, C.byref(p)))
        return p

    d_edges = [dalloc(e.nbytes) for e in sets]
    for d, e in zip(d_edges, sets):
        ctx._check(ppx.ppx_memcpy_h2d(ctx.handle, d, e.ctypes.data_as(C.c_void_p), e.nbytes))
    launches = 12
    d_valid = [dalloc(n) for _ in range(launches)]
    d_last = [dalloc(4 * n) for _ in range(launches)]
    ctx.set_option("rs_overlap", 1)
    try:
        for rep in range(2):
            for k in range(launches):
                ctx._check(ppx.ppx_check_edges_rs_dev(ctx.handle, n, d_edges[k % 3], 2.0, C.byref(vpc), d_valid[k], d_last[k], None))
            for k in reversed(range(launches)):
                v = np.zeros(n, np.uint8)
                l = np.zeros(n, np.float32)
                ctx._check(ppx.ppx_memcpy_d2h(ctx.handle, v.ctypes.data_as(C.c_void_p), d_valid[k], n))
                ctx._check(ppx.ppx_memcpy_d2h(ctx.handle, l.ctypes.data_as(C.c_void_p), d_last[k], 4 * n))
                assert np.array_equal(v.astype(bool), want[k % 3][0]), (rep, k)
                assert np.array_equal(l.view(np.uint32), want[k % 3][1].view(np.uint32)), (rep, k)
        # many tiny launches: dozens are resident at once, far more than the three of full-size launches -- the ring of
        # work-queue cursors (256 slots, each launch zeroes its successor's) must never hand two live launches one cursor
        tiny, count = 200, 600
        d_tv, d_tl = dalloc(tiny * count), dalloc(4 * tiny * count)
        for k in range(count):
            off = (k % (n // tiny)) * tiny
            ctx._check(ppx.ppx_check_edges_rs_dev(ctx.handle, tiny, C.c_void_p(d_edges[0].value + 72 * off), 2.0, C.byref(vpc),
                                                  C.c_void_p(d_tv.value + k * tiny), C.c_void_p(d_tl.value + 4 * k * tiny), None))
        v = np.zeros(tiny * count, np.uint8)
        l = np.zeros(tiny * count, np.float32)
        ctx._check(ppx.ppx_memcpy_d2h(ctx.handle, v.ctypes.data_as(C.c_void_p), d_tv, v.nbytes))
        ctx._check(ppx.ppx_memcpy_d2h(ctx.handle, l.ctypes.data_as(C.c_void_p), d_tl, l.nbytes))
        for k in range(count):
            off = (k % (n // tiny)) * tiny
            assert np.array_equal(v[k * tiny:(k + 1) * tiny].astype(bool), want[0][0][off:off + tiny]), k
            assert np.array_equal(l[k * tiny:(k + 1) * tiny].view(np.uint32), want[0][1][off:off + tiny].view(np.uint32)), k
        ppx.ppx_dev_free(ctx.handle, d_tv)
        ppx.ppx_dev_free(ctx.handle, d_tl)
    finally:
        ctx.set_option("rs_overlap", 0)
    # back to serial launches on the same context
    again = ctx.check_edges_rs(sets[0], 2.0, vp)
    assert np.array_equal(again[0], want[0][0]) and np.array_equal(again[1].view(np.uint32), want[0][1].view(np.uint32))
    for p in d_edges + d_valid + d_last:
        ppx.ppx_dev_free(ctx.handle, p)


def test_local_origin_nonzero(ctx, ref):
    """OccupancyMap::SetPosition != 0: ValidateBounds sees the LOCAL position, the step's distance to the border the
    WORLD position (state_validator_occupancy_map.cpp:18-20 vs :54-59).  The FP32 cell filter must certify against
    both frames: filter on / off bit-identical, and both equal to the live reference, near every border."""
    g = load_npz("edges_c2.npz")
    for pos in ((-6.0, 3.5), (7.25, -9.0)):
        w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
        for r in g["rects"][::3]:
            w.add_rectangle(*r)
        w.set_position(*pos)
        w.update()
        occ, sq = w.grids()
        desc = w.map_desc()
        assert tuple(desc.local_origin) == pos
        ctx.upload_map(desc, sq, occ, None)
        rng = np.random.default_rng(17)
        n = 40000
        starts = np.stack([rng.uniform(-21, 21, n), rng.uniform(-21, 21, n), rng.uniform(-np.pi, np.pi, n)], 1)
        d, a = rng.uniform(0.1, 15, n), rng.uniform(-np.pi, np.pi, n)
        goals = np.stack([starts[:, 0] + d * np.cos(a), starts[:, 1] + d * np.sin(a), rng.uniform(-np.pi, np.pi, n)], 1)
        rr = w.check_rs_edges(starts, goals, 2.0)
        assert 0.02 < rr["valid"].mean() < 0.98
        edges = pp.make_rs_edges(starts, rr["steer"], rr["dir"], rr["len"])
        vp = pp.ValidatorParams(1.0, 0.1)
        on = ctx.check_edges_rs(edges, 2.0, vp, want_samples=True)
        ctx.set_option("rs_filter", 0)
        try:
            off = ctx.check_edges_rs(edges, 2.0, vp, want_samples=True)
        finally:
            ctx.set_option("rs_filter", 1)
        for got in (on, off):
            assert np.array_equal(got[0], rr["valid"])
            assert np.array_equal(got[2].astype(np.int64), rr["samples"].astype(np.int64))
            assert_close_rel(got[1], rr["last"], REL, "last")
        assert np.array_equal(on[1].view(np.uint32), off[1].view(np.uint32))
        w.close()


def test_two_contexts_one_process(ctx, c2):
    """One context per GPU may live in one process, and the caller (or torch) may change the current device between
    calls: every entry point runs on its context's device and restores the caller's.  With a single GPU both contexts
    sit on device 0; the second one must not disturb the first one's scratch buffers or map."""
    import torch
    g = c2
    desc, occ, sq, pc = c2_world_desc(g)
    ctx.upload_map(desc, sq, occ, pc)  # earlier tests of this module left other maps on the shared context
    vp = pp.ValidatorParams(1.0, 0.1)
    edges = pp.make_rs_edges(g["starts"], g["steer"], g["dir"], g["len"])
    want = ctx.check_edges_rs(edges, 2.0, vp)
    ndev = torch.cuda.device_count()
    other = pp.Context(ndev - 1)
    try:
        other.upload_map(desc, sq, occ, pc)
        if ndev > 1:
            torch.cuda.set_device(0)
        got = other.check_edges_rs(edges, 2.0, vp)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32))
        if ndev > 1:
            assert torch.cuda.current_device() == 0
            torch.cuda.set_device(ndev - 1)
        again = ctx.check_edges_rs(edges, 2.0, vp)
        assert np.array_equal(again[0], want[0])
    finally:
        other.close()
        if ndev > 1:
            torch.cuda.set_device(0)


def test_rs_engines_identical(ctx, c3_edges):
    """k_check_rs2 (knot records in shared memory, bulk-copy staging, a sample per iteration; the default) against
    k_check_rs (round-1 micro-step machine): verdicts, `last` bits and sample counts identical -- for a batch that is
    not a multiple of the 32-edge grab, for a tiny batch, and for a device pointer that is NOT 16-byte aligned (the
    bulk copy needs an aligned source; the kernel then loads the records directly)."""
    import ctypes as C
    from pathplanning_b200.lib import ppx
    edges, rr = c3_edges
    ctx.upload_map(*rr["map"], None)
    vp = pp.ValidatorParams(1.0, 0.1)
    for n in (200000, 60001, 33, 5):
        e = np.ascontiguousarray(edges[:n])
        outs = []
        for engine in (0, 1):
            ctx.set_option("rs_engine", engine)
            try:
                outs.append(ctx.check_edges_rs(e, 2.0, vp, want_samples=True))
            finally:
                ctx.set_option("rs_engine", 1)
        a, b = outs
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2]), n
        assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), n
        assert np.array_equal(b[0], rr["valid"][:n])
    # misaligned device pointer: records start 72 bytes into the allocation
    n = 4097
    e = np.ascontiguousarray(edges[:n + 1])
    d = C.c_void_p()
    ctx._check(ppx.ppx_dev_alloc(ctx.handle, e.nbytes, C.byref(d)))
    d_valid, d_last = C.c_void_p(), C.c_void_p()
    ctx._check(ppx.ppx_dev_alloc(ctx.handle, n, C.byref(d_valid)))
    ctx._check(ppx.ppx_dev_alloc(ctx.handle, 4 * n, C.byref(d_last)))
    ctx._check(ppx.ppx_memcpy_h2d(ctx.handle, d, e.ctypes.data_as(C.c_void_p), e.nbytes))
    vpc = vp.to_c()
    ctx._check(ppx.ppx_check_edges_rs_dev(ctx.handle, n, C.c_void_p(d.value + 72), 2.0, C.byref(vpc), d_valid, d_last, None))
    v = np.zeros(n, np.uint8)
    l = np.zeros(n, np.float32)
    ctx._check(ppx.ppx_memcpy_d2h(ctx.handle, v.ctypes.data_as(C.c_void_p), d_valid, n))
    ctx._check(ppx.ppx_memcpy_d2h(ctx.handle, l.ctypes.data_as(C.c_void_p), d_last, 4 * n))
    assert np.array_equal(v.astype(bool), rr["valid"][1:n + 1])
    assert_close_rel(l, rr["last"][1:n + 1], REL, "last")
    for p in (d, d_valid, d_last):
        ppx.ppx_dev_free(ctx.handle, p)


@pytest.mark.parametrize("footprint", [(4.5, 2.0, 1.3, 3, 0.1), (4.5, 2.0, 0.0, 1, 0.1), (3.0, 1.6, 1.0, 5, 0.05)])
def test_rect_footprint_variant(ctx, ref, c3_edges, footprint):
    """BASELINE config 3's "rectangular car footprint".  The reference has none (one disc, two `TODO add offset`); the
    variant is DEFINED in include/ppx.h (covering discs along the heading, de-rated sphere tracing) and its checker is
    that definition restated over the reference's own Interpolate / IsStateValid / distance map -- "port", parity
    unpinned by definition, reported apart from the disc results.  Verdicts and sample counts identical, `last` 1e-6."""
    from pathplanning_b200 import synth
    edges, rr = c3_edges
    ctx.upload_map(*rr["map"], None)
    n = 60000
    cfg = synth.C3
    w = ref.World(cfg["lb"], cfg["ub"], cfg["resolution"])
    for r in synth.config_rects(cfg):
        w.add_rectangle(*r)
    w.update()
    want = w.check_rs_segments_rect(edges["start"][:n], rr["steer"][:n], rr["dir"][:n], rr["len"][:n], 2.0, footprint)
    w.close()
    v, l, s = ctx.check_edges_rs_rect(np.ascontiguousarray(edges[:n]), 2.0, footprint)
    assert 0.01 < want["valid"].mean() < 0.9
    diff = np.flatnonzero((v != want["valid"]) | (s.astype(np.int64) != want["samples"].astype(np.int64)))
    assert len(diff) <= 2, (len(diff), diff[:5])  # a disc centre within an ulp of a cell border may flip with the libm
    same = np.setdiff1d(np.arange(n), diff)
    assert_close_rel(l[same], want["last"][same], REL, "last")
    # the rectangle is stricter than the reference's single disc of the same half-width placed at the pose... only when the
    # pose disc lies inside the cover: with the centre at the pose and one disc (second case) rho >= width / 2
    if footprint[2] == 0.0 and footprint[3] == 1:
        vd, _ = ctx.check_edges_rs(np.ascontiguousarray(edges[:n]), 2.0, pp.ValidatorParams(footprint[1] / 2.0, footprint[4]))
        assert not (v & ~vd).any()
    poses = edges["start"][:5000]
    assert np.array_equal(ctx.is_state_valid_rect(poses, footprint), v[:5000] | (l[:5000] > 0) | (s[:5000] > 1))
