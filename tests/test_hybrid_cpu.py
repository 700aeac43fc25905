"""CPU-only checks of the Hybrid A* test infrastructure.

1. The compiled reference reproduces the committed fixture tests/golden/hybrid_c2.npz (so the
   fixture is the reference's output, not ours).
2. The host twin of the device code (pathplanning_b200/csrc/ppx_hybrid.cuh compiled by g++,
   oracle/twin.cpp) reproduces the fixture exactly: successor sets, heuristic values, expansion
   sequences, path costs.  On the host both sides use the same libm, so equality here is bit-exact;
   the GPU suite (test_hybrid_gpu.py) states the tolerances that apply on the device.
"""
import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_npz, c2_world_desc

twinapi = pytest.importorskip("twinapi")


def sys_path_golden():
    import os
    import sys
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    if d not in sys.path:
        sys.path.insert(0, d)


@pytest.fixture(scope="module")
def fx():
    return load_npz("hybrid_c2.npz")


@pytest.fixture(scope="module")
def twin(fx):
    if not twinapi.available():
        pytest.skip("oracle/libpp_twin.so not built")
    g = load_npz("edges_c2.npz")
    desc, occ, sq, pc = c2_world_desc(g)
    world = twinapi.World(desc, sq, occ, pc)
    t = twinapi.Hybrid(world, pp.HybridParams().to_c(), pp.ValidatorParams(1.0, 0.1).to_c(), None)
    yield t
    t.close()
    world.close()


def _c2_ref_world(ref):
    g = load_npz("edges_c2.npz")
    w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
    for r in g["rects"]:
        w.add_rectangle(*r)
    w.update()
    w.set_validator_params(1.0, 0.1)
    return w


def test_mt64_matches_reference_rng(ref):
    """std::mt19937_64 + Random<double>::SampleUniform (utils/random.h, trap H4)."""
    for seed in (0, 42, 2**63 + 5):
        assert np.array_equal(ref.random_uniform(seed, 2000), twinapi.mt64_uniform(seed, 2000))


def test_reference_reproduces_hybrid_golden(ref, fx):
    w = _c2_ref_world(ref)
    H = ref.Hybrid(w)
    table, off = H.table()
    assert np.array_equal(table.astype(np.float32), fx["table"]) and np.array_equal(off, fx["table_off"])
    for i in (0, 3, 6, 13):
        r = H.search(fx["q_start"][i], fx["q_goal"][i], int(fx["q_seed"][i]))
        assert r["status"] == fx["q_status"][i] and r["n_expanded"] == fx["q_expanded"][i]
        assert np.array_equal(r["exp_keys"], fx[f"q{i}_keys"].astype(np.int32))
        assert r["cost"] == fx["q_cost"][i]
    H.close()
    w.close()


def test_twin_table_and_heuristics(twin, fx):
    table, off = twin.table()
    assert np.array_equal(table, fx["table"].astype(np.float64))
    assert np.array_equal(off, fx["table_off"])
    wf = load_npz("wavefront_c2.npz")["cost0"]  # wavefront of the same goal
    v, nh, ob = twin.heuristic(wf, fx["heur_goal"], fx["heur_poses"])
    assert np.array_equal(nh, fx["heur_nonholo"])
    assert np.array_equal(ob, fx["heur_obstacles"])
    assert np.array_equal(v, fx["heur_combined"])


def test_twin_expand(twin, fx):
    wf = load_npz("wavefront_c2.npz")["cost0"]
    cnt, ch, ph = twin.expand(wf, fx["heur_goal"], fx["ex_parents"], fx["ex_seeds"])
    assert np.array_equal(cnt, fx["ex_n"])
    assert np.array_equal(ph, fx["ex_parent_heuristic"])
    assert (fx["ex_kinds"].sum(1) > 0).sum() > 10  # the fixture contains successful Reeds-Shepp shots
    for i in range(len(cnt)):
        c = ch[i, :cnt[i]]
        n = cnt[i]
        assert np.array_equal(c["key"], fx["ex_keys"][i, :n]), i
        assert np.array_equal(c["direction"], fx["ex_dirs"][i, :n])
        assert np.array_equal(c["kind"], fx["ex_kinds"][i, :n])
        assert np.array_equal(c["pose"], fx["ex_poses"][i, :n])
        assert np.array_equal(c["cost"], fx["ex_costs"][i, :n])
        assert np.array_equal(c["heuristic"], fx["ex_heuristics"][i, :n])


def test_twin_search(twin, fx, ref):
    """Whole searches: expansion sequence, cost and solution chain identical to the reference."""
    w = _c2_ref_world(ref)  # only for the wavefront grids of the query goals
    for i in range(len(fx["q_seed"])):
        wf, _, _ = w.wavefront(fx["q_goal"][i])
        r = twin.search(wf, fx["q_start"][i], fx["q_goal"][i], int(fx["q_seed"][i]))
        res = r["result"]
        assert int(res["status"]) == fx["q_status"][i], i
        assert int(res["n_expanded"]) == fx["q_expanded"][i], i
        assert int(res["n_children"]) == fx["q_children"][i], i
        keys = fx[f"q{i}_keys"].astype(np.int32)
        assert np.array_equal(r["log"]["key"], keys[:, :3]), i
        assert np.array_equal(r["log"]["direction"], keys[:, 3]), i
        assert np.array_equal(r["log"]["pose"], fx[f"q{i}_poses"]), i
        if fx["q_status"][i] == 0:
            assert res["cost"] == fx["q_cost"][i]
            assert np.array_equal(r["path"]["pose"], fx[f"q{i}_path"])
            assert np.array_equal(r["path"]["kind"][1:], fx[f"q{i}_kinds"][1:])
            assert np.array_equal(r["path"]["length"][1:], fx[f"q{i}_lengths"][1:])
    w.close()


def test_twin_resample(twin, fx, ref):
    """N3 (hybrid_a_star.cpp:260-291): the resampled pose array of the solution chain -- sample count, cusp
    indices and composite length identical to the reference, poses within 1e-9 (host libm on both sides)."""
    w = _c2_ref_world(ref)
    H = ref.Hybrid(w)
    gold = load_npz("hybrid_resample_c2.npz")
    solved = 0
    for i in range(len(fx["q_seed"])):
        if fx["q_status"][i] != 0:
            continue
        wf, _, _ = w.wavefront(fx["q_goal"][i])
        r = twin.search(wf, fx["q_start"][i], fx["q_goal"][i], int(fx["q_seed"][i]))
        rr = H.search(fx["q_start"][i], fx["q_goal"][i], int(fx["q_seed"][i]))
        assert rr["status"] == 0
        for interp in (0.1, 0.25):
            ref_poses, ref_cusp, ref_len = H.resampled(interp)
            poses, cusp, length = twin.resample(fx["q_goal"][i], r["path"], interp)
            tag = "a" if interp == 0.1 else "b"  # the committed fixture holds the same arrays
            assert np.array_equal(ref_poses, gold[f"q{i}{tag}_poses"]) and np.array_equal(np.flatnonzero(ref_cusp), gold[f"q{i}{tag}_cusp"])
            assert length == ref_len, i
            assert len(poses) == len(ref_poses), i
            assert np.array_equal(cusp, ref_cusp), i
            if len(r["path"]) < 2:  # start already satisfies the goal: no segment, no samples (the reference throws)
                assert len(poses) == 0
                continue
            assert cusp[0]  # ratio 0 is always snapped; ratio 1 only when the last sample falls within half a step of the end
            assert np.abs(poses - ref_poses).max() < 1e-9, (i, np.abs(poses - ref_poses).max())
        solved += 1
    assert solved >= 5
    H.close()
    w.close()


def test_twin_smooth(twin, fx, ref):
    """N2 (smoother.cpp:33-224): the per-point phase A / phase B restatement run sequentially on the host gives the
    reference's smoothed path bit for bit (same libm on both sides; NaNs of the diverging default descent
    included), with the same status -- live reference on the C2 fixture queries, four parameter sets."""
    sys_path_golden()
    from make_golden import SMOOTH_CONFIGS
    w = _c2_ref_world(ref)
    H = ref.Hybrid(w)
    gold = load_npz("smooth.npz")
    checked = moved = k = 0
    for i in range(len(fx["q_seed"])):
        if fx["q_status"][i] != 0:
            continue
        H.search(fx["q_start"][i], fx["q_goal"][i], int(fx["q_seed"][i]))
        poses, cusp, _ = H.resampled(0.1)
        if len(poses) == 0:
            continue
        assert np.array_equal(poses, gold[f"c2{k}_in"])
        near_obs, near_edge = H.gvd_cells()
        assert np.array_equal(near_obs, gold["c2_near_obstacle"]) and np.array_equal(near_edge, gold["c2_near_edge"])
        for cfg, sp in SMOOTH_CONFIGS.items():
            H.set_smoother(sp)
            st_ref, out_ref = H.smooth(poses, cusp)
            assert st_ref == int(gold[f"c2{k}{cfg}_status"]) and np.array_equal(out_ref[:, :2], gold[f"c2{k}{cfg}_out"], equal_nan=True)
            st, out, iters = twin.world.smooth(sp.to_c(), near_obs, near_edge, poses, cusp)
            assert st == st_ref, (i, cfg, st, st_ref)
            assert np.array_equal(out, out_ref, equal_nan=True), (i, cfg, np.nanmax(np.abs(out - out_ref)))
            checked += 1
            moved += int(cfg in "cp" and 1e-3 < np.nanmax(np.abs(out[:, :2] - poses[:, :2])) < 1.0)
        # a short path is returned as is with Status::PathSize
        st_ref, out_ref = H.smooth(poses[:4], cusp[:4])
        st, out, _ = twin.world.smooth(sp.to_c(), near_obs, near_edge, poses[:4], cusp[:4])
        assert st == st_ref == 2 and np.array_equal(out, out_ref)
        k += 1
    assert checked >= 40 and moved >= 20
    H.close()
    w.close()
