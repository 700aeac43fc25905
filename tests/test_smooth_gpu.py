"""GPU parity tests of the batched smoother (ppx_smoother_*) against tests/golden/smooth.npz, generated
from the compiled reference's Smoother::Smooth (algo/smoother.cpp:33-224) by tests/golden/make_golden.py.

Bars: statuses identical; positions within 1e-6 (relative to the map scale) wherever the reference's own number is
well conditioned.  Two upstream facts shape the cases:
* with the parameters HybridAStar::Initialize sets, the reference's descent DIVERGES (the curvature gradient has a
  1/sqrt(1 - cos^2) factor and the step grows ~100x per iteration until NaN): status is Failure for every fixture
  path, and SearchPath then returns the unsmoothed pose array.  Config "d" checks exactly that outcome; config "s"
  checks the first gradient step term by term, excluding the points whose reference value is the ill-conditioned
  quotient (they move by metres in one step; one ulp of acosf/cosf changes them at the 1e-3 level);
* without the curvature term (configs "c", "p") the descent is a contraction and 2000 steps agree to 1e-6.
"""
import os
import sys

import numpy as np
import pytest

import pathplanning_b200 as pp
from common import load_npz, c2_world_desc

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))

pytestmark = pytest.mark.gpu
TOL = 1e-6


@pytest.fixture(scope="module")
def gold():
    return load_npz("smooth.npz")


@pytest.fixture(scope="module")
def ctx():  # own context: these tests switch maps
    c = pp.Context(0)
    yield c
    c.close()


def _configs():
    mc = float(np.float32(0.5))
    return {"d": pp.SmootherParams(max_curvature=mc), "s": pp.SmootherParams(max_curvature=mc, max_iterations=1),
            "c": pp.SmootherParams(max_curvature=mc, curvature_weight=0.0),
            "p": pp.SmootherParams(max_curvature=mc, curvature_weight=0.0, path_weight=0.1, max_iterations=200)}


def _world(ctx, gold, tag):
    if tag == "c2":
        desc, occ, sq, pc = c2_world_desc(load_npz("edges_c2.npz"))
    else:
        occ = gold["open_occupancy"].astype(np.int32)
        desc = pp.MapDesc(rows=occ.shape[0], cols=occ.shape[1], resolution=float(gold["open_resolution"]),
                          world_origin=tuple(gold["open_world_origin"].tolist()), local_origin=(0.0, 0.0),
                          lb=tuple(gold["open_lb"].tolist()), ub=tuple(gold["open_ub"].tolist()))
        sq, pc = gold["open_sqdist"].astype(np.int32), gold["open_pathcost"].astype(np.float32)
    ctx.upload_map(desc, sq, occ, pc)


def _batch(gold, tag, max_points=640):
    n = int(gold[f"{tag}_n"])
    poses = np.zeros((n, max_points, 3))
    cusp = np.zeros((n, max_points), np.uint8)
    count = np.zeros(n, np.int32)
    for k in range(n):
        p = gold[f"{tag}{k}_in"]
        count[k] = len(p)
        poses[k, :len(p)] = p
        cusp[k, gold[f"{tag}{k}_cusp"]] = 1
    return poses, cusp, count


@pytest.mark.parametrize("tag", ["c2", "open"])
def test_smoother_golden(ctx, gold, tag):
    _world(ctx, gold, tag)
    poses, cusp, count = _batch(gold, tag)
    well = total = 0
    for cfg, sp in _configs().items():
        sm = pp.Smoother(ctx, sp)
        sm.set_gvd(gold[f"{tag}_near_obstacle"].astype(np.int32), gold[f"{tag}_near_edge"].astype(np.int32))
        out = sm.smooth(poses, cusp, count)
        sm.close()
        for k in range(len(count)):
            n = int(count[k])
            ref = gold[f"{tag}{k}{cfg}_out"]
            st_ref = int(gold[f"{tag}{k}{cfg}_status"])
            dev = out["poses"][k, :n]
            assert np.array_equal(dev[:, 2], poses[k, :n, 2])  # headings are not touched (smoother.cpp:152)
            if cfg == "d":  # diverging descent: the outcome is the status
                assert int(out["status"][k]) == st_ref == -1, (tag, k)
                continue
            if np.isnan(ref).any():  # 0 * inf of an inactive curvature term: Failure on both sides
                assert int(out["status"][k]) == st_ref == -1, (tag, k, cfg)
                continue
            d = np.abs(dev[:, :2] - ref).max(axis=1)
            if cfg == "s":
                ok = np.abs(ref - poses[k, :n, :2]).max(axis=1) < 1e-2  # well-conditioned points of the single step
                assert int(out["iterations"][k]) == 1
                assert d[ok].max() < TOL * 20.0, (tag, k, cfg, d[ok].max())
                well += int(ok.sum())
                total += n
                if ok.all():
                    assert int(out["status"][k]) == st_ref
            else:
                assert int(out["status"][k]) == st_ref, (tag, k, cfg, int(out["status"][k]), st_ref)
                assert d.max() < TOL * 20.0, (tag, k, cfg, d.max())
                assert int(out["iterations"][k]) == sp.max_iterations
    assert well > 0.9 * total


def test_smoother_edge_cases(ctx, gold):
    _world(ctx, gold, "open")
    poses, cusp, count = _batch(gold, "open")
    sm = pp.Smoother(ctx, _configs()["p"])
    with pytest.raises(pp.PPXError):
        sm.smooth(poses, cusp, count)  # no GVD grids yet
    sm.set_gvd(gold["open_near_obstacle"].astype(np.int32), gold["open_near_edge"].astype(np.int32))
    count2 = count.copy()
    count2[0] = 0  # no path: Failure, nothing written
    count2[1] = 4  # fewer than five points: Status::PathSize, path returned as is
    out = sm.smooth(poses, cusp, count2)
    assert int(out["status"][0]) == -1 and int(out["status"][1]) == 2
    assert np.array_equal(out["poses"][1, :4], poses[1, :4])
    full = sm.smooth(poses, cusp, count)
    for k in range(2, len(count)):  # the other paths do not depend on their neighbours in the batch
        assert np.array_equal(out["poses"][k, :count[k]], full["poses"][k, :count[k]])
    single = sm.smooth(poses[3:4], cusp[3:4], count[3:4])
    assert np.array_equal(single["poses"][0], full["poses"][3]) and single["status"][0] == full["status"][3]
    sm.close()


def test_search_resample_smooth_pipeline(ctx, gold):
    """Device search -> device resampling -> device smoothing on the C2 fixture queries, against the smoothed paths the
    reference produced from its own search (config "p")."""
    fx = load_npz("hybrid_c2.npz")
    _world(ctx, gold, "c2")
    h = pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(1.0, 0.1))
    res = h.search(fx["q_start"], fx["q_goal"], fx["q_seed"], max_path=256)
    rs = h.resample(fx["q_goal"], res["results"], res["paths"], 0.1, max_points=640)
    sm = pp.Smoother(ctx, _configs()["p"])
    sm.set_gvd(gold["c2_near_obstacle"].astype(np.int32), gold["c2_near_edge"].astype(np.int32))
    out = sm.smooth(rs["poses"], rs["cusp"], rs["count"])
    k = 0
    for i in range(len(fx["q_seed"])):
        n = int(rs["count"][i])
        if fx["q_status"][i] != 0 or n == 0:
            assert int(out["status"][i]) == -1
            continue
        ref = gold[f"c2{k}p_out"]
        assert n == len(ref)
        assert int(out["status"][i]) == int(gold[f"c2{k}p_status"])
        assert np.abs(out["poses"][i, :n, :2] - ref).max() < TOL * 20.0, (i, np.abs(out["poses"][i, :n, :2] - ref).max())
        k += 1
    assert k == int(gold["c2_n"])
    sm.close()
    h.close()
