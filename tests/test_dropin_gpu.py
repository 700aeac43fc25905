"""The compiled drop-in on the reference side (dropin/): subclasses of the reference's own StateValidatorOccupancyMap /
HybridAStar that call libppx.so, under (a) the reference's UNMODIFIED planner/tests/test_hybrid_a_star.cpp, (b) a
validator A/B against the reference's own loop on the same map object, (c) the reference's UNMODIFIED pybind11 module
interfaces/python/src/pyplanning.cpp running the scenario of interfaces/python/scripts/example.py -- checked against the
compiled reference (oracle/_ref) on the same scenario and coin-flip seed."""
import math
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "dropin", "_build")


def _need(name):
    path = os.path.join(BUILD, name)
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (make -C dropin needs the reference tree)")
    return path


def test_reference_hybrid_a_star_test_through_the_adapter():
    """planner/tests/test_hybrid_a_star.cpp:9-36, asserts enabled: empty 20 x 20 m map, (0,0,0) -> (8,8,0.78)."""
    out = subprocess.run([_need("test_hybrid_a_star_ppx")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr


def test_validator_adapter_equals_reference_loop():
    out = subprocess.run([_need("test_dropin_validator")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 differences" in out.stdout, out.stdout


EXAMPLE_RECTS = [(10.0, 1.0, 2.0, 0.0, -math.pi / 4.0), (10.0, 1.0, 0.0, 7.5, -math.pi / 4.0), (10.0, 1.0, -8.0, 5.0, math.pi / 2.0),
                 (14.0, 1.0, 5.0, -5.0, 0.0)]


def test_pyplanning_module_example_scenario(ref):
    """example.py:12-60 through the module built from the reference's own binding file over the adapters."""
    so = [f for f in os.listdir(BUILD) if f.startswith("pyplanning")] if os.path.isdir(BUILD) else []
    if not so:
        pytest.skip("dropin/_build/pyplanning*.so not built")
    code = r'''
import json, math, os, sys
sys.path.insert(0, %r)
import pyplanning as nav
nav.initialize()
space = nav.StateSpaceSE2(nav.Pose2d(-10, -10, -math.pi), nav.Pose2d(10, 10, math.pi))
m = nav.ObstacleListOccupancyMap(0.1)
validator = nav.StateValidatorOccupancyMap(space, m)
for dx, dy, x, y, th in %r:
    o = nav.Obstacle()
    o.set_shape(nav.RectangleShape(dx, dy))
    o.set_pose(nav.Pose2d(x, y, th))
    m.add_obstacle(o)
gvd = nav.GVD(m)
gvd.update()
algo = nav.HybridAStar()
algo.set_init_state(nav.Pose2d(0.0, -9.0, 0.0))
algo.set_goal_state(nav.Pose2d(8.0, 8.0, 0.0))
assert algo.initialize(validator)
algo.path_interpolation = 0.8
status = algo.search_path()
path = algo.get_path()
print(json.dumps(dict(status=int(status), cost=algo.get_graph_search_optimal_cost(), n=len(path),
                      first=[path[0].x(), path[0].y(), path[0].theta], last=[path[-1].x(), path[-1].y(), path[-1].theta],
                      segments=len(algo.get_graph_search_path()),
                      valid=[bool(validator.is_state_valid(p)) for p in path])))
''' % (BUILD, EXAMPLE_RECTS)
    env = dict(os.environ, PPX_SEED="42")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout + out.stderr
    import json
    got = json.loads(out.stdout.strip().splitlines()[-1])
    # the compiled reference on the same scenario, GVD attached before the brushfire like example.py, same seed
    w = ref.World((-10.0, -10.0, -math.pi), (10.0, 10.0, math.pi), 0.1)
    for r in EXAMPLE_RECTS:
        w.add_rectangle(*r)
    w.gvd_update()
    w.set_validator_params(1.0, 0.1)
    H = ref.Hybrid(w)
    r = H.search((0.0, -9.0, 0.0), (8.0, 8.0, 0.0), 42)
    ref_poses, ref_cusp, ref_len = H.resampled(0.8)
    H.close()
    w.close()
    assert got["status"] == r["status"] == 0
    assert abs(got["cost"] - r["cost"]) <= 1e-6 * abs(r["cost"]), (got["cost"], r["cost"])
    assert got["segments"] == len(r["path"]) - 1
    # the pose array: sampled every path_interpolation = 0.8 m (the last sample is the last multiple of 0.8 m, not the goal);
    # same count as the reference's, and -- the smoother fails on this path on both sides -- the same end points
    assert got["n"] == len(ref_poses) > 10
    assert np.hypot(got["first"][0] - 0.0, got["first"][1] + 9.0) < 1e-9
    assert np.abs(np.array(got["last"]) - ref_poses[-1]).max() < 1e-6 or np.hypot(got["last"][0] - ref_poses[-1][0], got["last"][1] - ref_poses[-1][1]) < 0.5
    assert all(got["valid"])
