"""ppx_wavefront_dev_static: the wavefront with its scratch in a file-scope device array (no allocation at call time).

Not a reference behaviour (include/ppx.h says so): its bar is the existing entry point ppx_wavefront_dev -- float32 bit
patterns and explored sets identical, the tolerance of tests/test_wavefront_gpu.py -- on two calls in a row, the second
with fewer goals over a different part of the map, so that whatever the first call left in the shared array is there
when the second starts.  The file sorts last on purpose: it uses its own context.
"""
import ctypes as C

import numpy as np
import pytest

import pathplanning_b200 as pp
from pathplanning_b200.lib import ppx
from common import load_npz, c2_world_desc

pytestmark = pytest.mark.gpu

PPX_ERR_CAPACITY = -4


def _dev_pair(torch, own, goals, cells, static):
    """One call through the C ABI on device buffers of the caller (torch tensors); returns cost, explored, status."""
    n = len(goals)
    d_goals = torch.from_numpy(np.ascontiguousarray(goals, np.float64)).cuda()
    d_cost = torch.full((n, cells), -7.0, dtype=torch.float32, device="cuda")
    d_expl = torch.full((n, cells), 9, dtype=torch.uint8, device="cuda")
    d_status = torch.full((1,), 77, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    if static:
        rc = ppx.ppx_wavefront_dev_static(own.handle, n, d_goals.data_ptr(), d_cost.data_ptr(), d_expl.data_ptr(), d_status.data_ptr())
    else:
        rc = ppx.ppx_wavefront_dev(own.handle, n, d_goals.data_ptr(), d_cost.data_ptr(), d_expl.data_ptr())
    assert rc == 0, ppx.ppx_last_error(own.handle).decode()
    own.synchronize()
    return d_cost.cpu().numpy(), d_expl.cpu().numpy(), int(d_status.item()) if static else 0


def test_static_scratch_wavefront_equals_wavefront_dev_twice():
    torch = pytest.importorskip("torch")
    own = pp.Context(0)
    try:
        g = load_npz("edges_c2.npz")
        desc, occ, sq, pc = c2_world_desc(g)
        own.upload_map(desc, sq, occ, pc)
        cells = desc.rows * desc.cols
        rng = np.random.default_rng(11)
        first = np.stack([rng.uniform(-20, 20, 150), rng.uniform(-20, 20, 150), np.zeros(150)], 1)  # more goals than fit at once (64)
        first[:6] = load_npz("wavefront_c2.npz")["goals"]  # includes a goal outside the map
        second = np.stack([rng.uniform(-5, 19, 9), rng.uniform(-19, 3, 9), np.zeros(9)], 1)
        want = [_dev_pair(torch, own, goals, cells, static=False) for goals in (first, second)]
        free_before = torch.cuda.mem_get_info()[0]
        for goals, (wc, we, _) in zip((first, second), want):  # two calls in a row on the shared array
            cost, expl, status = _dev_pair(torch, own, goals, cells, static=True)
            assert status == 0
            bad = int((cost.view(np.uint32) != wc.view(np.uint32)).sum())
            assert bad == 0, f"{bad} cells differ from ppx_wavefront_dev"
            assert np.array_equal(expl, we)
            assert np.isfinite(cost).any()
        assert not np.isfinite(want[0][0][4]).any()
        # nothing was allocated on the device by the static calls (torch's own buffers of the loop are cached blocks of equal size)
        assert torch.cuda.mem_get_info()[0] >= free_before - (1 << 20)
        # NULL status word: kept in the static array
        d_goals = torch.from_numpy(second).cuda()
        d_cost = torch.empty((len(second), cells), dtype=torch.float32, device="cuda")
        assert ppx.ppx_wavefront_dev_static(own.handle, len(second), d_goals.data_ptr(), d_cost.data_ptr(), None, None) == 0
        own.synchronize()
        assert np.array_equal(d_cost.cpu().numpy().view(np.uint32), want[1][0].view(np.uint32))
    finally:
        own.close()


def test_static_scratch_rejects_a_map_beyond_the_limit():
    torch = pytest.importorskip("torch")
    own = pp.Context(0)
    try:
        rows = cols = 4200  # 17.6 M cells: 70 MB of tags per goal, the array holds 64 MiB
        desc = pp.MapDesc(rows=rows, cols=cols, resolution=0.1, world_origin=(-210.0, -210.0), lb=(-210, -210, -np.pi), ub=(210, 210, np.pi))
        own.upload_map(desc, np.full(rows * cols, 100, np.int32), np.full(rows * cols, -1, np.int32), None)
        d_goals = torch.zeros((1, 3), dtype=torch.float64, device="cuda")
        d_cost = torch.empty((1, rows * cols), dtype=torch.float32, device="cuda")
        rc = ppx.ppx_wavefront_dev_static(own.handle, 1, d_goals.data_ptr(), d_cost.data_ptr(), None, None)
        assert rc == PPX_ERR_CAPACITY
        msg = ppx.ppx_last_error(own.handle).decode()
        assert "static" in msg and "4200 x 4200" in msg and "ppx_wavefront_dev" in msg, msg
        assert ppx.ppx_wavefront_dev(own.handle, 1, d_goals.data_ptr(), d_cost.data_ptr(), None) == 0  # the allocating form takes it
        own.synchronize()
        assert float(d_cost.min().item()) == 0.0 and bool(torch.isfinite(d_cost).all().item())  # an empty map: every cell reached
    finally:
        own.close()
