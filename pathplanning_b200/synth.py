"""Synthetic workloads of the BASELINE.json configs (maps, edges, queries).

Map construction is setup-time work outside the hot path: obstacle
outlines are rasterised here on the host the way the reference does (Bresenham,
state_validator/obstacle.cpp:7-61,77-94) and the squared obstacle distance is an exact
Euclidean distance transform.  Tests that compare against the compiled reference upload the
reference's own grids instead, so both sides always read identical inputs.
"""
import math

import numpy as np

INT_MAX = 2**31 - 1


def map_geometry(lb, ub, resolution):
    """OccupancyMap::InitializeSize as StateValidatorOccupancyMap's ctor calls it
    (state_validator_occupancy_map.cpp:10-12, occupancy_map.cpp:6-14): float width/height."""
    res = np.float32(resolution)
    width = np.float32(ub[0] - lb[0])
    height = np.float32(ub[1] - lb[1])
    origin = (-float(width) / 2.0, -float(height) / 2.0)
    rows = int(math.ceil(float(np.float32(width / res))))
    cols = int(math.ceil(float(np.float32(height / res))))
    return rows, cols, origin


def _trunc_cell(v, origin, res):
    return int((v - origin) / float(np.float32(res)))


def _bresenham(r0, c0, r1, c1, rows, cols, out):
    x0, y0, x1, y1 = r0, c0, r1, c1
    steep = abs(y1 - y0) > abs(x1 - x0)
    if steep:
        x0, y0 = y0, x0
        x1, y1 = y1, x1
    if x0 > x1:
        x0, x1 = x1, x0
        y0, y1 = y1, y0
    dx = x1 - x0
    dy = abs(y1 - y0)
    err = int(dx / 2)
    ystep = 1 if y0 < y1 else -1
    y = y0
    for x in range(x0, x1 + 1):
        r, c = (y, x) if steep else (x, y)
        if 0 <= r < rows and 0 <= c < cols:
            out.append((r, c))
        err -= dy
        if err < 0:
            y += ystep
            err += dx


def rasterize_rectangles(rects, lb, ub, resolution):
    """rects: iterable of (dx, dy, x, y, theta).  Returns the int32 occupancy grid (-1 free,
    obstacle id otherwise) like ObstacleListOccupancyMap::AddObstacle builds it."""
    rows, cols, origin = map_geometry(lb, ub, resolution)
    occ = np.full((rows, cols), -1, dtype=np.int32)
    for oid, (dx, dy, x, y, th) in enumerate(rects):
        dx2, dy2 = dx / 2.0, dy / 2.0
        c, s = math.cos(th), math.sin(th)
        verts = []
        for vx, vy in ((dx2, dy2), (-dx2, dy2), (-dx2, -dy2), (dx2, -dy2)):
            wx = (c * vx + (-s) * vy) + x
            wy = (s * vx + c * vy) + y
            verts.append((_trunc_cell(wx, origin[0], resolution), _trunc_cell(wy, origin[1], resolution)))
        cells = []
        for i in range(4):
            (r0, c0), (r1, c1) = verts[i], verts[(i + 1) % 4]
            _bresenham(r0, c0, r1, c1, rows, cols, cells)
        if cells:
            rr, cc = zip(*cells)
            occ[np.array(rr), np.array(cc)] = oid
    return occ


def exact_sqdist(occ):
    """Exact squared Euclidean cell distance to the nearest occupied cell (int32; INT_MAX when the
    map is empty, like GVD::ObstacleDistanceMap's initial value, gvd.cpp:22)."""
    from scipy import ndimage
    free = occ < 0
    if free.all():
        return np.full(occ.shape, INT_MAX, dtype=np.int32)
    idx = ndimage.distance_transform_edt(free, return_distances=False, return_indices=True)
    rr, cc = np.indices(occ.shape)
    dr = (idx[0] - rr).astype(np.int64)
    dc = (idx[1] - cc).astype(np.int64)
    return (dr * dr + dc * dc).astype(np.int32)


def parking_lot(lb, ub, prob, seed, margin=4.0, row_pitch=8.0, spot_pitch=3.0):
    """Rows of parked cars (4.5 x 2.0 m rectangles, nose-in at theta = pi/2), each spot occupied
    with probability `prob`."""
    rng = np.random.default_rng(seed)
    rects = []
    y = lb[1] + margin + 2.0
    while y < ub[1] - margin:
        x = lb[0] + margin
        while x < ub[0] - margin:
            if rng.random() < prob:
                rects.append((4.5, 2.0, float(x), float(y), math.pi / 2))
            x += spot_pitch
        y += row_pitch
    return rects


# ---- the named configs --------------------------------------------------------------------------
C2 = dict(name="C2", lb=(-20.0, -20.0, -math.pi), ub=(20.0, 20.0, math.pi), resolution=0.1, prob=0.75, seed=12345)
C3 = dict(name="C3", lb=(-100.0, -100.0, -math.pi), ub=(100.0, 100.0, math.pi), resolution=0.1, prob=0.5, seed=7)


def config_rects(cfg):
    return parking_lot(cfg["lb"], cfg["ub"], cfg["prob"], cfg["seed"])


def random_rs_queries(n, seed, extent=88.0, dmin=2.0, dmax=20.0):
    """BASELINE config C3 edge endpoints: start uniform in +-extent with theta~U(-pi,pi); goal at
    distance U(dmin,dmax) in a uniform direction with theta~U(-pi,pi)."""
    rng = np.random.default_rng(seed)
    starts = np.empty((n, 3))
    starts[:, 0] = rng.uniform(-extent, extent, n)
    starts[:, 1] = rng.uniform(-extent, extent, n)
    starts[:, 2] = rng.uniform(-math.pi, math.pi, n)
    d = rng.uniform(dmin, dmax, n)
    a = rng.uniform(-math.pi, math.pi, n)
    goals = np.empty((n, 3))
    goals[:, 0] = starts[:, 0] + d * np.cos(a)
    goals[:, 1] = starts[:, 1] + d * np.sin(a)
    goals[:, 2] = rng.uniform(-math.pi, math.pi, n)
    return starts, goals
