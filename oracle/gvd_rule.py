"""TEST INFRASTRUCTURE -- not part of the product (only tests/ may import this).

numpy restatement of the order-free characterisation of the reference's Voronoi-edge set that the
device kernel k_voronoi_edges (pathplanning_b200/csrc/ppx_gvd.cu) implements.

The reference marks edge cells while the obstacle brushfire runs (GVD::ObstacleDistanceMap::CheckVoro,
state_validator/gvd.cpp:105-131, called from Update :57-69) and unmarks a cell whenever it is popped again
(:53-55).  The claim pinned by tests/test_oracle_cpu.py::test_voronoi_rule_* is that the set left behind is
a function of the FINAL m_distance / m_obstacle grids only; `voronoi_edges` evaluates that function.
Parity pinned: against the compiled reference (oracle/_ref) on C2, C3 and random rotated clutter.
"""
import numpy as np

INT_MAX = 2**31 - 1


def voronoi_edges(occupancy, sqdist, nearest):
    """occupancy (R, C) int32 obstacle ids (-1 free); sqdist (R, C) int32 = m_distance; nearest (R, C, 2) int32 =
    m_obstacle, (-1, -1) where there is none.  Returns the bool edge grid."""
    R, C = occupancy.shape
    edge = np.zeros((R, C), bool)
    orow, ocol = nearest[..., 0].astype(np.int64), nearest[..., 1].astype(np.int64)
    oid = np.where(orow >= 0, occupancy[np.clip(orow, 0, R - 1), np.clip(ocol, 0, C - 1)], -2)
    dist = sqdist.astype(np.int64)
    for dr in (-1, 0, 1):
        for dc in (-1, 0, 1):
            if dr == 0 and dc == 0:
                continue
            rs, cs = slice(max(0, -dr), R - max(0, dr)), slice(max(0, -dc), C - max(0, dc))  # s
            rn, cn = slice(max(0, dr), R - max(0, -dr)), slice(max(0, dc), C - max(0, -dc))  # n = s + (dr, dc)
            sr, sc = np.mgrid[rs, cs]
            nr, nc = sr + dr, sc + dc
            os_r, os_c, on_r, on_c = orow[rs, cs], ocol[rs, cs], orow[rn, cn], ocol[rn, cn]
            ds, dn = dist[rs, cs], dist[rn, cn]
            ok = (os_r >= 0) & (on_r >= 0) & (oid[rs, cs] != oid[rn, cn])  # gvd.cpp:110-111
            ok &= (ds > 1) | (dn > 1)  # :113
            ok &= (np.abs(os_r - on_r) > 1) | (np.abs(os_c - on_c) > 1)  # :114
            s_stab = (sr - on_r) ** 2 + (sc - on_c) ** 2 - ds  # :115-118
            n_stab = (nr - os_r) ** 2 + (nc - os_c) ** 2 - dn
            ok &= (s_stab >= 0) & (n_stab >= 0)  # :119-120
            edge[rs, cs] |= ok & (s_stab <= n_stab)  # :121-124 (the n side is s's rule seen from n)
    return edge


def exact_edt(seed):
    """Exact squared distance to the nearest True cell and that cell ((R, C) int32, (R, C, 2) int32); INT_MAX / (-1, -1)
    when there is no seed.  Ties between equidistant seeds are broken by scipy, not like the device."""
    from scipy import ndimage
    R, C = seed.shape
    if not seed.any():
        return np.full((R, C), INT_MAX, np.int32), np.full((R, C, 2), -1, np.int32)
    idx = ndimage.distance_transform_edt(~seed, return_distances=False, return_indices=True)
    rr, cc = np.indices(seed.shape)
    sq = ((idx[0] - rr).astype(np.int64) ** 2 + (idx[1] - cc).astype(np.int64) ** 2).astype(np.int32)
    return sq, np.stack([idx[0], idx[1]], -1).astype(np.int32)


def path_cost(sqdist, voro_sqdist, resolution, alpha=20.0, d_max=30.0):
    """GVD::PathCostMap::Update, gvd.cpp:266-283, with the float / double mixing of the original expression."""
    res = np.float32(resolution)
    obst = (np.sqrt(sqdist.astype(np.float64)) * np.float64(res)).astype(np.float32)  # gvd.h:38
    voro = (np.sqrt(voro_sqdist.astype(np.float64)) * np.float64(res)).astype(np.float32)  # gvd.h:82
    alpha, d_max = np.float32(alpha), np.float32(d_max)
    with np.errstate(invalid="ignore", divide="ignore"):
        a = (alpha / (alpha + obst)) * (voro / (obst + voro))  # float
        t = (obst - d_max).astype(np.float64)
        q = (t * t) / (np.float64(d_max) * np.float64(d_max))
        out = (a.astype(np.float64) * q).astype(np.float32)
    return np.where((obst >= d_max) | np.isinf(voro), np.float32(0), out)
