"""numpy-level host API over the C ABI (include/ppx.h).

Mirrors, batched, the reference interfaces on the hot path:
  StateValidatorOccupancyMap::IsStateValid / IsPathValid  (state_validator_occupancy_map.cpp:15-71)
  ReedsShepp::Solver::GetOptimalPath / GetShortestPath     (geometry/reeds_shepp.cpp:608-683)
  PathConnectionReedsShepp::Connect + IsPathValid          (paths/path_reeds_shepp.cpp:174-179)
Everything here is argument marshalling; the arithmetic is in the CUDA kernels.
"""
import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from .lib import ppx, MapDescC, ValidatorParamsC, RSCostParamsC, HybridParamsC, HybridLimitsC, RRTParamsC, SmootherParamsC

RS_EDGE = np.dtype([("start", "<f8", (3,)), ("len", "<f8", (5,)), ("code", "<u8")], align=True)
CSTEER_EDGE = np.dtype([("start", "<f8", (3,)), ("length", "<f8"), ("beta", "<f8"), ("cos_beta", "<f8"),
                        ("dtheta_ddist", "<f8"), ("direction", "<i4"), ("_pad", "<i4")], align=True)
LINE_EDGE = np.dtype([("from", "<f8", (3,)), ("to", "<f8", (3,))], align=True)
assert RS_EDGE.itemsize == 72 and CSTEER_EDGE.itemsize == 64 and LINE_EDGE.itemsize == 48
HYBRID_CHILD = np.dtype([("pose", "<f8", (3,)), ("cost", "<f8"), ("heuristic", "<f8"), ("length", "<f8"),
                         ("key", "<i4", (3,)), ("direction", "<i4"), ("kind", "<i4"), ("last", "<f4")], align=True)
HYBRID_RESULT = np.dtype([("status", "<i4"), ("n_expanded", "<i4"), ("n_nodes", "<i4"), ("n_children", "<i4"),
                          ("n_rs_attempts", "<i4"), ("path_len", "<i4"), ("cost", "<f8"), ("expanded_hash", "<u8")], align=True)
HYBRID_PATH_STATE = np.dtype([("pose", "<f8", (3,)), ("length", "<f8"), ("last", "<f4"), ("kind", "i1"), ("direction", "i1"),
                              ("slot", "i1"), ("_pad", "i1")], align=True)
HYBRID_LOG_ENTRY = np.dtype([("key", "<i4", (3,)), ("direction", "<i4"), ("pose", "<f8", (3,))], align=True)
assert HYBRID_CHILD.itemsize == 72 and HYBRID_RESULT.itemsize == 40 and HYBRID_PATH_STATE.itemsize == 40
assert HYBRID_LOG_ENTRY.itemsize == 40
MAX_CHILDREN = 19
RRT_RESULT = np.dtype([("status", "<i4"), ("iterations", "<i4"), ("n_nodes", "<i4"), ("solution", "<i4"), ("path_len", "<i4"),
                       ("n_edges_checked", "<i4"), ("solution_cost", "<f8")], align=True)
assert RRT_RESULT.itemsize == 32


ASTAR_RESULT = np.dtype([("status", np.int32), ("n_explored", np.int32), ("n_nodes", np.int32), ("path_len", np.int32),
                         ("cost", np.float64)])


class PPXError(RuntimeError):
    pass


@dataclass
class MapDesc:
    rows: int
    cols: int
    resolution: float
    world_origin: tuple
    local_origin: tuple = (0.0, 0.0)
    lb: tuple = (-1.0, -1.0, -np.pi)
    ub: tuple = (1.0, 1.0, np.pi)

    def to_c(self):
        d = MapDescC()
        d.rows, d.cols, d.resolution = int(self.rows), int(self.cols), float(self.resolution)
        d.world_origin[:] = self.world_origin
        d.local_origin[:] = self.local_origin
        d.lb[:] = self.lb
        d.ub[:] = self.ub
        return d


@dataclass
class ValidatorParams:
    """StateValidatorOccupancyMap public fields (state_validator_occupancy_map.h:27-28)."""
    min_safe_radius: float = 1.0
    min_path_interpolation_distance: float = 0.1

    def to_c(self):
        return ValidatorParamsC(self.min_safe_radius, self.min_path_interpolation_distance)


@dataclass
class RSCostParams:
    min_turning_radius: float = 1.0
    reverse_cost_multiplier: float = 1.0
    forward_cost_multiplier: float = 1.0
    direction_switching_cost: float = 0.0
    shortest: int = 0

    def to_c(self):
        return RSCostParamsC(self.min_turning_radius, self.reverse_cost_multiplier, self.forward_cost_multiplier,
                             self.direction_switching_cost, self.shortest)


def make_rs_edges(starts, steer, direction, lengths):
    """Pack (n,3) starts, (n,5) steer / direction / length arrays into ppx_rs_edge records."""
    n = len(starts)
    e = np.zeros(n, dtype=RS_EDGE)
    e["start"] = starts
    e["len"] = lengths
    code = np.zeros(n, dtype=np.uint64)
    steer = np.asarray(steer, dtype=np.uint64)
    direction = np.asarray(direction, dtype=np.uint64)
    for i in range(5):
        code |= ((steer[:, i] & 3) | ((direction[:, i] & 3) << 2)) << np.uint64(4 * i)
    e["code"] = code
    return e


def make_csteer_edges(starts, steering, length, direction, wheelbase=2.6, rear_to_center=0.0):
    """Host half of KinematicBicycleModel::ConstantSteer (kinematic_bicycle_model.cpp:14-20):
    the per-steering constants are evaluated here with the host libm, like the reference does."""
    n = len(starts)
    e = np.zeros(n, dtype=CSTEER_EDGE)
    e["start"] = starts
    e["length"] = length
    tan_s = np.tan(np.asarray(steering, dtype=np.float64))
    beta = np.arctan(rear_to_center * tan_s / wheelbase)
    cos_beta = np.cos(beta)
    e["beta"] = beta
    e["cos_beta"] = cos_beta
    e["dtheta_ddist"] = cos_beta * tan_s / wheelbase
    e["direction"] = direction
    return e


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a, shape_last=3):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape_last and (a.ndim != 2 or a.shape[1] != shape_last):
        raise ValueError(f"expected (n,{shape_last}) float64, got {a.shape}")
    return a


class Context:
    """One ppx_ctx (one GPU, one stream)."""

    def __init__(self, device=0):
        h = C.c_void_p()
        rc = ppx.ppx_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise PPXError(f"ppx_ctx_create({device}) failed with {rc}: no usable CUDA device; "
                           "libppx has no CPU fallback")
        self._h = h
        self.device = device
        self.map = None

    def close(self):
        if getattr(self, "_h", None):
            ppx.ppx_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise PPXError(f"ppx error {rc}: {ppx.ppx_last_error(self._h).decode()}")

    @property
    def handle(self):
        return self._h

    def set_stream(self, cuda_stream_ptr):
        self._check(ppx.ppx_ctx_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def synchronize(self):
        self._check(ppx.ppx_ctx_synchronize(self._h))

    def set_option(self, name, value):
        self._check(ppx.ppx_ctx_set_option(self._h, name.encode(), int(value)))

    @property
    def launch_count(self):
        return int(ppx.ppx_ctx_launch_count(self._h))

    # ---- world model ------------------------------------------------------------------------
    def upload_map(self, desc: MapDesc, sqdist=None, occupancy=None, pathcost=None):
        cells = desc.rows * desc.cols
        keep = []
        for name, a, dt in (("sqdist", sqdist, np.int32), ("occupancy", occupancy, np.int32), ("pathcost", pathcost, np.float32)):
            if a is not None:
                a = np.ascontiguousarray(a, dtype=dt).reshape(-1)
                if a.size != cells:
                    raise ValueError(f"{name} has {a.size} cells, expected {cells}")
            keep.append(a)
        d = desc.to_c()
        self._check(ppx.ppx_map_upload(self._h, C.byref(d), _ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2])))
        self.map = desc

    def rasterize(self, desc: "MapDesc", polygons, occupancy=None, first_id=0):
        """Shape::RasterizeLine + ObstacleListOccupancyMap::AddObstacle (obstacle.cpp:7-85,
        obstacle_list_occupancy_map.cpp:29-45) for a list of polygons given by their world-frame vertices ((k, 2)
        arrays, as Obstacle::GetBoundaryWorldPosition returns them) -> occupancy (rows, cols) int32, -1 = free.
        first_id = -1 removes the outlines instead (RemoveObstacle, :47-61)."""
        occ = (np.full((desc.rows, desc.cols), -1, np.int32) if occupancy is None
               else np.ascontiguousarray(occupancy, np.int32).copy())
        polys = [np.ascontiguousarray(p, np.float64).reshape(-1, 2) for p in polygons]
        offsets = np.zeros(len(polys) + 1, np.int32)
        offsets[1:] = np.cumsum([len(p) for p in polys])
        verts = np.concatenate(polys) if polys else np.zeros((0, 2))
        assert first_id == -1 or occ.max(initial=-1) < first_id or not polys, "first_id must exceed the ids already in the grid"
        d = desc.to_c()
        self._check(ppx.ppx_map_rasterize(self.handle, C.byref(d), len(polys), _ptr(offsets), _ptr(verts), first_id, _ptr(occ)))
        return occ

    def astar_n2(self, starts, goals, max_path=0, max_nodes=0):
        """AStarN2 with Euclidean step cost and heuristic (a_star_n2.cpp:12-28, a_star.h:326-434) for (row, col) cell
        pairs on the uploaded occupancy grid -> dict(results (nq,) ASTAR_RESULT, paths (nq, max_path, 2) or None)"""
        s = np.ascontiguousarray(starts, np.int32).reshape(-1, 2)
        g = np.ascontiguousarray(goals, np.int32).reshape(-1, 2)
        nq = len(s)
        res = np.zeros(nq, ASTAR_RESULT)
        paths = np.zeros((nq, max_path, 2), np.int32) if max_path else None
        self._check(ppx.ppx_astar_n2(self.handle, nq, _ptr(s), _ptr(g), max_nodes, max_path, _ptr(res), _ptr(paths)))
        return dict(results=res, paths=paths)

    def astar_n2_ex(self, starts, goals, cost_kind=0, heuristic_kind=0, cell_cost=None, bidirectional=False, max_path=0, max_nodes=0):
        """ppx_astar_n2_ex: AStarN2 / BidirectionalAStarN2 with the device's functor table (see ppx.h)"""
        from .lib import AstarParamsC
        s = np.ascontiguousarray(starts, np.int32).reshape(-1, 2)
        g = np.ascontiguousarray(goals, np.int32).reshape(-1, 2)
        nq = len(s)
        res = np.zeros(nq, ASTAR_RESULT)
        paths = np.zeros((nq, max_path, 2), np.int32) if max_path else None
        cc = None if cell_cost is None else np.ascontiguousarray(cell_cost, np.float32)
        prm = AstarParamsC(cost_kind, heuristic_kind, None if cc is None else cc.ctypes.data, 1 if bidirectional else 0, 0)
        self._check(ppx.ppx_astar_n2_ex(self.handle, nq, _ptr(s), _ptr(g), C.byref(prm), max_nodes, max_path, _ptr(res), _ptr(paths)))
        return dict(results=res, paths=paths)

    def build_map(self, desc: MapDesc, occupancy, pathcost=None, want_sqdist=True):
        """ppx_map_build: occupancy grid in, obstacle squared-distance map computed on the device (N1)."""
        cells = desc.rows * desc.cols
        occ = np.ascontiguousarray(occupancy, dtype=np.int32).reshape(-1)
        if occ.size != cells:
            raise ValueError(f"occupancy has {occ.size} cells, expected {cells}")
        pc = None if pathcost is None else np.ascontiguousarray(pathcost, dtype=np.float32).reshape(-1)
        sq = np.empty(cells, dtype=np.int32) if want_sqdist else None
        d = desc.to_c()
        self._check(ppx.ppx_map_build(self._h, C.byref(d), _ptr(occ), _ptr(pc), _ptr(sq)))
        self.map = desc
        return None if sq is None else sq.reshape(desc.rows, desc.cols)

    def build_gvd(self, desc: MapDesc, occupancy, voronoi=True, want=("sqdist", "voro_sqdist", "pathcost")):
        """ppx_gvd_build: the whole GVD of the occupancy grid on the device (N1) -- obstacle distance map, Voronoi
        edges, Voronoi distance map, path-cost map; installed as the context's map.  Returns the grids named in
        `want` (also "nearest_obstacle", "nearest_edge": rows x cols x 2)."""
        cells = desc.rows * desc.cols
        occ = np.ascontiguousarray(occupancy, dtype=np.int32).reshape(-1)
        if occ.size != cells:
            raise ValueError(f"occupancy has {occ.size} cells, expected {cells}")
        shape = (desc.rows, desc.cols)
        out = {}
        if "sqdist" in want:
            out["sqdist"] = np.empty(shape, np.int32)
        if "voro_sqdist" in want:
            out["voro_sqdist"] = np.empty(shape, np.int32)
        if "pathcost" in want:
            out["pathcost"] = np.empty(shape, np.float32)
        if "nearest_obstacle" in want:
            out["nearest_obstacle"] = np.empty(shape + (2,), np.int32)
        if "nearest_edge" in want:
            out["nearest_edge"] = np.empty(shape + (2,), np.int32)
        d = desc.to_c()
        self._check(ppx.ppx_gvd_build(self._h, C.byref(d), _ptr(occ), 1 if voronoi else 0, _ptr(out.get("sqdist")),
                                      _ptr(out.get("voro_sqdist")), _ptr(out.get("pathcost")), _ptr(out.get("nearest_obstacle")),
                                      _ptr(out.get("nearest_edge"))))
        self.map = desc
        return out

    # ---- A7 -----------------------------------------------------------------------------------
    def is_state_valid(self, poses, vp: ValidatorParams):
        poses = _f64(poses)
        out = np.empty(len(poses), dtype=np.uint8)
        v = vp.to_c()
        self._check(ppx.ppx_is_state_valid(self._h, len(poses), _ptr(poses), C.byref(v), _ptr(out)))
        return out.astype(bool)

    def is_state_valid_rect(self, poses, footprint):
        """rectangular-footprint variant (ppx_footprint_params: length, width, rear_to_center, n_discs, min_step)"""
        from .lib import FootprintParamsC
        poses = _f64(poses)
        out = np.empty(len(poses), dtype=np.uint8)
        fp = FootprintParamsC(*footprint)
        self._check(ppx.ppx_is_state_valid_rect(self._h, len(poses), _ptr(poses), C.byref(fp), _ptr(out)))
        return out.astype(bool)

    def check_edges_rs_rect(self, edges, min_turning_radius, footprint):
        """-> (valid, last, samples) for the rectangular-footprint variant; not a reference behaviour (see ppx.h)"""
        from .lib import FootprintParamsC
        edges = np.ascontiguousarray(edges, dtype=RS_EDGE)
        n = len(edges)
        valid = np.empty(n, dtype=np.uint8)
        last = np.empty(n, dtype=np.float32)
        samples = np.empty(n, dtype=np.uint32)
        fp = FootprintParamsC(*footprint)
        self._check(ppx.ppx_check_edges_rs_rect(self._h, n, _ptr(edges), float(min_turning_radius), C.byref(fp), _ptr(valid), _ptr(last),
                                                _ptr(samples)))
        return valid.astype(bool), last, samples

    def check_edges_rs(self, edges, min_turning_radius, vp: ValidatorParams, want_samples=False):
        edges = np.ascontiguousarray(edges, dtype=RS_EDGE)
        n = len(edges)
        valid = np.empty(n, dtype=np.uint8)
        last = np.empty(n, dtype=np.float32)
        samples = np.empty(n, dtype=np.uint32) if want_samples else None
        v = vp.to_c()
        self._check(ppx.ppx_check_edges_rs(self._h, n, _ptr(edges), float(min_turning_radius), C.byref(v),
                                           _ptr(valid), _ptr(last), _ptr(samples)))
        return (valid.astype(bool), last, samples) if want_samples else (valid.astype(bool), last)

    def check_edges_rs_audit(self, edges, min_turning_radius, vp: ValidatorParams):
        """check_edges_rs plus the audit of the FP32 cell filter (ppx_check_edges_rs_audit)."""
        edges = np.ascontiguousarray(edges, dtype=RS_EDGE)
        n = len(edges)
        valid = np.empty(n, dtype=np.uint8)
        last = np.empty(n, dtype=np.float32)
        samples = np.empty(n, dtype=np.uint32)
        counters = np.zeros(3, dtype=np.uint64)
        dev = np.zeros(1, dtype=np.float32)
        v = vp.to_c()
        self._check(ppx.ppx_check_edges_rs_audit(self._h, n, _ptr(edges), float(min_turning_radius), C.byref(v),
                                                 _ptr(valid), _ptr(last), _ptr(samples), _ptr(counters), _ptr(dev)))
        return dict(valid=valid.astype(bool), last=last, samples=samples, certified=int(counters[0]),
                    parked=int(counters[1]), cell_mismatch=int(counters[2]), max_dev_cells=float(dev[0]))

    def check_edges_csteer(self, edges, vp: ValidatorParams, want_final=False):
        edges = np.ascontiguousarray(edges, dtype=CSTEER_EDGE)
        n = len(edges)
        valid = np.empty(n, dtype=np.uint8)
        last = np.empty(n, dtype=np.float32)
        final = np.empty((n, 3), dtype=np.float64) if want_final else None
        v = vp.to_c()
        self._check(ppx.ppx_check_edges_csteer(self._h, n, _ptr(edges), C.byref(v), _ptr(valid), _ptr(last), _ptr(final)))
        return (valid.astype(bool), last, final) if want_final else (valid.astype(bool), last)

    def check_edges_line(self, edges, vp: ValidatorParams):
        edges = np.ascontiguousarray(edges, dtype=LINE_EDGE)
        n = len(edges)
        valid = np.empty(n, dtype=np.uint8)
        last = np.empty(n, dtype=np.float32)
        v = vp.to_c()
        self._check(ppx.ppx_check_edges_line(self._h, n, _ptr(edges), C.byref(v), _ptr(valid), _ptr(last)))
        return valid.astype(bool), last

    # ---- A3 / A4 ------------------------------------------------------------------------------
    def rs_solve(self, starts, goals, cp: RSCostParams):
        """-> dict(edges, word, length, cost)"""
        starts, goals = _f64(starts), _f64(goals)
        n = len(starts)
        edges = np.zeros(n, dtype=RS_EDGE)
        word = np.empty(n, dtype=np.int32)
        length = np.empty(n, dtype=np.float64)
        cost = np.empty(n, dtype=np.float32)
        c = cp.to_c()
        self._check(ppx.ppx_rs_solve(self._h, n, _ptr(starts), _ptr(goals), C.byref(c), _ptr(edges), _ptr(word),
                                     _ptr(length), _ptr(cost)))
        return dict(edges=edges, word=word, length=length, cost=cost)

    def connect_check_rs(self, starts, goals, cp: RSCostParams, vp: ValidatorParams):
        """-> dict(valid, last, word, length, cost)"""
        starts, goals = _f64(starts), _f64(goals)
        n = len(starts)
        valid = np.empty(n, dtype=np.uint8)
        last = np.empty(n, dtype=np.float32)
        word = np.empty(n, dtype=np.int32)
        length = np.empty(n, dtype=np.float64)
        cost = np.empty(n, dtype=np.float32)
        c, v = cp.to_c(), vp.to_c()
        self._check(ppx.ppx_connect_check_rs(self._h, n, _ptr(starts), _ptr(goals), C.byref(c), C.byref(v), _ptr(valid),
                                             _ptr(last), _ptr(word), _ptr(length), _ptr(cost)))
        return dict(valid=valid.astype(bool), last=last, word=word, length=length, cost=cost)

    def rs_interpolate(self, edges, min_turning_radius, ratios):
        edges = np.ascontiguousarray(edges, dtype=RS_EDGE)
        ratios = np.ascontiguousarray(ratios, dtype=np.float64)
        n = len(edges)
        if ratios.ndim == 1:
            ratios = np.tile(ratios, (n, 1))
        nr = ratios.shape[1]
        poses = np.empty((n, nr, 3), dtype=np.float64)
        self._check(ppx.ppx_rs_interpolate(self._h, n, _ptr(edges), float(min_turning_radius), nr, _ptr(ratios), _ptr(poses)))
        return poses

    # ---- A8 -----------------------------------------------------------------------------------
    def wavefront(self, goals, want_explored=True):
        """ObstaclesHeuristic::Update for each goal -> (cost (n,rows,cols) f32, explored (n,rows,cols) bool)."""
        goals = _f64(goals)
        n = len(goals)
        rows, cols = self.map.rows, self.map.cols
        cost = np.empty((n, rows, cols), dtype=np.float32)
        explored = np.empty((n, rows, cols), dtype=np.uint8) if want_explored else None
        self._check(ppx.ppx_wavefront(self._h, n, _ptr(goals), _ptr(cost), _ptr(explored)))
        return (cost, explored.astype(bool)) if want_explored else cost

    # ---- A10 ----------------------------------------------------------------------------------
    def knn(self, points, queries, k):
        """Exact k nearest neighbours (Tree::GetNearestNodes) -> (indices (nq,k) int32, dist2 (nq,k))."""
        points = np.ascontiguousarray(points, np.float64)
        queries = np.ascontiguousarray(queries, np.float64)
        dim = queries.shape[1]
        nq = len(queries)
        idx = np.empty((nq, k), np.int32)
        d2 = np.empty((nq, k), np.float64)
        self._check(ppx.ppx_knn(self._h, dim, len(points), _ptr(points), nq, _ptr(queries), k, _ptr(idx), _ptr(d2)))
        return idx, d2

    # ---- A11 ----------------------------------------------------------------------------------
    def rrt_run(self, starts, goals, seeds, lb, ub, dim=2, connection=0, star=1, validator=0, max_iteration=100,
                max_nodes=10000, max_connection_distance=0.1, goal_bias=0.05, min_turning_radius=1.0, vp=None,
                want_trees=True):
        """RRT / RRT* (algo/rrt.h, algo/rrt_star.h), one tree per (start, goal, seed)."""
        starts, goals = _f64(starts), _f64(goals)
        seeds = np.ascontiguousarray(seeds, np.uint64)
        nt = len(starts)
        p = RRTParamsC(dim, connection, star, validator, max_iteration, max_nodes, max_connection_distance, goal_bias,
                       min_turning_radius)
        lb3 = list(lb) + [0.0] * (3 - len(lb))
        ub3 = list(ub) + [0.0] * (3 - len(ub))
        p.lb[:] = lb3
        p.ub[:] = ub3
        v = (vp or ValidatorParams()).to_c()
        res = np.zeros(nt, RRT_RESULT)
        cap = max_nodes + 2 if want_trees else 0
        nodes = np.zeros((nt, cap, 3)) if want_trees else None
        parents = np.zeros((nt, cap), np.int32) if want_trees else None
        costs = np.zeros((nt, cap)) if want_trees else None
        self._check(ppx.ppx_rrt_run(self._h, C.byref(p), C.byref(v), nt, _ptr(starts), _ptr(goals), _ptr(seeds), _ptr(res), cap,
                                    _ptr(nodes), _ptr(parents), _ptr(costs)))
        return dict(results=res, nodes=nodes, parents=parents, costs=costs)


@dataclass
class HybridParams:
    """HybridAStar::SearchParameters (algo/hybrid_a_star.h:29-50)."""
    wheelbase: float = 2.6
    min_turning_radius: float = 2.0
    direction_switching_cost: float = 0.0
    reverse_cost_multiplier: float = 1.0
    forward_cost_multiplier: float = 1.0
    voronoi_cost_multiplier: float = 1.0
    num_generated_motion: int = 5
    spatial_resolution: float = 1.0
    angular_resolution: float = 0.0872

    def to_c(self):
        return HybridParamsC(self.wheelbase, self.min_turning_radius, self.direction_switching_cost,
                             self.reverse_cost_multiplier, self.forward_cost_multiplier, self.voronoi_cost_multiplier,
                             self.num_generated_motion, 0, self.spatial_resolution, self.angular_resolution)


class HybridAStar:
    """ppx_hybrid: HybridAStar::Initialize + the batched graph search (algo/hybrid_a_star.cpp:206-257)."""

    def __init__(self, ctx: Context, params: HybridParams = None, vp: ValidatorParams = None):
        self.ctx = ctx
        self.params = params or HybridParams()
        self.vp = vp or ValidatorParams()
        h = C.c_void_p()
        p, v = self.params.to_c(), self.vp.to_c()
        ctx._check(ppx.ppx_hybrid_create(ctx.handle, C.byref(p), C.byref(v), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            ppx.ppx_hybrid_destroy(self._h)
            self._h = None

    @property
    def handle(self):
        return self._h

    def table(self):
        dims = np.zeros(3, np.int32)
        off = np.zeros(2, np.float64)
        self.ctx._check(ppx.ppx_hybrid_table(self._h, _ptr(dims), _ptr(off), None))
        vals = np.empty(tuple(dims.tolist()), np.float64)
        self.ctx._check(ppx.ppx_hybrid_table(self._h, _ptr(dims), _ptr(off), _ptr(vals)))
        return vals, off

    def heuristic(self, goal, poses):
        poses = _f64(poses)
        goal = np.ascontiguousarray(goal, np.float64)
        n = len(poses)
        v, nh, ob = np.empty(n), np.empty(n), np.empty(n)
        self.ctx._check(ppx.ppx_hybrid_heuristic(self._h, _ptr(goal), n, _ptr(poses), _ptr(v), _ptr(nh), _ptr(ob)))
        return v, nh, ob

    def expand(self, goal, parents, seeds):
        parents = _f64(parents)
        seeds = np.ascontiguousarray(seeds, np.uint64)
        goal = np.ascontiguousarray(goal, np.float64)
        n = len(parents)
        count = np.zeros(n, np.int32)
        children = np.zeros((n, MAX_CHILDREN), HYBRID_CHILD)
        ph = np.empty(n)
        self.ctx._check(ppx.ppx_hybrid_expand(self._h, _ptr(goal), n, _ptr(parents), _ptr(seeds), _ptr(count), _ptr(children), _ptr(ph)))
        return count, children, ph

    def timing(self):
        """(wavefront_ms, search_ms) of the last search call on this planner, device time"""
        a, b = C.c_double(), C.c_double()
        self.ctx._check(ppx.ppx_hybrid_timing(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def search(self, starts, goals, seeds, max_path=0, max_log=0, max_expansions=0, max_nodes=0):
        """-> dict(results (nq,) HYBRID_RESULT, paths (nq,max_path) or None, log (nq,max_log) or None)"""
        starts, goals = _f64(starts), _f64(goals)
        seeds = np.ascontiguousarray(seeds, np.uint64)
        nq = len(starts)
        lim = HybridLimitsC(max_nodes, max_expansions, max_path, max_log)
        res = np.zeros(nq, HYBRID_RESULT)
        paths = np.zeros((nq, max_path), HYBRID_PATH_STATE) if max_path else None
        log = np.zeros((nq, max_log), HYBRID_LOG_ENTRY) if max_log else None
        self.ctx._check(ppx.ppx_hybrid_search(self._h, nq, _ptr(starts), _ptr(goals), _ptr(seeds), C.byref(lim), _ptr(res),
                                              _ptr(paths), _ptr(log)))
        return dict(results=res, paths=paths, log=log)

    def resample(self, goals, results, paths, path_interpolation=0.1, max_points=4096):
        """The pose array HybridAStar::SearchPath hands to the smoother (hybrid_a_star.cpp:260-291) for the chains of
        search(): -> dict(poses (nq,max_points,3), cusp (nq,max_points) bool, count (nq,), length (nq,))"""
        goals = _f64(goals)
        results = np.ascontiguousarray(results, HYBRID_RESULT)
        paths = np.ascontiguousarray(paths, HYBRID_PATH_STATE)
        nq = len(goals)
        assert paths.ndim == 2 and len(paths) == nq and len(results) == nq
        poses = np.zeros((nq, max_points, 3), np.float64)
        cusp = np.zeros((nq, max_points), np.uint8)
        count = np.zeros(nq, np.int32)
        length = np.zeros(nq, np.float64)
        self.ctx._check(ppx.ppx_hybrid_resample(self._h, nq, _ptr(goals), _ptr(results), _ptr(paths), paths.shape[1],
                                                C.c_float(path_interpolation), max_points, _ptr(poses), _ptr(cusp), _ptr(count),
                                                _ptr(length)))
        return dict(poses=poses, cusp=cusp.astype(bool), count=count, length=length)


@dataclass
class SmootherParams:
    """Smoother::Parameters (algo/smoother.h:29-62) + the GVD constants and validator radius Smooth reads."""
    step_tolerance: float = 1e-3
    max_iterations: int = 2000
    learning_rate: float = 0.01
    path_weight: float = 0.0
    smooth_weight: float = 0.4
    voronoi_weight: float = 0.02
    collision_weight: float = 0.2
    curvature_weight: float = 0.4
    collision_ratio: float = 0.2
    max_curvature: float = 0.5  # 1 / minTurningRadius (hybrid_a_star.cpp:218)
    alpha: float = 20.0
    d_max: float = 30.0
    min_safe_radius: float = 1.0

    def to_c(self):
        return SmootherParamsC(self.step_tolerance, self.max_iterations, self.learning_rate, self.path_weight,
                               self.smooth_weight, self.voronoi_weight, self.collision_weight, self.curvature_weight,
                               self.collision_ratio, self.max_curvature, self.alpha, self.d_max, self.min_safe_radius)


class Smoother:
    """ppx_smoother: Smoother::Initialize / Smooth / GetPath (algo/smoother.cpp:18-158) for batches of pose arrays."""

    def __init__(self, ctx: Context, params: SmootherParams = None):
        self.ctx = ctx
        self.params = params or SmootherParams()
        h = C.c_void_p()
        p = self.params.to_c()
        ctx._check(ppx.ppx_smoother_create(ctx.handle, C.byref(p), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            ppx.ppx_smoother_destroy(self._h)
            self._h = None

    def set_gvd(self, nearest_obstacle, nearest_edge):
        """(rows, cols, 2) int32 grids of GVD::GetNearestObstacleCell / GetNearestVoronoiEdgeCell"""
        a = np.ascontiguousarray(nearest_obstacle, np.int32)
        b = np.ascontiguousarray(nearest_edge, np.int32)
        self.ctx._check(ppx.ppx_smoother_set_gvd(self._h, _ptr(a), _ptr(b)))

    def use_gvd(self):
        """take the two cell grids from the GVD the context built on the device (Context.build_gvd)"""
        self.ctx._check(ppx.ppx_smoother_use_gvd(self._h))

    def smooth(self, poses, cusp, count):
        """poses (nq, max_points, 3), cusp (nq, max_points), count (nq,) as returned by HybridAStar.resample
        -> dict(poses, status (nq,), iterations (nq,))"""
        poses = np.ascontiguousarray(poses, np.float64)
        cusp = np.ascontiguousarray(cusp, np.uint8)
        count = np.ascontiguousarray(count, np.int32)
        nq, max_points = poses.shape[0], poses.shape[1]
        assert poses.shape == (nq, max_points, 3) and cusp.shape == (nq, max_points) and count.shape == (nq,)
        out = np.zeros_like(poses)
        status = np.zeros(nq, np.int32)
        iters = np.zeros(nq, np.int32)
        self.ctx._check(ppx.ppx_smoother_smooth(self._h, nq, _ptr(poses), _ptr(cusp), _ptr(count), max_points, _ptr(out),
                                                _ptr(status), _ptr(iters)))
        return dict(poses=out, status=status, iterations=iters)
