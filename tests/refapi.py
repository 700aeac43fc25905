"""ctypes wrapper of the compiled reference harness (oracle/ref_harness.cpp ->
oracle/_ref/libpp_ref.so).  TEST INFRASTRUCTURE: imported only by tests/, smoke() and bench.py's
cpu_baseline / --impl reference legs."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libpp_ref.so")
_lib = None


def available():
    return os.path.exists(REF_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(REF_PATH)
        _lib.ref_world_create.restype = C.c_void_p
        _lib.ref_hybrid_create.restype = C.c_void_p
        for name in ("ref_check_rs_edges", "ref_check_rs_segments", "ref_wavefront", "ref_nonholo_build"):
            getattr(_lib, name).restype = C.c_double
        _lib.ref_hybrid_expansions.restype = C.c_long
        _lib.ref_hybrid_solution.restype = C.c_long
        _lib.ref_hybrid_resampled.restype = C.c_long
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _d3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


def selfcheck():
    w = np.zeros(75, dtype=np.int32)
    bad = lib().ref_selfcheck(_p(w))
    return bad, dict(zip(range(-37, 38), w.tolist()))


# ---- Reeds-Shepp ------------------------------------------------------------------------------
def rs_optimal_batch(starts, goals, rmin=1.0, rev=1.0, fwd=1.0, sw=0.0):
    starts = np.ascontiguousarray(starts, dtype=np.float64)
    goals = np.ascontiguousarray(goals, dtype=np.float64)
    n = len(starts)
    words = np.empty(n, np.int32)
    steer = np.empty((n, 5), np.int32)
    direction = np.empty((n, 5), np.int32)
    length = np.empty((n, 5), np.float64)
    total = np.empty(n, np.float64)
    cost = np.empty(n, np.float32)
    lib().ref_rs_optimal_batch(C.c_long(n), _p(starts), _p(goals), C.c_double(rmin), C.c_float(rev), C.c_float(fwd),
                               C.c_float(sw), _p(words), _p(steer), _p(direction), _p(length), _p(total), _p(cost))
    return dict(word=words, steer=steer, dir=direction, len=length, length=total, cost=cost)


def rs_shortest(start, goal, rmin=1.0):
    steer = np.empty(5, np.int32)
    direction = np.empty(5, np.int32)
    length = np.empty(5, np.float64)
    total = C.c_double()
    final = np.empty(3, np.float64)
    w = lib().ref_rs_shortest(_d3(start), _d3(goal), C.c_double(rmin), _p(steer), _p(direction), _p(length),
                              C.byref(total), _p(final))
    return dict(word=w, steer=steer, dir=direction, len=length, length=total.value, final=final)


def rs_interpolate(start, steer, direction, length, rmin, ratios, truncate=-1.0):
    ratios = np.ascontiguousarray(ratios, dtype=np.float64)
    steer = np.ascontiguousarray(steer, dtype=np.int32)
    direction = np.ascontiguousarray(direction, dtype=np.int32)
    length = np.ascontiguousarray(length, dtype=np.float64)
    poses = np.empty((len(ratios), 3), np.float64)
    out_len = C.c_double()
    final = np.empty(3, np.float64)
    lib().ref_rs_interpolate(_d3(start), _p(steer), _p(direction), _p(length), C.c_double(rmin), C.c_double(truncate),
                             C.c_int(len(ratios)), _p(ratios), _p(poses), C.byref(out_len), _p(final))
    return poses, out_len.value, final


def csteer_interpolate(start, steering, length, direction, wheelbase, ratios):
    ratios = np.ascontiguousarray(ratios, dtype=np.float64)
    poses = np.empty((len(ratios), 3), np.float64)
    lib().ref_csteer_interpolate(_d3(start), C.c_double(steering), C.c_double(length), C.c_int(direction),
                                 C.c_double(wheelbase), C.c_int(len(ratios)), _p(ratios), _p(poses))
    return poses


def random_uniform(seed, n, lb=0.0, ub=1.0):
    out = np.empty(n, np.float64)
    lib().ref_random_uniform(C.c_uint64(seed), C.c_long(n), C.c_double(lb), C.c_double(ub), _p(out))
    return out


# ---- World ------------------------------------------------------------------------------------
class World:
    """StateSpaceSE2 + ObstacleListOccupancyMap + StateValidatorOccupancyMap of the reference."""

    def __init__(self, lb, ub, resolution):
        self.lb, self.ub, self.resolution = tuple(lb), tuple(ub), float(resolution)
        self.h = C.c_void_p(lib().ref_world_create(_d3(lb), _d3(ub), C.c_float(resolution)))
        self.min_safe_radius = 1.0
        self.min_step = 0.1

    def close(self):
        if self.h:
            lib().ref_world_destroy(self.h)
            self.h = None

    def add_rectangle(self, dx, dy, x, y, theta):
        return lib().ref_world_add_rectangle(self.h, C.c_double(dx), C.c_double(dy), C.c_double(x), C.c_double(y),
                                             C.c_double(theta))

    def add_shape(self, kind, a, b, x, y, theta):
        """kind 0 RectangleShape(a, b), 1 RegularPolygonShape(a, int(b)), 2 CircleShape(a, int(b)) at Pose2d(x, y, theta)"""
        return lib().ref_world_add_shape(self.h, C.c_int(kind), C.c_double(a), C.c_double(b), C.c_double(x), C.c_double(y),
                                         C.c_double(theta))

    def astar_n2(self, start, goal, cap=4096):
        """AStarN2 with Euclidean cost / heuristic functors -> dict(status, cost, explored, seconds, path (k,2))"""
        cost, sec = C.c_double(), C.c_double()
        explored, plen = C.c_long(), C.c_long()
        path = np.zeros((cap, 2), np.int32)
        st = lib().ref_astar_n2(self.h, C.c_int(int(start[0])), C.c_int(int(start[1])), C.c_int(int(goal[0])), C.c_int(int(goal[1])),
                                C.byref(cost), C.byref(explored), C.byref(sec), C.c_long(cap), _p(path), C.byref(plen))
        return dict(status=int(st), cost=cost.value, explored=explored.value, seconds=sec.value, path=path[:min(plen.value, cap)].copy(),
                    path_len=plen.value)

    def astar_n2_ex(self, start, goal, cost_kind=0, heuristic_kind=0, cell_cost=None, bidirectional=False, cap=8192):
        """AStarN2 / BidirectionalAStarN2 with the functor table of ppx_astar_params"""
        cost = C.c_double()
        explored, plen = C.c_long(), C.c_long()
        path = np.zeros((cap, 2), np.int32)
        cc = None if cell_cost is None else np.ascontiguousarray(cell_cost, np.float32)
        st = lib().ref_astar_n2_ex(self.h, C.c_int(int(start[0])), C.c_int(int(start[1])), C.c_int(int(goal[0])), C.c_int(int(goal[1])),
                                   C.c_int(cost_kind), C.c_int(heuristic_kind), None if cc is None else _p(cc), C.c_int(int(bidirectional)),
                                   C.byref(cost), C.byref(explored), C.c_long(cap), _p(path), C.byref(plen))
        return dict(status=int(st), cost=cost.value, explored=explored.value, path=path[:min(plen.value, cap)].copy(), path_len=plen.value)

    def remove_obstacle(self, index):
        """RemoveObstacle of the index-th obstacle added to this world"""
        return lib().ref_world_remove_obstacle(self.h, C.c_int(index))

    def set_cells(self, rows, cols, obstacle_id=0):
        rows = np.ascontiguousarray(rows, np.int32)
        cols = np.ascontiguousarray(cols, np.int32)
        lib().ref_world_set_cells(self.h, C.c_long(len(rows)), _p(rows), _p(cols), C.c_int(obstacle_id))

    def set_position(self, x, y):
        """OccupancyMap::SetPosition: the local frame's origin in world coordinates (occupancy_map.h:60-61)"""
        lib().ref_world_set_position(self.h, C.c_double(x), C.c_double(y))

    def update(self):
        lib().ref_world_update(self.h)

    def info(self):
        rows, cols, res = C.c_int(), C.c_int(), C.c_float()
        org = (C.c_double * 4)()
        lib().ref_world_info(self.h, C.byref(rows), C.byref(cols), org, C.byref(res))
        return dict(rows=rows.value, cols=cols.value, world_origin=(org[0], org[1]), local_origin=(org[2], org[3]),
                    resolution=res.value)

    def grids(self):
        i = self.info()
        occ = np.empty((i["rows"], i["cols"]), np.int32)
        sq = np.empty((i["rows"], i["cols"]), np.int32)
        lib().ref_world_copy_grids(self.h, _p(occ), _p(sq))
        return occ, sq

    def gvd_update(self):
        i = self.info()
        pc = np.empty((i["rows"], i["cols"]), np.float32)
        vd = np.empty((i["rows"], i["cols"]), np.int32)
        lib().ref_world_gvd_update(self.h, _p(pc), _p(vd))
        return pc, vd

    def set_gvd_distances(self, sqdist, voro_sqdist):
        """overwrite the GVD's two distance grids (the potential is recomputed from them at every GVD::Update);
        returns how many (obstacle, Voronoi) cells differed from the reference's own"""
        a = np.ascontiguousarray(sqdist, np.int32)
        b = np.ascontiguousarray(voro_sqdist, np.int32)
        diff = (C.c_long * 2)()
        lib().ref_world_set_gvd_distances(self.h, _p(a), _p(b), diff)
        return int(diff[0]), int(diff[1])

    def gvd_cells(self):
        """(rows, cols, 2) nearest-obstacle and nearest-Voronoi-edge cells of the world's GVD (after gvd_update)"""
        i = self.info()
        a = np.zeros((i["rows"], i["cols"], 2), np.int32)
        b = np.zeros((i["rows"], i["cols"], 2), np.int32)
        lib().ref_world_gvd_cells(self.h, _p(a), _p(b))
        return a, b

    def set_validator_params(self, min_safe_radius, min_step):
        self.min_safe_radius, self.min_step = float(min_safe_radius), float(min_step)
        lib().ref_world_set_validator_params(self.h, C.c_float(min_safe_radius), C.c_float(min_step))

    def map_desc(self):
        """ppx MapDesc equivalent of this world (what the C++ host mirror passes to ppx_map_upload)."""
        from pathplanning_b200 import MapDesc
        i = self.info()
        return MapDesc(rows=i["rows"], cols=i["cols"], resolution=i["resolution"], world_origin=i["world_origin"],
                       local_origin=i["local_origin"], lb=self.lb, ub=self.ub)

    def is_state_valid(self, poses):
        poses = np.ascontiguousarray(poses, np.float64)
        out = np.empty(len(poses), np.uint8)
        lib().ref_is_state_valid_batch(self.h, C.c_long(len(poses)), _p(poses), _p(out))
        return out.astype(bool)

    def check_rs_edges(self, starts, goals, rmin=1.0, rev=1.0, fwd=1.0, sw=0.0, full=True):
        """Solve + build + IsPathValid.  Returns dict incl. wall seconds."""
        starts = np.ascontiguousarray(starts, np.float64)
        goals = np.ascontiguousarray(goals, np.float64)
        n = len(starts)
        valid = np.empty(n, np.uint8)
        last = np.empty(n, np.float32)
        if full:
            samples = np.empty(n, np.int32)
            words = np.empty(n, np.int32)
            steer = np.empty((n, 5), np.int32)
            direction = np.empty((n, 5), np.int32)
            length = np.empty((n, 5), np.float64)
            total = np.empty(n, np.float64)
            cost = np.empty(n, np.float32)
        else:
            samples = words = steer = direction = length = total = cost = None
        sec = lib().ref_check_rs_edges(self.h, C.c_long(n), _p(starts), _p(goals), C.c_double(rmin), C.c_float(rev),
                                       C.c_float(fwd), C.c_float(sw), _p(valid), _p(last), _p(samples), _p(words),
                                       _p(steer), _p(direction), _p(length), _p(total), _p(cost))
        return dict(valid=valid.astype(bool), last=last, samples=samples, word=words, steer=steer, dir=direction,
                    len=length, length=total, cost=cost, seconds=sec)

    def check_rs_segments(self, starts, steer, direction, length, rmin=1.0):
        starts = np.ascontiguousarray(starts, np.float64)
        steer = np.ascontiguousarray(steer, np.int32)
        direction = np.ascontiguousarray(direction, np.int32)
        length = np.ascontiguousarray(length, np.float64)
        n = len(starts)
        valid = np.empty(n, np.uint8)
        last = np.empty(n, np.float32)
        samples = np.empty(n, np.int32)
        sec = lib().ref_check_rs_segments(self.h, C.c_long(n), _p(starts), _p(steer), _p(direction), _p(length),
                                          C.c_double(rmin), _p(valid), _p(last), _p(samples))
        return dict(valid=valid.astype(bool), last=last, samples=samples, seconds=sec)

    def check_rs_segments_rect(self, starts, steer, direction, length, rmin, footprint):
        """rectangular-footprint variant restated over the reference's Interpolate / IsStateValid / distance map ("port");
        footprint = (length, width, rear_to_center, n_discs, min_step)"""
        starts = np.ascontiguousarray(starts, np.float64)
        steer = np.ascontiguousarray(steer, np.int32)
        direction = np.ascontiguousarray(direction, np.int32)
        length = np.ascontiguousarray(length, np.float64)
        n = len(starts)
        valid = np.empty(n, np.uint8)
        last = np.empty(n, np.float32)
        samples = np.empty(n, np.int32)
        fp = np.array(footprint, np.float64)
        f = lib().ref_check_rs_segments_rect
        f.restype = C.c_double
        sec = f(self.h, C.c_long(n), _p(starts), _p(steer), _p(direction), _p(length), C.c_double(rmin), _p(fp), _p(valid), _p(last),
                _p(samples))
        return dict(valid=valid.astype(bool), last=last, samples=samples, seconds=sec)

    def check_csteer(self, starts, steering, length, direction, wheelbase=2.6):
        starts = np.ascontiguousarray(starts, np.float64)
        steering = np.ascontiguousarray(steering, np.float64)
        length = np.ascontiguousarray(length, np.float64)
        direction = np.ascontiguousarray(direction, np.int32)
        n = len(starts)
        valid = np.empty(n, np.uint8)
        last = np.empty(n, np.float32)
        final = np.empty((n, 3), np.float64)
        flen = np.empty(n, np.float64)
        lib().ref_check_csteer(self.h, C.c_long(n), _p(starts), _p(steering), _p(length), _p(direction),
                               C.c_double(wheelbase), _p(valid), _p(last), _p(final), _p(flen))
        return dict(valid=valid.astype(bool), last=last, final=final, final_len=flen)

    def check_line_edges(self, a, b):
        a = np.ascontiguousarray(a, np.float64)
        b = np.ascontiguousarray(b, np.float64)
        n = len(a)
        valid = np.empty(n, np.uint8)
        last = np.empty(n, np.float32)
        lib().ref_check_line_edges(self.h, C.c_long(n), _p(a), _p(b), _p(valid), _p(last))
        return dict(valid=valid.astype(bool), last=last)

    def wavefront(self, goal, rev=1.0, fwd=1.0):
        i = self.info()
        cost = np.empty((i["rows"], i["cols"]), np.float32)
        explored = np.empty((i["rows"], i["cols"]), np.uint8)
        sec = lib().ref_wavefront(self.h, _d3(goal), C.c_double(rev), C.c_double(fwd), _p(cost), _p(explored))
        return cost, explored.astype(bool), sec


def nonholo_build(lb, ub, spatial_res=1.0, ang_res=0.0872, rmin=2.0, rev=1.0, fwd=1.0, sw=0.0):
    dims = np.zeros(3, np.int32)
    off = np.zeros(2, np.float64)
    lib().ref_nonholo_build(_d3(lb), _d3(ub), C.c_double(spatial_res), C.c_double(ang_res), C.c_double(rmin),
                            C.c_double(rev), C.c_double(fwd), C.c_double(sw), _p(dims), _p(off), None)
    vals = np.empty(tuple(dims.tolist()), np.float64)
    sec = lib().ref_nonholo_build(_d3(lb), _d3(ub), C.c_double(spatial_res), C.c_double(ang_res), C.c_double(rmin),
                                  C.c_double(rev), C.c_double(fwd), C.c_double(sw), _p(dims), _p(off), _p(vals))
    return vals, off, sec


# ---- Hybrid A* --------------------------------------------------------------------------------
DEFAULT_HYBRID = dict(min_turning_radius=2.0, direction_switching_cost=0.0, reverse_cost_multiplier=1.0,
                      forward_cost_multiplier=1.0, voronoi_cost_multiplier=1.0, num_generated_motion=5,
                      spatial_resolution=1.0, angular_resolution=0.0872)


def shape_vertices(kind, a, b, x, y, theta, cap=256):
    """Obstacle::GetBoundaryWorldPosition of the shape at the pose -> (k, 2) world coordinates"""
    out = np.zeros((cap, 2), np.float64)
    n = lib().ref_shape_vertices(C.c_int(kind), C.c_double(a), C.c_double(b), C.c_double(x), C.c_double(y), C.c_double(theta),
                                 _p(out), C.c_int(cap))
    return out[:n].copy()


class Hybrid:
    """HybridAStar of the reference over a World (Initialize runs in the constructor)."""

    def __init__(self, world, **kw):
        p = dict(DEFAULT_HYBRID)
        p.update(kw)
        self.params = p
        self.world = world
        L = lib()
        L.ref_hybrid_create.restype = C.c_void_p
        self.h = C.c_void_p(L.ref_hybrid_create(
            world.h, C.c_double(p["min_turning_radius"]), C.c_double(p["direction_switching_cost"]),
            C.c_double(p["reverse_cost_multiplier"]), C.c_double(p["forward_cost_multiplier"]),
            C.c_double(p["voronoi_cost_multiplier"]), C.c_uint(p["num_generated_motion"]),
            C.c_double(p["spatial_resolution"]), C.c_double(p["angular_resolution"])))
        if not self.h:
            raise RuntimeError("ref_hybrid_create failed")

    def close(self):
        if self.h:
            lib().ref_hybrid_destroy(self.h)
            self.h = None

    def table(self):
        dims = np.zeros(3, np.int32)
        off = np.zeros(2, np.float64)
        lib().ref_hybrid_table(self.h, _p(dims), _p(off), None)
        vals = np.empty(tuple(dims.tolist()), np.float64)
        lib().ref_hybrid_table(self.h, _p(dims), _p(off), _p(vals))
        return vals, off

    def search(self, start, goal, seed, cap=200000):
        """Wavefront + GVD update + graph search (hybrid_a_star.cpp:248-257)."""
        cost = C.c_double()
        nexp, nchild, nsamp = C.c_long(), C.c_long(), C.c_long()
        times = (C.c_double * 3)()
        st = lib().ref_hybrid_search(self.h, _d3(start), _d3(goal), C.c_uint64(seed), C.c_int(1), C.byref(cost),
                                     C.byref(nexp), C.byref(nchild), C.byref(nsamp), times)
        n = nexp.value
        keys = np.empty((max(n, 1), 4), np.int32)
        poses = np.empty((max(n, 1), 3), np.float64)
        lib().ref_hybrid_expansions(self.h, C.c_long(n), _p(keys), _p(poses))
        sp = np.empty((4096, 3), np.float64)
        kinds = np.empty(4096, np.int32)
        lens = np.empty(4096, np.float64)
        m = lib().ref_hybrid_solution(self.h, C.c_long(4096), _p(sp), _p(kinds), _p(lens)) if st == 0 else 0
        return dict(status=st, cost=cost.value, n_expanded=n, n_children=nchild.value, n_samples=nsamp.value,
                    times=list(times), exp_keys=keys[:n], exp_poses=poses[:n], path=sp[:m].copy(), kinds=kinds[:m].copy(),
                    lengths=lens[:m].copy())

    def resampled(self, path_interpolation=0.1, cap=8192):
        """hybrid_a_star.cpp:260-291 on the last search: resampled graph-search path + cusp flags + composite length."""
        poses = np.empty((cap, 3), np.float64)
        cusp = np.zeros(cap, np.uint8)
        length = C.c_double()
        n = lib().ref_hybrid_resampled(self.h, C.c_float(path_interpolation), C.c_long(cap), _p(poses), _p(cusp), C.byref(length))
        return poses[:n].copy(), cusp[:n].astype(bool), length.value

    def gvd_cells(self):
        """(rows, cols, 2) nearest-obstacle and nearest-Voronoi-edge cells of the planner's GVD (after a search)"""
        i = self.world.info()
        rows, cols = i["rows"], i["cols"]
        a = np.zeros((rows, cols, 2), np.int32)
        b = np.zeros((rows, cols, 2), np.int32)
        lib().ref_hybrid_gvd_cells(self.h, _p(a), _p(b))
        return a, b

    def set_smoother(self, sp):
        """Smoother::SetParameters from a pathplanning_b200.SmootherParams"""
        v = np.array([sp.step_tolerance, sp.max_iterations, sp.learning_rate, sp.path_weight, sp.smooth_weight, sp.voronoi_weight,
                      sp.collision_weight, sp.curvature_weight, sp.collision_ratio, sp.max_curvature], np.float64)
        lib().ref_hybrid_smoother_set(self.h, _p(v))

    def smooth(self, poses, cusp):
        """Smoother::Smooth of the planner's smoother -> (Smoother::Status, GetPath())"""
        poses = np.ascontiguousarray(poses, np.float64)
        cusp = np.ascontiguousarray(cusp, np.uint8)
        out = np.zeros_like(poses)
        st = lib().ref_hybrid_smooth(self.h, C.c_long(len(poses)), _p(poses), _p(cusp), _p(out))
        return int(st), out

    def search_full(self, start, goal, seed):
        sec = C.c_double()
        plen = C.c_long()
        st = lib().ref_hybrid_search_full(self.h, _d3(start), _d3(goal), C.c_uint64(seed), C.byref(sec), C.byref(plen))
        return st, sec.value, plen.value

    def prepare_goal(self, goal):
        lib().ref_hybrid_prepare_goal(self.h, _d3(goal))

    def expand(self, parent, seed):
        poses = np.empty((11, 3), np.float64)
        keys = np.empty((11, 3), np.int32)
        dirs = np.empty(11, np.int32)
        costs = np.empty(11, np.float64)
        kinds = np.empty(11, np.int32)
        heur = np.empty(11, np.float64)
        ph = C.c_double()
        n = lib().ref_hybrid_expand(self.h, _d3(parent), C.c_uint64(seed), _p(poses), _p(keys), _p(dirs), _p(costs),
                                    _p(kinds), _p(heur), C.byref(ph))
        return dict(n=n, poses=poses[:n], keys=keys[:n], dirs=dirs[:n], costs=costs[:n], kinds=kinds[:n],
                    heuristics=heur[:n], parent_heuristic=ph.value)

    def heuristic(self, poses):
        poses = np.ascontiguousarray(poses, np.float64)
        n = len(poses)
        v, nh, ob = np.empty(n), np.empty(n), np.empty(n)
        lib().ref_hybrid_heuristic(self.h, C.c_long(n), _p(poses), _p(v), _p(nh), _p(ob))
        return v, nh, ob


# ---- Tree / RRT ---------------------------------------------------------------------------------
def tree_knn_r2(points, queries, k):
    """Tree<Point2d,2>::GetNearestNodes of the reference over the brute-force flann stand-in."""
    points = np.ascontiguousarray(points, np.float64)
    queries = np.ascontiguousarray(queries, np.float64)
    out = np.empty((len(queries), k), np.int32)
    lib().ref_tree_knn_r2(C.c_long(len(points)), _p(points), C.c_long(len(queries)), _p(queries), C.c_int(k), _p(out))
    return out


def rrt_r2_run(lb, ub, start, goal, seed, max_iteration, max_nodes, max_connection_distance, goal_bias, star):
    """RRTR2 / RRTStarR2 (the reference templates) in free space; returns the whole tree."""
    cap = max_nodes + 8
    n, plen = C.c_long(), C.c_long()
    xy = np.zeros((cap, 2), np.float64)
    parents = np.zeros(cap, np.int32)
    costs = np.zeros(cap, np.float64)
    pxy = np.zeros((cap, 2), np.float64)
    st = lib().ref_rrt_r2_run((C.c_double * 2)(*lb), (C.c_double * 2)(*ub), (C.c_double * 2)(*start), (C.c_double * 2)(*goal),
                              C.c_uint64(seed), C.c_uint(max_iteration), C.c_uint(max_nodes), C.c_double(max_connection_distance),
                              C.c_double(goal_bias), C.c_int(star), C.c_long(cap), C.byref(n), _p(xy), _p(parents), _p(costs),
                              C.byref(plen), _p(pxy))
    return dict(status=st, n_nodes=n.value, nodes=xy[:n.value], parents=parents[:n.value], costs=costs[:n.value], path=pxy[:plen.value])


def rrtstar_se2_restated(world, lb, ub, start, goal, seed, max_iteration, max_nodes, max_connection_distance, goal_bias,
                         reeds_shepp, rmin=2.0):
    """RRT* over Pose2d: restated loop + the reference's own sampling / connection / validator (see ref_harness.cpp)."""
    cap = max_nodes + 8
    n, it, edges = C.c_long(), C.c_long(), C.c_long()
    sec = C.c_double()
    xyt = np.zeros((cap, 3), np.float64)
    parents = np.zeros(cap, np.int32)
    costs = np.zeros(cap, np.float64)
    st = lib().ref_rrtstar_se2_restated(world.h if world is not None else None, _d3(lb), _d3(ub), _d3(start), _d3(goal),
                                        C.c_uint64(seed), C.c_uint(max_iteration), C.c_uint(max_nodes),
                                        C.c_double(max_connection_distance), C.c_double(goal_bias), C.c_int(reeds_shepp),
                                        C.c_double(rmin), C.c_long(cap), C.byref(n), _p(xyt), _p(parents), _p(costs),
                                        C.byref(it), C.byref(edges), C.byref(sec))
    return dict(status=st, n_nodes=n.value, iterations=it.value, edges_checked=edges.value, seconds=sec.value,
                nodes=xyt[:n.value], parents=parents[:n.value], costs=costs[:n.value])
