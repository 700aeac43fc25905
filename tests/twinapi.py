"""ctypes wrapper of the host twin (oracle/twin.cpp -> oracle/libpp_twin.so).  TEST INFRASTRUCTURE:
the kernel arithmetic headers compiled for the host, used by `-m "not gpu"` tests to exercise the
kernels' control logic against golden vectors without a GPU.  Never imported by the product."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TWIN_PATH = os.path.join(ROOT, "oracle", "libpp_twin.so")
_lib = None


def available():
    return os.path.exists(TWIN_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(TWIN_PATH)
        _lib.twin_world_create.restype = C.c_void_p
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class World:
    def __init__(self, desc, sqdist, occupancy=None, pathcost=None):
        self.desc = desc
        d = desc.to_c()
        sq = np.ascontiguousarray(sqdist, np.int32)
        occ = None if occupancy is None else np.ascontiguousarray(occupancy, np.int32)
        pc = None if pathcost is None else np.ascontiguousarray(pathcost, np.float32)
        self.h = C.c_void_p(lib().twin_world_create(C.byref(d), _p(sq), _p(occ), _p(pc)))

    def close(self):
        if self.h:
            lib().twin_world_destroy(self.h)
            self.h = None

    def smooth(self, params_c, near_obstacle, near_edge, poses, cusp):
        """-> (status, smoothed poses (n,3), iterations)"""
        poses = np.ascontiguousarray(poses, np.float64)
        cusp = np.ascontiguousarray(cusp, np.uint8)
        a = np.ascontiguousarray(near_obstacle, np.int32)
        b = np.ascontiguousarray(near_edge, np.int32)
        out = np.zeros_like(poses)
        it = C.c_int()
        st = lib().twin_smooth(self.h, C.byref(params_c), _p(a), _p(b), C.c_long(len(poses)), _p(poses), _p(cusp), _p(out),
                               C.byref(it))
        return int(st), out, it.value

    def check_rs(self, edges, rmin, radius, min_step, impl=1):
        n = len(edges)
        edges = np.ascontiguousarray(edges)
        valid = np.empty(n, np.uint8)
        last = np.empty(n, np.float32)
        samples = np.empty(n, np.uint32)
        micro = np.empty(n, np.uint32)
        lib().twin_check_rs(self.h, C.c_long(n), _p(edges), C.c_double(rmin), C.c_float(radius), C.c_float(min_step),
                            C.c_int(impl), _p(valid), _p(last), _p(samples), _p(micro))
        return dict(valid=valid.astype(bool), last=last, samples=samples, microsteps=micro)


class Hybrid:
    """The Hybrid A* device code (ppx_hybrid.cuh) run sequentially on the host."""

    def __init__(self, world, params, vp, table=None):
        from pathplanning_b200.lib import HybridParamsC, ValidatorParamsC
        self.world = world
        self.params = HybridParamsC(**params) if isinstance(params, dict) else params
        self.vp = vp
        L = lib()
        L.twin_hybrid_create.restype = C.c_void_p
        t = None if table is None else np.ascontiguousarray(table, np.float64)
        self.h = C.c_void_p(L.twin_hybrid_create(world.h, C.byref(self.params), C.byref(vp), _p(t)))
        assert self.h

    def close(self):
        if self.h:
            lib().twin_hybrid_destroy(self.h)
            self.h = None

    def table(self):
        dims = np.zeros(4, np.int32)
        off = np.zeros(2, np.float64)
        lib().twin_hybrid_info(self.h, _p(dims), _p(off), None, None)
        vals = np.empty(tuple(dims[:3].tolist()), np.float64)
        lib().twin_hybrid_info(self.h, _p(dims), _p(off), None, _p(vals))
        return vals, off

    def heuristic(self, wf, goal, poses):
        poses = np.ascontiguousarray(poses, np.float64)
        wf = np.ascontiguousarray(wf, np.float32)
        goal = np.ascontiguousarray(goal, np.float64)
        n = len(poses)
        v, nh, ob = np.empty(n), np.empty(n), np.empty(n)
        lib().twin_hybrid_heuristic(self.h, _p(wf), _p(goal), C.c_long(n), _p(poses), _p(v), _p(nh), _p(ob))
        return v, nh, ob

    def expand(self, wf, goal, parents, seeds):
        from pathplanning_b200.api import HYBRID_CHILD
        parents = np.ascontiguousarray(parents, np.float64)
        seeds = np.ascontiguousarray(seeds, np.uint64)
        wf = np.ascontiguousarray(wf, np.float32)
        goal = np.ascontiguousarray(goal, np.float64)
        n = len(parents)
        count = np.zeros(n, np.int32)
        children = np.zeros((n, 19), HYBRID_CHILD)
        ph = np.empty(n)
        lib().twin_hybrid_expand(self.h, _p(wf), _p(goal), C.c_long(n), _p(parents), _p(seeds), _p(count), _p(children), _p(ph))
        return count, children, ph

    def search(self, wf, start, goal, seed, max_path=2048, max_log=100000, max_expansions=0):
        from pathplanning_b200.api import HYBRID_RESULT, HYBRID_PATH_STATE, HYBRID_LOG_ENTRY
        from pathplanning_b200.lib import HybridLimitsC
        wf = np.ascontiguousarray(wf, np.float32)
        lim = HybridLimitsC(0, max_expansions, max_path, max_log)
        res = np.zeros(1, HYBRID_RESULT)
        path = np.zeros(max_path, HYBRID_PATH_STATE)
        log = np.zeros(max_log, HYBRID_LOG_ENTRY)
        s = np.ascontiguousarray(start, np.float64)
        g = np.ascontiguousarray(goal, np.float64)
        gf = np.zeros((max_log, 2), np.float64)
        lib().twin_hybrid_search(self.h, _p(wf), _p(s), _p(g), C.c_uint64(seed), C.byref(lim), _p(res), _p(path), _p(log), _p(gf))
        r = res[0]
        n = min(int(r["n_expanded"]), max_log)
        return dict(result=r, path=path[:int(r["path_len"])], log=log[:n], gf=gf[:n])


    def resample(self, goal, path, path_interpolation=0.1, cap=8192):
        """-> (poses (n,3), cusp (n,) bool, composite length)"""
        path = np.ascontiguousarray(path)
        g = np.ascontiguousarray(goal, np.float64)
        poses = np.zeros((cap, 3), np.float64)
        cusp = np.zeros(cap, np.uint8)
        length = np.zeros(1, np.float64)
        fn = lib().twin_hybrid_resample
        fn.restype = C.c_long
        n = fn(self.h, _p(g), _p(path), C.c_int(len(path)), C.c_float(path_interpolation), C.c_long(cap), _p(poses), _p(cusp),
               _p(length))
        assert n >= 0, n
        return poses[:n].copy(), cusp[:n].astype(bool), float(length[0])


def mt64_uniform(seed, n):
    out = np.empty(n, np.float64)
    lib().twin_mt64_uniform(C.c_uint64(seed), C.c_long(n), _p(out))
    return out
