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
