"""Shared helpers of the test-suite."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


def load_kat():
    return json.load(open(os.path.join(GOLDEN, "rs_kat.json")))


def load_npz(name):
    return np.load(os.path.join(GOLDEN, name))


def c2_world_desc(g):
    """MapDesc + grids of the golden C2 parking-lot fixture."""
    from pathplanning_b200 import MapDesc
    occ = g["occupancy"].astype(np.int32)
    desc = MapDesc(rows=occ.shape[0], cols=occ.shape[1], resolution=float(g["resolution"]),
                   world_origin=tuple(g["world_origin"].tolist()), local_origin=(0.0, 0.0),
                   lb=tuple(g["lb"].tolist()), ub=tuple(g["ub"].tolist()))
    return desc, occ, g["sqdist"].astype(np.int32), g["pathcost"].astype(np.float32)


def assert_close_rel(a, b, rel, what=""):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    err = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
    err = np.where(both_inf | (a == b), 0.0, err)
    bad = ~(err <= rel)
    assert not bad.any(), f"{what}: {bad.sum()} of {bad.size} beyond rel {rel}; worst {np.nanmax(err)}"
