"""B200-native collision-checked expansion path for lfilipozzi/PathPlanning.

The product is ``libppx.so`` (hand-written sm_100a CUDA behind the C ABI in ``include/ppx.h``).
This package is the thin Python host side used by tests and bench.py: a ctypes binding
(:mod:`pathplanning_b200.lib`) and numpy-level wrappers (:mod:`pathplanning_b200.api`) that
mirror the reference's validator / solver / heuristic interfaces.  There is no CPU fallback:
importing :mod:`pathplanning_b200.lib` raises if the library is missing, and creating a
:class:`pathplanning_b200.api.Context` raises without a CUDA device.
"""
from . import lib  # noqa: F401  (raises ImportError loudly when libppx.so is not built)
from .api import (  # noqa: F401
    Context, MapDesc, ValidatorParams, RSCostParams, PPXError,
    RS_EDGE, CSTEER_EDGE, LINE_EDGE, make_rs_edges, make_csteer_edges, HybridParams, HybridAStar,
    SmootherParams, Smoother,
)
