"""ctypes binding of libppx.so -- one declaration per entry point of include/ppx.h."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libppx.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(nvcc, sm_100a).  pathplanning_b200 has no CPU fallback.")

ppx = C.CDLL(LIB_PATH)


class MapDescC(C.Structure):
    _fields_ = [("rows", C.c_int32), ("cols", C.c_int32), ("resolution", C.c_float), ("_pad", C.c_float),
                ("world_origin", C.c_double * 2), ("local_origin", C.c_double * 2),
                ("lb", C.c_double * 3), ("ub", C.c_double * 3)]


class ValidatorParamsC(C.Structure):
    _fields_ = [("min_safe_radius", C.c_float), ("min_path_interpolation_distance", C.c_float)]


class FootprintParamsC(C.Structure):
    _fields_ = [("length", C.c_double), ("width", C.c_double), ("rear_to_center", C.c_double), ("n_discs", C.c_int32),
                ("min_step", C.c_float)]


class AstarParamsC(C.Structure):
    _fields_ = [("cost_kind", C.c_int32), ("heuristic_kind", C.c_int32), ("cell_cost", C.c_void_p), ("bidirectional", C.c_int32),
                ("_pad", C.c_int32)]


class RSCostParamsC(C.Structure):
    _fields_ = [("min_turning_radius", C.c_double), ("reverse_cost_multiplier", C.c_float),
                ("forward_cost_multiplier", C.c_float), ("direction_switching_cost", C.c_float),
                ("shortest", C.c_int32)]


class HybridParamsC(C.Structure):
    _fields_ = [("wheelbase", C.c_double), ("min_turning_radius", C.c_double), ("direction_switching_cost", C.c_double),
                ("reverse_cost_multiplier", C.c_double), ("forward_cost_multiplier", C.c_double),
                ("voronoi_cost_multiplier", C.c_double), ("num_generated_motion", C.c_uint32), ("_pad", C.c_uint32),
                ("spatial_resolution", C.c_double), ("angular_resolution", C.c_double)]


class RRTParamsC(C.Structure):
    _fields_ = [("dim", C.c_int32), ("connection", C.c_int32), ("star", C.c_int32), ("validator", C.c_int32),
                ("max_iteration", C.c_uint32), ("max_nodes", C.c_uint32), ("max_connection_distance", C.c_double),
                ("goal_bias", C.c_double), ("min_turning_radius", C.c_double), ("lb", C.c_double * 3), ("ub", C.c_double * 3)]


class HybridLimitsC(C.Structure):
    _fields_ = [("max_nodes", C.c_int32), ("max_expansions", C.c_int32), ("max_path", C.c_int32), ("max_log", C.c_int32)]


class SmootherParamsC(C.Structure):
    _fields_ = [("step_tolerance", C.c_float), ("max_iterations", C.c_int32), ("learning_rate", C.c_float),
                ("path_weight", C.c_float), ("smooth_weight", C.c_float), ("voronoi_weight", C.c_float),
                ("collision_weight", C.c_float), ("curvature_weight", C.c_float), ("collision_ratio", C.c_float),
                ("max_curvature", C.c_float), ("alpha", C.c_float), ("d_max", C.c_float), ("min_safe_radius", C.c_float)]


_p = C.c_void_p
_i64 = C.c_int64
_SIGS = {
    "ppx_version": (C.c_int, []),
    "ppx_ctx_create": (C.c_int, [C.c_int, C.POINTER(_p)]),
    "ppx_ctx_destroy": (C.c_int, [_p]),
    "ppx_ctx_set_stream": (C.c_int, [_p, _p]),
    "ppx_ctx_synchronize": (C.c_int, [_p]),
    "ppx_ctx_set_option": (C.c_int, [_p, C.c_char_p, _i64]),
    "ppx_last_error": (C.c_char_p, [_p]),
    "ppx_ctx_launch_count": (C.c_int64, [_p]),
    "ppx_dev_alloc": (C.c_int, [_p, C.c_uint64, C.POINTER(_p)]),
    "ppx_dev_free": (C.c_int, [_p, _p]),
    "ppx_host_alloc": (C.c_int, [_p, C.c_uint64, C.POINTER(_p)]),
    "ppx_host_free": (C.c_int, [_p, _p]),
    "ppx_memcpy_h2d": (C.c_int, [_p, _p, _p, C.c_uint64]),
    "ppx_memcpy_d2h": (C.c_int, [_p, _p, _p, C.c_uint64]),
    "ppx_map_upload": (C.c_int, [_p, C.POINTER(MapDescC), _p, _p, _p]),
    "ppx_map_build": (C.c_int, [_p, C.POINTER(MapDescC), _p, _p, _p]),
    "ppx_gvd_build": (C.c_int, [_p, C.POINTER(MapDescC), _p, C.c_int32, _p, _p, _p, _p, _p]),
    "ppx_is_state_valid": (C.c_int, [_p, _i64, _p, C.POINTER(ValidatorParamsC), _p]),
    "ppx_check_edges_rs": (C.c_int, [_p, _i64, _p, C.c_double, C.POINTER(ValidatorParamsC), _p, _p, _p]),
    "ppx_check_edges_rs_audit": (C.c_int, [_p, _i64, _p, C.c_double, C.POINTER(ValidatorParamsC), _p, _p, _p, _p, _p]),
    "ppx_check_edges_rs_dev": (C.c_int, [_p, _i64, _p, C.c_double, C.POINTER(ValidatorParamsC), _p, _p, _p]),
    "ppx_check_edges_csteer": (C.c_int, [_p, _i64, _p, C.POINTER(ValidatorParamsC), _p, _p, _p]),
    "ppx_check_edges_line": (C.c_int, [_p, _i64, _p, C.POINTER(ValidatorParamsC), _p, _p]),
    "ppx_rs_solve": (C.c_int, [_p, _i64, _p, _p, C.POINTER(RSCostParamsC), _p, _p, _p, _p]),
    "ppx_rs_solve_dev": (C.c_int, [_p, _i64, _p, _p, C.POINTER(RSCostParamsC), _p, _p, _p, _p]),
    "ppx_connect_check_rs": (C.c_int, [_p, _i64, _p, _p, C.POINTER(RSCostParamsC), C.POINTER(ValidatorParamsC),
                                       _p, _p, _p, _p, _p]),
    "ppx_connect_check_rs_dev": (C.c_int, [_p, _i64, _p, _p, C.POINTER(RSCostParamsC), C.POINTER(ValidatorParamsC),
                                           _p, _p, _p, _p, _p]),
    "ppx_rs_interpolate": (C.c_int, [_p, _i64, _p, C.c_double, C.c_int32, _p, _p]),
    "ppx_wavefront": (C.c_int, [_p, _i64, _p, _p, _p]),
    "ppx_wavefront_dev": (C.c_int, [_p, _i64, _p, _p, _p]),
    "ppx_wavefront_dev_static": (C.c_int, [_p, _i64, _p, _p, _p, _p]),
    "ppx_knn": (C.c_int, [_p, C.c_int32, _i64, _p, _i64, _p, C.c_int32, _p, _p]),
    "ppx_knn_dev": (C.c_int, [_p, C.c_int32, _i64, _p, _i64, _p, C.c_int32, _p, _p]),
    "ppx_rrt_run": (C.c_int, [_p, C.POINTER(RRTParamsC), _p, _i64, _p, _p, _p, _p, _i64, _p, _p, _p]),
    "ppx_hybrid_create": (C.c_int, [_p, C.POINTER(HybridParamsC), C.POINTER(ValidatorParamsC), C.POINTER(_p)]),
    "ppx_hybrid_destroy": (C.c_int, [_p]),
    "ppx_hybrid_table": (C.c_int, [_p, _p, _p, _p]),
    "ppx_hybrid_heuristic": (C.c_int, [_p, _p, _i64, _p, _p, _p, _p]),
    "ppx_hybrid_expand": (C.c_int, [_p, _p, _i64, _p, _p, _p, _p, _p]),
    "ppx_hybrid_search": (C.c_int, [_p, _i64, _p, _p, _p, C.POINTER(HybridLimitsC), _p, _p, _p]),
    "ppx_hybrid_resample": (C.c_int, [_p, _i64, _p, _p, _p, C.c_int32, C.c_float, C.c_int32, _p, _p, _p, _p]),
    "ppx_smoother_create": (C.c_int, [_p, C.POINTER(SmootherParamsC), C.POINTER(_p)]),
    "ppx_smoother_destroy": (C.c_int, [_p]),
    "ppx_smoother_set_gvd": (C.c_int, [_p, _p, _p]),
    "ppx_smoother_use_gvd": (C.c_int, [_p]),
    "ppx_smoother_smooth": (C.c_int, [_p, _i64, _p, _p, _p, C.c_int32, _p, _p, _p]),
    "ppx_map_rasterize": (C.c_int, [_p, C.POINTER(MapDescC), C.c_int32, _p, _p, C.c_int32, _p]),
    "ppx_astar_n2": (C.c_int, [_p, _i64, _p, _p, C.c_int32, C.c_int32, _p, _p]),
    "ppx_astar_n2_ex": (C.c_int, [_p, _i64, _p, _p, C.POINTER(AstarParamsC), C.c_int32, C.c_int32, _p, _p]),
    "ppx_hybrid_search_dev": (C.c_int, [_p, _i64, _p, _p, _p, C.POINTER(HybridLimitsC), _p, _p, _p]),
    "ppx_hybrid_timing": (C.c_int, [_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "ppx_is_state_valid_rect": (C.c_int, [_p, _i64, _p, C.POINTER(FootprintParamsC), _p]),
    "ppx_check_edges_rs_rect": (C.c_int, [_p, _i64, _p, C.c_double, C.POINTER(FootprintParamsC), _p, _p, _p]),
    "ppx_check_edges_rs_rect_dev": (C.c_int, [_p, _i64, _p, C.c_double, C.POINTER(FootprintParamsC), _p, _p, _p]),
    "ppx_comm_unique_id": (C.c_int, [_p]),
    "ppx_comm_init": (C.c_int, [_p, _p, C.c_int32, C.c_int32]),
    "ppx_comm_destroy": (C.c_int, [_p]),
    "ppx_comm_broadcast_map": (C.c_int, [_p, C.c_int32]),
    "ppx_comm_gather_results": (C.c_int, [_p, _p, _i64, C.c_int32, _p, _i64, _p]),
    "ppx_comm_gather_paths": (C.c_int, [_p, _i64, _p, _p, C.c_int32, _p, _p, _i64, _p, _i64, _p]),
}

for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(ppx, _name)  # AttributeError here = header and library out of sync
    _fn.restype = _res
    _fn.argtypes = _args

EXPORTED = tuple(_SIGS)
