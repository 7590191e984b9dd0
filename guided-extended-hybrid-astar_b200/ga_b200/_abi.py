"""ctypes binding of libgridaccel.so (include/gridaccel.h).  No fallback of any kind: if the
library is missing or a call fails, an exception is raised."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libgridaccel.so")


class GaError(RuntimeError):
    pass


class CollisionParams(C.Structure):
    _fields_ = [("min_thresh", C.c_int), ("max_thresh", C.c_int), ("disk_r_c", C.c_int),
                ("yaw_res_deg", C.c_int), ("nb_disks", C.c_int), ("pool_win", C.c_int)]


class ExpandParams(C.Structure):
    _fields_ = [("max_steer", C.c_double), ("wheel_base", C.c_double), ("arc_l", C.c_double),
                ("yaw_res_deg", C.c_double), ("turn_on_point_angle_deg", C.c_double),
                ("nb_steer", C.c_int)]


class ProxParams(C.Structure):
    _fields_ = [("lf", C.c_double), ("lb", C.c_double), ("width", C.c_double), ("con2star", C.c_double),
                ("h_prox_cost", C.c_double), ("interp_res", C.c_double)]


class AstarParams(C.Structure):
    _fields_ = [("astar_movement_cost", C.c_double), ("astar_prox_cost", C.c_double),
                ("astar_lane_movement_cost", C.c_double), ("unknown_cost_w", C.c_double),
                ("alpha", C.c_double), ("do_max", C.c_double), ("do_min", C.c_double),
                ("motion_res_min", C.c_double), ("motion_res_max", C.c_double),
                ("dist_val_min", C.c_double), ("dist_val_max", C.c_double),
                ("vor_dim", C.c_int), ("dt_variant", C.c_int)]


class HeurQuery(C.Structure):
    _fields_ = [("goal_x", C.c_int), ("goal_y", C.c_int), ("start_x", C.c_int), ("start_y", C.c_int),
                ("flags", C.c_uint), ("max_extra_nodes", C.c_uint)]


HEUR_FIXEDPOINT, HEUR_REFORDER, HEUR_EARLY_EXIT, HEUR_ONLY_NEAR = 0, 1, 2, 4

NODE_DTYPE = np.dtype([("x", "f8"), ("y", "f8"), ("yaw", "f8"), ("motion_res", "f8"),
                       ("with_rear_axis", "i4"), ("pad_", "i4")])
SUCC_DTYPE = np.dtype([("x", "f8"), ("y", "f8"), ("yaw", "f8"), ("xi", "i4"), ("yi", "i4"),
                       ("yawi", "i4"), ("n_poses", "i4"), ("valid", "u1"),
                       ("near_cell_edge", "u1"), ("collision_free", "u1"), ("collision_ambiguous", "u1"),
                       ("pad_", "u1", (4,))])
SUCC_COST_DTYPE = np.dtype([("heur_cost", "f8"), ("heur_dist", "f8"), ("movement_cost", "f8"), ("motion_res", "f8")])
OUT_OF_HEURISTIC = 999999999.0
assert NODE_DTYPE.itemsize == 40 and SUCC_DTYPE.itemsize == 48 and SUCC_COST_DTYPE.itemsize == 32

# ga_grid_id
GRID_PATCH, GRID_SAFETY, GRID_ASTAR, GRID_GEOFENCE, GRID_MOVEMENT_COST, GRID_H_PROX, \
    GRID_MOTION_RES, GRID_OBSTACLE_DIST, GRID_VORONOI_DIST, GRID_CART, GRID_PMAP = range(11)
GRID_HEUR_COST, GRID_HEUR_DIST, GRID_HEUR_PARENT, GRID_HEUR_FLAGS, GRID_HEUR_PACKED, GRID_HEUR_REC = 16, 18, 20, 22, 24, 26
# GA_GRID_HEUR_REC: 12 bytes per cell -- cost_ (two 32-bit words, low first), then dir:3 | unknown:1 | straight:14 | diagonal:14
HEUR_REC_DTYPE = np.dtype([("cost_lo", np.uint32), ("cost_hi", np.uint32), ("w", np.uint32)])
assert HEUR_REC_DTYPE.itemsize == 12
GRID_DTYPES = {GRID_PATCH: np.uint8, GRID_SAFETY: np.uint8, GRID_ASTAR: np.uint8,
               GRID_GEOFENCE: np.uint8, GRID_MOVEMENT_COST: np.float64, GRID_H_PROX: np.float64,
               GRID_MOTION_RES: np.float64, GRID_OBSTACLE_DIST: np.float32,
               GRID_VORONOI_DIST: np.float32, GRID_CART: np.uint8, GRID_PMAP: np.uint8}
for _s in (0, 1):
    GRID_DTYPES[GRID_HEUR_COST + _s] = np.float64
    GRID_DTYPES[GRID_HEUR_DIST + _s] = np.float64
    GRID_DTYPES[GRID_HEUR_PARENT + _s] = np.int32
    GRID_DTYPES[GRID_HEUR_FLAGS + _s] = np.uint8
    GRID_DTYPES[GRID_HEUR_PACKED + _s] = np.uint8
    GRID_DTYPES[GRID_HEUR_REC + _s] = HEUR_REC_DTYPE

K_PACK, K_DILATE, K_HEUR_INIT, K_HEUR_RELAX, K_HEUR_FINISH, K_COLLISION, K_EXPAND, K_DT, \
    K_POTENTIAL, K_CARTOGRAPH, K_RAYTRACE, K_PROX, K_COUNT = range(13)
DT_FIXEDPOINT, DT_IPPFLOAT = 0, 1
H_CLOSED, H_UNKNOWN = 1, 2

# every symbol include/gridaccel.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "ga_abi_version": (C.c_int, []),
    "ga_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "ga_destroy": (C.c_int, [_P]),
    "ga_last_error": (C.c_char_p, [_P]),
    "ga_sync": (C.c_int, [_P]),
    "ga_stream": (_P, [_P]),
    "ga_set_collision_params": (C.c_int, [_P, C.POINTER(CollisionParams), _P]),
    "ga_patch_reset": (C.c_int, [_P, C.c_int]),
    "ga_patch_insert": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int]),
    "ga_build_safety": (C.c_int, [_P]),
    "ga_check_poses": (C.c_int, [_P, _P, _P, _P, C.c_int, _P, C.c_int, C.c_int, _P]),
    "ga_set_expand_params": (C.c_int, [_P, C.POINTER(ExpandParams)]),
    "ga_expand_stride": (C.c_int, [_P]),
    "ga_expand": (C.c_int, [_P, _P, C.c_int, _P]),
    "ga_expand_submit": (C.c_int, [_P, _P, C.c_int, C.POINTER(C.c_int)]),
    "ga_expand_collect": (C.c_int, [_P, C.c_int, _P, C.POINTER(C.c_int)]),
    "ga_expand_collect_flags": (C.c_int, [_P, C.c_int, _P, C.POINTER(C.c_int)]),
    "ga_expand_ex": (C.c_int, [_P, _P, C.c_int, C.c_int, _P, _P, _P]),
    "ga_set_astar_params": (C.c_int, [_P, C.POINTER(AstarParams)]),
    "ga_cart_reset": (C.c_int, [_P, C.c_int]),
    "ga_cartograph": (C.c_int, [_P, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "ga_cart_pass_local_map": (C.c_int, [_P, C.c_int, C.c_int, C.c_int]),
    "ga_cart_load_prev": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int]),
    "ga_cart_dim": (C.c_int, [_P]),
    "ga_raytrace": (C.c_int, [_P, C.c_void_p, C.c_void_p]),
    "ga_cartograph_raytraced": (C.c_int, [_P, C.c_int, C.c_int]),
    "ga_set_prox_params": (C.c_int, [_P, C.POINTER(ProxParams)]),
    "ga_prox_of_corners": (C.c_int, [_P, C.c_void_p, C.c_int, C.c_void_p]),
    "ga_valid_close_pose": (C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int)]),
    "ga_valid_close_pose_noalloc": (C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int)]),
    "ga_path_prox_costs": (C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "ga_build_heuristic": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "ga_build_heuristic_batch": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ga_potential_sites": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_int, C.POINTER(C.c_int)]),
    "ga_build_potential": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_int, C.POINTER(C.c_int)]),
    "ga_upload": (C.c_int, [_P, C.c_int, _P, C.c_size_t]),
    "ga_download": (C.c_int, [_P, C.c_int, _P, C.c_size_t]),
    "ga_download_async": (C.c_int, [_P, C.c_int, _P, C.c_size_t]),
    "ga_download_window": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "ga_grid_bytes": (C.c_int, [_P, C.c_int, C.POINTER(C.c_size_t)]),
    "ga_device_ptr": (C.c_int, [_P, C.c_int, C.POINTER(_P)]),
    "ga_patch_dim": (C.c_int, [_P]),
    "ga_astar_dim": (C.c_int, [_P]),
    "ga_prof_enable": (C.c_int, [_P, C.c_int]),
    "ga_prof_reset": (C.c_int, [_P]),
    "ga_prof_get": (C.c_int, [_P, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]),
    "ga_launch_count": (C.c_longlong, [_P]),
    "ga_heur_stats": (C.c_int, [_P, C.POINTER(C.c_longlong)]),
    "ga_set_heur_knobs": (C.c_int, [_P, _P]),
    "ga_set_safety_staging": (C.c_int, [_P, C.c_int]),
    "ga_build_heuristic_ex": (C.c_int, [_P, C.c_int, C.POINTER(HeurQuery), C.POINTER(C.c_int)]),
    "ga_heur_near_nodes": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, C.POINTER(C.c_int)]),
    "ga_heur_closing_order": (C.c_int, [_P, _P, C.c_int, C.POINTER(C.c_int)]),
    "ga_heur_reforder_host": (C.c_int, [C.POINTER(AstarParams), C.c_int, _P, _P, _P, _P, C.POINTER(HeurQuery), _P, _P, _P,
                                        _P, _P, C.POINTER(C.c_int)]),
}

_lib = None


def load():
    """dlopen libgridaccel.so and type every entry point.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GaError("libgridaccel.so is not built (%s): run python -c 'import __graft_entry__ as g; "
                      "g.build()' or make -C guided-extended-hybrid-astar_b200/csrc" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)
