"""ctypes binding of oracle/libga_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  Nothing under guided-extended-hybrid-astar_b200/ does."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libga_oracle.so")
_lib = None

H_CLOSED, H_UNKNOWN = 1, 2
DT_FIXEDPOINT, DT_IPPFLOAT = 0, 1


class DisksInfo(C.Structure):
    _fields_ = [("nb_disks", C.c_int), ("nb_yaws", C.c_int), ("yaw_res_deg", C.c_int),
                ("disk_r", C.c_double), ("disk_r_c", C.c_int)]


class ExpandParams(C.Structure):
    _fields_ = [("max_steer", C.c_double), ("wheel_base", C.c_double), ("arc_l", C.c_double),
                ("yaw_res_astar_deg", C.c_double), ("nb_steer", C.c_int), ("with_rear_axis", C.c_int),
                ("turn_on_point_angle_deg", C.c_double), ("astar_dim", C.c_int)]


class HeurParams(C.Structure):
    _fields_ = [("astar_dim", C.c_int), ("astar_prox_cost", C.c_double),
                ("astar_lane_movement_cost", C.c_double), ("unknown_cost_w", C.c_double),
                ("heuristic_early_exit", C.c_int), ("max_extra_nodes", C.c_uint),
                ("get_only_near", C.c_int)]


class PotParams(C.Structure):
    _fields_ = [("alpha", C.c_double), ("do_max", C.c_double), ("do_min", C.c_double),
                ("motion_res_min", C.c_double), ("motion_res_max", C.c_double),
                ("dist_val_min", C.c_double), ("dist_val_max", C.c_double),
                ("vor_dim", C.c_int), ("dt_variant", C.c_int)]


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
        _lib.gao_angle_0_2pi.restype = C.c_double
        _lib.gao_angle_mpi_pi.restype = C.c_double
        _lib.gao_angle_diff.restype = C.c_double
        for f in ("gao_angle_0_2pi", "gao_angle_mpi_pi"):
            getattr(_lib, f).argtypes = [C.c_double]
        _lib.gao_angle_diff.argtypes = [C.c_double, C.c_double]
        _lib.gao_yaw_idx.argtypes = [C.c_double, C.c_int]
        _lib.gao_threshold.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int]
        _lib.gao_check_pose.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double,
                                        C.c_double, C.c_int]
        _lib.gao_expand_node.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_double,
                                         C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int]
        _lib.gao_calc_disks.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_double, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_int]
        _lib.gao_move_car_some_steps.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int,
                                                 C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        _lib.gao_turn_on_rear_axis.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_int]
        _lib.gao_steering_inputs.argtypes = [C.c_double, C.c_int, C.c_void_p]
        _lib.gao_bilinear.argtypes = [C.c_double] * 4 + [C.c_int] * 4 + [C.c_double] * 2
        _lib.gao_bilinear.restype = C.c_double
        _lib.gao_bilin_at.argtypes = [C.c_double, C.c_double, C.c_void_p, C.c_int]
        _lib.gao_bilin_at.restype = C.c_double
        _lib.gao_prox_of_corners.argtypes = [C.c_void_p, C.c_int, C.c_void_p] + [C.c_double] * 3
        _lib.gao_prox_of_corners.restype = C.c_double
        _lib.gao_path_prox_costs.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        _lib.gao_costmap_value.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        _lib.gao_check_grid.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def ellipse_rows(r):
    hw = np.zeros(2 * r + 1, np.int32)
    se = np.zeros((2 * r + 1, 2 * r + 1), np.uint8)
    taps = lib().gao_ellipse_rows(r, _p(hw), _p(se))
    return hw, se, taps


def threshold(patch, lo, hi):
    patch = np.ascontiguousarray(patch, np.uint8)
    out = np.empty_like(patch)
    lib().gao_threshold(_p(patch), _p(out), patch.size, lo, hi)
    return out


def dilate_disk(img, r):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    lib().gao_dilate_disk(_p(img), _p(out), img.shape[0], r)
    return out


def process_safety_patch(patch, lo, hi, r):
    patch = np.ascontiguousarray(patch, np.uint8)
    out = np.empty_like(patch)
    lib().gao_process_safety_patch(_p(patch), _p(out), patch.shape[0], lo, hi, r)
    return out


def pool_max(safety, win=4):
    safety = np.ascontiguousarray(safety, np.uint8)
    pd = safety.shape[0]
    ad = pd // win
    out = np.empty((ad, ad), np.uint8)
    lib().gao_pool_max(_p(safety), _p(out), pd, ad, win)
    return out


def pass_local_map(patch, local_map, ox, oy):
    local_map = np.ascontiguousarray(local_map, np.uint8)
    lib().gao_pass_local_map(_p(patch), patch.shape[0], _p(local_map), int(ox), int(oy), local_map.shape[0])


def cartograph(cart, local_map, ox, oy):
    local_map = np.ascontiguousarray(local_map, np.uint8)
    lib().gao_cartograph(_p(cart), cart.shape[0], _p(local_map), int(ox), int(oy), local_map.shape[0])


def copy_patch_to_patch(prev, prev_o, nxt, o):
    lib().gao_copy_patch_to_patch(_p(prev), prev.shape[0], int(prev_o[0]), int(prev_o[1]), _p(nxt), nxt.shape[0],
                                  int(o[0]), int(o[1]))


class ProxParams(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("lf", "lb", "width", "con2star", "h_prox_cost", "interp_res")]


def bilinear(q11, q12, q21, q22, x1, x2, y1, y2, x, y):
    return lib().gao_bilinear(q11, q12, q21, q22, x1, x2, y1, y2, x, y)


def bilin_at(x, y, grid):
    grid = np.ascontiguousarray(grid, np.float64)
    return lib().gao_bilin_at(x, y, _p(grid), grid.shape[0])


def prox_of_corners(grid, params, poses):
    """poses [n, 3] in metres / rad -> getProxOfCorners(pose * con2star_) per pose"""
    grid = np.ascontiguousarray(grid, np.float64)
    p = ProxParams(*params)
    return np.array([lib().gao_prox_of_corners(_p(grid), grid.shape[0], C.byref(p), float(x), float(y), float(w))
                     for x, y, w in np.asarray(poses, np.float64).reshape(-1, 3)], np.float64)


def path_prox_costs(grid, params, poses, offsets):
    grid = np.ascontiguousarray(grid, np.float64)
    poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
    offsets = np.ascontiguousarray(offsets, np.int32)
    p = ProxParams(*params)
    out = np.zeros(len(offsets) - 1, np.float64)
    lib().gao_path_prox_costs(_p(grid), grid.shape[0], C.byref(p), _p(poses), _p(offsets), len(offsets) - 1, _p(out))
    return out


def valid_close_pose(safety, centers, yaw_res_deg, prox, params, near_cells, center):
    """step 3 of getValidClosePose; centers [nb_yaws, nb_disks, 2]; returns the pose or None"""
    safety = np.ascontiguousarray(safety, np.uint8)
    centers = np.ascontiguousarray(centers, np.int32)
    prox = np.ascontiguousarray(prox, np.float64)
    member = np.zeros(prox.shape[0] * prox.shape[0], np.uint8)
    member[np.asarray(near_cells, np.int64)] = 1
    p = ProxParams(*params)
    c = np.ascontiguousarray(center, np.float64)
    out = np.zeros(3, np.float64)
    lib().gao_valid_close_pose.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p]
    ok = lib().gao_valid_close_pose(_p(safety), safety.shape[0], _p(centers), centers.shape[-2], int(yaw_res_deg), _p(prox),
                                    prox.shape[0], C.byref(p), _p(member), _p(c), _p(out))
    return out if ok else None


GM_DIM = 801


def precast_lists():
    off = np.zeros(181, np.int32)
    ent = np.zeros(2 * GM_DIM * GM_DIM + 4096, np.int32)
    n = lib().gao_precast_lists(_p(ent), ent.size, _p(off))
    assert n <= ent.size
    return ent[:n].copy(), off


_PRECAST = None


def raytrace(grid_map_gt):
    """GridMapSim::raytracing (gridmap_sim.cpp:42-80) on an 801 x 801 window"""
    global _PRECAST
    if _PRECAST is None:
        _PRECAST = precast_lists()
    gt = np.ascontiguousarray(grid_map_gt, np.uint8)
    out = np.empty((GM_DIM, GM_DIM), np.uint8)
    lib().gao_raytrace(_p(gt), _p(_PRECAST[0]), _p(_PRECAST[1]), _p(out))
    return out


def calc_disks(length, width, lb, len_per_disk=2.0, double_rows=False, safety=0.1, yaw_res_deg=3):
    info = DisksInfo()
    cen = np.zeros(360 * 32 * 2, np.int32)
    rc = lib().gao_calc_disks(length, width, lb, len_per_disk, int(double_rows), safety, yaw_res_deg, C.byref(info),
                              _p(cen), cen.size)
    assert rc == 0
    return info, cen[:info.nb_yaws * info.nb_disks * 2].reshape(info.nb_yaws, info.nb_disks, 2).copy()


def yaw_idx(yaw, yaw_res_deg=3):
    return lib().gao_yaw_idx(yaw, yaw_res_deg)


def check_pose(safety, centers, yaw_res_deg, x, y, yaw, variant=0):
    return lib().gao_check_pose(_p(safety), safety.shape[0], _p(centers), centers.shape[1], yaw_res_deg, x, y, yaw,
                                variant)


def check_segments(safety, centers, yaw_res_deg, x, y, yaw, seg_off, variant=0):
    x, y, yaw = (np.ascontiguousarray(v, np.float64) for v in (x, y, yaw))
    seg = np.ascontiguousarray(seg_off, np.int32)
    centers = np.ascontiguousarray(centers, np.int32)
    hit = np.empty(len(seg) - 1, np.int32)
    lib().gao_check_segments(_p(safety), safety.shape[0], _p(centers), centers.shape[1], yaw_res_deg, _p(x), _p(y),
                             _p(yaw), _p(seg), len(seg) - 1, variant, _p(hit))
    return hit


def move_car_some_steps(x, y, yaw, arc_l, motion_res, direction, steer, wb):
    xs, ys, yaws = (np.empty(512) for _ in range(3))
    n = lib().gao_move_car_some_steps(x, y, yaw, arc_l, motion_res, direction, steer, wb, _p(xs), _p(ys), _p(yaws), 512)
    return xs[:n].copy(), ys[:n].copy(), yaws[:n].copy()


def turn_on_rear_axis(yaw, delta, yaw_res_deg):
    yaws = np.empty(1024)
    n = lib().gao_turn_on_rear_axis(yaw, delta, yaw_res_deg, _p(yaws), 1024)
    return yaws[:n].copy()


def steering_inputs(max_steer, nb):
    out = np.empty(nb)
    lib().gao_steering_inputs(max_steer, nb, _p(out))
    return out


def expand_node(safety, centers, yaw_res_coll_deg, ep, x, y, yaw, motion_res, cap=64):
    centers = np.ascontiguousarray(centers, np.int32)
    valid = np.zeros(cap, np.uint8)
    disc = np.zeros((cap, 3), np.int32)
    end = np.zeros((cap, 3), np.float64)
    npos = np.zeros(cap, np.int32)
    k = lib().gao_expand_node(_p(safety), safety.shape[0], _p(centers), centers.shape[1], yaw_res_coll_deg,
                              C.cast(C.byref(ep), C.c_void_p), x, y, yaw, motion_res, _p(valid), _p(disc), _p(end),
                              _p(npos), cap)
    return valid[:k], disc[:k], end[:k], npos[:k]


def heuristic_reforder(hp, grid, fence, mc, prox, goal, start):
    n = grid.size
    cost, dist = np.empty(n), np.empty(n)
    parent = np.empty(n, np.int32)
    flags = np.empty(n, np.uint8)
    order = np.empty(n, np.int32)
    near_ids = np.empty(n, np.int32)
    near_d = np.empty(n)
    n_near = C.c_int()
    fence_p = _p(fence) if fence is not None else None
    k = lib().gao_heuristic_reforder(C.byref(hp), _p(grid), fence_p, _p(mc), _p(prox), int(goal[0]), int(goal[1]), int(start[0]),
                                     int(start[1]), _p(cost), _p(dist), _p(parent), _p(flags), _p(order), _p(near_ids),
                                     _p(near_d), C.byref(n_near))
    sh = grid.shape
    return dict(n_closed=k, cost=cost.reshape(sh), dist=dist.reshape(sh), parent=parent.reshape(sh),
                flags=flags.reshape(sh), order=order[:k].copy(), near_ids=near_ids[:n_near.value].copy(),
                near_dist=near_d[:n_near.value].copy())


def heuristic_fixedpoint(hp, grid, fence, mc, prox, goal):
    n = grid.size
    cost, dist = np.empty(n), np.empty(n)
    parent = np.empty(n, np.int32)
    flags = np.empty(n, np.uint8)
    fence_p = _p(fence) if fence is not None else None
    k = lib().gao_heuristic_fixedpoint(C.byref(hp), _p(grid), fence_p, _p(mc), _p(prox), int(goal[0]), int(goal[1]), _p(cost),
                                       _p(dist), _p(parent), _p(flags))
    sh = grid.shape
    return dict(n_closed=k, cost=cost.reshape(sh), dist=dist.reshape(sh), parent=parent.reshape(sh),
                flags=flags.reshape(sh))


def astar_path(parent, flags, x, y):
    dim = parent.shape[0]
    xs = np.empty(dim * dim, np.int32)
    ys = np.empty(dim * dim, np.int32)
    n = lib().gao_astar_path(dim, _p(parent), _p(flags), int(x), int(y), _p(xs), _p(ys), dim * dim)
    return xs[:n].copy(), ys[:n].copy()


def distance_transform(src, variant):
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    out = np.empty((h, w), np.float32)
    lib().gao_distance_transform(_p(src), w, w, h, _p(out), variant)
    return out


def voronoi_potential_field(pp, grid, ego, vor_xy, obstacle_dist, voronoi_dist, h_prox, motion_res):
    v = np.ascontiguousarray(vor_xy, np.int32).reshape(-1, 2)
    return lib().gao_voronoi_potential_field(C.byref(pp), _p(grid), grid.shape[0], int(ego[0]), int(ego[1]), _p(v), len(v),
                                             _p(obstacle_dist), _p(voronoi_dist), _p(h_prox), _p(motion_res))
