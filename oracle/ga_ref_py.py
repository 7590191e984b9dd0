"""ctypes binding of oracle/_ref/libga_ref.so -- TEST INFRASTRUCTURE ONLY.

libga_ref.so holds the reference's OWN translation units (collision_checking.cpp, a_star.cpp,
vehicle.cpp, util1/2.cpp, transforms.cpp, VoronoiDiagramGenerator.cpp), compiled unmodified from
/root/reference by `make -C oracle ref` behind the declaration shims in oracle/ref_shim/.  It pins
the plain-C oracle (tests/test_ref_pins_oracle.py), generates tests/golden/ref_golden.npz and is
the "reference" CPU arm of bench.py.  The reference keeps all state in static class members, so
this module is ONE planner per process, exactly like the reference."""
import ctypes as C
import os
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libga_ref.so")
_lib = None
_cfg_file = None

GRID_DTYPE = {0: np.uint8, 1: np.uint8, 2: np.uint8, 3: np.uint8, 4: np.float64, 5: np.float64, 6: np.float64,
              7: np.float32, 8: np.float32}
PATCH, SAFETY, ASTAR, GEOFENCE, MOVEMENT_COST, H_PROX, MOTION_RES, OBSTACLE_DIST, VORONOI_DIST = range(9)

# the keys of LIB/config/config.yml that AStar::initialize / CollisionChecker::initialize read
# (a_star.cpp:15-38, collision_checking.cpp:14-26), shipped values
DEFAULT_CONFIG = {
    "YAW_RES_COLL": 3, "SAFETY_DISTANCE_M": 0.1, "SEARCH_DIST": 5.0, "MIN_THRESH": 100, "MAX_THRESH": 130,
    "LEN_PER_DISK": 2.0, "DOUBLE_DISK_ROWS": False, "MAX_PATCH_INS_DIST": 40,
    "ASTAR_UNKNOWN_COST": 0, "HEURISTIC_EARLY_EXIT": True, "MAX_EXTRA_NODES_ASTAR": 10000,
    "MOTION_RES_MIN": 0.1, "MOTION_RES_MAX": 1.0, "MIN_DIST_VAL": 3, "MAX_DIST_VAL": 10,
}


def available():
    return os.path.exists(LIB)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libga_ref.so is not built (make -C oracle ref needs /root/reference)")
        _lib = C.CDLL(LIB, mode=C.RTLD_GLOBAL)
        L = _lib
        L.gar_last_error.restype = C.c_char_p
        L.gar_disk_r.restype = C.c_double
        L.gar_init.argtypes = [C.c_char_p, C.c_int] + [C.c_double] * 5 + [C.c_int] + [C.c_double] * 6
        L.gar_reinit.argtypes = [C.c_int]
        L.gar_disk_centers.argtypes = [C.c_void_p]
        L.gar_dil_kernel.argtypes = [C.c_void_p]
        L.gar_pass_local_map.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.gar_get_grid.argtypes = [C.c_int, C.c_void_p, C.c_size_t]
        L.gar_set_grid.argtypes = [C.c_int, C.c_void_p, C.c_size_t]
        L.gar_set_heuristic_config.argtypes = [C.c_int, C.c_uint, C.c_double]
        L.gar_calc_distance_heuristic.argtypes = [C.c_int] * 6
        L.gar_dump_closed_set.argtypes = [C.c_int] + [C.c_void_p] * 4
        L.gar_nodes_near_goal.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.gar_astar_path.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.gar_calc_voronoi_potential_field.argtypes = [C.c_int, C.c_int, C.c_int]
        L.gar_yaw_idx.argtypes = [C.c_double]
        L.gar_cartograph.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.gar_cart_load_prev.argtypes = [C.c_double] * 4
        L.gar_cart_get.argtypes = [C.c_void_p, C.c_size_t]
        L.gar_con2gm.restype = C.c_double
        L.gar_raytrace.argtypes = [C.c_void_p, C.c_void_p]
        L.gar_precast_lists.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.gar_gm_last_error.restype = C.c_char_p
        L.gar_voronoi_endpoints.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_int]
        L.gar_check_grid.argtypes = [C.c_int] * 4
        L.gar_check_pose.argtypes = [C.c_double] * 3 + [C.c_int]
        L.gar_check_segments.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int, C.c_void_p]
        L.gar_move_car_some_steps.argtypes = [C.c_double] * 5 + [C.c_int, C.c_double] + [C.c_void_p] * 3 + [C.c_int]
        L.gar_turn_on_rear_axis.argtypes = [C.c_double] * 5 + [C.c_void_p, C.c_int]
        L.gar_bilinear.argtypes = [C.c_double] * 4 + [C.c_int] * 4 + [C.c_double] * 2
        L.gar_bilinear.restype = C.c_double
        L.gar_bilin_interp.argtypes = [C.c_double] * 2
        L.gar_bilin_interp.restype = C.c_double
        L.gar_prox_of_corners.argtypes = [C.c_double] * 3
        L.gar_prox_of_corners.restype = C.c_double
        L.gar_con2star.restype = C.c_double
        L.gar_valid_close_pose.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p]
        L.gar_path_prox_costs.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p]
    return _lib


def _check(rc):
    if rc < 0:
        raise RuntimeError("reference: " + lib().gar_last_error().decode())
    return rc


def init(patch_dim, veh=(0.55, 3.165, 3.993, 1.17, 2.13, False), weights=(5.0, 5.0, 3.0), vor=(100.0, 7.0, 3.0),
         config=None):
    """HybridAStar::initialize's share for this path + create_patch's reinit.
    veh = (max_steer, wb, lf, lb, width, is_ushift); weights = (movement, prox, lane) costs;
    vor = (alpha, do_max, do_min)."""
    global _cfg_file
    cfg = dict(DEFAULT_CONFIG)
    cfg.update(config or {})
    f = tempfile.NamedTemporaryFile("w", suffix=".yml", delete=False)
    for k, v in cfg.items():
        f.write("%s: %s\n" % (k, v))
    f.close()
    _cfg_file = f.name
    max_steer, wb, lf, lb, width, ushift = veh
    _check(lib().gar_init(_cfg_file.encode(), int(patch_dim), max_steer, wb, lf, lb, width, int(ushift),
                          weights[0], weights[1], weights[2], vor[0], vor[1], vor[2]))
    os.unlink(_cfg_file)


def reinit(patch_dim):
    _check(lib().gar_reinit(int(patch_dim)))


def dims():
    L = lib()
    return L.gar_patch_dim(), L.gar_astar_dim()


def disks():
    L = lib()
    nd, ny = L.gar_nb_disks(), L.gar_nb_yaws()
    cen = np.zeros((ny, nd, 2), np.int32)
    L.gar_disk_centers(_p(cen))
    return dict(nb_disks=nd, nb_yaws=ny, disk_r=L.gar_disk_r(), disk_r_c=L.gar_disk_r_c(), centers=cen)


def dil_kernel():
    r = lib().gar_disk_r_c()
    k = np.zeros((2 * r + 1, 2 * r + 1), np.uint8)
    lib().gar_dil_kernel(_p(k))
    return k


def _shape(gid):
    pd, ad = dims()
    return (pd, pd) if gid < 2 else ((ad, ad) if gid < 7 else (200, 200))


def get_grid(gid):
    out = np.empty(_shape(gid), GRID_DTYPE[gid])
    if lib().gar_get_grid(gid, _p(out), out.nbytes) != 0:
        raise RuntimeError("reference: bad grid id / size")
    return out


def set_grid(gid, arr):
    arr = np.ascontiguousarray(arr, GRID_DTYPE[gid])
    assert arr.shape == _shape(gid), (arr.shape, _shape(gid))
    if lib().gar_set_grid(gid, _p(arr), arr.nbytes) != 0:
        raise RuntimeError("reference: bad grid id / size")


def pass_local_map(win, ox, oy):
    win = np.ascontiguousarray(win, np.uint8)
    _check(lib().gar_pass_local_map(_p(win), int(ox), int(oy), win.shape[0]))


def process_safety_patch():
    _check(lib().gar_process_safety_patch())


def calc_astar_grid():
    _check(lib().gar_calc_astar_grid())


def set_heuristic_config(early_exit=True, max_extra_nodes=10000, unknown_cost_w=0.0):
    lib().gar_set_heuristic_config(int(early_exit), int(max_extra_nodes), float(unknown_cost_w))


def calc_distance_heuristic(goal, start, for_path=True, get_only_near=False):
    """AStar::calcDistanceHeuristic; returns the dense dump of the selected closed set."""
    n = _check(lib().gar_calc_distance_heuristic(int(goal[0]), int(goal[1]), int(start[0]), int(start[1]),
                                                 int(for_path), int(get_only_near)))
    _, ad = dims()
    cost, dist = np.empty((ad, ad)), np.empty((ad, ad))
    parent = np.empty((ad, ad), np.int32)
    flags = np.empty((ad, ad), np.uint8)
    k = _check(lib().gar_dump_closed_set(int(for_path), _p(cost), _p(dist), _p(parent), _p(flags)))
    assert k == n
    return dict(n_closed=n, cost=cost, dist=dist, parent=parent, flags=flags)


def nodes_near_goal(cap=4096):
    d = np.empty(cap)
    xy = np.empty((cap, 2), np.int32)
    n = lib().gar_nodes_near_goal(_p(d), _p(xy), cap)
    n = min(n, cap)
    return d[:n].copy(), xy[:n].copy()


def astar_path(x, y):
    _, ad = dims()
    xs, ys = np.empty(ad * ad, np.int32), np.empty(ad * ad, np.int32)
    n = _check(lib().gar_astar_path(int(x), int(y), _p(xs), _p(ys), ad * ad))
    return xs[:n].copy(), ys[:n].copy()


def calc_voronoi_potential_field(ego, dt_variant=0):
    _check(lib().gar_calc_voronoi_potential_field(int(ego[0]), int(ego[1]), int(dt_variant)))


def voronoi_endpoints(sites_xy, vor_dim=200, min_dist=3.0):
    """cells the reference's Fortune sweep marks for these sites (row-major (x, y) pairs)"""
    s = np.ascontiguousarray(sites_xy, np.int32).reshape(-1, 2)
    out = np.empty((vor_dim * vor_dim, 2), np.int32)
    n = _check(lib().gar_voronoi_endpoints(_p(s), len(s), vor_dim, float(min_dist), _p(out), len(out)))
    return out[:n].copy()


def cart_reset(patch_dim):
    _check(lib().gar_cart_reset(int(patch_dim)))


def cartograph(win, ox, oy):
    win = np.ascontiguousarray(win, np.uint8)
    _check(lib().gar_cartograph(_p(win), int(ox), int(oy), win.shape[0]))


def cart_pass_local_map(ox, oy, dim):
    _check(lib().gar_cart_pass_local_map(int(ox), int(oy), int(dim)))


def cart_load_prev(prev_utm, utm):
    _check(lib().gar_cart_load_prev(float(prev_utm[0]), float(prev_utm[1]), float(utm[0]), float(utm[1])))


def cart_get():
    d = lib().gar_cart_dim()
    out = np.empty((d, d), np.uint8)
    _check(lib().gar_cart_get(_p(out), out.nbytes))
    return out


def con2gm():
    return lib().gar_con2gm()


def con2star():
    return lib().gar_con2star()


def bilinear(q11, q12, q21, q22, x1, x2, y1, y2, x, y):
    """util::bilinearInterpolation (util1.cpp:152-200)"""
    return lib().gar_bilinear(q11, q12, q21, q22, x1, x2, y1, y2, x, y)


def bilin_interp(x, y):
    """util::getBilinInterp on AStar::h_prox_arr_ (set_grid(5, ...))"""
    return lib().gar_bilin_interp(x, y)


def path_prox_costs(poses_m, offsets, h_prox_cost=5.0, interp_res=0.1):
    """the prox_cost loop of getRSPathCosts / getRAPathCosts over a ragged batch; poses in metres"""
    poses_m = np.ascontiguousarray(poses_m, np.float64).reshape(-1, 3)
    offsets = np.ascontiguousarray(offsets, np.int32)
    out = np.zeros(len(offsets) - 1, np.float64)
    lib().gar_path_prox_costs(_p(poses_m), _p(offsets), len(offsets) - 1, h_prox_cost, interp_res, _p(out))
    return out


def valid_close_pose(center, near_cells, yaw_res_coll=3.0):
    """step 3 of HybridAStar::getValidClosePose composed from the reference's parts; the pose or None"""
    c = np.ascontiguousarray(center, np.float64)
    cells = np.ascontiguousarray(near_cells, np.int32)
    out = np.zeros(3, np.float64)
    ok = lib().gar_valid_close_pose(_p(c), _p(cells), len(cells), float(yaw_res_coll), _p(out))
    return out if ok else None


def prox_of_corners(x, y, yaw):
    """HybridAStar::getProxOfCorners composed from the reference's parts; pose in astar cells"""
    return lib().gar_prox_of_corners(x, y, yaw)


GM_DIM = 801  # gridmap_sim.hpp:17


def raytrace(grid_map_gt):
    """GridMapSim::raytracing on an 801 x 801 ground-truth window (0 free / 255 occupied)"""
    gt = np.ascontiguousarray(grid_map_gt, np.uint8)
    assert gt.shape == (GM_DIM, GM_DIM)
    out = np.empty((GM_DIM, GM_DIM), np.uint8)
    if lib().gar_raytrace(_p(gt), _p(out)) < 0:
        raise RuntimeError("reference: " + lib().gar_gm_last_error().decode())
    return out


def precast_lists():
    """the reference's 180 distance-sorted precast lists, flattened: (entries = ix * 801 + iy, offsets[181])"""
    off = np.zeros(181, np.int32)
    ent = np.zeros(2 * GM_DIM * GM_DIM + 4096, np.int32)
    n = lib().gar_precast_lists(_p(ent), ent.size, _p(off))
    assert n <= ent.size
    return ent[:n].copy(), off


def yaw_idx(yaw):
    return lib().gar_yaw_idx(float(yaw))


def check_grid(x, y, yaw_idx_, variant=0):
    return lib().gar_check_grid(int(x), int(y), int(yaw_idx_), int(variant))


def check_pose(x, y, yaw, variant=0):
    return lib().gar_check_pose(float(x), float(y), float(yaw), int(variant))


def check_segments(x, y, yaw, seg_off, variant=0):
    x, y, yaw = (np.ascontiguousarray(v, np.float64) for v in (x, y, yaw))
    seg = np.ascontiguousarray(seg_off, np.int32)
    hit = np.empty(len(seg) - 1, np.int32)
    lib().gar_check_segments(_p(x), _p(y), _p(yaw), _p(seg), len(seg) - 1, int(variant), _p(hit))
    return hit


def move_car_some_steps(x, y, yaw, arc_l, motion_res, direction, steer):
    xs, ys, yaws = (np.empty(512) for _ in range(3))
    n = lib().gar_move_car_some_steps(x, y, yaw, arc_l, motion_res, int(direction), steer, _p(xs), _p(ys), _p(yaws), 512)
    return xs[:n].copy(), ys[:n].copy(), yaws[:n].copy()


def turn_on_rear_axis(x, y, yaw, delta, yaw_res_deg):
    yaws = np.empty(1024)
    n = lib().gar_turn_on_rear_axis(x, y, yaw, delta, yaw_res_deg, _p(yaws), 1024)
    return yaws[:n].copy()
