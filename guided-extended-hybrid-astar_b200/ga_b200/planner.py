"""Host mirror of the reference's CollisionChecker / AStar / Vehicle interfaces.

The reference keeps all state in static class members (one planner per process,
collision_checking.hpp:50-76, a_star.hpp:38-99).  Here the state lives in a GridContext (one
ga_ctx: device arenas + one CUDA stream), and CollisionChecker / AStar are views on it that keep
the reference's method names, argument meaning and error behaviour
(wrapper_hybrid_a_star_lib.cpp:176-199, 327-346).  Every grid computation runs in
libgridaccel.so; nothing here computes a grid on the CPU.
"""
import ctypes as C
import math

import numpy as np

from . import _abi
from ._abi import GaError

GM_RES = 0.15625    # params.hpp:8
ASTAR_RES = 0.625   # params.hpp:7
CON2GM = 1 / GM_RES
CON2STAR = 1 / ASTAR_RES
TO_RAD = math.pi / 180
FREE, UNKNOWN, OCC = 0, 1, 2   # GRID_VAL, collision_checking.hpp:39-47
NB_STEER = 13                  # hybrid_a_star_lib.hpp:48


class VehConfig:
    """Default vehicle (python/path_planner_lib/data_structures.py:19-34)."""

    def __init__(self, max_steer=0.55, width=2.13, wb=3.165, lf=3.993, lb=1.17, is_ushift=False):
        self.max_steer, self.width, self.wb, self.lf, self.lb = max_steer, width, wb, lf, lb
        self.is_ushift = is_ushift


class LibConfig:
    """The keys of config/config.yml and config/params.yml the grid path reads, shipped values."""

    def __init__(self, **kw):
        self.YAW_RES_COLL = 3
        self.DOUBLE_DISK_ROWS = False
        self.LEN_PER_DISK = 2.0
        self.SAFETY_DISTANCE_M = 0.1
        self.MIN_THRESH = 100
        self.MAX_THRESH = 130
        self.ASTAR_UNKNOWN_COST = 0.0
        self.MAX_EXTRA_NODES_ASTAR = 10000
        self.HEURISTIC_EARLY_EXIT = True
        self.TURN_ON_POINT_ANGLE = 90.0
        self.YAW_RES = 10
        self.MOTION_RES_MIN = 0.1
        self.MOTION_RES_MAX = 1.0
        self.MIN_DIST_VAL = 3.0
        self.MAX_DIST_VAL = 10.0
        self.RA_FREQ = 5
        # params.yml
        self.astar_prox_cost_ = 5.0
        self.astar_movement_cost_ = 5.0
        self.astar_lane_movement_cost_ = 3.0
        self.alpha_ = 100.0
        self.do_max_ = 7.0
        self.do_min_ = 3.0
        # a_star.hpp:36; IPP float chamfer = what the reference's Docker OpenCV (WITH_IPP) runs
        self.VOR_DIM = 200
        self.DT_VARIANT = _abi.DT_IPPFLOAT
        for k, v in kw.items():
            if not hasattr(self, k):
                raise KeyError(k)
            setattr(self, k, v)

    @classmethod
    def from_yaml(cls, path2config, path2params=None):
        import yaml
        cfg = cls()
        for path in (path2config, path2params):
            if path is None:
                continue
            with open(path) as f:
                for k, v in (yaml.safe_load(f) or {}).items():
                    if hasattr(cfg, k):
                        setattr(cfg, k, v)
        return cfg


class Vehicle:
    """Vehicle::initialize (vehicle.cpp:11-32)."""

    def __init__(self, max_steer, w_b, l_f, l_b, width, is_ushift):
        self.max_steer_, self.w_b_, self.lf_, self.lb_ = max_steer, w_b, l_f, l_b
        self.length_ = l_f + l_b
        self.width_ = width
        self.is_ushift_ = is_ushift

    @classmethod
    def from_config(cls, v):
        return cls(v.max_steer, v.wb, v.lf, v.lb, v.width, v.is_ushift)

    # host rollouts with the C library's sin / cos / tan / fmod, the reference's operand order: what the device
    # results are re-checked against when a pose lies on a cell / yaw-bin edge (ga_succ.near_cell_edge)
    @staticmethod
    def _mpi_pi(a):  # util::constrainAngleMinPIPlusPi (util1.cpp:17-24)
        while a < 0:
            a += 2 * math.pi
        return math.fmod(a + math.pi, 2.0 * math.pi) - math.pi

    def move_car_some_steps(self, pose, arc_l, motion_res, direction, steer):
        """Vehicle::move_car_some_steps (vehicle.cpp:100-122) -> lists x, y, yaw"""
        n = _cround(arc_l / motion_res)
        x, y, yaw = pose
        xs, ys, yaws = [], [], []
        d = motion_res * direction
        for _ in range(n):
            x += d * math.cos(yaw)  # Vehicle::move (vehicle.cpp:34-43)
            y += d * math.sin(yaw)
            yaw += d * math.tan(steer) / self.w_b_
            yaw = self._mpi_pi(yaw)
            xs.append(x)
            ys.append(y)
            yaws.append(yaw)
        return xs, ys, yaws

    def turn_on_rear_axis(self, pose, delta_angle, yaw_res_coll_deg):
        """Vehicle::turn_on_rear_axis (vehicle.cpp:45-98) -> lists x, y, yaw"""
        x, y, yaw = pose
        goal_yaw = self._mpi_pi(yaw + delta_angle)
        res = yaw_res_coll_deg * TO_RAD
        sg = (0.0 < delta_angle) - (delta_angle < 0.0)
        yaws = []
        while True:
            yaw = self._mpi_pi(yaw + sg * res)
            yaws.append(yaw)
            diff = self._mpi_pi(self._mpi_pi(goal_yaw) - self._mpi_pi(yaw))  # util::getAngleDiff (util1.cpp:26-34)
            if not abs(diff) >= res:
                break
        if yaws[-1] != goal_yaw:
            yaws[-1] = goal_yaw
        return [x] * len(yaws), [y] * len(yaws), yaws


def _cround(v):
    """std::round: half away from zero."""
    return int(math.floor(v + 0.5)) if v >= 0 else -int(math.floor(-v + 0.5))


def calc_disks(veh, cfg):
    """CollisionChecker::calculateDisks / getDiskRadius / getDiskPositions
    (collision_checking.cpp:44-145): footprint disks and the per-yaw centre LUT, in fp64 on the
    host exactly as the reference initialises them.  Returns (nb_disks, disk_r, disk_r_c,
    centers[nb_yaws, nb_disks, 2] int32)."""
    eff_nb = int(math.ceil(veh.length_ / cfg.LEN_PER_DISK))
    mult = 2 if cfg.DOUBLE_DISK_ROWS else 1
    nb_disks = mult * eff_nb
    eff_width = veh.width_ / mult
    radius = math.sqrt(((veh.length_ * veh.length_) / (eff_nb * eff_nb) + (eff_width * eff_width)) / 4)
    dist = 2 * math.sqrt((radius * radius) - (eff_width * eff_width) / 4)
    first = dist / 2 - veh.lb_
    positions = [first + i * dist for i in range(eff_nb)]
    disk_r = radius + cfg.SAFETY_DISTANCE_M
    disk_r_c = int(math.ceil(disk_r * CON2GM))
    nb_yaws = 360 // cfg.YAW_RES_COLL
    yaw_res_rad = cfg.YAW_RES_COLL * TO_RAD
    centers = np.zeros((nb_yaws, nb_disks, 2), np.int32)
    for yi in range(nb_yaws):
        yaw = yi * yaw_res_rad
        c, s = math.cos(yaw), math.sin(yaw)
        k = 0
        for pos in positions:
            offs = (eff_width / 2.0, -eff_width / 2.0) if cfg.DOUBLE_DISK_ROWS else (0.0,)
            for yo in offs:
                rx = pos * c - yo * s
                ry = pos * s + yo * c
                centers[yi, k, 0] = _cround(rx * CON2GM)
                centers[yi, k, 1] = _cround(ry * CON2GM)
                k += 1
    return nb_disks, disk_r, disk_r_c, centers


class GridContext:
    """One planner instance = one ga_ctx."""

    def __init__(self, device=0):
        self.lib = _abi.load()
        h = C.c_void_p()
        rc = self.lib.ga_create(device, C.byref(h))
        if rc != 0:
            raise GaError("ga_create(device=%d) failed with %d: no usable CUDA device "
                          "(this backend has no CPU path)" % (device, rc))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.ga_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != 0:
            raise GaError("%s (status %d)" % (self.lib.ga_last_error(self.h).decode(), rc))

    def sync(self):
        self.check(self.lib.ga_sync(self.h))

    def stream(self):
        return self.lib.ga_stream(self.h)

    def grid_shape(self, gid):
        if gid in (_abi.GRID_PATCH, _abi.GRID_SAFETY):
            d = self.lib.ga_patch_dim(self.h)
        elif gid == _abi.GRID_CART:
            d = self.lib.ga_cart_dim(self.h)
        elif gid == _abi.GRID_PMAP:
            d = 801
        elif gid in (_abi.GRID_OBSTACLE_DIST, _abi.GRID_VORONOI_DIST):
            nb = C.c_size_t()
            self.check(self.lib.ga_grid_bytes(self.h, gid, C.byref(nb)))
            d = int(round(math.sqrt(nb.value // 4)))
        else:
            d = self.lib.ga_astar_dim(self.h)
        return (d, d)

    def download(self, gid):
        out = np.empty(self.grid_shape(gid), _abi.GRID_DTYPES[gid])
        self.check(self.lib.ga_download(self.h, gid, _abi.ptr(out), out.nbytes))
        return out

    def upload(self, gid, arr):
        a = np.ascontiguousarray(arr, _abi.GRID_DTYPES[gid])
        if a.shape != self.grid_shape(gid):
            raise GaError("grid %d expects shape %s, got %s" % (gid, self.grid_shape(gid), a.shape))
        self.check(self.lib.ga_upload(self.h, gid, _abi.ptr(a), a.nbytes))
        self.sync()  # `a` may be a temporary

    def device_ptr(self, gid):
        p = C.c_void_p()
        self.check(self.lib.ga_device_ptr(self.h, gid, C.byref(p)))
        return p.value

    # measurement
    def prof_enable(self, on=True):
        self.check(self.lib.ga_prof_enable(self.h, int(on)))

    def prof_reset(self):
        self.check(self.lib.ga_prof_reset(self.h))

    def prof_get(self, k):
        ms, n = C.c_double(), C.c_longlong()
        self.check(self.lib.ga_prof_get(self.h, k, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def launch_count(self):
        return self.lib.ga_launch_count(self.h)


class CollisionChecker:
    """Mirror of the reference's CollisionChecker (collision_checking.hpp:36-120)."""

    FREE, UNKNOWN, OCC = FREE, UNKNOWN, OCC

    def __init__(self, ctx, vehicle, cfg=None):
        self.ctx, self.vehicle, self.cfg = ctx, vehicle, cfg or LibConfig()
        self.patch_dim_ = 0
        self.params_set_ = False

    # -- initialize / calculateDisks / resetPatch (collision_checking.cpp:12-42, 75-145)
    def initialize(self, patch_dim, path2config=None):
        if path2config is not None:
            self.cfg = LibConfig.from_yaml(path2config)
        self.calculateDisks()
        self.resetPatch(patch_dim)

    def calculateDisks(self):
        cfg = self.cfg
        self.nb_disks_, self.disk_r_, self.disk_r_c_, self.disk_centers_ = calc_disks(self.vehicle, cfg)
        self.yaw_res_coll_ = cfg.YAW_RES_COLL
        self.nb_disc_yaws_ = 360 // cfg.YAW_RES_COLL
        p = _abi.CollisionParams(int(cfg.MIN_THRESH), int(cfg.MAX_THRESH), self.disk_r_c_,
                                 int(cfg.YAW_RES_COLL), self.nb_disks_, 4)
        cen = np.ascontiguousarray(self.disk_centers_, np.int32)
        self.ctx.check(self.ctx.lib.ga_set_collision_params(self.ctx.h, C.byref(p), _abi.ptr(cen)))
        self.params_set_ = True

    def resetPatch(self, patch_dim):
        self.ctx.check(self.ctx.lib.ga_patch_reset(self.ctx.h, int(patch_dim)))
        self.patch_dim_ = int(patch_dim)

    # -- passLocalMap (collision_checking.cpp:176-191, 269-313)
    def passLocalMap(self, local_map, origin, dim):
        a = np.ascontiguousarray(local_map, np.uint8)
        if a.size != dim * dim:
            raise GaError("local map must hold dim*dim cells")
        self.ctx.check(self.ctx.lib.ga_patch_insert(self.ctx.h, _abi.ptr(a), int(origin[0]), int(origin[1]), int(dim)))
        self.ctx.sync()

    passLocalMapData = passLocalMap

    # -- processSafetyPatch (collision_checking.cpp:193-224); the pooled astar grid comes with it
    def processSafetyPatch(self):
        self.ctx.check(self.ctx.lib.ga_build_safety(self.ctx.h))

    @property
    def patch_arr_(self):
        return self.ctx.download(_abi.GRID_PATCH)

    @property
    def patch_safety_arr_(self):
        return self.ctx.download(_abi.GRID_SAFETY)

    # -- checks (collision_checking.cpp:333-452)
    def getYawIdx(self, yaw):
        a = yaw
        while a < 0:
            a += 2 * math.pi
        a = math.fmod(a, 2 * math.pi)
        return min(int(a / (self.yaw_res_coll_ * TO_RAD)), self.nb_disc_yaws_ - 1)

    def returnDiskPositions(self, yaw):
        return [tuple(int(v) for v in c) for c in self.disk_centers_[self.getYawIdx(yaw)]]

    def _check(self, x, y, yaw, seg_off, variant):
        x = np.ascontiguousarray(x, np.float64)
        y = np.ascontiguousarray(y, np.float64)
        yaw = np.ascontiguousarray(yaw, np.float64)
        if not (x.shape == y.shape == yaw.shape and x.ndim == 1):
            raise GaError("x, y, yaw must be 1-D and of equal length")
        seg = np.ascontiguousarray(seg_off, np.int32)
        hit = np.empty(len(seg) - 1, np.int32)
        self.ctx.check(self.ctx.lib.ga_check_poses(self.ctx.h, _abi.ptr(x), _abi.ptr(y), _abi.ptr(yaw), len(x),
                                                   _abi.ptr(seg), len(seg) - 1, int(variant), _abi.ptr(hit)))
        return hit

    def checkPose(self, pose):
        return bool(self._check([pose[0]], [pose[1]], [pose[2]], [0, 1], 0)[0] < 0)

    def checkPoseVariant(self, pose):
        return bool(self._check([pose[0]], [pose[1]], [pose[2]], [0, 1], 1)[0] < 0)

    def checkGrid(self, pose):
        """pose = (x, y, yaw_idx) in gm cells / yaw bins (collision_checking.cpp:344-362)."""
        return self.checkPose(((pose[0] + 0.5) * GM_RES, (pose[1] + 0.5) * GM_RES,
                               (pose[2] + 0.5) * self.yaw_res_coll_ * TO_RAD)) if pose[0] >= 0 and pose[1] >= 0 \
            else False

    checkGridVariant = checkGrid  # the pybind name is bound to checkGrid (wrapper:193)

    def getPathCollisionIndex(self, x_list, y_list, yaw_list):
        return int(self._check(x_list, y_list, yaw_list, [0, len(x_list)], 0)[0])

    def checkPathCollision(self, x_list, y_list, yaw_list):
        return self.getPathCollisionIndex(x_list, y_list, yaw_list) < 0

    # batched forms (one launch per expansion batch)
    def checkSegments(self, x, y, yaw, seg_off, variant=0):
        return self._check(x, y, yaw, seg_off, variant)

    def setExpandParams(self, arc_l=ASTAR_RES * 1.5, nb_steer=NB_STEER):
        cfg, v = self.cfg, self.vehicle
        p = _abi.ExpandParams(v.max_steer_, v.w_b_, arc_l, float(cfg.YAW_RES), float(cfg.TURN_ON_POINT_ANGLE), nb_steer)
        self.ctx.check(self.ctx.lib.ga_set_expand_params(self.ctx.h, C.byref(p)))
        self.expand_stride_ = self.ctx.lib.ga_expand_stride(self.ctx.h)
        self.expand_params_ = dict(arc_l=arc_l, nb_steer=nb_steer, yaw_res=float(cfg.YAW_RES),
                                   turn_on_point_angle=float(cfg.TURN_ON_POINT_ANGLE))

    def expandNodes(self, nodes, exact=True):
        """nodes: structured array _abi.NODE_DTYPE -> successors [n_nodes, stride] _abi.SUCC_DTYPE.
        exact: primitives the device flags near_cell_edge (a pose within 1e-9 of a cell / yaw-bin boundary, where the
        device's libm and glibc may round to different cells) are rolled out again on the host with the reference's
        arithmetic and re-checked from those poses, so that valid / xi / yi / yawi equal the reference's for EVERY
        primitive (Vehicle::move_car_some_steps / turn_on_rear_axis, vehicle.cpp:45-122)."""
        nodes = np.ascontiguousarray(nodes, _abi.NODE_DTYPE)
        out = np.zeros((len(nodes), self.expand_stride_), _abi.SUCC_DTYPE)
        self.ctx.check(self.ctx.lib.ga_expand(self.ctx.h, _abi.ptr(nodes), len(nodes), _abi.ptr(out)))
        if exact:
            self._recheck_flagged(nodes, out)
        return out

    def _recheck_flagged(self, nodes, out):
        flagged = np.argwhere((out["near_cell_edge"] != 0) & (out["n_poses"] > 0))
        if len(flagged) == 0:
            return
        ep = self.expand_params_
        n_steer = ep["nb_steer"]
        delta_steer = (2 * self.vehicle.max_steer_) / float(n_steer - 1)
        step = ep["turn_on_point_angle"] * TO_RAD
        ra, d = [], -2 * math.pi + step
        while d <= 2 * math.pi - step:  # the accumulating loop of setNeighbors (hybrid_a_star_lib.cpp:519-531)
            if d != 0:
                ra.append(d)
            d += step
        xs, ys, yaws, seg = [], [], [], [0]
        for i, k in flagged:
            nd = nodes[i]
            pose = (float(nd["x"]), float(nd["y"]), float(nd["yaw"]))
            if k < 2 * n_steer:
                steer = (k // 2 - n_steer // 2) * delta_steer
                lx, ly, lyaw = self.vehicle.move_car_some_steps(pose, ep["arc_l"], float(nd["motion_res"]), -1 if k & 1 else 1, steer)
            else:
                lx, ly, lyaw = self.vehicle.turn_on_rear_axis(pose, ra[k - 2 * n_steer], float(self.cfg.YAW_RES_COLL))
            xs += lx
            ys += ly
            yaws += lyaw
            seg.append(len(xs))
        hits = self.checkSegments(np.array(xs), np.array(ys), np.array(yaws), np.array(seg, np.int32))
        adim = self.ctx.lib.ga_astar_dim(self.ctx.h)
        inv = 1.0 / (ep["yaw_res"] * TO_RAD)
        for j, (i, k) in enumerate(flagged):
            x, y, yaw = xs[seg[j + 1] - 1], ys[seg[j + 1] - 1], yaws[seg[j + 1] - 1]
            s = out[i, k]
            s["x"], s["y"], s["yaw"] = x, y, yaw
            s["xi"], s["yi"], s["yawi"] = int(x * CON2STAR), int(y * CON2STAR), _cround(yaw * inv)  # mapCont2Disc
            ok = hits[j] < 0
            on_grid = (k >= 2 * n_steer) or (-1 < s["xi"] < adim and -1 < s["yi"] < adim)
            s["valid"] = 1 if (ok and on_grid) else 0
            s["collision_free"] = 1 if ok else 0
            s["near_cell_edge"] = 0
            out[i, k] = s

    def expandNodesWithCosts(self, nodes, for_path=True):
        """ga_expand_ex: successors plus the grid reads the search makes per successor (heuristic cost_ / cost_dist_
        of closed_set_path_ or closed_set_guidance_, movement_cost_map_, motion_res_map_) and getProxOfCorners of every
        parent pose -- the planes stay on the device.  Needs HybridAStar(ctx, vehicle) (prox parameters)."""
        nodes = np.ascontiguousarray(nodes, _abi.NODE_DTYPE)
        out = np.zeros((len(nodes), self.expand_stride_), _abi.SUCC_DTYPE)
        cost = np.zeros((len(nodes), self.expand_stride_), _abi.SUCC_COST_DTYPE)
        prox = np.zeros(len(nodes))
        self.ctx.check(self.ctx.lib.ga_expand_ex(self.ctx.h, _abi.ptr(nodes), len(nodes), 1 if for_path else 0, _abi.ptr(out),
                                                 _abi.ptr(cost), _abi.ptr(prox)))
        return out, cost, prox


class Cartographing:
    """Mirror of the reference's Cartographing (LIB/include/cartographing_lib/cartographing.hpp:19-41,
    pybind: wrapper_hybrid_a_star_lib.cpp) on top of ga_cart_*: the accumulated map of the
    unknown-environment loop stays on the device, only sensor windows are uploaded."""

    def __init__(self, ctx):
        self.ctx = ctx

    def resetPatch(self, patch_dim):
        self.ctx.check(self.ctx.lib.ga_cart_reset(self.ctx.h, int(patch_dim)))

    def cartograph(self, local_map, origin, dim=None):
        w = np.ascontiguousarray(local_map, np.uint8)
        dim = w.shape[0] if dim is None else int(dim)
        self.ctx.check(self.ctx.lib.ga_cartograph(self.ctx.h, _abi.ptr(w), int(origin[0]), int(origin[1]), dim))

    def cartographRaytraced(self, origin):
        """cartograph of the window ga_raytrace left on the device (no host round trip)"""
        self.ctx.check(self.ctx.lib.ga_cartograph_raytraced(self.ctx.h, int(origin[0]), int(origin[1])))

    def passLocalMap(self, origin, dim):
        self.ctx.check(self.ctx.lib.ga_cart_pass_local_map(self.ctx.h, int(origin[0]), int(origin[1]), int(dim)))

    @staticmethod
    def _to_int(v):
        # Point<double>::toInt = std::round, half away from zero (data_structures1.hpp:57-60)
        return int(math.copysign(math.floor(abs(v) + 0.5), v))

    def loadPrevPatch(self, prev_origin_utm, origin_utm):
        c2g = 1.0 / GM_RES  # grid_tf::con2gm_
        po = [self._to_int(v * c2g) for v in prev_origin_utm]
        no = [self._to_int(v * c2g) for v in origin_utm]
        self.ctx.check(self.ctx.lib.ga_cart_load_prev(self.ctx.h, po[0], po[1], no[0], no[1]))

    def getMap(self):
        return self.ctx.download(_abi.GRID_CART)


class GridMapSim:
    """Mirror of the reference's GridMapSim (LIB/include/gridmap_sim_lib/gridmap_sim.hpp:92-112):
    the simulated lidar ray-cast on an 801 x 801 ground-truth window."""
    GM_DIM = 801

    def __init__(self, ctx):
        self.ctx = ctx

    def gmSimulation(self, grid_map_gt, fetch=True):
        gt = np.ascontiguousarray(grid_map_gt, np.uint8)
        if gt.shape != (self.GM_DIM, self.GM_DIM):
            raise ValueError("grid_map_gt must be 801 x 801 (GM_DIM)")
        out = np.empty_like(gt) if fetch else None
        self.ctx.check(self.ctx.lib.ga_raytrace(self.ctx.h, _abi.ptr(gt), _abi.ptr(out) if fetch else None))
        return out


class HybridAStar:
    """The grid reads of the reference's HybridAStar that score sampled poses
    (LIB/src/hybridastar_planning_lib/hybrid_a_star_lib.cpp:265-289, 588-594, 670-676; pybind:
    wrapper_hybrid_a_star_lib.cpp) on top of ga_prox_*: AStar::h_prox_arr_ stays on the device, a whole
    batch of analytic-expansion candidates is scored per call."""

    def __init__(self, ctx, vehicle, h_prox_cost=5.0, interp_res=0.1):
        self.ctx, self.vehicle = ctx, vehicle
        self.h_prox_cost_, self.interp_res_ = float(h_prox_cost), float(interp_res)  # params.yml:8, config.yml:22
        p = _abi.ProxParams(vehicle.lf_, vehicle.lb_, vehicle.width_, CON2STAR, self.h_prox_cost_, self.interp_res_)
        ctx.check(ctx.lib.ga_set_prox_params(ctx.h, C.byref(p)))

    def getProxOfCornersBatch(self, poses):
        """poses [n, 3] (x, y, yaw) in metres / rad -> getProxOfCorners(pose * con2star_) per pose"""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
        out = np.zeros(len(poses), np.float64)
        self.ctx.check(self.ctx.lib.ga_prox_of_corners(self.ctx.h, _abi.ptr(poses), len(poses), _abi.ptr(out)))
        return out

    def getProxOfCorners(self, pose):
        return float(self.getProxOfCornersBatch([pose])[0])

    def getValidClosePose(self, center_pose, near_cells, ego_pose=None):
        """step 3 of HybridAStar::getValidClosePose (hybrid_a_star_lib.cpp:1604-1668): the cheapest
        collision-free pose around center_pose whose astar cell is in near_cells (nodes_near_goal_ as
        cell indices y * astar_dim + x); ego_pose (the reference's default) when there is none"""
        c = np.ascontiguousarray(center_pose, np.float64)
        cells = np.ascontiguousarray(near_cells, np.int32)
        out, found = np.zeros(3, np.float64), C.c_int()
        self.ctx.check(self.ctx.lib.ga_valid_close_pose(self.ctx.h, _abi.ptr(c), _abi.ptr(cells), len(cells), _abi.ptr(out),
                                                        C.byref(found)))
        return out if found.value else ego_pose

    def getPathProxCosts(self, poses, path_offsets):
        """the prox_cost term of getRSPathCosts / getRAPathCosts for every path of a ragged batch"""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
        off = np.ascontiguousarray(path_offsets, np.int32)
        if len(off) < 1 or off[-1] != len(poses):
            raise ValueError("path_offsets must end at len(poses)")
        out = np.zeros(len(off) - 1, np.float64)
        self.ctx.check(self.ctx.lib.ga_path_prox_costs(self.ctx.h, _abi.ptr(poses), _abi.ptr(off), len(off) - 1, _abi.ptr(out)))
        return out


def unpack_heur_records(rec):
    """GA_GRID_HEUR_REC (12 B per cell) -> the four result planes (cost, cost_dist, parent, flags), bit for bit what
    GA_GRID_HEUR_COST / _DIST / _PARENT / _FLAGS hold after a device build."""
    rec = np.ascontiguousarray(rec)
    dim = rec.shape[0]
    cost = (rec["cost_lo"].astype(np.uint64) | (rec["cost_hi"].astype(np.uint64) << np.uint64(32))).view(np.float64)
    w = rec["w"]
    closed = np.isfinite(cost)
    straight, diag = ((w >> 4) & 16383).astype(np.float64), ((w >> 18) & 16383).astype(np.float64)
    dist = np.where(closed, straight * 0.625 + diag * (0.625 * math.sqrt(2.0)), np.inf)   # finish C: two products, one sum
    # motion_ (a_star.hpp:50-52): parent = cell - motion_[dir]
    mdx = np.array([1, 0, -1, 0, -1, -1, 1, 1]); mdy = np.array([0, 1, 0, -1, -1, 1, -1, 1])
    d = (w & 7).astype(np.int64)
    ys, xs = np.mgrid[0:dim, 0:dim]
    parent = ((ys - mdy[d]) * dim + (xs - mdx[d])).astype(np.int32)
    root = closed & (((w >> 4) & 0x0FFFFFFF) == 0)
    parent[root] = -1
    parent[~closed] = -2
    flags = (closed.astype(np.uint8) * _abi.H_CLOSED) | (((w >> 3) & 1).astype(np.uint8) * _abi.H_UNKNOWN * closed)
    return cost, dist, parent, flags.astype(np.uint8)


class NodeDisc:
    """data_structures2.hpp:47-67"""
    __slots__ = ("pos", "cost_", "cost_dist_", "parent_index_", "is_unknown_")

    def __init__(self, pos, cost, cost_dist, parent, unk):
        self.pos, self.cost_, self.cost_dist_, self.parent_index_, self.is_unknown_ = pos, cost, cost_dist, parent, unk


class ClosedSet:
    """Read API of std::unordered_map<size_t, NodeDisc> (closed_set_guidance_ / closed_set_path_,
    a_star.hpp:69-72) over the dense result planes."""

    def __init__(self, cost, dist, parent, flags):
        self.cost, self.dist, self.parent, self.flags = cost, dist, parent, flags
        self.dim = cost.shape[0]

    def contains(self, idx):
        return 0 <= idx < self.flags.size and bool(self.flags.flat[idx] & _abi.H_CLOSED)

    __contains__ = contains

    def at(self, idx):
        if not self.contains(idx):
            raise KeyError(idx)
        return NodeDisc((idx % self.dim, idx // self.dim), float(self.cost.flat[idx]), float(self.dist.flat[idx]),
                        int(self.parent.flat[idx]), bool(self.flags.flat[idx] & _abi.H_UNKNOWN))

    __getitem__ = at

    def size(self):
        return int(np.count_nonzero(self.flags & _abi.H_CLOSED))

    __len__ = size

    def keys(self):
        return np.flatnonzero(self.flags.ravel() & _abi.H_CLOSED)


def heuristic_reference_order_host(cfg, grid, fence, mc, prox, goal, start, get_only_near=False):
    """ga_heur_reforder_host: the library's GA_HEUR_REFORDER replay on host planes (no device)."""
    lib = _abi.load()
    d = grid.shape[0]
    ap = _abi.AstarParams(cfg.astar_movement_cost_, cfg.astar_prox_cost_, cfg.astar_lane_movement_cost_,
                          float(cfg.ASTAR_UNKNOWN_COST), cfg.alpha_, cfg.do_max_, cfg.do_min_, cfg.MOTION_RES_MIN,
                          cfg.MOTION_RES_MAX, float(cfg.MIN_DIST_VAL), float(cfg.MAX_DIST_VAL), int(cfg.VOR_DIM),
                          int(cfg.DT_VARIANT))
    flags = _abi.HEUR_REFORDER | (_abi.HEUR_EARLY_EXIT if cfg.HEURISTIC_EARLY_EXIT else 0) | \
        (_abi.HEUR_ONLY_NEAR if get_only_near else 0)
    q = _abi.HeurQuery(int(goal[0]), int(goal[1]), int(start[0]), int(start[1]), flags, int(cfg.MAX_EXTRA_NODES_ASTAR))
    g, f = np.ascontiguousarray(grid, np.uint8), np.ascontiguousarray(fence, np.uint8)
    m, p = np.ascontiguousarray(mc, np.float64), np.ascontiguousarray(prox, np.float64)
    cost, dist = np.empty((d, d)), np.empty((d, d))
    parent, fl, order = np.empty((d, d), np.int32), np.empty((d, d), np.uint8), np.empty(d * d, np.int32)
    n = C.c_int()
    rc = lib.ga_heur_reforder_host(C.byref(ap), d, _abi.ptr(g), _abi.ptr(f), _abi.ptr(m), _abi.ptr(p), C.byref(q),
                                   _abi.ptr(cost), _abi.ptr(dist), _abi.ptr(parent), _abi.ptr(fl), _abi.ptr(order), C.byref(n))
    if rc != 0:
        raise _abi.GaError("ga_heur_reforder_host failed: %d" % rc)
    return dict(n_closed=n.value, cost=cost, dist=dist, parent=parent, flags=fl, order=order[:n.value].copy())


class AStar:
    """Mirror of the reference's AStar grid builder (a_star.hpp:32-140)."""

    def __init__(self, ctx, cfg=None):
        self.ctx, self.cfg = ctx, cfg or LibConfig()
        self.closed_sets_ = [None, None]
        self.set_params()

    def set_params(self):
        c = self.cfg
        p = _abi.AstarParams(c.astar_movement_cost_, c.astar_prox_cost_, c.astar_lane_movement_cost_,
                             float(c.ASTAR_UNKNOWN_COST), c.alpha_, c.do_max_, c.do_min_, c.MOTION_RES_MIN,
                             c.MOTION_RES_MAX, float(c.MIN_DIST_VAL), float(c.MAX_DIST_VAL), int(c.VOR_DIM),
                             int(c.DT_VARIANT))
        self.ctx.check(self.ctx.lib.ga_set_astar_params(self.ctx.h, C.byref(p)))

    @property
    def astar_dim_(self):
        return self.ctx.lib.ga_astar_dim(self.ctx.h)

    # grids (a_star.hpp:73-99): value copies, like Vec2DFlat::getNumpyArr
    astar_grid_ = property(lambda s: s.ctx.download(_abi.GRID_ASTAR))
    h_prox_arr_ = property(lambda s: s.ctx.download(_abi.GRID_H_PROX))
    motion_res_map_ = property(lambda s: s.ctx.download(_abi.GRID_MOTION_RES))
    movement_cost_map_ = property(lambda s: s.ctx.download(_abi.GRID_MOVEMENT_COST))
    restr_geofence_grid_ = property(lambda s: s.ctx.download(_abi.GRID_GEOFENCE))
    obstacle_dist_ = property(lambda s: s.ctx.download(_abi.GRID_OBSTACLE_DIST))
    voronoi_dist_ = property(lambda s: s.ctx.download(_abi.GRID_VORONOI_DIST))

    def calcAstarGridCuda(self):
        """Pooling::execute is fused into processSafetyPatch's kernel (ga_build_safety): the pooled
        grid is already current.  Kept for call-order compatibility (a_star.cpp:536-542)."""
        return None

    def calcIndex(self, x, y):
        return y * self.astar_dim_ + x  # a_star.cpp:303-316

    def potentialSites(self, ego_index):
        V = self.cfg.VOR_DIM
        xy = np.empty((V * V, 2), np.int32)
        n = C.c_int()
        self.ctx.check(self.ctx.lib.ga_potential_sites(self.ctx.h, int(ego_index[0]), int(ego_index[1]), _abi.ptr(xy),
                                                       V * V, C.byref(n)))
        return xy[:n.value].copy()

    def calcVoronoiPotentialField(self, ego_index, voronoi_endpoints):
        """a_star.cpp:440-479.  voronoi_endpoints: [n,2] window cells marked by the host-side Fortune
        sweep (generateVoronoi_wrapper) for the sites potentialSites() returned."""
        v = np.ascontiguousarray(voronoi_endpoints, np.int32).reshape(-1, 2)
        had = C.c_int()
        self.ctx.check(self.ctx.lib.ga_build_potential(self.ctx.h, int(ego_index[0]), int(ego_index[1]),
                                                       _abi.ptr(v) if len(v) else None, len(v), C.byref(had)))
        return bool(had.value)

    def astar_params(self):
        c = self.cfg
        return _abi.AstarParams(c.astar_movement_cost_, c.astar_prox_cost_, c.astar_lane_movement_cost_,
                                float(c.ASTAR_UNKNOWN_COST), c.alpha_, c.do_max_, c.do_min_, c.MOTION_RES_MIN,
                                c.MOTION_RES_MAX, float(c.MIN_DIST_VAL), float(c.MAX_DIST_VAL), int(c.VOR_DIM),
                                int(c.DT_VARIANT))

    def calcDistanceHeuristic(self, goal_pos, start_pos=None, for_path=True, get_only_near=False, reference_order=None):
        """a_star.cpp:110-295.  reference_order=False (default without a start cell): the GPU fixed point --
        every cell reachable from the goal is closed, the start cell is not needed.  reference_order=True
        (default when start_pos is given): the reference's own closing order with its early exit
        (cfg.HEURISTIC_EARLY_EXIT, cfg.MAX_EXTRA_NODES_ASTAR) and get_only_near -- bit-exact planes and
        nodes_near_goal_ (GA_HEUR_REFORDER, replayed on the host by the library)."""
        slot = 1 if for_path else 0
        n = C.c_int()
        if reference_order is None:
            reference_order = start_pos is not None
        if not reference_order:
            if get_only_near and start_pos is None:
                raise ValueError("get_only_near needs the start cell")
            self.ctx.check(self.ctx.lib.ga_build_heuristic(self.ctx.h, slot, int(goal_pos[0]), int(goal_pos[1]), C.byref(n)))
        else:
            if start_pos is None:
                raise ValueError("the reference's closing order depends on the start cell")
            flags = _abi.HEUR_REFORDER | (_abi.HEUR_EARLY_EXIT if self.cfg.HEURISTIC_EARLY_EXIT else 0) | \
                (_abi.HEUR_ONLY_NEAR if get_only_near else 0)
            q = _abi.HeurQuery(int(goal_pos[0]), int(goal_pos[1]), int(start_pos[0]), int(start_pos[1]), flags,
                               int(self.cfg.MAX_EXTRA_NODES_ASTAR))
            self.ctx.check(self.ctx.lib.ga_build_heuristic_ex(self.ctx.h, slot, C.byref(q), C.byref(n)))
        self.closed_sets_[slot] = None
        self.near_query_ = (slot, start_pos) if get_only_near else None
        return n.value

    @property
    def nodes_near_goal_(self):
        """a_star.hpp:82: (dist, cell id) of the closed cells within 10 cells of the start of the last
        get_only_near build."""
        if getattr(self, "near_query_", None) is None:
            return []
        slot, start = self.near_query_
        ids, dist, n = np.zeros(400, np.int32), np.zeros(400), C.c_int()
        self.ctx.check(self.ctx.lib.ga_heur_near_nodes(self.ctx.h, slot, int(start[0]), int(start[1]), _abi.ptr(ids),
                                                       _abi.ptr(dist), 400, C.byref(n)))
        return list(zip(dist[:n.value].tolist(), ids[:n.value].tolist()))

    def closing_order(self):
        cap = self.astar_dim_ ** 2
        ids, n = np.zeros(cap, np.int32), C.c_int()
        self.ctx.check(self.ctx.lib.ga_heur_closing_order(self.ctx.h, _abi.ptr(ids), cap, C.byref(n)))
        return ids[:n.value].copy()

    @staticmethod
    def calcDistanceHeuristicBatch(astars, goals, for_path=True):
        """calcDistanceHeuristic for a batch of independent planners (one AStar / context each, same
        device) as one kernel chain (ga_build_heuristic_batch); returns the closed-cell counts."""
        slot = 1 if for_path else 0
        n = len(astars)
        hs = (C.c_void_p * n)(*[a.ctx.h for a in astars])
        g = np.ascontiguousarray(goals, np.int32).reshape(n, 2)
        out = np.zeros(n, np.int32)
        lead = astars[0].ctx
        lead.check(lead.lib.ga_build_heuristic_batch(hs, n, slot, g.ctypes.data_as(C.POINTER(C.c_int)),
                                                     out.ctypes.data_as(C.POINTER(C.c_int))))
        for a in astars:
            a.closed_sets_[slot] = None
        return out

    def _closed(self, slot):
        if self.closed_sets_[slot] is None:
            d = self.ctx.download
            self.closed_sets_[slot] = ClosedSet(d(_abi.GRID_HEUR_COST + slot), d(_abi.GRID_HEUR_DIST + slot),
                                                d(_abi.GRID_HEUR_PARENT + slot), d(_abi.GRID_HEUR_FLAGS + slot))
        return self.closed_sets_[slot]

    closed_set_guidance_ = property(lambda s: s._closed(0))
    closed_set_path_ = property(lambda s: s._closed(1))

    def findValidNeighborIndex(self, index, cs=None):
        cs = cs or self.closed_set_guidance_
        for b in (index + 1, index - 1, index + self.astar_dim_, index - self.astar_dim_):  # a_star.cpp:351-378
            if cs.contains(b):
                return b
        return -1

    def getAstarPath(self, x_index, y_index):
        """a_star.cpp:386-421: parent walk in closed_set_guidance_."""
        cs = self.closed_set_guidance_
        idx = self.calcIndex(x_index, y_index)
        if not cs.contains(idx):
            idx = self.findValidNeighborIndex(idx, cs)
            if idx == -1:
                return [], []
        xs, ys = [], []
        par = cs.parent.ravel()
        while idx != -1:
            xs.append(idx % cs.dim)
            ys.append(idx // cs.dim)
            idx = int(par[idx])
        return xs, ys
