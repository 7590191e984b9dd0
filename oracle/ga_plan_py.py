"""ctypes driver of the plan-level harness (oracle/ref_shim/plan_api.cpp) -- TEST / BENCH INFRASTRUCTURE ONLY.

Two libraries export the same gap_* entry points (oracle/Makefile):
  arm "ref"   oracle/_ref/libga_ref.so   -- the reference's serial Hybrid A* on its own AStar / CollisionChecker
  arm "b200"  oracle/_ref/libga_planb.so -- the SAME hybrid_a_star_lib.cpp on the drop-in classes over libgridaccel.so
The reference keeps all planner state in static members, so ONE planner per process: use run_query_subprocess()
(a fresh process per query) or import this module in a process that only ever loads one arm.

CLI:  python oracle/ga_plan_py.py --arm ref|b200 --scenario Valet_parking_015 --out result.npz
      python oracle/ga_plan_py.py --arm ref --synthetic 2048 --seed 50000 --queries 4 --out result.npz
"""
import argparse
import ctypes as C
import hashlib
import os
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIBS = {"ref": os.path.join(HERE, "_ref", "libga_ref.so"), "b200": os.path.join(HERE, "_ref", "libga_planb.so")}
GOLDEN = os.path.join(ROOT, "tests", "golden")

GM_RES = 0.15625
ASTAR_RES = 0.625

# LIB/config/config.yml, shipped values (every key HybridAStar / AStar / CollisionChecker::initialize read)
CONFIG = {
    "ONLY_FORWARD": False, "IS_USHIFT": False, "COMFORT_BRAKE_ACC": 1, "WAYPOINT_TYPE": 0, "MAX_DIST4WAYPOINTS": 50,
    "WAYPOINT_DIST": 40, "INTERP_RES": 0.1, "MIN_THRESH": 100, "MAX_THRESH": 130, "MAX_PATCH_INS_DIST": 40,
    "PATCH_PADDING_DIM": 100, "YAW_RES_COLL": 3, "DOUBLE_DISK_ROWS": False, "LEN_PER_DISK": 2.0,
    "SAFETY_DISTANCE_M": 0.1, "SEARCH_DIST": 5.0, "ASTAR_UNKNOWN_COST": 0, "MAX_EXTRA_NODES_ASTAR": 10000,
    "HEURISTIC_EARLY_EXIT": True, "TURN_ON_POINT_ANGLE": 90, "YAW_RES": 10, "MOTION_RES_MIN": 0.1,
    "MOTION_RES_MAX": 1.0, "MIN_DIST_VAL": 3, "MAX_DIST_VAL": 10, "EXTRA_STEER_COST_ANALYTIC": 5,
    "REAR_AXIS_COST": 70, "TURN_ON_POINT_HORIZON": 15, "TIMEOUT": 2000, "NON_H_NO_OBS_PATCH_DIM": 101, "RA_FREQ": 5,
    "MAX_EXTRA_NODES_HASTAR": 50, "DIST_THRESH_ANALYTIC_M": 10, "RS_2ND_STEER": 0.5, "APPROX_GOAL_DIST": 2,
    "APPROX_GOAL_ANGLE": 25,
}
# LIB/config/params.yml in gap_init's order
PARAMS12 = [100.0, 7.0, 3.0, 5.0, 5.0, 3.0, 20.0, 1.0, 3.0, 5.0, 10.0, 0.9]
# VehConfig defaults (python/path_planner_lib/data_structures.py:19-34)
VEHICLE = dict(max_steer=0.55, wb=3.165, lf=3.993, lb=1.17, width=2.13)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _share_dir(cfg):
    """lib_share_dir with config/config.yml written from cfg and a persistent data/ (nonh_noobs.data cache)."""
    text = "".join("%s: %s\n" % (k, v) for k, v in cfg.items())
    tag = hashlib.sha1(text.encode()).hexdigest()[:12]
    base = os.path.join(HERE, "_ref", "share")
    d = os.path.join(base, tag)
    os.makedirs(os.path.join(d, "config"), exist_ok=True)
    os.makedirs(os.path.join(base, "data"), exist_ok=True)
    link = os.path.join(d, "data")
    if not os.path.exists(link):
        try:
            os.symlink(os.path.join("..", "data"), link)
        except FileExistsError:
            pass
    path = os.path.join(d, "config", "config.yml")
    if not os.path.exists(path):
        tmp = path + ".%d" % os.getpid()
        with open(tmp, "w") as f:
            f.write(text)
        os.replace(tmp, path)
    return d


class Planner:
    """One planner process-wide (static state in the library)."""

    def __init__(self, arm, is_ushift=False, timeout_ms=10**9, config=None, is_sim=True, dt_variant=1):
        self.arm = arm
        path = LIBS[arm]
        if not os.path.exists(path):
            raise RuntimeError("%s is not built (make -C oracle %s)" % (path, "ref" if arm == "ref" else "planb"))
        self.L = L = C.CDLL(path, mode=C.RTLD_GLOBAL)
        L.gap_last_error.restype = C.c_char_p
        L.gap_init.argtypes = [C.c_char_p] + [C.c_double] * 5 + [C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.gap_create_patch.argtypes = [C.c_double] * 5 + [C.c_void_p]
        L.gap_reinit.argtypes = [C.c_double, C.c_double, C.c_int]
        L.gap_insert_window_sim.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double]
        L.gap_pass_local_map.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.gap_check_pose.argtypes = [C.c_double] * 3
        L.gap_recalc_env.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.gap_plan.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.gap_get_path.argtypes = [C.c_void_p] * 6
        L.gap_get_trace.argtypes = [C.c_void_p, C.c_void_p]
        L.gap_get_expanded.argtypes = [C.c_void_p]
        L.gap_get_closed.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.gap_get_heuristic.argtypes = [C.c_int] + [C.c_void_p] * 4
        L.gap_get_map.argtypes = [C.c_int, C.c_void_p]
        if arm == "b200":
            L.gap_b200_set_mode.argtypes = [C.c_int]
            L.gap_b200_set_lookahead.argtypes = [C.c_int]
            L.gap_b200_stats.argtypes = [C.c_void_p, C.c_void_p]
        cfg = dict(CONFIG)
        cfg.update(config or {})
        cfg["IS_USHIFT"] = bool(is_ushift)
        self.cfg = cfg
        p = np.array(PARAMS12, np.float64)
        v = VEHICLE
        t0 = time.perf_counter()
        self._ck(L.gap_init(_share_dir(cfg).encode(), v["max_steer"], v["wb"], v["lf"], v["lb"], v["width"],
                            int(bool(is_ushift)), _p(p), int(min(timeout_ms, 2**31 - 1)), int(is_sim)))
        self.init_s = time.perf_counter() - t0
        # cv::distanceTransform arithmetic: 1 = IPP float (OpenCV built WITH_IPP, the reference's Docker image), 0 = 16.16
        L.gap_set_dt_variant.argtypes = [C.c_int]
        self._ck(L.gap_set_dt_variant(int(dt_variant)))
        self.origin = np.zeros(2)
        self.patch_dim = 0

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError("gap: " + self.L.gap_last_error().decode())

    def set_heuristic_mode(self, mode):
        """arm b200 only: 0 = the reference's pop order (bit-exact sequences), 1 = GPU fixed point."""
        if self.arm == "b200":
            self._ck(self.L.gap_b200_set_mode(int(mode)))

    def create_patch(self, ego_utm, goal_utm):
        out = np.zeros(3)
        self._ck(self.L.gap_create_patch(ego_utm[0], ego_utm[1], goal_utm[0], goal_utm[1],
                                         float(self.cfg["PATCH_PADDING_DIM"]), _p(out)))
        self.origin = out[:2].copy()
        self.patch_dim = int(out[2])
        return self.origin, self.patch_dim

    def reinit(self, origin, patch_dim):
        self._ck(self.L.gap_reinit(float(origin[0]), float(origin[1]), int(patch_dim)))
        self.origin = np.array(origin, np.float64)
        self.patch_dim = int(patch_dim)

    def to_patch(self, pose_utm):
        return np.array([pose_utm[0] - self.origin[0], pose_utm[1] - self.origin[1], pose_utm[2]])

    def insert_window_sim(self, window, ego_patch):
        w = np.ascontiguousarray(window, np.uint8)
        self._ck(self.L.gap_insert_window_sim(_p(w), w.shape[0], float(ego_patch[0]), float(ego_patch[1])))

    def pass_local_map(self, window, ox, oy):
        w = np.ascontiguousarray(window, np.uint8)
        self._ck(self.L.gap_pass_local_map(_p(w), int(ox), int(oy), w.shape[0]))

    def process_safety(self):
        self._ck(self.L.gap_process_safety())

    def check_pose(self, pose):
        r = self.L.gap_check_pose(float(pose[0]), float(pose[1]), float(pose[2]))
        if r < 0:
            self._ck(r)
        return bool(r)

    def recalc_env(self, ego_patch, goal_patch):
        e, g = np.array(ego_patch, np.float64), np.array(goal_patch, np.float64)
        ms = C.c_double(0)
        n = self.L.gap_recalc_env(_p(e), _p(g), C.byref(ms))
        if n < 0:
            self._ck(n)
        return n, ms.value

    def plan(self, ego_patch, start_patch, goal_patch, do_analytic=True, trace=True):
        e, s, g = (np.array(a, np.float64) for a in (ego_patch, start_patch, goal_patch))
        out = np.zeros(8)
        self._ck(self.L.gap_plan(_p(e), _p(s), _p(g), int(do_analytic), int(trace), _p(out)))
        n = int(out[3])
        x, y, yaw, delta = (np.zeros(n) for _ in range(4))
        d, t = np.zeros(n, np.int32), np.zeros(n, np.int32)
        self.L.gap_get_path(_p(x), _p(y), _p(yaw), _p(delta), _p(d), _p(t))
        tr, h = np.zeros(3, np.uint64), np.zeros(1, np.uint64)
        self.L.gap_get_trace(_p(tr), _p(h))
        exp = np.zeros((int(tr[0]), 3))
        self.L.gap_get_expanded(_p(exp))
        nc = int(out[1])
        key5, cost2 = np.zeros((nc, 5), np.int64), np.zeros((nc, 2))
        self.L.gap_get_closed(_p(key5), _p(cost2), nc)
        extra = {}
        if self.arm == "b200":
            st, ns = np.zeros(8, np.uint64), np.zeros(3, np.uint64)
            self.L.gap_b200_stats(_p(st), _p(ns))
            extra = dict(b200_stats=st, b200_ns=ns)  # ns: submit, wait + copy, answer table  # fans checked synchronously, cached / direct path checks, poses, look-ahead launches / hits / re-checks, kernel launches
        return dict(extra, found=int(out[0]), n_closed=nc, n_open=int(out[2]), path_cost=out[4], idx_analytic=int(out[5]),
                    is_analytic=int(out[6]), search_ms=out[7], path_x=x, path_y=y, path_yaw=yaw, path_delta=delta,
                    path_dir=d, path_type=t, n_expanded=int(tr[0]), n_checks=int(tr[1]), n_check_poses=int(tr[2]),
                    check_hash=h[0], expanded=exp, closed_keys=key5, closed_cost=cost2)

    def heuristic(self, which):
        d = self.L.gap_astar_dim()
        cost, dist = np.zeros((d, d)), np.zeros((d, d))
        par, fl = np.zeros((d, d), np.int32), np.zeros((d, d), np.uint8)
        n = self.L.gap_get_heuristic(int(which), _p(cost), _p(dist), _p(par), _p(fl))
        return dict(n=n, cost=cost, dist=dist, parent=par, flags=fl)

    def get_map(self, which):
        d = self.L.gap_astar_dim()
        a = np.zeros((d, d), np.uint8 if which == 0 else np.float64)
        self._ck(self.L.gap_get_map(int(which), _p(a)))
        return a


# ---- workloads --------------------------------------------------------------------------------

def load_scenario(name):
    z = np.load(os.path.join(GOLDEN, "scenario_%s.npz" % name))
    return z["global_map_all_visible"], z["start_pose"], z["goal_pose"]


def local_window(global_map, ego_utm, dim=801):
    """simulation.py:262-277 + util.get_local_map (util.py:145-157): window around the rounded ego cell."""
    ex, ey = int(round(ego_utm[0] / GM_RES)), int(round(ego_utm[1] / GM_RES))
    off = dim // 2
    ox, oy = max(ex - off, 0), max(ey - off, 0)
    return np.ascontiguousarray(global_map[oy:oy + dim, ox:ox + dim])


def run_scenario(pl, name, trace=True, do_plan=True):
    """BASELINE config 1: one planning query on a reference map, ALL_VISIBLE = true (first cycle of do_planning)."""
    gmap, start, goal = load_scenario(name)
    t0 = time.perf_counter()
    pl.create_patch(start, goal)
    ego, g = pl.to_patch(start), pl.to_patch(goal)
    pl.insert_window_sim(local_window(gmap, start), ego)
    pl.process_safety()
    t1 = time.perf_counter()
    ok = (pl.check_pose(ego), pl.check_pose(g))
    n_guid, env_ms = pl.recalc_env(ego, g)
    t2 = time.perf_counter()
    res = dict(patch_dim=pl.patch_dim, origin=pl.origin, poses_ok=np.array(ok), n_guidance_path=n_guid, env_ms=env_ms,
               insert_safety_ms=(t1 - t0) * 1e3, env_total_ms=(t2 - t1) * 1e3)
    if do_plan:
        res.update(pl.plan(ego, ego, g, True, trace))
        res["query_ms"] = (time.perf_counter() - t0) * 1e3
    return res


def run_patch_move(pl, name):
    """config 3's patch move: a first cycle on the scenario (environment built), then the ego has moved and a new patch is
    created around it (planning.py:668-715 -> create_patch -> HybridAStar::reinit -> AStar::reinit, a_star.cpp:70-99):
    h_prox_arr_ / motion_res_map_ are carried over the move (util::saveTemp + copyPatch2Patch), the next environment
    build and heuristic run on the carried field."""
    gmap, start, goal = load_scenario(name)
    first = run_scenario(pl, name, trace=False, do_plan=False)
    ego2 = np.array([start[0] - 9.0, start[1] + 13.0, start[2]])
    pl.create_patch(ego2, goal)
    ego, g = pl.to_patch(ego2), pl.to_patch(goal)
    out = dict(first_patch_dim=first["patch_dim"], patch_dim=pl.patch_dim, origin=pl.origin,
               carried_h_prox=pl.get_map(1), carried_motion_res=pl.get_map(2))
    if pl.arm == "b200":
        d = pl.L.gap_astar_dim()
        for which, key in ((1, "device_h_prox"), (2, "device_motion_res")):
            a = np.zeros((d, d))
            pl._ck(pl.L.gap_b200_device_map(which, _p(a)))
            out[key] = a
    pl.insert_window_sim(local_window(gmap, ego2), ego)
    pl.process_safety()
    n_guid, env_ms = pl.recalc_env(ego, g)
    h = pl.heuristic(0)
    out.update(n_guidance_path=n_guid, h_prox=pl.get_map(1), motion_res=pl.get_map(2), hg_cost=h["cost"], hg_parent=h["parent"],
               hg_flags=h["flags"], hg_dist=h["dist"])
    return out


def sample_queries(pl, dim, seed, n, margin_cells=101):
    """start / goal poses on a synthetic map: uniform in metres, yaw U[-pi, pi), until checkPose passes
    (SURVEY.md 8(d)); the ego keeps 101 astar cells from the far border (Voronoi window)."""
    rng = np.random.default_rng(seed + 1)
    side = dim * GM_RES
    lo, hi = 12.0, side - margin_cells * ASTAR_RES
    out = []
    while len(out) < n:
        pair = []
        for _ in range(2):
            for _try in range(10000):
                p = np.array([rng.uniform(lo, hi), rng.uniform(lo, hi), rng.uniform(-np.pi, np.pi)])
                if pl.check_pose(p):
                    break
            pair.append(p)
        out.append(pair)
    return out


def run_synthetic(pl, dim, seed, n_queries, trace=True):
    """BASELINE config 5's unit: queries on one synthetic dim x dim map that IS the patch."""
    sys.path.insert(0, os.path.join(ROOT, "guided-extended-hybrid-astar_b200"))
    from ga_b200.synth import gen_map
    m = gen_map(dim, seed)
    pl.reinit((0.0, 0.0), dim)
    pl.pass_local_map(m, 0, 0)
    pl.process_safety()
    qs = sample_queries(pl, dim, seed, n_queries)
    results = []
    for ego, goal in qs:
        t0 = time.perf_counter()
        n_guid, env_ms = pl.recalc_env(ego, goal)
        r = pl.plan(ego, ego, goal, True, trace)
        r.update(env_ms=env_ms, n_guidance_path=n_guid, query_ms=(time.perf_counter() - t0) * 1e3, ego=ego, goal=goal)
        results.append(r)
    return results


def run_query_range(pl, dim, first_id, last_id, trace=False):
    """BASELINE config 5 proper: the queries [first_id, last_id) of the 4096-query job.  Query q lives on the synthetic
    map of seed 50000 + q // 64 (64 start / goal pairs per map) and is the (q % 64)-th pair sampled on it;
    a planner process walks its block of ids map by map (ga_b200/sharding.py decides who gets which block)."""
    sys.path.insert(0, os.path.join(ROOT, "guided-extended-hybrid-astar_b200"))
    from ga_b200.synth import gen_map
    from ga_b200.sharding import QUERIES_PER_MAP, map_seed
    results, cur_seed, pairs = [], None, None
    for q in range(first_id, last_id):
        seed = map_seed(q)
        t0 = time.perf_counter()
        if seed != cur_seed:
            pl.reinit((0.0, 0.0), dim)
            pl.pass_local_map(gen_map(dim, seed), 0, 0)
            pl.process_safety()
            pairs = sample_queries(pl, dim, seed, QUERIES_PER_MAP)
            cur_seed = seed
            t0 = time.perf_counter()  # map generation and the sampling of the pairs are the job's input, not planner work
        ego, goal = pairs[q % QUERIES_PER_MAP]
        n_guid, env_ms = pl.recalc_env(ego, goal)
        r = pl.plan(ego, ego, goal, True, trace)
        r.update(env_ms=env_ms, n_guidance_path=n_guid, query_ms=(time.perf_counter() - t0) * 1e3, ego=ego, goal=goal,
                 query_id=q)
        results.append(r)
    return results


def _flatten(prefix, r, out):
    for k, v in r.items():
        out[prefix + k] = np.asarray(v)


def run_query_subprocess(arm, out_path, scenario=None, synthetic=None, seed=0, queries=1, ushift=False, mode=0,
                         timeout=600, extra=()):
    """fresh process per run (the reference is one planner per process and leaks search state across plans)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--arm", arm, "--out", out_path, "--mode", str(mode)]
    if scenario:
        cmd += ["--scenario", scenario]
    if synthetic:
        cmd += ["--synthetic", str(synthetic), "--seed", str(seed), "--queries", str(queries)]
    if ushift:
        cmd += ["--ushift"]
    cmd += list(extra)
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError("plan runner failed (%s): %s\n%s" % (arm, p.stdout[-2000:], p.stderr[-4000:]))
    return dict(np.load(out_path, allow_pickle=False))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", choices=["ref", "b200"], required=True)
    ap.add_argument("--scenario")
    ap.add_argument("--synthetic", type=int)
    ap.add_argument("--seed", type=int, default=50000)
    ap.add_argument("--queries", type=int, default=1)
    ap.add_argument("--query-ids", type=int, nargs=2, help="config 5: run the queries [a, b) of the sharded job on --synthetic maps")
    ap.add_argument("--ushift", action="store_true")
    ap.add_argument("--mode", type=int, default=0, help="arm b200: 0 reference-order heuristic, 1 GPU fixed point")
    ap.add_argument("--timeout-ms", type=int, default=10**9)
    ap.add_argument("--no-trace", action="store_true")
    ap.add_argument("--heuristic", action="store_true", help="also store the heuristic planes")
    ap.add_argument("--no-lookahead", action="store_true", help="arm b200: check every fan synchronously")
    ap.add_argument("--patch-move", action="store_true", help="with --scenario: first cycle, then a patch move and a second environment build")
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    pl = Planner(a.arm, is_ushift=a.ushift, timeout_ms=a.timeout_ms)
    pl.set_heuristic_mode(a.mode)
    if a.arm == "b200" and a.no_lookahead:
        pl._ck(pl.L.gap_b200_set_lookahead(0))
    out = dict(init_s=np.array(pl.init_s))
    if a.scenario and a.patch_move:
        if a.arm == "b200":
            pl.L.gap_b200_device_map.argtypes = [C.c_int, C.c_void_p]
        _flatten("", run_patch_move(pl, a.scenario), out)
    elif a.scenario:
        r = run_scenario(pl, a.scenario, trace=not a.no_trace)
        _flatten("", r, out)
        if a.heuristic:
            for w, nm in ((0, "hg_"), (1, "hp_")):
                _flatten(nm, pl.heuristic(w), out)
                if a.arm == "b200":  # the device's own planes beside what the closed set's accessors return
                    d = pl.L.gap_astar_dim()
                    dev = dict(cost=np.zeros((d, d)), dist=np.zeros((d, d)), parent=np.zeros((d, d), np.int32), flags=np.zeros((d, d), np.uint8))
                    pl.L.gap_b200_heur_planes.argtypes = [C.c_int] + [C.c_void_p] * 4
                    pl._ck(pl.L.gap_b200_heur_planes(w, _p(dev["cost"]), _p(dev["dist"]), _p(dev["parent"]), _p(dev["flags"])))
                    _flatten("dev_" + nm, dev, out)
            out["h_prox"] = pl.get_map(1)
            out["motion_res"] = pl.get_map(2)
            out["astar_grid"] = pl.get_map(0)
    else:
        if a.query_ids:
            rs = run_query_range(pl, a.synthetic, a.query_ids[0], a.query_ids[1], trace=not a.no_trace)
        else:
            rs = run_synthetic(pl, a.synthetic, a.seed, a.queries, trace=not a.no_trace)
        out["n_queries"] = np.array(len(rs))
        for i, r in enumerate(rs):
            _flatten("q%d_" % i, r, out)
    np.savez_compressed(a.out, **out)
    brief = {k: (v.tolist() if v.size < 4 else v.shape) for k, v in out.items() if not k.startswith(("q", "hp_", "hg_"))}
    print(brief)


if __name__ == "__main__":
    main()
