#!/usr/bin/env python
"""bench.py -- headline measurement of the B200 grid backend (BASELINE.json configs[1]).

Workload ("synthetic-4096"): synthetic 4096 x 4096 occupancy patch (SURVEY.md 8(d) gen_map, seed
20240 + i) -> threshold + 21x21 disk dilation + 4x4 max-pool (ga_build_safety) -> obstacle-aware
2D heuristic grid from a goal cell (ga_build_heuristic) on the 1024 x 1024 astar grid.
One step = one such environment build.  metric = heuristic-grid Mcells/s = astar cells closed by
the heuristic / step time (the dilation + pool that feed it are inside the step).

value : inputs resident in HBM, device-timed with CUDA events on the library's stream.
e2e   : same step through the reference-facing host API with HOST buffers: pinned patch -> H2D,
        build, D2H of the closed-set planes (cost, cost_dist, parent, flags) the serial search reads.
N > 1 : one process per GPU (torchrun), every rank builds its own environments (independent
        queries, no data-path collective), value = all ranks' cells / max-over-ranks time.
--impl reference : the reference's own CPU code for the same step (oracle/_ref when built, else
        the oracle port), timed on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "guided-extended-hybrid-astar_b200"))

PATCH_DIM = 4096
N_ROT = 4            # distinct environments rotated through so consecutive steps never share L2 contents
METRIC = "heuristic_grid_mcells_per_s"
UNIT = "Mcells/s"
BYTES_PER_CELL_HEUR = 39.0   # read 1+1+8+8, write 8+8+4+1 per astar cell
BYTES_PER_CELL_SAFETY = 2.0 + 1.0 / 16
# dram__bytes_read.sum + dram__bytes_write.sum of one heur_sweep_kernel launch on the 1024^2 grid of
# this workload, from the `ncu --set full` capture summarised in profiles/ (None until captured)
NCU_TRAFFIC_BYTES = 17946900 + 148480  # profiles/r1_ncu_full_summary.csv (18.1 MB: the planes finish writes are not this kernel's, improved cells stay in L2)


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6400.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line):
    NVML every 2 ms when nvidia_ml_py is importable, the nvidia-smi query otherwise."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz, self.active = None, False
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _nvml(self):
        nv = self.nv
        bits = ((nv.nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown"),
                (nv.nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                (nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_thermal_slowdown"),
                (nv.nvmlClocksThrottleReasonSwPowerCap, "sw_power_cap"))
        while not self.stop_flag:
            if self.active:
                try:
                    self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in bits:
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(0.002)

    def _smi(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [v.strip() for v in out.strip().split(",")]
                if self.active:
                    self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:]):
                    if self.active and v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.02)

    def run(self):
        (self._nvml if self.nv else self._smi)()

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "source": "nvml" if self.nv else "nvidia-smi"}


def make_env(ga, device, seed, torch):
    """one planner context with its 4096^2 patch resident, plus pinned host buffers for e2e"""
    from ga_b200 import _abi
    ctx = ga.GridContext(device)
    cfg = ga.LibConfig()
    astar = ga.AStar(ctx, cfg)
    cc = ga.CollisionChecker(ctx, ga.Vehicle.from_config(ga.VehConfig()), cfg)
    cc.initialize(PATCH_DIM)
    patch = ga.gen_map(PATCH_DIM, seed)
    h_patch = torch.from_numpy(patch).pin_memory()
    ctx.upload(_abi.GRID_PATCH, h_patch.numpy())
    cc.processSafetyPatch()
    grid = astar.astar_grid_
    goal = ga.sample_free_cells(grid, 1, seed + 1)[0]
    adim = astar.astar_dim_
    outs = {gid: torch.empty((adim, adim), dtype=dt).pin_memory() for gid, dt in
            ((_abi.GRID_HEUR_COST, torch.float64), (_abi.GRID_HEUR_DIST, torch.float64),
             (_abi.GRID_HEUR_PARENT, torch.int32), (_abi.GRID_HEUR_FLAGS, torch.uint8))}
    return dict(ctx=ctx, cc=cc, astar=astar, patch=patch, h_patch=h_patch, goal=(int(goal[0]), int(goal[1])), outs=outs)


def step_resident(env, n_closed=None):
    lib, h = env["ctx"].lib, env["ctx"].h
    env["ctx"].check(lib.ga_build_safety(h))
    env["ctx"].check(lib.ga_build_heuristic(h, 0, env["goal"][0], env["goal"][1], n_closed))


def step_e2e(env):
    from ga_b200 import _abi
    ctx = env["ctx"]
    lib, h = ctx.lib, ctx.h
    hp = env["h_patch"]
    ctx.check(lib.ga_upload(h, _abi.GRID_PATCH, C.c_void_p(hp.data_ptr()), hp.numel()))
    ctx.check(lib.ga_build_safety(h))
    ctx.check(lib.ga_build_heuristic(h, 0, env["goal"][0], env["goal"][1], None))
    for gid, t in env["outs"].items():
        ctx.check(lib.ga_download(h, gid, C.c_void_p(t.data_ptr()), t.numel() * t.element_size()))


def step_e2e_async(env):
    """step_e2e without the host waiting for the result: the caller syncs the context before it touches
    the environment's buffers again"""
    from ga_b200 import _abi
    ctx = env["ctx"]
    lib, h = ctx.lib, ctx.h
    hp = env["h_patch"]
    ctx.check(lib.ga_upload(h, _abi.GRID_PATCH, C.c_void_p(hp.data_ptr()), hp.numel()))
    ctx.check(lib.ga_build_safety(h))
    ctx.check(lib.ga_build_heuristic(h, 0, env["goal"][0], env["goal"][1], None))
    for gid, t in env["outs"].items():
        ctx.check(lib.ga_download_async(h, gid, C.c_void_p(t.data_ptr()), t.numel() * t.element_size()))


def _ref_goal(patch):
    import ga_oracle_py as go
    thr = go.threshold(patch, 100, 130)
    free = np.argwhere(thr[::4, ::4] == 0)
    return int(free[len(free) // 2][1]), int(free[len(free) // 2][0])


def cpu_reference_steps(dim, seeds, disk_r_c):
    """the same step on the host cores, one thread (the reference keeps all planner state in static
    members: one planner per process).  oracle/_ref (the reference's own translation units) when it
    was built, else the plain-C oracle port.  Returns (seconds, cells closed, kind)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ga_b200
    import ga_oracle_py as go
    import ga_ref_py as gr
    total, cells = 0.0, 0
    if gr.available():
        gr.init(dim)
        gr.set_heuristic_config(early_exit=False, max_extra_nodes=1 << 30, unknown_cost_w=0.0)  # whole reachable grid, as the GPU builds
        L = gr.lib()
        for seed in seeds:
            patch = ga_b200.gen_map(dim, seed)
            goal = _ref_goal(patch)
            t0 = time.perf_counter()
            gr.set_grid(gr.PATCH, patch)       # CollisionChecker::patch_arr_
            gr.process_safety_patch()          # collision_checking.cpp:193-224
            gr.calc_astar_grid()               # a_star.cpp:536-542 (CPU max in place of cuDNN)
            n = L.gar_calc_distance_heuristic(goal[0], goal[1], goal[0], goal[1], 0, 0)  # a_star.cpp:110-295
            total += time.perf_counter() - t0
            cells += n
        return total, cells, "reference"
    for seed in seeds:
        patch = ga_b200.gen_map(dim, seed)
        goal = _ref_goal(patch)
        t0 = time.perf_counter()
        safety = go.process_safety_patch(patch, 100, 130, disk_r_c)
        grid = go.pool_max(safety)
        ad = grid.shape[0]
        hp = go.HeurParams(ad, 5.0, 3.0, 0.0, 0, 1 << 30, 0)
        res = go.heuristic_reforder(hp, grid, np.zeros((ad, ad), np.uint8), np.full((ad, ad), 5.0), np.zeros((ad, ad)), goal, goal)
        total += time.perf_counter() - t0
        cells += res["n_closed"]
    return total, cells, "port"


def _ref_worker(a):
    dim, seeds, r_c = a
    return cpu_reference_steps(dim, seeds, r_c)  # (seconds inside the reference's calls, cells, kind)


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the step, one planner process per
    host core (its only way to use more than one core), each step = cores x one 2048^2 sample."""
    if rank != 0:
        return
    import multiprocessing as mp
    import ga_b200
    dim = args.ref_dim
    if dim <= 0:
        # bounded sample: the whole --steps/--warmup run stays within ~2 minutes of host time
        budget = 120.0 / (args.steps + args.warmup)
        dim = 2048 if budget >= 0.9 else 1024 if budget >= 0.25 else 512
    veh = ga_b200.Vehicle.from_config(ga_b200.VehConfig())
    _, _, r_c, _ = ga_b200.calc_disks(veh, ga_b200.LibConfig())
    cores = max(1, min(len(os.sched_getaffinity(0)), args.ref_procs))
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        def one_step(i):
            # the processes run concurrently; the step lasts as long as the slowest planner's calls
            # (map generation and process start-up are not the reference's work and stay outside)
            out = pool.map(_ref_worker, [(dim, [20240 + (i * cores + w) % 64], r_c) for w in range(cores)])
            return max(o[0] for o in out), sum(o[1] for o in out), out[0][2]
        for i in range(args.warmup):
            one_step(i)
        total, cells, kind = 0.0, 0, "port"
        for i in range(args.steps):
            dt, n, kind = one_step(args.warmup + i)
            total += dt
            cells += n
    val = cells / total / 1e6
    sample = "%d step(s); one step = %d planner processes x (threshold + dilate + pool + heuristic) on a %dx%d patch" % (
        args.steps, cores, dim, dim)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic-4096", "patch_dim": PATCH_DIM, "astar_dim": PATCH_DIM // 4,
                       "sample_patch_dim": dim},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_batch_leg(ga, torch, device, rank, args, sampler, hbm_peak):
    """BASELINE config 5's shape on one GPU: a batch of independent queries on 2048^2 maps (64 queries
    share a map), one heuristic grid per query, all grids relaxed by ONE kernel chain
    (ga_build_heuristic_batch).  Reports grids/s, Mcells/s and the algorithmic-bytes rate."""
    from ga_b200 import _abi
    nq, dim = args.batch, 2048
    maps = {}
    ctxs, astars, goals = [], [], []
    for qi in range(nq):
        seed = 50000 + rank * 1000 + qi // 64
        if seed not in maps:
            maps[seed] = ga.gen_map(dim, seed)
        ctx = ga.GridContext(device)
        cfg = ga.LibConfig()
        astar = ga.AStar(ctx, cfg)
        cc = ga.CollisionChecker(ctx, ga.Vehicle.from_config(ga.VehConfig()), cfg)
        cc.initialize(dim)
        ctx.upload(_abi.GRID_PATCH, maps[seed])
        cc.processSafetyPatch()
        if qi % 64 == 0:
            grid = astar.astar_grid_
            free = ga.sample_free_cells(grid, 64, seed + 1)
        ctxs.append(ctx)
        astars.append(astar)
        goals.append(free[qi % 64])
    closed = ga.AStar.calcDistanceHeuristicBatch(astars, goals, for_path=False)
    lead = ctxs[0]
    hs = (C.c_void_p * nq)(*[c.h for c in ctxs])
    g = np.ascontiguousarray(goals, np.int32).reshape(nq, 2)
    gp = g.ctypes.data_as(C.POINTER(C.c_int))
    reps = max(3, min(20, args.steps // 10))
    for _ in range(2):
        lead.check(lead.lib.ga_build_heuristic_batch(hs, nq, 0, gp, None))
    lead.sync()
    lead.prof_reset()
    lead.prof_enable(True)
    s = torch.cuda.ExternalStream(lead.stream())
    a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active = True
    a.record(s)
    for _ in range(reps):
        lead.check(lead.lib.ga_build_heuristic_batch(hs, nq, 0, gp, None))
    b_.record(s)
    lead.sync()
    sampler.active = False
    lead.prof_enable(False)
    ms = a.elapsed_time(b_) / reps
    relax_ms = lead.prof_get(_abi.K_HEUR_RELAX)[0] / reps
    adim = dim // 4
    alg = BYTES_PER_CELL_HEUR * adim * adim * nq
    out = {"workload": "%d queries on 2048x2048 maps (64 per map), one 512x512 heuristic grid each, one kernel chain" % nq,
           "grids_per_s": nq / (ms * 1e-3), "mcells_per_s": float(closed.sum()) / (ms * 1e-3) / 1e6, "ms_per_batch": ms,
           "sweep_kernel_ms": relax_ms, "sweep_kernel_algorithmic_gbs": alg / (relax_ms * 1e-3) / 1e9,
           "sweep_kernel_roofline_frac": alg / (relax_ms * 1e-3) / 1e9 / hbm_peak}
    for c in ctxs:
        c.close()
    return out


def run_replan_leg(ga, device, args):
    """BASELINE config 3's shape: the unknown-environment replanning cycle on the default scenario's
    patch size (1492^2, astar 373^2).  One cycle = ray-cast the 801^2 ground-truth window (host buffer
    in), merge it into the accumulated map, hand the window to the collision checker, rebuild the
    safety + pooled grids and the guidance heuristic.  Timed end to end on the host clock; the
    reference's own code for the same cycle (gridmap_sim.cpp, cartographing.cpp,
    collision_checking.cpp, a_star.cpp) on one core beside it."""
    from ga_b200 import _abi
    dim = 1492
    world = ga.gen_map(2600, 31)
    world = np.where(world == 255, 255, 0).astype(np.uint8)
    origin = np.array([500, 400])
    cycles = [(origin + np.array([420 + 7 * k, 430 + 4 * k])) for k in range(30)]
    wins = [np.ascontiguousarray(world[e[1] - 400:e[1] + 401, e[0] - 400:e[0] + 401]) for e in cycles]
    ctx = ga.GridContext(device)
    cfg = ga.LibConfig()
    astar = ga.AStar(ctx, cfg)
    cc = ga.CollisionChecker(ctx, ga.Vehicle.from_config(ga.VehConfig()), cfg)
    cc.initialize(dim)
    cart, sim = ga.Cartographing(ctx), ga.GridMapSim(ctx)
    cart.resetPatch(dim)
    goal = (300, 60)

    def cycle(k):
        wo = cycles[k] - origin - 400
        sim.gmSimulation(wins[k], fetch=False)
        cart.cartographRaytraced(wo)
        cart.passLocalMap(wo, 801)
        cc.processSafetyPatch()
        ctx.check(ctx.lib.ga_build_heuristic(ctx.h, 0, goal[0], goal[1], None))

    for k in range(3):
        cycle(k)
    ctx.sync()
    t0 = time.perf_counter()
    for k in range(3, len(cycles)):
        cycle(k)
    ctx.sync()
    dt = (time.perf_counter() - t0) / (len(cycles) - 3)
    out = {"workload": "unknown-environment cycle on a 1492^2 patch: 801^2 ray-cast + map merge + hand-over + safety/pool + 373^2 heuristic",
           "cycles_per_s": 1.0 / dt, "ms_per_cycle": dt * 1e3, "h2d_bytes_per_cycle": 801 * 801}
    ctx.close()
    if not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import ga_ref_py as gr
        if gr.available():
            gr.init(dim)
            gr.cart_reset(dim)
            gr.set_heuristic_config(early_exit=False, max_extra_nodes=1 << 30, unknown_cost_w=0.0)
            L = gr.lib()
            ts = []
            for k in range(6):
                wo = cycles[k] - origin - 400
                t0 = time.perf_counter()
                pm = gr.raytrace(wins[k])
                gr.cartograph(pm, wo[0], wo[1])
                gr.cart_pass_local_map(wo[0], wo[1], 801)
                gr.process_safety_patch()
                gr.calc_astar_grid()
                L.gar_calc_distance_heuristic(goal[0], goal[1], goal[0], goal[1], 0, 0)
                ts.append(time.perf_counter() - t0)
            out["cpu_reference_ms_per_cycle"] = 1e3 * float(np.median(ts[1:]))
            out["cpu_reference_sample"] = "5 cycles of the same chain through the reference's own code, 1 core, median"
    return out


def run_collision_leg(env, args, ga):
    """BASELINE metric M3 (collision checks/s) on environment 0 (4096^2 patch, 3-disk footprint):
    (a) pose checks through ga_check_poses: host pose arrays in, first-hit index per segment out, one
        launch per batch (H2D + kernel + D2H inside the timing: the API takes host buffers);
    (b) expansion batches through ga_expand: every motion primitive of every node (13 steers x 2
        directions + the 6 U-Shift rear-axis rotations, config 4) rolled out and checked in one launch;
    next to the reference's own checkPose loop on one host core."""
    from ga_b200 import _abi
    ctx, cc = env["ctx"], env["cc"]
    rng = np.random.default_rng(7)
    n, seg_len = 1 << 20, 8
    ext = PATCH_DIM * 0.15625
    x, y = rng.uniform(0, ext, n), rng.uniform(0, ext, n)
    yaw = rng.uniform(-3.2, 3.2, n)
    seg = np.arange(0, n + 1, seg_len, dtype=np.int32)
    cc.checkSegments(x, y, yaw, seg)
    ctx.sync()
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        hits = cc.checkSegments(x, y, yaw, seg)
    ctx.sync()
    dt = (time.perf_counter() - t0) / reps
    out = {"pose_checks_per_s": n / dt, "poses_per_launch": n, "segments": len(seg) - 1, "disks": 3,
           "free_segments": float((hits < 0).mean())}
    # expansion batches
    cc.setExpandParams()
    nodes = np.zeros(4096, _abi.NODE_DTYPE)
    nodes["x"], nodes["y"] = rng.uniform(20, ext - 20, 4096), rng.uniform(20, ext - 20, 4096)
    nodes["yaw"], nodes["motion_res"], nodes["with_rear_axis"] = rng.uniform(-3.14, 3.14, 4096), 0.3125, 1
    succ = cc.expandNodes(nodes)
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        succ = cc.expandNodes(nodes)
    ctx.sync()
    dt = (time.perf_counter() - t0) / reps
    poses = int(succ["n_poses"].sum())
    out.update({"expand_nodes_per_s": len(nodes) / dt, "expand_primitives_per_s": int((succ["n_poses"] > 0).sum()) / dt,
                "expand_pose_checks_per_s": poses / dt, "expand_batch_nodes": len(nodes), "ushift": True})
    # (c) proximity scoring of sampled analytic-expansion paths: host pose lists
    #     in, one cost per path out; h_prox_arr_ stays on the device
    adim = PATCH_DIM // 4
    prox = rng.random((adim, adim)) * (rng.random((adim, adim)) < 0.5)
    ctx.upload(_abi.GRID_H_PROX, prox)
    hy = ga.HybridAStar(ctx, cc.vehicle)
    n_paths, path_len = 8192, 128  # 12.8 m sampled at INTERP_RES = 0.1
    pp = np.stack([rng.uniform(10, ext - 10, n_paths * path_len), rng.uniform(10, ext - 10, n_paths * path_len),
                   rng.uniform(-3.14, 3.14, n_paths * path_len)], 1)
    poff = np.arange(0, n_paths * path_len + 1, path_len, dtype=np.int32)
    hy.getPathProxCosts(pp, poff)
    t0 = time.perf_counter()
    for _ in range(reps):
        costs = hy.getPathProxCosts(pp, poff)
    dt = (time.perf_counter() - t0) / reps
    out.update({"prox_poses_per_s": len(pp) / dt, "prox_paths_per_launch": n_paths, "prox_path_len": path_len})
    # (d) getValidClosePose step 3: 64 x 64 x 114 candidate poses per call
    grid = ctx.download(_abi.GRID_ASTAR)
    free = np.argwhere(grid == 0)
    cyx = free[len(free) // 2]
    ys, xs = np.mgrid[0:adim, 0:adim]
    near = (np.sqrt((xs - cyx[1]) ** 2.0 + (ys - cyx[0]) ** 2.0) < 10.0) & (grid != 2)
    cells = (ys[near] * adim + xs[near]).astype(np.int32)
    center = (cyx[1] * 0.625 + 0.3, cyx[0] * 0.625 + 0.05, 0.4)
    close = hy.getValidClosePose(center, cells)
    t0 = time.perf_counter()
    for _ in range(reps):
        close = hy.getValidClosePose(center, cells)
    dt = (time.perf_counter() - t0) / reps
    out.update({"close_pose_ms": dt * 1e3, "close_pose_candidates": 64 * 64 * 114})
    if not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import ga_ref_py as gr
        if gr.available():
            m = 200000
            gr.init(PATCH_DIM)
            gr.set_grid(gr.H_PROX, prox)
            t0 = time.perf_counter()
            ref_costs = gr.path_prox_costs(pp[:m], poff[:m // path_len + 1])
            out["cpu_reference_prox_poses_per_s"] = (m // path_len * path_len) / (time.perf_counter() - t0)
            out["prox_max_diff_vs_reference"] = float(np.max(np.abs(ref_costs - costs[:len(ref_costs)])) / np.max(np.abs(ref_costs)))
            gr.set_grid(gr.SAFETY, ctx.download(_abi.GRID_SAFETY))
            t0 = time.perf_counter()
            ref_close = gr.valid_close_pose(center, cells)
            out["cpu_reference_close_pose_ms"] = (time.perf_counter() - t0) * 1e3
            out["close_pose_equal_to_reference"] = bool((ref_close is None) == (close is None) and
                                                        (close is None or np.array_equal(ref_close, close)))
            t0 = time.perf_counter()
            gr.check_segments(x[:m], y[:m], yaw[:m], seg[:m // seg_len + 1])
            out["cpu_reference_pose_checks_per_s"] = m / (time.perf_counter() - t0)
            out["cpu_reference_sample"] = "%d poses, CollisionChecker::getPathCollisionIndex per segment, 1 core" % m
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ref-dim", type=int, default=0, help="patch size of the bounded CPU sample (reference arm); 0 = from --steps")
    ap.add_argument("--ref-procs", type=int, default=64, help="planner processes of the reference arm (capped by the host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-collision", action="store_true", help="skip the collision-checks/s leg")
    ap.add_argument("--no-replan", action="store_true", help="skip the unknown-environment cycle leg")
    ap.add_argument("--batch", type=int, default=128, help="queries of the batched-heuristic leg (0 = skip)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import ga_b200
    from ga_b200 import _abi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 backend has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    envs = [make_env(ga_b200, local_rank, 20240 + rank * N_ROT + i, torch) for i in range(N_ROT)]
    hbm_peak, peak_src = load_peaks()

    def barrier():
        for e in envs:
            e["ctx"].sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # cells closed per environment (fixed by the map), also the parity hook of smoke()
    closed = []
    for e in envs:
        n = C.c_int()
        step_resident(e, C.byref(n))
        closed.append(n.value)

    def timed(step_fn, steps, profile):
        for i in range(args.warmup):
            step_fn(envs[i % N_ROT])
        barrier()
        if profile:
            for e in envs:
                e["ctx"].prof_reset()
                e["ctx"].prof_enable(True)
        launches0 = sum(e["ctx"].launch_count() for e in envs)
        sampler.active = True
        t0 = time.perf_counter()
        # steps run back to back on the device: step i+1's stream waits for step i's end event, so the
        # span first-start .. last-end is exactly K serialised steps without any host synchronisation
        first = last = None
        for i in range(steps):
            e = envs[i % N_ROT]
            s = torch.cuda.ExternalStream(e["ctx"].stream())
            if last is not None:
                s.wait_event(last)
            a = torch.cuda.Event(enable_timing=True)
            a.record(s)
            first = first or a
            step_fn(e)
            last = torch.cuda.Event(enable_timing=True)
            last.record(s)
        barrier()
        wall = time.perf_counter() - t0
        sampler.active = False
        dev_ms = first.elapsed_time(last)
        launches = sum(e["ctx"].launch_count() for e in envs) - launches0
        cells = sum(closed[i % N_ROT] for i in range(steps))
        if profile:
            for e in envs:
                e["ctx"].prof_enable(False)
        return dev_ms, wall * 1e3, launches, cells

    def timed_pipelined(steps):
        """the e2e step with the N_ROT environments (one context, stream and pinned buffer set each) in
        flight together: environment k's upload and downloads overlap the kernels of the others.  Every
        step still copies its own patch in and its four result planes out inside the timed region."""
        for i in range(args.warmup):
            step_e2e_async(envs[i % N_ROT])
        for e in envs:
            e["ctx"].sync()
        barrier()
        sampler.active = True
        t0 = time.perf_counter()
        for i in range(steps):
            e = envs[i % N_ROT]
            if i >= N_ROT:
                e["ctx"].sync()  # this environment's previous step is complete: its buffers are free again
            step_e2e_async(e)
        for e in envs:
            e["ctx"].sync()
        barrier()
        wall = time.perf_counter() - t0
        sampler.active = False
        return wall * 1e3, sum(closed[i % N_ROT] for i in range(steps))

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev_ms, wall_ms, launches, cells = timed(step_resident, args.steps, False)
    # per-kernel device time: the same K steps once more with CUDA events around every launch on the
    # library's stream (kept out of the run `value` comes from: the event records cost a little)
    timed(step_resident, args.steps, True)
    kt = {}
    for k, name in ((_abi.K_DILATE, "safety"), (_abi.K_HEUR_INIT, "heur_init"),
                    (_abi.K_HEUR_RELAX, "heur_relax"), (_abi.K_HEUR_FINISH, "heur_finish")):
        ms = n = 0
        for e in envs:
            m_, n_ = e["ctx"].prof_get(k)
            ms += m_
            n += n_
        kt[name] = (ms, n)
    # the steps are serialised on each context's stream; K steps' device time = sum of per-step event spans
    e2e_dev_ms, e2e_serial_ms, _, e2e_cells = timed(step_e2e, args.steps, False)
    e2e_wall_ms, e2e_cells = timed_pipelined(args.steps)
    batch = run_batch_leg(ga_b200, torch, local_rank, rank, args, sampler, hbm_peak) if args.batch > 0 else None
    collision = run_collision_leg(envs[0], args, ga_b200) if (rank == 0 and not args.no_collision) else None
    replan = run_replan_leg(ga_b200, local_rank, args) if (rank == 0 and not args.no_replan) else None
    sampler.stop_flag = True

    # max over ranks
    t = torch.tensor([dev_ms, e2e_wall_ms, float(cells), float(e2e_cells), float(launches), e2e_serial_ms], dtype=torch.float64,
                     device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms, e2e_wall_ms, e2e_serial_ms = float(tmax[0]), float(tmax[1]), float(tmax[5])
        cells, e2e_cells, launches = float(tsum[2]), float(tsum[3]), int(tsum[4])
    value = cells / (dev_ms * 1e-3) / 1e6
    e2e_value = e2e_cells / (e2e_wall_ms * 1e-3) / 1e6

    if rank == 0:
        adim = PATCH_DIM // 4
        relax_ms, relax_n = kt["heur_relax"]
        dil_ms, dil_n = kt["safety"]
        # dominant kernel = the wavefront relaxation; algorithmic bytes = 39 B x astar cells of one grid per launch
        alg_bytes = BYTES_PER_CELL_HEUR * adim * adim
        achieved = alg_bytes / (relax_ms / max(relax_n, 1) * 1e-3) / 1e9
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "synthetic-4096", "patch_dim": PATCH_DIM, "astar_dim": adim,
                           "disk_r_c": envs[0]["cc"].disk_r_c_, "l2": "inputs rotate over %d environments per rank (~130 MB each) > L2" % N_ROT,
                           "cells_closed_per_step": cells / args.steps / world},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": PATCH_DIM * PATCH_DIM,
                        "d2h_bytes_per_step": adim * adim * 21, "ms_per_step": e2e_wall_ms / args.steps,
                        "pipeline": "%d planner contexts in flight (one stream and pinned buffer set each): a step's copies "
                                    "overlap the other contexts' kernels" % N_ROT,
                        "serial_value": e2e_cells / (e2e_serial_ms * 1e-3) / 1e6,
                        "serial_ms_per_step": e2e_serial_ms / args.steps},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "kernel": "heur_sweep_kernel", "achieved": achieved, "peak": hbm_peak,
                             "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": NCU_TRAFFIC_BYTES, "peak_source": peak_src,
                             "note": "dependency-depth bound, not bandwidth bound: one grid's wavefront keeps ~5% of the "
                                     "worker warps busy; see `batch` for the same kernel fed with independent grids",
                             "algorithmic_bytes_per_launch": alg_bytes, "avg_launch_ms": relax_ms / max(relax_n, 1)},
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kt.items()},
                "safety_roofline": {"achieved": BYTES_PER_CELL_SAFETY * PATCH_DIM * PATCH_DIM /
                                    (dil_ms / max(dil_n, 1) * 1e-3) / 1e9, "unit": "GB/s",
                                    "note": "safety_kernel (threshold + dilation + pool fused), 2+1/16 B per gm cell"},
                "clocks": sampler.result()}
        if batch:
            line["batch"] = batch
        if collision:
            line["collision"] = collision
        if replan:
            line["replan_cycle"] = replan
        if not args.no_cpu_baseline:
            # bounded sample of the same step on the host: a 2048^2 patch seeded like environment 0
            dt, n, kind = cpu_reference_steps(2048, [20240], envs[0]["cc"].disk_r_c_)
            line["cpu_baseline"] = {"value": n / dt / 1e6, "unit": UNIT, "cores": 1, "kind": kind,
                                    "sample": "1 step (threshold + dilate + pool + heuristic) on a 2048x2048 patch, seed 20240 (%.1f s)" % dt}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    for e in envs:
        e["ctx"].close()


if __name__ == "__main__":
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    main()
