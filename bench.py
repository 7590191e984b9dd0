#!/usr/bin/env python
"""bench.py -- headline measurement of the B200 grid backend (BASELINE.json configs[1]).

Workload ("synthetic-4096"): one ENVIRONMENT BUILD of the planner on a synthetic 4096 x 4096 occupancy patch
(SURVEY.md 8(d) gen_map, seed 20240 + i), in the reference's call order (planning.py:756-783, hybrid_a_star_lib.cpp:1225-1234):
  passLocalMap(801^2 sensor window)  ->  processSafetyPatch (threshold + 21x21 disk dilation, whole patch) + 4x4 max-pool
  ->  calcVoronoiPotentialField(ego) (two distance transforms + setCostmapValue on the 200^2 window; the Fortune sweep
  between them is the reference's host code in BOTH arms)  ->  calcDistanceHeuristic(goal, ego) on the 1024 x 1024
  astar grid, with the proximity term the real planner carries (PROX instantiation).
One step = one such build.  metric = heuristic-grid Mcells/s = astar cells closed by the heuristic / step time.

value : inputs resident in HBM (patch, window, sweep end points), device-timed with CUDA events on the library's stream:
        safety + pool, potential field, heuristic (GPU fixed point: every reachable cell closed).
e2e   : the same build through the host API with HOST buffers, as the drop-in planner classes issue it: pinned 801^2
        window -> ga_patch_insert (H2D), safety, ga_potential_sites (D2H), Fortune sweep on the host, ga_build_potential
        (H2D), heuristic, D2H of what the serial search reads (h_prox / motion_res windows + the closed set as 12-byte
        records, GA_GRID_HEUR_REC; e2e.planes_*: the same with the four planes, 21 B per cell).
        e2e.lean_* : the same with the packed 1-byte plane instead of the four planes (planner that takes its per-successor
        costs from ga_expand's epilogue).
N > 1 : one process per GPU (torchrun), every rank builds its own environments (independent queries, no data-path
        collective), value = all ranks' cells / max-over-ranks time.
plan  : metric M1, plan queries/s: the reference's serial Hybrid A* (hybrid_a_star_lib.cpp, unmodified) on the drop-in
        classes over this library, P planner processes per GPU, next to the pure reference planner on the host cores.
--impl reference : the reference's own CPU code for the same step at the SAME patch size (oracle/_ref), one planner
        process per host core, per-stage times and both heuristic configurations (shipped early exit / whole grid).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "guided-extended-hybrid-astar_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

PATCH_DIM = 4096
WIN_DIM = 801        # sensor window of one cycle (simulation.py:53)
VOR_DIM = 200
N_ROT = 4            # distinct environments rotated through so consecutive steps never share L2 contents
METRIC = "heuristic_grid_mcells_per_s"
UNIT = "Mcells/s"
BYTES_PER_CELL_HEUR = 39.0   # read 1+1+8+8, write 8+8+4+1 per astar cell
BYTES_PER_CELL_SAFETY = 2.0 + 1.0 / 16
# dram__bytes_read.sum + dram__bytes_write.sum of one heur_sweep_kernel launch on the 1024^2 grid of this workload,
# from the `ncu --set full` capture summarised in profiles/ (None until captured for this round's kernel)
NCU_TRAFFIC_BYTES = 18417700 + 33536   # profiles/r2_ncu_full_summary.csv, heur_sweep_kernel


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def pick_points(grid, seed):
    """ego and goal astar cells of an environment, from the pooled grid (bit-identical in both arms): the ego keeps
    110 cells from every border (Voronoi window), the goal is any free cell at least a third of the
    grid away."""
    ad = grid.shape[0]
    rng = np.random.default_rng(seed)
    ys, xs = np.nonzero(grid == 0)
    inner = (xs >= 110) & (xs < ad - 110) & (ys >= 110) & (ys < ad - 110)
    k = rng.integers(0, int(inner.sum()))
    ego = (int(xs[inner][k]), int(ys[inner][k]))
    far = np.hypot(xs - ego[0], ys - ego[1]) > ad / 3
    k = rng.integers(0, int(far.sum()))
    return ego, (int(xs[far][k]), int(ys[far][k]))


def host_sweep(sites):
    """the Fortune sweep between the two distance transforms (deps_lib/VoronoiDiagramGenerator.cpp): serial fp64 host
    code that stays the reference's in both arms; here through oracle/_ref (the reference's own TU)."""
    import ga_ref_py as gr
    if len(sites) == 0:
        return np.zeros((0, 2), np.int32)
    return np.ascontiguousarray(gr.voronoi_endpoints(sites, VOR_DIM, 3.0), np.int32).reshape(-1, 2)


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
    """one planner context with its 4096^2 patch resident, the cycle's 801^2 sensor window in pinned host memory,
    ego / goal cells, the sweep end points of its potential window, and pinned result buffers for e2e"""
    from ga_b200 import _abi
    ctx = ga.GridContext(device)
    cfg = ga.LibConfig()
    astar = ga.AStar(ctx, cfg)
    cc = ga.CollisionChecker(ctx, ga.Vehicle.from_config(ga.VehConfig()), cfg)
    cc.initialize(PATCH_DIM)
    patch = ga.gen_map(PATCH_DIM, seed)
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    grid = astar.astar_grid_
    ego, goal = pick_points(grid, seed + 1)
    adim = astar.astar_dim_
    # the window the sensor delivers this cycle: 801^2 cells around the ego (same content as the patch there)
    wo = (min(max(ego[0] * 4 - WIN_DIM // 2, 0), PATCH_DIM - WIN_DIM), min(max(ego[1] * 4 - WIN_DIM // 2, 0), PATCH_DIM - WIN_DIM))
    h_win = torch.from_numpy(np.ascontiguousarray(patch[wo[1]:wo[1] + WIN_DIM, wo[0]:wo[0] + WIN_DIM])).pin_memory()
    sites = astar.potentialSites(ego)
    vor = host_sweep(sites)
    astar.calcVoronoiPotentialField(ego, vor)
    outs = {gid: torch.empty((adim, adim), dtype=dt).pin_memory() for gid, dt in
            ((_abi.GRID_HEUR_COST, torch.float64), (_abi.GRID_HEUR_DIST, torch.float64),
             (_abi.GRID_HEUR_PARENT, torch.int32), (_abi.GRID_HEUR_FLAGS, torch.uint8))}
    rec = torch.empty((adim * adim * 12,), dtype=torch.uint8).pin_memory()
    wins = {gid: torch.empty((VOR_DIM, VOR_DIM), dtype=torch.float64).pin_memory() for gid in (_abi.GRID_H_PROX, _abi.GRID_MOTION_RES)}
    h_sites = np.empty((VOR_DIM * VOR_DIM, 2), np.int32)
    return dict(ctx=ctx, cc=cc, astar=astar, patch=patch, h_win=h_win, win_origin=wo, ego=ego, goal=goal, vor=vor,
                n_sites=len(sites), outs=outs, rec=rec, wins=wins, h_sites=h_sites,
                vor_origin=(max(ego[0] - VOR_DIM // 2, 0), max(ego[1] - VOR_DIM // 2, 0)))


def step_resident(env, n_closed=None):
    """device-resident build: safety + pool, potential field (end points of the sweep given), heuristic with prox"""
    from ga_b200 import _abi
    ctx = env["ctx"]
    lib, h = ctx.lib, ctx.h
    v = env["vor"]
    ctx.check(lib.ga_build_safety(h))
    ctx.check(lib.ga_build_potential(h, env["ego"][0], env["ego"][1], _abi.ptr(v) if len(v) else None, len(v), None))
    ctx.check(lib.ga_build_heuristic(h, 0, env["goal"][0], env["goal"][1], n_closed))


def _e2e_front(env, sweep_s):
    """window in, safety, sites out, host sweep, potential field, heuristic -- everything up to the result copies"""
    from ga_b200 import _abi
    ctx = env["ctx"]
    lib, h = ctx.lib, ctx.h
    hw = env["h_win"]
    ctx.check(lib.ga_patch_insert(h, C.c_void_p(hw.data_ptr()), env["win_origin"][0], env["win_origin"][1], WIN_DIM))
    ctx.check(lib.ga_build_safety(h))
    n = C.c_int()
    ctx.check(lib.ga_potential_sites(h, env["ego"][0], env["ego"][1], _abi.ptr(env["h_sites"]), VOR_DIM * VOR_DIM, C.byref(n)))
    t0 = time.perf_counter()
    v = host_sweep(env["h_sites"][:n.value])
    sweep_s[0] += time.perf_counter() - t0
    ctx.check(lib.ga_build_potential(h, env["ego"][0], env["ego"][1], _abi.ptr(v) if len(v) else None, len(v), None))
    ctx.check(lib.ga_build_heuristic(h, 0, env["goal"][0], env["goal"][1], None))


def step_e2e(env, sweep_s, wait=True, lean=False, planes=False):
    from ga_b200 import _abi
    ctx = env["ctx"]
    lib, h = ctx.lib, ctx.h
    _e2e_front(env, sweep_s)
    ox, oy = env["vor_origin"]
    for gid, t in env["wins"].items():  # the window of h_prox_arr_ / motion_res_map_ the search reads cell by cell
        ctx.check(lib.ga_download_window(h, gid, ox, oy, VOR_DIM, VOR_DIM, C.c_void_p(t.data_ptr())))
    if lean:
        t = env["outs"][_abi.GRID_HEUR_FLAGS]
        ctx.check(lib.ga_download_async(h, _abi.GRID_HEUR_PACKED, C.c_void_p(t.data_ptr()), t.numel()))
    elif planes:
        for gid, t in env["outs"].items():
            ctx.check(lib.ga_download_async(h, gid, C.c_void_p(t.data_ptr()), t.numel() * t.element_size()))
    else:  # what the drop-in's closed sets read in place: the four planes as 12-byte records
        t = env["rec"]
        ctx.check(lib.ga_download_async(h, _abi.GRID_HEUR_REC, C.c_void_p(t.data_ptr()), t.numel()))
    if wait:
        ctx.sync()


E2E_H2D = WIN_DIM * WIN_DIM          # + the sweep's end points (a few KB, counted per run)
E2E_D2H_FULL = 2 * VOR_DIM * VOR_DIM * 8 + (PATCH_DIM // 4) ** 2 * 12
E2E_D2H_PLANES = 2 * VOR_DIM * VOR_DIM * 8 + (PATCH_DIM // 4) ** 2 * 21
E2E_D2H_LEAN = 2 * VOR_DIM * VOR_DIM * 8 + (PATCH_DIM // 4) ** 2


def cpu_reference_steps(dim, seeds, full_grid=True, repeats=1):
    """the same environment build through the reference's OWN code (oracle/_ref: collision_checking.cpp, a_star.cpp,
    VoronoiDiagramGenerator.cpp ... compiled unmodified; cv::dilate / distanceTransform / cuDNN pooling restated in
    oracle/ref_shim), one thread -- the reference keeps its planner state in static members, one planner per process.
    full_grid: HEURISTIC_EARLY_EXIT off, the whole reachable grid as the GPU builds; else the shipped config.yml
    (early exit, MAX_EXTRA_NODES_ASTAR 10000).  Returns dict(seconds, cells, stage seconds)."""
    import ga_b200
    import ga_ref_py as gr
    if not gr.available():
        raise RuntimeError("oracle/_ref/libga_ref.so is not built")
    gr.init(dim)
    if full_grid:
        gr.set_heuristic_config(early_exit=False, max_extra_nodes=1 << 30, unknown_cost_w=0.0)
    else:
        gr.set_heuristic_config(early_exit=True, max_extra_nodes=10000, unknown_cost_w=0.0)
    st = dict(insert=0.0, dilate=0.0, pool=0.0, potential=0.0, heuristic=0.0)
    cells = 0
    for seed in seeds:
        patch = ga_b200.gen_map(dim, seed)
        gr.set_grid(gr.PATCH, patch)
        gr.process_safety_patch()
        gr.calc_astar_grid()
        ego, goal = pick_points(gr.get_grid(gr.ASTAR), seed + 1)
        wo = (min(max(ego[0] * 4 - WIN_DIM // 2, 0), dim - WIN_DIM), min(max(ego[1] * 4 - WIN_DIM // 2, 0), dim - WIN_DIM))
        win = np.ascontiguousarray(patch[wo[1]:wo[1] + WIN_DIM, wo[0]:wo[0] + WIN_DIM])
        for _ in range(repeats):
            t0 = time.perf_counter()
            gr.pass_local_map(win, wo[0], wo[1])            # collision_checking.cpp:269-313
            t1 = time.perf_counter()
            gr.process_safety_patch()                       # :193-224
            t2 = time.perf_counter()
            gr.calc_astar_grid()                            # a_star.cpp:536-542 (CPU max in place of cuDNN)
            t3 = time.perf_counter()
            gr.calc_voronoi_potential_field(ego, 1)         # :440-479, incl. the Fortune sweep
            t4 = time.perf_counter()
            r = gr.calc_distance_heuristic(goal, ego, for_path=False)  # :110-295
            t5 = time.perf_counter()
            for k, d in zip(("insert", "dilate", "pool", "potential", "heuristic"), (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4)):
                st[k] += d
            cells += r["n_closed"]
    return dict(seconds=sum(st.values()), cells=cells, stages=st)


def _ref_worker(a):
    dim, seeds, full_grid = a
    return cpu_reference_steps(dim, seeds, full_grid)


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the step at the same patch size, one planner
    process per host core (its only way to use more than one core); each step = cores x one environment build."""
    if rank != 0:
        return
    import multiprocessing as mp
    dim = args.ref_dim if args.ref_dim > 0 else PATCH_DIM
    # the whole --steps/--warmup run has to end within a few minutes: ~1.5 s per 4096^2 build
    est = {4096: 1.6, 2048: 0.4, 1024: 0.1}.get(dim, 1.6)
    if args.ref_dim <= 0 and est * (args.steps + args.warmup) > 420:
        dim = 2048 if 0.4 * (args.steps + args.warmup) <= 420 else 1024
    cores = max(1, min(len(os.sched_getaffinity(0)), args.ref_procs))
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        def one_step(i, full):
            # the processes run concurrently; the step lasts as long as the slowest planner's calls (map generation
            # and process start-up are not the reference's work and stay outside)
            out = pool.map(_ref_worker, [(dim, [20240 + (i * cores + w) % 64], full) for w in range(cores)])
            return max(o["seconds"] for o in out), sum(o["cells"] for o in out), out[0]["stages"]
        for i in range(args.warmup):
            one_step(i, True)
        total, cells = 0.0, 0
        stages = None
        for i in range(args.steps):
            dt, n, stg = one_step(args.warmup + i, True)
            total += dt
            cells += n
            stages = stg if stages is None else {k: stages[k] + stg[k] for k in stg}
        # the shipped configuration (early exit after the ego cell + 10 000 nodes) on a few builds, beside it
        ship_t, ship_n, ship_stg = 0.0, 0, None
        for i in range(min(args.steps, 3)):
            dt, n, stg = one_step(1000 + i, False)
            ship_t += dt
            ship_n += n
            ship_stg = stg if ship_stg is None else {k: ship_stg[k] + stg[k] for k in stg}
    val = cells / total / 1e6
    ms = lambda d, n: {k + "_ms": 1e3 * v / n for k, v in d.items()}
    sample = "%d step(s); one step = %d planner processes x one environment build (window insert + threshold + dilate + pool " \
             "+ potential field incl. Fortune sweep + heuristic, whole reachable grid) on a %dx%d patch" % (args.steps, cores, dim, dim)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic-4096", "patch_dim": PATCH_DIM, "astar_dim": PATCH_DIM // 4,
                       "sample_patch_dim": dim, "same_config": dim == PATCH_DIM},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample,
                             "cpu_model": cpu_model(),
                             "stages_per_build_one_process": ms(stages, args.steps),
                             "dilate_share": stages["dilate"] / max(sum(stages.values()), 1e-12),
                             "shipped_config": {"note": "HEURISTIC_EARLY_EXIT True, MAX_EXTRA_NODES_ASTAR 10000 (config.yml:47-49)",
                                                "value": ship_n / max(ship_t, 1e-12) / 1e6, "ms_per_step": 1e3 * ship_t / max(min(args.steps, 3), 1),
                                                "cells_per_build": ship_n / max(min(args.steps, 3), 1) / cores,
                                                "stages_per_build_one_process": ms(ship_stg, max(min(args.steps, 3), 1))}},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_plan_leg(args, local_rank, rank):
    """metric M1, plan queries/s (BASELINE config 5's unit): independent start/goal queries on synthetic 2048^2 maps.
    Arm B = the reference's unmodified hybrid_a_star_lib.cpp on the drop-in classes over libgridaccel.so
    (oracle/_ref/libga_planb.so), P planner processes on this rank's GPU; on rank 0 also arm A = the pure reference
    planner, P processes on the host cores.  One planner per process (static state), so processes, not threads.
    A query = recalculateEnv (pool + potential field + guidance heuristic) + hybridAStarPlanning's search; the shipped
    2000 ms timeout applies.  queries/s = queries / the slowest process's summed query time."""
    import ga_plan_py as gp
    if not (os.path.exists(gp.LIBS["b200"]) and os.path.exists(gp.LIBS["ref"])):
        return {"unavailable": "oracle/_ref plan libraries not built (need /root/reference at build time)"}
    from ga_b200.sharding import shard_queries
    procs, nq = args.plan_procs, args.plan_queries
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # the job: --plan-total queries (default: a bounded sample of procs x nq per GPU; 4096 = BASELINE config 5), whole
    # maps per rank (ga_b200/sharding.py), each rank's block split over its planner processes
    total = args.plan_total if args.plan_total > 0 else world * procs * nq
    ids = shard_queries(total, world, rank) if args.plan_total > 0 else np.arange(rank * procs * nq, (rank + 1) * procs * nq)
    blocks = [b for b in np.array_split(ids, procs) if len(b)]
    tmp = tempfile.mkdtemp(prefix="ga_plan_")

    # Several planner PROCESSES on one GPU only run concurrently under CUDA MPS (otherwise the driver time-slices whole
    # contexts and every launch waits for a slice).  One MPS control daemon per rank, restricted to the rank's GPU.
    mps_env, mps_note = None, ""
    if not args.no_mps and procs > 1:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = vis.split(",")[local_rank] if vis else str(local_rank)
        cand = dict(os.environ, CUDA_VISIBLE_DEVICES=phys, CUDA_MPS_PIPE_DIRECTORY=os.path.join(tmp, "mps_pipe"),
                    CUDA_MPS_LOG_DIRECTORY=os.path.join(tmp, "mps_log"))
        os.makedirs(cand["CUDA_MPS_PIPE_DIRECTORY"], exist_ok=True)
        os.makedirs(cand["CUDA_MPS_LOG_DIRECTORY"], exist_ok=True)
        try:
            r = subprocess.run(["nvidia-cuda-mps-control", "-d"], env=cand, capture_output=True, timeout=20)
            if r.returncode == 0:
                mps_env = cand
            else:
                mps_note = r.stderr.decode()[-200:]
        except (OSError, subprocess.TimeoutExpired):
            mps_env = None

    def run_arm(arm, tag):
        ps = []
        if arm == "b200" and mps_env is not None:
            # an MPS client names devices relative to the set its control daemon was started with (one GPU here)
            env = dict(mps_env, CUDA_VISIBLE_DEVICES="0", GA_DEVICE="0")
        else:
            env = dict(os.environ, GA_DEVICE=str(local_rank))
        for w, blk in enumerate(blocks):
            out = os.path.join(tmp, "%s_%d_%d.npz" % (tag, rank, w))
            cmd = [sys.executable, os.path.join(ROOT, "oracle", "ga_plan_py.py"), "--arm", arm, "--synthetic", "2048",
                   "--query-ids", str(int(blk[0])), str(int(blk[-1]) + 1), "--no-trace", "--timeout-ms", "2000", "--out", out]
            ps.append((subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE), out))
        tot, found, exp, env_ms, search_ms, n = [], 0, 0, 0.0, 0.0, 0
        for p, out in ps:
            err = p.communicate()[1]
            if p.returncode != 0:
                return {"error": err.decode()[-400:]}
            z = np.load(out)
            k = int(z["n_queries"])
            tot.append(sum(float(z["q%d_query_ms" % i]) for i in range(k)))
            for i in range(k):
                found += int(z["q%d_found" % i])
                exp += int(z["q%d_n_closed" % i])
                env_ms += float(z["q%d_env_ms" % i])
                search_ms += float(z["q%d_search_ms" % i])
            n += k
        return {"queries_per_s": n / (max(tot) * 1e-3), "queries": n, "processes": procs, "found": found,
                "expanded_nodes": exp, "env_ms_per_query": env_ms / n, "search_ms_per_query": search_ms / n}

    out = {"workload": "queries %d..%d of a %d-query job on synthetic 2048^2 maps (64 queries per map, shipped config), %d planner "
                       "processes on this rank" % (int(ids[0]) if len(ids) else 0, int(ids[-1]) if len(ids) else 0, total, len(blocks)),
           "b200": run_arm("b200", "b")}
    out["b200"]["mps"] = mps_env is not None
    if mps_note:
        out["b200"]["mps_note"] = mps_note
    if mps_env is not None:
        try:
            subprocess.run(["nvidia-cuda-mps-control"], input=b"quit\n", env=mps_env, capture_output=True, timeout=20)
        except (OSError, subprocess.TimeoutExpired):
            pass
    if rank == 0 and not args.no_cpu_baseline:
        out["cpu_reference"] = run_arm("ref", "a")
        out["cpu_reference"]["cores"] = procs
    return out


def run_batch_leg(ga, torch, device, rank, args, sampler, hbm_peak, nq=None):
    """BASELINE config 5's shape on one GPU: a batch of independent queries on 2048^2 maps (64 queries
    share a map), one heuristic grid per query, all grids relaxed by ONE kernel chain
    (ga_build_heuristic_batch).  Reports grids/s, Mcells/s and the algorithmic-bytes rate."""
    from ga_b200 import _abi
    nq, dim = (nq or args.batch), 2048
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
    reps = max(3, min(20, args.steps // 10)) if nq <= 256 else 3
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
            for tag, early in (("", False), ("_shipped_config", True)):
                gr.init(dim)
                gr.cart_reset(dim)
                gr.set_heuristic_config(early_exit=early, max_extra_nodes=10000 if early else 1 << 30, unknown_cost_w=0.0)
                L = gr.lib()
                ts = []
                for k in range(6):
                    wo = cycles[k] - origin - 400
                    ego = ((cycles[k][0] - origin[0]) // 4, (cycles[k][1] - origin[1]) // 4)
                    t0 = time.perf_counter()
                    pm = gr.raytrace(wins[k])
                    gr.cartograph(pm, wo[0], wo[1])
                    gr.cart_pass_local_map(wo[0], wo[1], 801)
                    gr.process_safety_patch()
                    gr.calc_astar_grid()
                    L.gar_calc_distance_heuristic(goal[0], goal[1], int(ego[0]), int(ego[1]), 0, 0)
                    ts.append(time.perf_counter() - t0)
                out["cpu_reference_ms_per_cycle" + tag] = 1e3 * float(np.median(ts[1:]))
            out["cpu_reference_sample"] = "5 cycles of the same chain through the reference's own code, 1 core, median; whole " \
                                          "reachable grid (as the GPU builds) and the shipped early-exit configuration"
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
    # kernel-level figure of the same batch: device time of check_poses_kernel alone (CUDA events around the launch)
    # against the bytes it must move -- 24 B pose in, one 32 B sector per disk lookup (random gathers), 4 B per segment out
    ctx.prof_reset()
    ctx.prof_enable(True)
    for _ in range(reps):
        cc.checkSegments(x, y, yaw, seg)
    ctx.sync()
    ctx.prof_enable(False)
    k_ms = ctx.prof_get(_abi.K_COLLISION)[0] / reps
    sector_bytes = n * (24 + 3 * 32) + 4 * (len(seg) - 1)
    hbm_peak = load_peaks()[0]
    out["check_poses_kernel"] = {"ms": k_ms, "pose_checks_per_s_device": n / (k_ms * 1e-3),
                                 "algorithmic_bytes_per_pose": 24 + 3 + 4 / seg_len, "sector_bytes_per_pose": 24 + 3 * 32 + 4 / seg_len,
                                 "achieved_gbs_sectors": sector_bytes / (k_ms * 1e-3) / 1e9,
                                 "frac_of_hbm_peak": sector_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak,
                                 "note": "L2 / DRAM random-access bound: 3 scattered byte gathers per pose, each a 32 B sector; the host "
                                         "API figure above is PCIe bound (24 B per pose over the link)"}
    # expansion batches
    cc.setExpandParams()
    nodes = np.zeros(4096, _abi.NODE_DTYPE)
    nodes["x"], nodes["y"] = rng.uniform(20, ext - 20, 4096), rng.uniform(20, ext - 20, 4096)
    nodes["yaw"], nodes["motion_res"], nodes["with_rear_axis"] = rng.uniform(-3.14, 3.14, 4096), 0.3125, 1
    succ = cc.expandNodes(nodes, exact=False)
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        succ = cc.expandNodes(nodes, exact=False)
    ctx.sync()
    dt = (time.perf_counter() - t0) / reps
    poses = int(succ["n_poses"].sum())
    ctx.prof_reset()
    ctx.prof_enable(True)
    for _ in range(reps):
        cc.expandNodes(nodes, exact=False)
    ctx.sync()
    ctx.prof_enable(False)
    e_ms = ctx.prof_get(_abi.K_EXPAND)[0] / reps
    out["expand_kernel"] = {"ms": e_ms, "nodes_per_s_device": len(nodes) / (e_ms * 1e-3), "pose_checks_per_s_device": poses / (e_ms * 1e-3),
                            "note": "fp64 rollout (sin / cos / tan per step) + 3 gathers per pose; compute bound on the fp64 pipe"}
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
    ap.add_argument("--ref-dim", type=int, default=0, help="patch size of the reference arm (0 = 4096, shrunk only if --steps makes the run too long)")
    ap.add_argument("--ref-procs", type=int, default=64, help="planner processes of the reference arm (capped by the host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-collision", action="store_true", help="skip the collision-checks/s leg")
    ap.add_argument("--no-replan", action="store_true", help="skip the unknown-environment cycle leg")
    ap.add_argument("--no-plan", action="store_true", help="skip the plan queries/s leg")
    ap.add_argument("--plan-procs", type=int, default=4, help="planner processes per GPU of the plan leg (fixed per GPU: weak scaling)")
    ap.add_argument("--plan-queries", type=int, default=6, help="queries per planner process of the plan leg")
    ap.add_argument("--plan-total", type=int, default=0, help="plan leg: total queries of the job over all ranks (4096 = BASELINE config 5); 0 = procs x queries per GPU")
    ap.add_argument("--e2e-threads", type=int, default=0, help="planner contexts / host threads of the e2e leg (0 = as many as --e2e-per-core allows)")
    ap.add_argument("--e2e-per-core", type=float, default=1.0, help="e2e planner threads per host core of this rank")
    ap.add_argument("--no-mps", action="store_true", help="plan leg: do not start a CUDA MPS daemon for the planner processes")
    ap.add_argument("--plan-only", action="store_true", help="run only the plan leg (development)")
    ap.add_argument("--batch", type=int, default=128, help="queries of the batched-heuristic leg (0 = skip)")
    ap.add_argument("--batch-large", type=int, default=1024, help="queries of the large batched-heuristic run on rank 0 (0 = skip)")
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

    # one host thread per planner context; the Fortune sweep (7 ms of host work per step) scales with the cores a rank has
    n_thr = max(N_ROT, min(args.e2e_threads if args.e2e_threads > 0 else 1 << 30, int(args.e2e_per_core * len(os.sched_getaffinity(0))) // max(world, 1)))
    if args.plan_only:
        plan = run_plan_leg(args, local_rank, rank)
        if world > 1:
            allp = [None] * world
            dist.all_gather_object(allp, plan)
            plan = {"per_rank": allp}
            dist.destroy_process_group()
        if rank == 0:
            print(json.dumps(plan), flush=True)
        return
    envs_e2e = [make_env(ga_b200, local_rank, 20240 + rank * 64 + i, torch) for i in range(n_thr)]
    envs = envs_e2e[:N_ROT]
    hbm_peak, peak_src = load_peaks()

    def barrier():
        for e in envs_e2e:
            e["ctx"].sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # cells closed per environment (fixed by the map)
    closed_e2e = []
    for e in envs_e2e:
        n = C.c_int()
        step_resident(e, C.byref(n))
        closed_e2e.append(n.value)
    closed = closed_e2e[:N_ROT]

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
        # steps run back to back on the device: step i+1's stream waits for step i's end event, so the span
        # first-start .. last-end is exactly K serialised steps
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

    def timed_e2e(steps, threaded, lean, planes=False):
        """host-clocked: every step copies its window in and its results out inside the timed region and waits for them.
        threaded: one host thread per planner context (its own stream and pinned buffers), as a planner service with a
        thread pool per GPU runs it: the Fortune sweeps of different contexts run on different host cores and one
        context's copies overlap the other contexts' kernels."""
        sweep_s = [0.0]
        for e in envs_e2e:
            step_e2e(e, sweep_s, wait=True, lean=lean, planes=planes)
        barrier()
        nenv = len(envs_e2e) if threaded else N_ROT
        per = [steps // nenv + (1 if i < steps % nenv else 0) for i in range(nenv)]
        sweeps = [[0.0] for _ in range(nenv)]

        def worker(i):
            for _ in range(per[i]):
                step_e2e(envs_e2e[i], sweeps[i], wait=True, lean=lean, planes=planes)

        sampler.active = True
        t0 = time.perf_counter()
        if threaded:
            ths = [threading.Thread(target=worker, args=(i,)) for i in range(nenv)]
            for th in ths:
                th.start()
            for th in ths:
                th.join()
        else:
            for i in range(steps):
                step_e2e(envs_e2e[i % N_ROT], sweeps[i % N_ROT], wait=True, lean=lean, planes=planes)
        barrier()
        wall = time.perf_counter() - t0
        sampler.active = False
        cells = sum(closed_e2e[i] * per[i] for i in range(nenv)) if threaded else sum(closed_e2e[i % N_ROT] for i in range(steps))
        return wall * 1e3, cells, sum(x[0] for x in sweeps) * 1e3

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev_ms, wall_ms, launches, cells = timed(step_resident, args.steps, False)
    # per-kernel device time: the same K steps once more with CUDA events around every launch on the library's
    # stream (kept out of the run `value` comes from: the event records cost a little)
    timed(step_resident, args.steps, True)
    kt = {}
    for k, name in ((_abi.K_DILATE, "safety"), (_abi.K_DT, "distance_transforms"), (_abi.K_POTENTIAL, "costmap"),
                    (_abi.K_HEUR_INIT, "heur_init"), (_abi.K_HEUR_RELAX, "heur_relax"), (_abi.K_HEUR_FINISH, "heur_finish")):
        ms = n = 0
        for e in envs:
            m_, n_ = e["ctx"].prof_get(k)
            ms += m_
            n += n_
        kt[name] = (ms, n)
    e2e_steps = max(20, args.steps // 2)
    e2e_steps = max(e2e_steps, 4 * n_thr)
    e2e_serial_ms, e2e_serial_cells, sweep_ms = timed_e2e(e2e_steps, False, False)
    e2e_wall_ms, e2e_cells, sweep_thr_ms = timed_e2e(e2e_steps, True, False)
    lean_wall_ms, lean_cells, _ = timed_e2e(e2e_steps, True, True)
    planes_wall_ms, planes_cells, _ = timed_e2e(e2e_steps, True, False, planes=True)
    batch = run_batch_leg(ga_b200, torch, local_rank, rank, args, sampler, hbm_peak) if args.batch > 0 else None
    # the same with 1024 queries in one chain (rank 0 only: it is a statement about the kernel, not about scaling)
    batch_large = (run_batch_leg(ga_b200, torch, local_rank, rank, args, sampler, hbm_peak, nq=args.batch_large)
                   if (rank == 0 and args.batch > 0 and args.batch_large > 0) else None)
    collision = run_collision_leg(envs[0], args, ga_b200) if (rank == 0 and not args.no_collision) else None
    replan = run_replan_leg(ga_b200, local_rank, args) if (rank == 0 and not args.no_replan) else None
    sampler.stop_flag = True
    for e in envs_e2e:
        e["ctx"].sync()
    plan = None
    if not args.no_plan:
        if world > 1:
            dist.barrier()
        plan = run_plan_leg(args, local_rank, rank)
        if world > 1:
            allp = [None] * world
            dist.all_gather_object(allp, plan)
            if rank == 0:
                plan["per_rank_b200"] = [p.get("b200") for p in allp]

    # max over ranks of the times, sum over ranks of the work
    plan_qps = plan["b200"].get("queries_per_s", 0.0) if plan and "b200" in plan else 0.0
    plan_t = (plan["b200"]["queries"] / plan_qps) if plan_qps else 0.0
    plan_q = float(plan["b200"]["queries"]) if plan_qps else 0.0
    batch_ms = batch["ms_per_batch"] if batch else 0.0
    t = torch.tensor([dev_ms, e2e_wall_ms, float(cells), float(e2e_cells), float(launches), e2e_serial_ms, lean_wall_ms,
                      float(lean_cells), plan_t, plan_q, batch_ms, float(args.batch if batch else 0), planes_wall_ms, float(planes_cells)],
                     dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms, e2e_wall_ms, e2e_serial_ms, lean_wall_ms, plan_t, batch_ms = (float(tmax[i]) for i in (0, 1, 5, 6, 8, 10))
        cells, e2e_cells, launches, lean_cells, plan_q = float(tsum[2]), float(tsum[3]), int(tsum[4]), float(tsum[7]), float(tsum[9])
        batch_n = float(tsum[11])
        planes_wall_ms, planes_cells = float(tmax[12]), float(tsum[13])
    else:
        batch_n = float(args.batch if batch else 0)
    value = cells / (dev_ms * 1e-3) / 1e6
    e2e_value = e2e_cells / (e2e_wall_ms * 1e-3) / 1e6

    if rank == 0:
        adim = PATCH_DIM // 4
        relax_ms, relax_n = kt["heur_relax"]
        dil_ms, dil_n = kt["safety"]
        # dominant kernel = the wavefront relaxation; algorithmic bytes = 39 B x astar cells of one grid per launch
        alg_bytes = BYTES_PER_CELL_HEUR * adim * adim
        achieved = alg_bytes / (relax_ms / max(relax_n, 1) * 1e-3) / 1e9
        n_vor = int(np.mean([len(e["vor"]) for e in envs]))
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "synthetic-4096", "patch_dim": PATCH_DIM, "astar_dim": adim,
                           "step": "safety + pool (4096^2) -> potential field (200^2 window, PROX) -> heuristic (1024^2, fixed point)",
                           "disk_r_c": envs[0]["cc"].disk_r_c_, "l2": "inputs rotate over %d environments per rank (~130 MB each) > L2" % N_ROT,
                           "cells_closed_per_step": cells / args.steps / world, "voronoi_end_points": n_vor,
                           "potential_sites": int(np.mean([e["n_sites"] for e in envs]))},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": E2E_H2D + 8 * n_vor,
                        "d2h_bytes_per_step": E2E_D2H_FULL + 8 * int(np.mean([e["n_sites"] for e in envs])),
                        "ms_per_step": e2e_wall_ms / e2e_steps, "steps": e2e_steps,
                        "what": "801^2 window in (ga_patch_insert), safety, sites out, Fortune sweep on the host, potential field, "
                                "heuristic, h_prox / motion_res windows + the closed set out as 12-byte records (GA_GRID_HEUR_REC: "
                                "cost, parent direction, unknown flag and step counts of every cell -- what the drop-in's closed "
                                "sets read in place; lossless against the four planes)",
                        "pipeline": "%d planner contexts, one host thread each (own stream and pinned buffers)" % n_thr,
                        "host_sweep_ms_per_step": sweep_ms / e2e_steps,
                        "host_sweep_ms_per_step_under_load": sweep_thr_ms / e2e_steps,
                        "serial_value": e2e_serial_cells / (e2e_serial_ms * 1e-3) / 1e6,
                        "serial_ms_per_step": e2e_serial_ms / e2e_steps,
                        "planes_value": planes_cells / (planes_wall_ms * 1e-3) / 1e6, "planes_ms_per_step": planes_wall_ms / e2e_steps,
                        "planes_d2h_bytes_per_step": E2E_D2H_PLANES,
                        "planes_what": "the same with the four planes (cost, cost_dist, parent, flags: 21 B per cell) instead of the records",
                        "lean_value": lean_cells / (lean_wall_ms * 1e-3) / 1e6, "lean_ms_per_step": lean_wall_ms / e2e_steps,
                        "lean_d2h_bytes_per_step": E2E_D2H_LEAN,
                        "lean_what": "the packed 1-byte plane (GA_GRID_HEUR_PACKED) instead of the four planes: a planner that "
                                     "takes per-successor costs from ga_expand's epilogue"},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "kernel": "heur_sweep_kernel", "achieved": achieved, "peak": hbm_peak,
                             "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": NCU_TRAFFIC_BYTES, "peak_source": peak_src,
                             "note": "dependency-depth bound, not bandwidth bound: one grid's wavefront keeps a few % of the "
                                     "worker warps busy; see `batch` for the same kernel fed with independent grids",
                             "algorithmic_bytes_per_launch": alg_bytes, "avg_launch_ms": relax_ms / max(relax_n, 1)},
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kt.items()},
                "safety_roofline": {"achieved": BYTES_PER_CELL_SAFETY * PATCH_DIM * PATCH_DIM /
                                    (dil_ms / max(dil_n, 1) * 1e-3) / 1e9, "unit": "GB/s", "peak": hbm_peak,
                                    "frac": BYTES_PER_CELL_SAFETY * PATCH_DIM * PATCH_DIM / (dil_ms / max(dil_n, 1) * 1e-3) / 1e9 / hbm_peak,
                                    "note": "safety_kernel (threshold + dilation + pool fused), 2+1/16 B per gm cell"},
                "clocks": sampler.result()}
        if batch:
            batch["grids_per_s_all_ranks"] = batch_n / (batch_ms * 1e-3)
            line["batch"] = batch
        if batch_large:
            batch_large.pop("grids_per_s_all_ranks", None)
            line["batch_large"] = batch_large
        if collision:
            line["collision"] = collision
        if replan:
            line["replan_cycle"] = replan
        if plan:
            if plan_t:
                plan["queries_per_s_all_ranks"] = plan_q / plan_t
            line["plan"] = plan
        if not args.no_cpu_baseline:
            # bounded sample of the same step on the host, one core: two builds at the full patch size
            r = cpu_reference_steps(PATCH_DIM, [20240], full_grid=True, repeats=2)
            line["cpu_baseline"] = {"value": r["cells"] / r["seconds"] / 1e6, "unit": UNIT, "cores": 1, "kind": "reference",
                                    "cpu_model": cpu_model(),
                                    "stages_ms_per_build": {k: 1e3 * v / 2 for k, v in r["stages"].items()},
                                    "sample": "2 environment builds on the 4096x4096 patch of seed 20240 through the reference's own "
                                              "code, whole reachable grid (%.1f s)" % r["seconds"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    for e in envs_e2e:
        e["ctx"].close()


if __name__ == "__main__":
    main()
