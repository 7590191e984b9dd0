import ctypes as C, os, sys, threading, time
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [R, R + '/guided-extended-hybrid-astar_b200', R + '/oracle']
import numpy as np, torch
import ga_b200
from ga_b200 import _abi
import bench
NT = int(sys.argv[1]) if len(sys.argv) > 1 else 16
envs = [bench.make_env(ga_b200, 0, 20240 + i, torch) for i in range(NT)]
names = ["insert+safety", "sites(sync)", "sweep", "potential+heur launch", "windows(sync)", "rec launch", "final sync"]
acc = [[0.0] * len(names) for _ in range(NT)]
def step(env, a):
    ctx = env["ctx"]; lib, h = ctx.lib, ctx.h
    t = [time.perf_counter()]
    hw = env["h_win"]
    ctx.check(lib.ga_patch_insert(h, C.c_void_p(hw.data_ptr()), env["win_origin"][0], env["win_origin"][1], bench.WIN_DIM))
    ctx.check(lib.ga_build_safety(h)); t.append(time.perf_counter())
    n = C.c_int()
    ctx.check(lib.ga_potential_sites(h, env["ego"][0], env["ego"][1], _abi.ptr(env["h_sites"]), bench.VOR_DIM ** 2, C.byref(n))); t.append(time.perf_counter())
    v = bench.host_sweep(env["h_sites"][:n.value]); t.append(time.perf_counter())
    ctx.check(lib.ga_build_potential(h, env["ego"][0], env["ego"][1], _abi.ptr(v) if len(v) else None, len(v), None))
    ctx.check(lib.ga_build_heuristic(h, 0, env["goal"][0], env["goal"][1], None)); t.append(time.perf_counter())
    ox, oy = env["vor_origin"]
    for gid, tt in env["wins"].items():
        ctx.check(lib.ga_download_window(h, gid, ox, oy, bench.VOR_DIM, bench.VOR_DIM, C.c_void_p(tt.data_ptr())))
    t.append(time.perf_counter())
    tt = env["rec"]
    ctx.check(lib.ga_download_async(h, _abi.GRID_HEUR_REC, C.c_void_p(tt.data_ptr()), tt.numel())); t.append(time.perf_counter())
    ctx.sync(); t.append(time.perf_counter())
    for i in range(len(names)):
        a[i] += t[i + 1] - t[i]
for e, a in zip(envs, acc):
    step(e, [0.0] * len(names))
PER = 8
def worker(i):
    for _ in range(PER):
        step(envs[i], acc[i])
t0 = time.perf_counter()
ths = [threading.Thread(target=worker, args=(i,)) for i in range(NT)]
[t.start() for t in ths]; [t.join() for t in ths]
wall = time.perf_counter() - t0
print("threads %d: %.3f ms per build overall, per thread %.2f ms per build" % (NT, wall * 1e3 / (NT * PER), wall * 1e3 / PER))
tot = np.array(acc).sum(0) / (NT * PER) * 1e3
for nme, v in zip(names, tot):
    print("  %-24s %.3f ms" % (nme, v))
