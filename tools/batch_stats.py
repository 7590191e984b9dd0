"""Batched heuristic build: per-kernel device time and work counters (run on the GPU box).
usage: batch_stats.py [n_queries] [patch_dim]"""
import ctypes as C, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R + '/guided-extended-hybrid-astar_b200']
import numpy as np
import ga_b200
from ga_b200 import _abi
nq = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
ctxs, astars, goals = [], [], []
for qi in range(nq):
    seed = 50000 + qi // 64
    ctx = ga_b200.GridContext(0)
    cfg = ga_b200.LibConfig()
    astar = ga_b200.AStar(ctx, cfg)
    cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
    cc.initialize(dim)
    if qi % 64 == 0:
        patch = ga_b200.gen_map(dim, seed)
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    if qi % 64 == 0:
        free = ga_b200.sample_free_cells(astar.astar_grid_, 64, seed + 1)
    ctxs.append(ctx); astars.append(astar); goals.append(free[qi % 64])
closed = ga_b200.AStar.calcDistanceHeuristicBatch(astars, goals, for_path=False)
lead = ctxs[0]
hs = (C.c_void_p * nq)(*[c.h for c in ctxs])
g = np.ascontiguousarray(goals, np.int32).reshape(nq, 2)
gp = g.ctypes.data_as(C.POINTER(C.c_int))
for _ in range(2):
    lead.check(lead.lib.ga_build_heuristic_batch(hs, nq, 0, gp, None))
lead.sync(); lead.prof_reset(); lead.prof_enable(True)
reps = 5
for _ in range(reps):
    lead.check(lead.lib.ga_build_heuristic_batch(hs, nq, 0, gp, None))
lead.sync(); lead.prof_enable(False)
st = (C.c_longlong * 7)()
lead.check(lead.lib.ga_heur_stats(lead.h, st))
t = {k: lead.prof_get(k)[0] / reps for k in (_abi.K_HEUR_INIT, _abi.K_HEUR_RELAX, _abi.K_HEUR_FINISH)}
tiles = nq * ((dim // 4 + 29) // 30) ** 2
v = max(st[0], 1)
print("nq %d dim %d | tiles %d visits %d (%.1f/tile) sweeps %d improving %d | pop %.1f stage %.2f relax %.2f wb %.2f us/visit | "
      "init %.3f relax %.3f finish %.3f ms | %.0f grids/s %.0f Mcells/s alg %.0f GB/s" %
      (nq, dim, tiles, st[0], st[0] / tiles, st[1], st[2], *[st[k] / v / 1965.0 for k in (3, 4, 5, 6)],
       t[_abi.K_HEUR_INIT], t[_abi.K_HEUR_RELAX], t[_abi.K_HEUR_FINISH], nq / sum(t.values()) * 1e3,
       closed.sum() / sum(t.values()) / 1e3, 39.0 * nq * (dim // 4) ** 2 / t[_abi.K_HEUR_RELAX] / 1e6), flush=True)
for c in ctxs:
    c.close()
