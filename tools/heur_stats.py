"""Work counters and per-kernel device time of the heuristic build (run on the GPU box)."""
import ctypes as C
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R + '/guided-extended-hybrid-astar_b200']
import numpy as np
import ga_b200
from ga_b200 import _abi

cases = ((4096, 20240), (4096, 20241), (2048, 50000), (1492, 5))
for dim, seed in cases:
    ctx = ga_b200.GridContext(0)
    cfg = ga_b200.LibConfig()
    astar = ga_b200.AStar(ctx, cfg)
    cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
    cc.initialize(dim)
    ctx.upload(_abi.GRID_PATCH, ga_b200.gen_map(dim, seed))
    cc.processSafetyPatch()
    goal = ga_b200.sample_free_cells(astar.astar_grid_, 1, seed + 1)[0]
    for rep in range(3):
        n = astar.calcDistanceHeuristic(goal, None, for_path=False)
    ctx.prof_reset()
    ctx.prof_enable(True)
    reps = 10
    for rep in range(reps):
        n = astar.calcDistanceHeuristic(goal, None, for_path=False)
    ctx.prof_enable(False)
    st = (C.c_longlong * 7)()
    ctx.check(ctx.lib.ga_heur_stats(ctx.h, st))
    t = {k: ctx.prof_get(k)[0] / reps for k in (_abi.K_HEUR_INIT, _abi.K_HEUR_RELAX, _abi.K_HEUR_FINISH)}
    v = max(st[0], 1)
    print("   per visit: pop %.2f us stage %.2f us relax %.2f us wb %.2f us (at 1.965 GHz); relax per inner iter %.3f us" %
          tuple([st[k] / v / 1965.0 for k in (3, 4, 5, 6)] + [st[5] / max(st[1], 1) / 1965.0]))
    print("dim %d seed %d goal %s closed %d | visits %d inner %d improving %d | init %.3f relax %.3f finish %.3f ms" %
          (dim, seed, tuple(goal), n, st[0], st[1], st[2], t[_abi.K_HEUR_INIT], t[_abi.K_HEUR_RELAX], t[_abi.K_HEUR_FINISH]),
          flush=True)
    ctx.close()
