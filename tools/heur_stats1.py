"""one-case variant of heur_stats.py (development)"""
import ctypes as C, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R + '/guided-extended-hybrid-astar_b200']
import ga_b200
from ga_b200 import _abi
dim, seed = int(sys.argv[1]) if len(sys.argv) > 1 else 4096, 20240
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
ctx.prof_reset(); ctx.prof_enable(True)
reps = 10
for rep in range(reps):
    n = astar.calcDistanceHeuristic(goal, None, for_path=False)
ctx.prof_enable(False)
st = (C.c_longlong * 7)()
ctx.check(ctx.lib.ga_heur_stats(ctx.h, st))
t = {k: ctx.prof_get(k)[0] / reps for k in (_abi.K_HEUR_INIT, _abi.K_HEUR_RELAX, _abi.K_HEUR_FINISH)}
v = max(st[0], 1)
print("%s W=%s C=%s | visits %d inner %d improving %d | pop %.2f stage %.2f relax %.2f wb %.2f us/visit, %.3f us/inner | relax %.3f finish %.3f ms" %
      (os.environ.get("GA_HEUR_IMPL", "sweep"), os.environ.get("GA_SW_WARPS"), os.environ.get("GA_SW_CTAS"), st[0], st[1], st[2],
       *[st[k] / v / 1965.0 for k in (3, 4, 5, 6)], st[5] / max(st[1], 1) / 1965.0, t[_abi.K_HEUR_RELAX], t[_abi.K_HEUR_FINISH]), flush=True)
ctx.close()
