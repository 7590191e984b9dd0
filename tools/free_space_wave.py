"""Wave speed of the sweep kernel without obstacles (development): relax time on an empty 1024^2 grid
for a corner goal and a centre goal = tiles crossed x the per-tile hop cost with no bends."""
import ctypes as C, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R + '/guided-extended-hybrid-astar_b200']
import numpy as np
import ga_b200
from ga_b200 import _abi
dim = 4096
ctx = ga_b200.GridContext(0)
cfg = ga_b200.LibConfig()
astar = ga_b200.AStar(ctx, cfg)
cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
cc.initialize(dim)
for name, patch in (("empty", np.zeros((dim, dim), np.uint8)), ("synthetic", ga_b200.gen_map(dim, 20240))):
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    grid = astar.astar_grid_
    free = np.argwhere(grid == 0)
    for gname, goal in (("corner", free[0][::-1]), ("centre", free[len(free) // 2][::-1])):
        for rep in range(3):
            astar.calcDistanceHeuristic(goal, None, for_path=False)
        ctx.prof_reset(); ctx.prof_enable(True)
        for rep in range(10):
            n = astar.calcDistanceHeuristic(goal, None, for_path=False)
        ctx.prof_enable(False)
        st = (C.c_longlong * 7)()
        ctx.check(ctx.lib.ga_heur_stats(ctx.h, st))
        ms = ctx.prof_get(_abi.K_HEUR_RELAX)[0] / 10
        span = max(goal[0], 1023 - goal[0], goal[1], 1023 - goal[1]) / 30.0
        print("%s %s goal %s: relax %.3f ms, closed %d, visits %d sweeps %d | %.1f tiles to the far edge -> %.2f us per tile" %
              (name, gname, tuple(int(v) for v in goal), ms, n, st[0] // 13, st[1] // 13, span, ms * 1e3 / span), flush=True)
ctx.close()
