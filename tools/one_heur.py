"""One 4096^2 environment, a few heuristic builds (ncu target)."""
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R + '/guided-extended-hybrid-astar_b200']
import ga_b200
from ga_b200 import _abi

dim, seed = 4096, 20240
ctx = ga_b200.GridContext(0)
cfg = ga_b200.LibConfig()
astar = ga_b200.AStar(ctx, cfg)
cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
cc.initialize(dim)
ctx.upload(_abi.GRID_PATCH, ga_b200.gen_map(dim, seed))
cc.processSafetyPatch()
goal = ga_b200.sample_free_cells(astar.astar_grid_, 1, seed + 1)[0]
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    n = astar.calcDistanceHeuristic(goal, None, for_path=False)
print("closed", n)
ctx.close()
