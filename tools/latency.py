"""Per-call latency of the collision entry points at the batch sizes the serial search produces:
ga_check_poses for one expansion fan (26 segments) and ga_expand / ga_expand_ex for 1 .. 4096 nodes, next to the
reference's own loop on one host core when oracle/_ref is present.  python tools/latency.py [patch_dim]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "guided-extended-hybrid-astar_b200"), os.path.join(ROOT, "oracle")]
import ga_b200  # noqa: E402
from ga_b200 import _abi  # noqa: E402

dim = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
ctx = ga_b200.GridContext(0)
cfg = ga_b200.LibConfig()
astar = ga_b200.AStar(ctx, cfg)
veh = ga_b200.Vehicle.from_config(ga_b200.VehConfig())
cc = ga_b200.CollisionChecker(ctx, veh, cfg)
cc.initialize(dim)
patch = ga_b200.gen_map(dim, 50000)
ctx.upload(_abi.GRID_PATCH, patch)
cc.processSafetyPatch()
astar.calcDistanceHeuristic(ga_b200.sample_free_cells(astar.astar_grid_, 1, 3)[0], None, for_path=True)
hy = ga_b200.HybridAStar(ctx, veh)
cc.setExpandParams()
rng = np.random.default_rng(1)
ext = dim * 0.15625


def bench(fn, reps):
    fn()
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e6


print("patch %d^2" % dim)
for n_seg, per in ((1, 1), (26, 3), (26, 9), (32, 30), (256, 9), (2048, 9)):
    n = n_seg * per
    x, y, yaw = rng.uniform(20, ext - 20, n), rng.uniform(20, ext - 20, n), rng.uniform(-3, 3, n)
    seg = np.arange(0, n + 1, per, dtype=np.int32)
    us = bench(lambda: cc.checkSegments(x, y, yaw, seg), 300 if n < 5000 else 50)
    print("ga_check_poses  %5d segments x %2d poses: %8.1f us / call  (%.1f M pose checks/s)" % (n_seg, per, us, n / us))
for nodes_n in (1, 8, 64, 512, 4096):
    nodes = np.zeros(nodes_n, _abi.NODE_DTYPE)
    nodes["x"], nodes["y"] = rng.uniform(20, ext - 20, nodes_n), rng.uniform(20, ext - 20, nodes_n)
    nodes["yaw"], nodes["motion_res"], nodes["with_rear_axis"] = rng.uniform(-3, 3, nodes_n), 0.3125, 0
    us = bench(lambda: cc.expandNodes(nodes), 200 if nodes_n < 1000 else 30)
    us2 = bench(lambda: cc.expandNodesWithCosts(nodes), 200 if nodes_n < 1000 else 30)
    print("ga_expand %5d nodes: %8.1f us / call = %6.2f us / node;  ga_expand_ex: %8.1f us = %6.2f us / node"
          % (nodes_n, us, us / nodes_n, us2, us2 / nodes_n))
try:
    import ga_ref_py as gr
    if gr.available():
        gr.init(dim)
        gr.set_grid(gr.SAFETY, ctx.download(_abi.GRID_SAFETY))
        n = 26 * 3
        x, y, yaw = rng.uniform(20, ext - 20, n), rng.uniform(20, ext - 20, n), rng.uniform(-3, 3, n)
        seg = np.arange(0, n + 1, 3, dtype=np.int32)
        t0 = time.perf_counter()
        for _ in range(2000):
            gr.check_segments(x, y, yaw, seg)
        print("reference, 1 core: 26 segments x 3 poses through CollisionChecker::getPathCollisionIndex: %.1f us / fan"
              % ((time.perf_counter() - t0) / 2000 * 1e6))
except Exception as e:  # noqa: BLE001
    print("reference arm unavailable:", e)
ctx.close()
