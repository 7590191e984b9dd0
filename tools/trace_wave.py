"""Per-tile event timeline of one heuristic build (development; needs a library built with
`make -C guided-extended-hybrid-astar_b200/csrc EXTRA=-DSW_TRACE`)."""
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
ctx.upload(_abi.GRID_PATCH, np.zeros((dim, dim), np.uint8) if len(sys.argv) < 2 else ga_b200.gen_map(dim, int(sys.argv[1])))
cc.processSafetyPatch()
goal = (0, 0)
lib = ctx.lib
lib.ga_heur_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint, C.POINTER(C.c_uint), C.c_int]
n = C.c_uint()
for rep in range(3):
    astar.calcDistanceHeuristic(goal, None, for_path=False)
lib.ga_heur_trace(ctx.h, None, 0, C.byref(n), 1)
astar.calcDistanceHeuristic(goal, None, for_path=False)
buf = np.zeros(1 << 20, np.uint64)
lib.ga_heur_trace(ctx.h, buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(n), 1)
ev = buf[:n.value]
t = (ev >> np.uint64(24)).astype(np.int64)
tile = ((ev >> np.uint64(12)) & np.uint64(0xFFF)).astype(int)
kind = ((ev >> np.uint64(8)) & np.uint64(15)).astype(int)
arg = (ev & np.uint64(255)).astype(int)
t0 = t.min()
order = np.argsort(t, kind="stable")
names = {1: "visit", 2: "staged", 3: "sweep", 4: "publish", 5: "refresh", 6: "wait", 7: "release", 8: "waited"}
print("events", n.value, "span %.1f us" % ((t.max() - t0) / 1e3))
tx = 35
for k in range(0, 6):
    tl = k * tx + k
    print("--- tile (%d,%d)" % (k, k))
    for i in order:
        if tile[i] == tl:
            a = arg[i]
            extra = {2: "run=%d fresh=%d" % (a & 1, a >> 1), 3: "dir=%d changed=%d" % (a & 3, a >> 2),
                     5: "from=%d improves=%d" % ((a & 7) - 1, a >> 3), 8: "news=%d" % a, 7: "any=%d" % a}.get(kind[i], "")
            print("  %8.2f us  %-8s %s" % ((t[i] - t0) / 1e3, names[kind[i]], extra))
# when did each diagonal tile make its LAST changing sweep?
last = {}
for i in order:
    if kind[i] == 3 and (arg[i] >> 2):
        last[tile[i]] = (t[i] - t0) / 1e3
print("last changing sweep of the diagonal tiles:", ["%.1f" % last.get(k * tx + k, -1) for k in range(0, 35, 2)])
ctx.close()
