"""Single-grid relaxation time of the bench's environments under different scheduling knobs (run on the GPU box).
usage: knob_sweep.py  -- prints ms per build of heur_sweep_kernel for each knob set, 4 environments of 4096^2 with prox."""
import ctypes as C, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R, R + '/guided-extended-hybrid-astar_b200', R + '/oracle']
import numpy as np
import torch
import ga_b200
from ga_b200 import _abi
import bench

envs = [bench.make_env(ga_b200, 0, 20240 + i, torch) for i in range(4)]
sets = [(-1, -1, -1, -1, -1, -1, -1)] + [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for k in sets:
    k7 = np.array(k, np.int32)
    tot, n = 0.0, 0
    for e in envs:
        ctx = e["ctx"]
        ctx.check(ctx.lib.ga_set_heur_knobs(ctx.h, _abi.ptr(k7)))
        for _ in range(3):
            bench.step_resident(e)
        ctx.sync(); ctx.prof_reset(); ctx.prof_enable(True)
        for _ in range(10):
            bench.step_resident(e)
        ctx.sync(); ctx.prof_enable(False)
        ms, cnt = ctx.prof_get(_abi.K_HEUR_RELAX)
        tot += ms; n += cnt
    print(k, "relax %.4f ms" % (tot / n), flush=True)
