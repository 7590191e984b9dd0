"""Runs one workload through both plan arms (fresh process each) and prints where they agree / differ.
python tools/plan_compare.py [--scenario Valet_parking_015 | --synthetic 1024 --seed 50000 --queries 3] [--ushift] [--mode 0|1]"""
import argparse
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ga_plan_py as gp  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scenario")
ap.add_argument("--synthetic", type=int)
ap.add_argument("--seed", type=int, default=50000)
ap.add_argument("--queries", type=int, default=3)
ap.add_argument("--ushift", action="store_true")
ap.add_argument("--mode", type=int, default=0)
a = ap.parse_args()
if not a.scenario and not a.synthetic:
    a.scenario = "Valet_parking_015"
d = tempfile.mkdtemp()
kw = dict(scenario=a.scenario, synthetic=a.synthetic, seed=a.seed, queries=a.queries, ushift=a.ushift)
ra = gp.run_query_subprocess("ref", os.path.join(d, "a.npz"), extra=["--heuristic"] if a.scenario else [], **kw)
rb = gp.run_query_subprocess("b200", os.path.join(d, "b.npz"), mode=a.mode, extra=["--heuristic"] if a.scenario else [], **kw)
if "b200_stats" in rb:
    print("b200 host ms in submit / collect (wait + copy) / answer table:", (rb["b200_ns"] / 1e6).tolist())
    print("b200_stats (sync fans, cached checks, direct checks, poses, look-ahead launches, hits, re-checks, launches):", rb["b200_stats"].tolist())
for k in sorted(ra):
    if k not in rb:
        print("missing in b200:", k)
        continue
    x, y = ra[k], rb[k]
    if x.shape != y.shape:
        print("%-24s SHAPE %s vs %s" % (k, x.shape, y.shape))
    elif x.size <= 3:
        print("%-24s %s | %s %s" % (k, x.tolist(), y.tolist(), "" if np.array_equal(x, y, equal_nan=x.dtype.kind == "f") else "  <-- differs"))
    else:
        same = np.array_equal(x, y, equal_nan=x.dtype.kind == "f")
        extra = ""
        if not same:
            nd = np.count_nonzero(~((x == y) | ((x != x) & (y != y)) if x.dtype.kind == "f" else (x == y)))
            extra = "  <-- %d of %d differ" % (nd, x.size)
            if k.endswith("expanded"):
                first = int(np.argmax(np.any(x != y, axis=1)))
                extra += ", first at row %d" % first
        print("%-24s %s%s" % (k, "equal" if same else "DIFFERENT", extra))
