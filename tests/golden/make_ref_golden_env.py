"""Golden vectors of the unknown-environment rows from the REFERENCE's own
code (oracle/_ref/libga_ref.so = cartographing.cpp + gridmap_sim.cpp compiled unmodified).  Run in the
build container (needs /root/reference): python tests/golden/make_ref_golden_env.py
Inputs are regenerated from seeds by the tests (ga_b200.gen_map), only outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "guided-extended-hybrid-astar_b200")]
import ga_ref_py as gr  # noqa: E402
from ga_b200.synth import gen_map  # noqa: E402

WORLD_DIM, WORLD_SEED, DIM = 2600, 31, 1492
ORIGIN = (500, 400)


def cycle_inputs(world, cyc):
    """ego cell in the world, ground-truth window, window origin inside the patch"""
    ego = (ORIGIN[0] + 420 + 35 * cyc, ORIGIN[1] + 430 + 20 * cyc)
    gt = np.ascontiguousarray(world[ego[1] - 400:ego[1] + 401, ego[0] - 400:ego[0] + 401])
    return gt, (ego[0] - ORIGIN[0] - 400, ego[1] - ORIGIN[1] - 400)


def make_world():
    w = gen_map(WORLD_DIM, WORLD_SEED)
    return np.where(w == 255, 255, 0).astype(np.uint8)


if __name__ == "__main__":
    world = make_world()
    gr.init(DIM)
    gr.cart_reset(DIM)
    out = {"meta": np.array([WORLD_DIM, WORLD_SEED, DIM, ORIGIN[0], ORIGIN[1]])}
    for cyc in range(3):
        gt, wo = cycle_inputs(world, cyc)
        pm = gr.raytrace(gt)
        gr.cartograph(pm, wo[0], wo[1])
        gr.cart_pass_local_map(wo[0], wo[1], 801)
        out["pmap_%d" % cyc] = pm
    out["cart"] = gr.cart_get()
    out["patch"] = gr.get_grid(gr.PATCH)
    # a patch move: new origin 12.3 m / -7.9 m away, previous map carried over
    prev_utm, utm = (1000.0, 2000.0), (1012.3, 1992.1)
    gr.reinit(DIM)
    gr.cart_reset(DIM)
    gr.cart_load_prev(prev_utm, utm)
    out["moved_utm"] = np.array([prev_utm, utm])
    out["cart_moved"] = gr.cart_get()
    np.savez_compressed(os.path.join(HERE, "ref_golden_env.npz"), **out)
    print("written", os.path.getsize(os.path.join(HERE, "ref_golden_env.npz")) // 1024, "KiB")
