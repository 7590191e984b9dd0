"""Golden vectors of the proximity scoring from the REFERENCE's own code
(oracle/_ref/libga_ref.so: util::bilinearInterpolation, util::getBilinInterp, Vehicle's corners and
AStar::h_prox_arr_ compiled unmodified; HybridAStar::getProxOfCorners composed from them in
oracle/ref_shim/ref_api.cpp).  Run in the build container (needs /root/reference):
python tests/golden/make_ref_golden_prox.py
Inputs are regenerated from seeds by the tests, only outputs are stored."""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "guided-extended-hybrid-astar_b200")]

VEH = (0.55, 3.165, 3.993, 1.17, 2.13, False)
PATCH_DIM, ASTAR_DIM, SEED = 1000, 250, 4242
CON2STAR = 1 / 0.625
H_PROX_COST, INTERP_RES = 5.0, 0.1  # params.yml:8, config.yml:22
N_POSES, N_PATHS = 4000, 300


def inputs():
    """prox grid, poses [n, 3] (metres / rad), ragged path offsets (some paths empty)"""
    rng = np.random.default_rng(SEED)
    grid = rng.random((ASTAR_DIM, ASTAR_DIM))
    grid[rng.random((ASTAR_DIM, ASTAR_DIM)) < 0.3] = 0.0
    grid[3, 5] = np.nan
    lo, span = 6.0 / CON2STAR, (ASTAR_DIM - 13) / CON2STAR  # every corner inside the grid
    poses = np.stack([lo + rng.random(N_POSES) * span, lo + rng.random(N_POSES) * span,
                      (rng.random(N_POSES) - 0.5) * 4 * math.pi], 1)
    poses[:200, :2] = np.round(poses[:200, :2] * CON2STAR) / CON2STAR  # on cell centres: the no-interpolation branches
    cuts = np.sort(rng.integers(0, N_POSES + 1, N_PATHS - 1))
    cuts[10:14] = cuts[10]  # empty paths
    offsets = np.concatenate([[0], np.sort(cuts), [N_POSES]]).astype(np.int32)
    return grid, poses, offsets


CP_SEED, CP_CENTERS = 77, 8


def close_pose_inputs(astar):
    """centres (metres / rad) and their nodes_near_goal_ cell lists for the getValidClosePose rows: the
    non-occupied cells closer than 10 cells to the centre cell (a_star.cpp:182-193)"""
    rng = np.random.default_rng(CP_SEED)
    ad = astar.shape[0]
    free = np.argwhere(astar == 0)
    picks = [free[i] for i in rng.integers(0, len(free), CP_CENTERS - 2)] + [np.array([2, 3]), np.array([ad - 3, ad - 2])]
    ys, xs = np.mgrid[0:ad, 0:ad]
    out = []
    for cy, cx in picks:
        near = (np.sqrt((xs - cx) ** 2.0 + (ys - cy) ** 2.0) < 10.0) & (astar != 2)
        out.append(((cx * 0.625 + 0.1, cy * 0.625 + 0.2, float(rng.uniform(-3, 3))), (ys[near] * ad + xs[near]).astype(np.int32)))
    return out


if __name__ == "__main__":
    import ga_ref_py as gr
    from ga_b200.synth import gen_map
    grid, poses, offsets = inputs()
    gr.init(PATCH_DIM, veh=VEH)
    assert gr.dims()[1] == ASTAR_DIM and gr.con2star() == CON2STAR
    gr.set_grid(gr.H_PROX, grid)
    v = np.array([gr.prox_of_corners(x * CON2STAR, y * CON2STAR, w) for x, y, w in poses])
    costs = np.zeros(len(offsets) - 1)
    for i in range(len(costs)):
        c = 0.0
        for j in range(offsets[i], offsets[i + 1]):
            c += v[j] * H_PROX_COST * INTERP_RES  # hybrid_a_star_lib.cpp:592
        costs[i] = c
    # getValidClosePose step 3 on a synthetic map (safety patch through the reference's own processSafetyPatch)
    gr.pass_local_map(gen_map(PATCH_DIM, CP_SEED), 0, 0)
    gr.process_safety_patch()
    gr.calc_astar_grid()
    close = []
    for center, cells in close_pose_inputs(gr.get_grid(gr.ASTAR)):
        r = gr.valid_close_pose(center, cells)
        close.append(np.concatenate([[0.0], np.zeros(3)]) if r is None else np.concatenate([[1.0], r]))
    np.savez_compressed(os.path.join(HERE, "ref_golden_prox.npz"), prox=v, costs=costs, close=np.array(close))
    print("close poses found:", int(np.array(close)[:, 0].sum()), "of", len(close))
    print("ref_golden_prox.npz: %d poses, %d paths, max %.3f" % (len(v), len(costs), v.max()))
