"""Generates tests/golden/ref_golden.npz by RUNNING THE REFERENCE'S OWN CODE
(oracle/_ref/libga_ref.so: collision_checking.cpp, a_star.cpp, vehicle.cpp, util*.cpp,
VoronoiDiagramGenerator.cpp compiled unmodified, see oracle/Makefile `ref`) on seeded synthetic
patches.  Run in the container that mounts /root/reference; the .npz is committed so that the GPU
box (no /root/reference) can check the CUDA path against outputs of the reference itself.
Inputs are regenerated from seeds by the tests (ga_b200.synth.gen_map), only outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "guided-extended-hybrid-astar_b200")]
import ga_ref_py as gr  # noqa: E402
from ga_b200.synth import gen_map  # noqa: E402

out = {}

# ---- case A: 600^2 patch, seed 31 -> safety, pooled grid, heuristic (both exits), pose checks
DIM, SEED = 600, 31
gr.init(DIM)
patch = gen_map(DIM, SEED)
gr.pass_local_map(patch, 0, 0)
gr.process_safety_patch()
gr.calc_astar_grid()
safety, grid = gr.get_grid(gr.SAFETY), gr.get_grid(gr.ASTAR)
out["A_dim_seed"] = np.array([DIM, SEED])
out["A_safety"] = safety
out["A_astar"] = grid
d = gr.disks()
out["A_centers"] = d["centers"]
out["A_disk_r_c"] = np.array(d["disk_r_c"])
rng = np.random.default_rng(SEED + 1)
free = np.argwhere(grid == 0)
(gy, gx), (sy, sx) = free[rng.integers(0, len(free))], free[rng.integers(0, len(free))]
out["A_goal_start"] = np.array([gx, gy, sx, sy])
for tag, early in (("full", False), ("early", True)):
    gr.set_heuristic_config(early, 500, 0.0)
    r = gr.calc_distance_heuristic((gx, gy), (sx, sy), for_path=False)
    for k in ("cost", "dist", "parent", "flags"):
        out["A_heur_%s_%s" % (tag, k)] = r[k]
    xs, ys = gr.astar_path(sx, sy)
    out["A_heur_%s_path" % tag] = np.stack([xs, ys], 1)
n = 6000
x, y = rng.uniform(-2, DIM * 0.15625 + 2, n), rng.uniform(-2, DIM * 0.15625 + 2, n)
yaw = rng.uniform(-7, 7, n)
seg = np.concatenate([[0], np.sort(rng.choice(np.arange(1, n), 800, replace=False)), [n]]).astype(np.int32)
out["A_pose_xyyaw"] = np.stack([x, y, yaw], 1)
out["A_seg"] = seg
out["A_first_hit_v0"] = gr.check_segments(x, y, yaw, seg, 0)
out["A_first_hit_v1"] = gr.check_segments(x, y, yaw, seg, 1)

# ---- case B: 1000^2 patch, seed 32 -> Voronoi potential field (both DT variants), twice
DIM, SEED = 1000, 32
for variant in (0, 1):
    gr.init(DIM)
    patch = gen_map(DIM, SEED)
    gr.pass_local_map(patch, 0, 0)
    gr.process_safety_patch()
    gr.calc_astar_grid()
    grid = gr.get_grid(gr.ASTAR)
    for rep, ego in enumerate([(120, 110), (131, 140)]):
        ox, oy = max(ego[0] - 100, 0), max(ego[1] - 100, 0)
        wy, wx = np.nonzero(grid[oy:oy + 200, ox:ox + 200] == 2)
        vor = gr.voronoi_endpoints(np.stack([wx, wy], 1))
        gr.calc_voronoi_potential_field(ego, variant)
        t = "B_v%d_r%d_" % (variant, rep)
        out[t + "ego"] = np.array(ego)
        out[t + "vor_xy"] = vor
        out[t + "obstacle_dist"] = gr.get_grid(gr.OBSTACLE_DIST)
        out[t + "voronoi_dist"] = gr.get_grid(gr.VORONOI_DIST)
        out[t + "h_prox"] = gr.get_grid(gr.H_PROX)
        out[t + "motion_res"] = gr.get_grid(gr.MOTION_RES)
out["B_dim_seed"] = np.array([DIM, SEED])
out["B_astar"] = grid

# ---- case C: motion primitives of the default vehicle
rng = np.random.default_rng(33)
steers = np.linspace(-0.55, 0.55, 13)
prim_in, prim_out = [], []
for i in range(64):
    x, y, yaw = rng.uniform(0, 60), rng.uniform(0, 60), rng.uniform(-np.pi, np.pi)
    mres, st, dr = rng.uniform(0.1, 1.0), steers[rng.integers(0, 13)], int(rng.choice([-1, 1]))
    xs, ys, ws = gr.move_car_some_steps(x, y, yaw, 0.9375, mres, dr, st)
    prim_in.append([x, y, yaw, mres, st, dr, len(xs)])
    prim_out.append(np.stack([xs, ys, ws], 1))
out["C_prim_in"] = np.array(prim_in)
out["C_prim_out"] = np.concatenate(prim_out)
ra_in, ra_out = [], []
for i in range(32):
    yaw = rng.uniform(-np.pi, np.pi)
    delta = np.radians(float(rng.choice([90, -90, 180, -180, 270, -270])))
    ws = gr.turn_on_rear_axis(1.0, 2.0, yaw, delta, 3.0)
    ra_in.append([yaw, delta, len(ws)])
    ra_out.append(ws)
out["C_ra_in"] = np.array(ra_in)
out["C_ra_out"] = np.concatenate(ra_out)

np.savez_compressed(os.path.join(HERE, "ref_golden.npz"), **out)
print("written", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "ref_golden.npz")) // 1024, "KiB")
