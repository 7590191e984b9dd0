"""Generates tests/golden/scenario_<name>.npz: the ground-truth global map of a reference scenario,
preprocessed the way the reference's simulator does it, so that the GPU box (no /root/reference)
can run BASELINE config 1 (default SCENARIO_NAME map, ALL_VISIBLE = true).

Restates ros2/scripts/sim_util.py:20-81 (load_png -> pad_map(800) -> transform_binary_to_gm) and the
preset poses of ros2/scripts/simulation.py:236-260.  Run where /root/reference is mounted."""
import os
import sys

import cv2
import numpy as np
import scipy.signal

HERE = os.path.dirname(os.path.abspath(__file__))
ENV = "/root/reference/src/freespace_planner/ros2/scripts/sim_data/env_models"
UNKNOWN, OCC = 127, 255

PRESETS = {  # simulation.py:236-260 (yaw in degrees there)
    "Valet_parking_015": ((165.734262, 128.0, 0.0), (161.325933, 161.124533, 90.0)),
    "Free_015": ((165.734262, 128.0, 0.0), (161.325933, 161.124533, 90.0)),
    "Maze": ((130.0, 130.0, 0.0), (181.0, 226.0, 0.0)),
}


def load_png(name):  # sim_util.py:31-65
    img = cv2.imread(os.path.join(ENV, name + ".png"), cv2.IMREAD_UNCHANGED)
    if img is None:
        raise SystemExit("cannot read " + name)
    if name == "Maze":
        img = np.flipud(img[:, :, -1]).copy()
        img[0, :] = OCC
        img[-1, :] = OCC
        img[:, 0] = OCC
        img[:, -1] = OCC
        img[img >= UNKNOWN] = OCC
        img[img < UNKNOWN] = 0
        img = cv2.resize(img, (int(img.shape[0] * 1.5), int(img.shape[0] * 1.5)), interpolation=cv2.INTER_NEAREST)
        img = OCC - img
    elif img.ndim > 2:
        img = np.flipud(cv2.cvtColor(img[:, :, :3], cv2.COLOR_BGR2GRAY))
    else:
        img = np.flipud(img)
    return OCC - img


def pad_map(gt, pad):  # sim_util.py:68-81
    w = gt.shape[0]
    first = UNKNOWN * np.ones((pad, w))
    second = UNKNOWN * np.ones((w + 2 * pad, pad))
    return np.hstack((second, np.vstack((first, gt, first)), second)).astype(np.uint8)


def transform_binary_to_gm(binary):  # sim_util.py:20-28
    g = scipy.signal.convolve2d(binary, np.ones((3, 3)) * (1 / 9), fillvalue=UNKNOWN)
    g[g > 100] = 1.0
    g[g > 1] = 2
    g *= UNKNOWN
    g[g > UNKNOWN] = OCC
    return g


def main():
    for name in sys.argv[1:] or ["Valet_parking_015"]:
        gm = pad_map(load_png(name), 800)
        vis = transform_binary_to_gm(gm)  # ALL_VISIBLE = true (simulation.py:58-60); 'full' convolution: +2 cells
        assert set(np.unique(vis)) <= {0.0, 127.0, 255.0}
        start, goal = PRESETS[name]
        np.savez_compressed(os.path.join(HERE, "scenario_%s.npz" % name),
                            global_map_all_visible=vis.astype(np.uint8), global_map_gt=gm,
                            start_pose=np.array([start[0], start[1], np.deg2rad(start[2])]),
                            goal_pose=np.array([goal[0], goal[1], np.deg2rad(goal[2])]))
        print(name, gm.shape, vis.shape, np.unique(vis, return_counts=True))


if __name__ == "__main__":
    main()
