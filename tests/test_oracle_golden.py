"""CPU: pins the oracle against the committed cv2 golden vectors (tests/golden/make_golden.py) and
against the reference's own known-answer tests for this path."""
import math
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv_golden.npz"))


@pytest.mark.parametrize("r", [1, 2, 3, 4, 7, 10, 11, 12, 16])
def test_ellipse_structuring_element(oracle, r):
    hw, se, taps = oracle.ellipse_rows(r)
    assert np.array_equal(se, G["se_%d" % r])
    assert taps == int(G["se_%d" % r].sum())
    if r == 10:  # 21x21 disk, 333 taps
        assert taps == 333
        assert list(2 * hw + 1) == [1, 9, 13, 15, 17, 19, 19, 21, 21, 21, 21, 21, 21, 21, 19, 19, 17, 15, 13, 9, 1]


@pytest.mark.parametrize("i", range(4))
def test_dilate_matches_cv2(oracle, i):
    out = oracle.dilate_disk(G["dil_in_%d" % i], int(G["dil_r_%d" % i]))
    assert np.array_equal(out, G["dil_out_%d" % i])


@pytest.mark.parametrize("i", range(4))
def test_distance_transform_variants_match_cv2(oracle, i):
    """16.16 variant: bit-exact everywhere.  IPP float variant: bit-exact for every distance below
    32 cells (20 m; float chamfer sums are exact there up to single roundings that any visiting
    order shares).  Beyond 32 cells the closed-source IPP kernel resolves a few exact-tie sums one
    float ulp higher than a plain raster chamfer (11 of 40 000 cells in golden 1, 36 in golden 3);
    nothing the planner reads reaches that far for obstacle distances (do_max 7 m, MAX_DIST_VAL
    10 m), the Voronoi distance can, with a relative effect of 1e-7 on h_prox."""
    src = G["dt_src_%d" % i]
    assert np.array_equal(oracle.distance_transform(src, oracle.DT_FIXEDPOINT), G["dt_fix_%d" % i])
    got, want = oracle.distance_transform(src, oracle.DT_IPPFLOAT), G["dt_ipp_%d" % i]
    near = want < 32.0
    assert np.array_equal(got[near], want[near])
    far_bad = got != want
    assert far_bad.mean() < 2e-3
    assert np.all(np.abs(got[far_bad] - want[far_bad]) <= np.spacing(want[far_bad]))


def test_threshold_order(oracle):
    v = np.arange(256, dtype=np.uint8)
    t = oracle.threshold(v, 100, 130)
    assert np.all(t[:100] == 0) and np.all(t[100:131] == 1) and np.all(t[131:] == 2)
    t = oracle.threshold(v, 140, 120)  # overlapping masks: setTo(OCC) runs after setTo(FREE)
    assert np.all(t[:121] == 0) and np.all(t[121:] == 2)


def test_pool_floor_semantics(oracle):
    s = np.zeros((10, 10), np.uint8)
    s[9, 9] = 2   # trailing 10 mod 4 rows / cols are ignored
    s[3, 7] = 1
    a = oracle.pool_max(s)
    assert a.shape == (2, 2) and a.tolist() == [[0, 1], [0, 0]]


def test_pass_local_map_drops_and_skips(oracle):
    p = np.full((20, 20), 127, np.uint8)
    w = np.arange(25, dtype=np.uint8).reshape(5, 5)
    oracle.pass_local_map(p, w, 17, 3)      # overhangs in x: dropped
    assert np.all(p == 127)
    oracle.pass_local_map(p, w, 3, 18)      # overhangs in y: rows skipped
    assert np.array_equal(p[18:20, 3:8], w[:2]) and np.all(p[:18] == 127)
    oracle.pass_local_map(p, w, 3, -2)
    assert np.array_equal(p[0:3, 3:8], w[2:])


def test_reference_collision_known_answers(oracle):
    """LIB/test/cpp/collision_checker_lib_test/test_collision_checking.cpp:9-30"""
    info, cen = oracle.calc_disks(3.0 + 1.5, 2.0, 1.5)
    safety = np.full((1000, 1000), 1, np.uint8)
    lib = oracle.lib()
    chk = lambda x, y, yi: lib.gao_check_grid(safety.ctypes.data, 1000, cen.ctypes.data, info.nb_disks, x, y, yi, 0)
    assert chk(500, 500, 0) == 1
    assert chk(0, 0, 0) == 0
    assert chk(999, 999, 0) == 0


def test_reference_angle_known_answers(oracle):
    """LIB/test/cpp/util_lib_test/test_util.cpp:10-155 (subset feeding getYawIdx)"""
    lib = oracle.lib()
    PI = math.pi
    assert lib.gao_angle_0_2pi(0.0) == 0.0
    assert abs(lib.gao_angle_0_2pi(-PI / 2) - 3 * PI / 2) < 1e-12
    assert abs(lib.gao_angle_0_2pi(5 * PI) - PI) < 1e-12
    assert abs(lib.gao_angle_mpi_pi(3 * PI / 2) + PI / 2) < 1e-12
    assert lib.gao_angle_mpi_pi(PI) == -PI
    assert abs(lib.gao_angle_diff(0.1, -0.1) + 0.2) < 1e-12
    assert oracle.yaw_idx(0.0) == 0 and oracle.yaw_idx(-1e-9) == 119 and oracle.yaw_idx(2 * PI - 1e-12) == 119
    assert oracle.yaw_idx(math.radians(3.0) + 1e-9) == 1


def test_default_footprint(oracle):
    """default vehicle -> 3 disks, disk_r_c 10, 120 yaw bins"""
    info, cen = oracle.calc_disks(3.993 + 1.17, 2.13, 1.17)
    assert (info.nb_disks, info.nb_yaws, info.disk_r_c) == (3, 120, 10)
    assert abs(info.disk_r - 1.4691914584892793) < 1e-15
    assert cen[0, :, 1].tolist() == [0, 0, 0]
    assert np.array_equal(cen[60, :, 0], -cen[0, :, 0])  # yaw = pi


def test_heuristic_reference_order_vs_fixed_point(oracle):
    """the restated reference order and the fixed point agree on an obstacle-free uniform grid and
    the fixed point is never worse anywhere"""
    import ga_b200
    dim = 96
    rng = np.random.default_rng(3)
    grid = np.zeros((dim, dim), np.uint8)
    for _ in range(25):
        x, y = rng.integers(0, dim, 2)
        grid[y:y + rng.integers(2, 12), x:x + rng.integers(2, 12)] = 2
    grid[36:50, 36:50] = 0
    mc = np.full((dim, dim), 5.0)
    prox = np.zeros((dim, dim))
    fence = np.zeros((dim, dim), np.uint8)
    hp = oracle.HeurParams(dim, 5.0, 3.0, 0.0, 0, 10000, 0)
    ref = oracle.heuristic_reforder(hp, grid, fence, mc, prox, (40, 40), (5, 5))
    fix = oracle.heuristic_fixedpoint(hp, grid, fence, mc, prox, (40, 40))
    m = ref["flags"] & 1 > 0
    assert np.array_equal(m, fix["flags"] & 1 > 0)
    assert np.all(fix["cost"][m] <= ref["cost"][m])
    assert (np.abs(fix["cost"][m] - ref["cost"][m]) > 1e-5 * np.maximum(fix["cost"][m], 1)).mean() < 0.05
    # early exit closes a subset
    hp2 = oracle.HeurParams(dim, 5.0, 3.0, 0.0, 1, 50, 0)
    ref2 = oracle.heuristic_reforder(hp2, grid, fence, mc, prox, (40, 40), (45, 45))
    m2 = ref2["flags"] & 1 > 0
    assert m2.sum() < m.sum() and np.all(m[m2])
    xs, ys = oracle.astar_path(ref["parent"], ref["flags"], 5, 5)
    assert (xs[-1], ys[-1]) == (40, 40) and len(xs) >= 36


def test_primitives(oracle):
    xs, ys, yaws = oracle.move_car_some_steps(10.0, 10.0, 0.3, 0.9375, 0.1, 1, 0.2, 3.165)
    assert len(xs) == 9
    x, y, yaw = 10.0, 10.0, 0.3
    for i in range(9):
        x += 0.1 * math.cos(yaw)
        y += 0.1 * math.sin(yaw)
        yaw += 0.1 * math.tan(0.2) / 3.165
        yaw = math.fmod(yaw + math.pi, 2.0 * math.pi) - math.pi  # constrainAngleMinPIPlusPi, util1.cpp:17-24
        assert (xs[i], ys[i], yaws[i]) == (x, y, yaw)
    y90 = oracle.turn_on_rear_axis(0.0, math.pi / 2, 3.0)
    assert len(y90) in (29, 30) and y90[-1] == math.pi / 2 and abs(y90[0] - math.radians(3.0)) < 1e-12
    st = oracle.steering_inputs(0.55, 13)
    assert st[0] == -0.55 and st[6] == 0.0 and st[12] == 0.55
