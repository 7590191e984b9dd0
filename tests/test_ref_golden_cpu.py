"""CPU: the plain-C oracle against tests/golden/ref_golden.npz -- outputs of the reference's OWN
code (tests/golden/make_ref_golden.py ran oracle/_ref/libga_ref.so).  Works without
/root/reference, so the oracle stays pinned on the GPU box too."""
import os

import numpy as np
import pytest

from ga_b200.synth import gen_map

R = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_golden.npz"))


def test_safety_pool_disks(oracle):
    dim, seed = (int(v) for v in R["A_dim_seed"])
    info, cen = oracle.calc_disks(3.993 + 1.17, 2.13, 1.17)
    assert np.array_equal(cen, R["A_centers"]) and info.disk_r_c == int(R["A_disk_r_c"])
    s = oracle.process_safety_patch(gen_map(dim, seed), 100, 130, info.disk_r_c)
    assert np.array_equal(s, R["A_safety"])
    assert np.array_equal(oracle.pool_max(s), R["A_astar"])


@pytest.mark.parametrize("tag,early", [("full", 0), ("early", 1)])
def test_heuristic_reference_order(oracle, tag, early):
    grid = R["A_astar"]
    ad = grid.shape[0]
    gx, gy, sx, sy = (int(v) for v in R["A_goal_start"])
    hp = oracle.HeurParams(ad, 5.0, 3.0, 0.0, early, 500, 0)
    got = oracle.heuristic_reforder(hp, grid, np.zeros((ad, ad), np.uint8), np.full((ad, ad), 5.0), np.zeros((ad, ad)),
                                    (gx, gy), (sx, sy))
    for k in ("cost", "dist", "parent", "flags"):
        assert np.array_equal(got[k], R["A_heur_%s_%s" % (tag, k)]), k
    xs, ys = oracle.astar_path(got["parent"], got["flags"], sx, sy)
    assert np.array_equal(np.stack([xs, ys], 1), R["A_heur_%s_path" % tag])


def test_pose_checks(oracle):
    p, seg = R["A_pose_xyyaw"], R["A_seg"]
    for v in (0, 1):
        got = oracle.check_segments(R["A_safety"], R["A_centers"], 3, p[:, 0], p[:, 1], p[:, 2], seg, v)
        assert np.array_equal(got, R["A_first_hit_v%d" % v])


@pytest.mark.parametrize("variant", [0, 1])
def test_potential_field(oracle, variant):
    grid = R["B_astar"]
    ad = grid.shape[0]
    pp = oracle.PotParams(100.0, 7.0, 3.0, 0.1, 1.0, 3.0, 10.0, 200, variant)
    od, vd = np.zeros((200, 200), np.float32), np.zeros((200, 200), np.float32)
    prox, mres = np.zeros((ad, ad)), np.full((ad, ad), 1.0)
    for rep in (0, 1):
        t = "B_v%d_r%d_" % (variant, rep)
        oracle.voronoi_potential_field(pp, grid, tuple(int(v) for v in R[t + "ego"]), R[t + "vor_xy"], od, vd, prox, mres)
        assert np.array_equal(od, R[t + "obstacle_dist"])
        assert np.array_equal(vd, R[t + "voronoi_dist"])
        assert np.array_equal(prox, R[t + "h_prox"], equal_nan=True)
        assert np.array_equal(mres, R[t + "motion_res"])


def test_motion_primitives(oracle):
    o = 0
    for x, y, yaw, mres, st, dr, n in R["C_prim_in"]:
        xs, ys, ws = oracle.move_car_some_steps(x, y, yaw, 0.9375, mres, int(dr), st, 3.165)
        n = int(n)
        assert len(xs) == n and np.array_equal(np.stack([xs, ys, ws], 1), R["C_prim_out"][o:o + n])
        o += n
    o = 0
    for yaw, delta, n in R["C_ra_in"]:
        n = int(n)
        assert np.array_equal(oracle.turn_on_rear_axis(yaw, delta, 3.0), R["C_ra_out"][o:o + n])
        o += n


def test_unknown_environment_rows(oracle):
    """tests/golden/ref_golden_env.npz (the reference's own gridmap_sim.cpp + cartographing.cpp, see
    make_ref_golden_env.py) against the oracle: ray-cast windows, accumulated map, collision-checker
    patch and the map carried over a patch move, bit for bit."""
    import importlib.util
    import os
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    spec = importlib.util.spec_from_file_location("make_ref_golden_env", os.path.join(here, "make_ref_golden_env.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    g = np.load(os.path.join(here, "ref_golden_env.npz"))
    world = gen.make_world()
    dim = gen.DIM
    cart = np.full((dim, dim), 127, np.uint8)
    patch = np.full((dim, dim), 127, np.uint8)
    for cyc in range(3):
        gt, wo = gen.cycle_inputs(world, cyc)
        pm = oracle.raytrace(gt)
        assert np.array_equal(pm, g["pmap_%d" % cyc])
        oracle.cartograph(cart, pm, wo[0], wo[1])
        oracle.pass_local_map(patch, cart[wo[1]:wo[1] + 801, wo[0]:wo[0] + 801], wo[0], wo[1])
    assert np.array_equal(cart, g["cart"]) and np.array_equal(patch, g["patch"])
    prev_utm, utm = g["moved_utm"]
    rnd = lambda v: int(np.copysign(np.floor(abs(v) + 0.5), v))  # noqa: E731  Point<double>::toInt = std::round
    nxt = np.full((dim, dim), 127, np.uint8)
    oracle.copy_patch_to_patch(cart, [rnd(v * 6.4) for v in prev_utm], nxt, [rnd(v * 6.4) for v in utm])
    assert np.array_equal(nxt, g["cart_moved"])


def _prox_golden():
    import importlib.util
    import os
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    spec = importlib.util.spec_from_file_location("make_ref_golden_prox", os.path.join(here, "make_ref_golden_prox.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    return gen, np.load(os.path.join(here, "ref_golden_prox.npz"))


def test_prox_scoring_rows(oracle):
    """tests/golden/ref_golden_prox.npz (the reference's own bilinear interpolation, corners and h_prox_arr_,
    see make_ref_golden_prox.py) and the known answers of LIB/test/cpp/util_lib_test/test_util.cpp:342-440
    against the oracle, bit for bit."""
    kat = [((0, 0, 0, 0), (0.0, 0.0), 0.0), ((0, 0, 2, 2), (1.0, 0.0), 2.0), ((0, 0, 2, 2), (0.5, 0.5), 1.0),
           ((0, 2, 0, 2), (0.0, 1.0), 2.0), ((0, 2, 0, 2), (1.0, 0.0), 0.0), ((0, 2, 0, 2), (0.5, 0.5), 1.0),
           ((0, 1, 1, 2), (0.0, 0.0), 0.0), ((0, 1, 1, 2), (1.0, 1.0), 2.0), ((0, 1, 1, 2), (0.5, 0.5), 1.0)]
    for q, (x, y), want in kat:
        assert oracle.bilinear(*q, 0, 1, 0, 1, x, y) == want
    gen, g = _prox_golden()
    grid, poses, offsets = gen.inputs()
    params = (gen.VEH[2], gen.VEH[3], gen.VEH[4], gen.CON2STAR, gen.H_PROX_COST, gen.INTERP_RES)
    assert np.array_equal(oracle.prox_of_corners(grid, params, poses), g["prox"])
    assert np.array_equal(oracle.path_prox_costs(grid, params, poses, offsets), g["costs"])


def test_valid_close_pose_rows(oracle):
    """the getValidClosePose rows of ref_golden_prox.npz (step 3 through the reference's own checkPose /
    calcIndex / getBilinInterp on the reference's own safety patch) against the oracle chain
    threshold -> dilation -> pool -> gao_valid_close_pose: the chosen poses bit for bit."""
    gen, g = _prox_golden()
    grid, _, _ = gen.inputs()
    info, cen = oracle.calc_disks(gen.VEH[2] + gen.VEH[3], gen.VEH[4], gen.VEH[3])
    safety = oracle.process_safety_patch(gen_map(gen.PATCH_DIM, gen.CP_SEED), 100, 130, info.disk_r_c)
    astar = oracle.pool_max(safety)
    params = (gen.VEH[2], gen.VEH[3], gen.VEH[4], gen.CON2STAR, gen.H_PROX_COST, gen.INTERP_RES)
    for (center, cells), want in zip(gen.close_pose_inputs(astar), g["close"]):
        got = oracle.valid_close_pose(safety, cen, 3, grid, params, cells, center)
        assert (got is not None) == bool(want[0])
        if got is not None:
            assert np.array_equal(got, want[1:])
