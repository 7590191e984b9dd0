"""Parity of the CUDA path (through the C ABI) with the CPU oracle, on a B200."""
import ctypes as C
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import ga_b200
from ga_b200 import _abi


def _setup(planner, patch_dim, patch):
    ctx, cc, astar = planner
    cc.initialize(patch_dim)
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    return ctx, cc, astar


# ---- K1 + K2: threshold, disk dilation, 4x4 pool -- bit-exact ---------------------------------

@pytest.mark.parametrize("dim,seed", [(64, 1), (333, 2), (500, 3), (1000, 4), (1492, 5), (2048, 6), (2050, 7), (4096, 20240)])
def test_safety_and_pool_bit_exact(planner, oracle, dim, seed):
    patch = ga_b200.gen_map(dim, seed)
    # sprinkle every u8 value so both thresholds are exercised at their edges
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, dim * dim, dim)
    patch.ravel()[idx] = rng.integers(0, 256, dim).astype(np.uint8)
    ctx, cc, astar = _setup(planner, dim, patch)
    want = oracle.process_safety_patch(patch, 100, 130, cc.disk_r_c_)
    got = cc.patch_safety_arr_
    assert np.array_equal(got, want)
    assert np.array_equal(astar.astar_grid_, oracle.pool_max(want))


@pytest.mark.parametrize("r", [0, 1, 3, 10, 16])
def test_dilation_radii(planner, oracle, r):
    ctx, cc, astar = planner
    dim = 300
    cc.calculateDisks()
    p = _abi.CollisionParams(100, 130, r, 3, cc.nb_disks_, 4)
    cen = np.ascontiguousarray(cc.disk_centers_, np.int32)
    ctx.check(ctx.lib.ga_set_collision_params(ctx.h, C.byref(p), _abi.ptr(cen)))
    cc.resetPatch(dim)
    rng = np.random.default_rng(r)
    patch = np.where(rng.random((dim, dim)) < 0.003, 255, 0).astype(np.uint8)
    patch[rng.random((dim, dim)) < 0.003] = 127
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    assert np.array_equal(cc.patch_safety_arr_, oracle.process_safety_patch(patch, 100, 130, r))


@pytest.mark.parametrize("lo,hi", [(0, 130), (0, 255), (1, 127), (100, 127), (100, 128), (128, 128), (129, 200), (200, 255), (255, 255)])
@pytest.mark.parametrize("dim", [320, 300, 301])
def test_threshold_edges(planner, oracle, lo, hi, dim):
    """MIN_THRESH / MAX_THRESH at the edges of the SWAR byte compare (0: nothing is below; 127 / 128: the carry bit of
    the compare; 255: nothing is above) on uniformly random bytes, for the 16-byte, 4-byte and byte-wise row modes
    (dim % 16 == 0, % 4 == 0, neither) -- collision_checking.cpp:213-220."""
    ctx, cc, astar = planner
    cc.calculateDisks()
    p = _abi.CollisionParams(lo, hi, 2, 3, cc.nb_disks_, 4)
    cen = np.ascontiguousarray(cc.disk_centers_, np.int32)
    ctx.check(ctx.lib.ga_set_collision_params(ctx.h, C.byref(p), _abi.ptr(cen)))
    cc.resetPatch(dim)
    rng = np.random.default_rng(lo * 256 + hi + dim)
    patch = rng.integers(0, 256, (dim, dim)).astype(np.uint8)
    patch[rng.random((dim, dim)) < 0.9] = max(lo - 1, 0)   # mostly FREE (when anything is), so that dilation has room
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    want = oracle.process_safety_patch(patch, lo, hi, 2)
    assert np.array_equal(cc.patch_safety_arr_, want)
    assert np.array_equal(astar.astar_grid_, oracle.pool_max(want))


@pytest.mark.parametrize("dim", [64, 1008, 2048])
def test_safety_staging_modes(planner, oracle, dim):
    """the tile's raw bytes by one bulk tensor copy (TMA, zero fill outside the image) or by per-thread cp.async: same bytes"""
    patch = ga_b200.gen_map(dim, dim)
    rng = np.random.default_rng(dim)
    patch[0, :] = rng.integers(0, 256, dim)      # the image borders are where the two stagings differ in mechanism
    patch[-1, :] = rng.integers(0, 256, dim)
    patch[:, 0] = rng.integers(0, 256, dim)
    patch[:, -1] = rng.integers(0, 256, dim)
    ctx, cc, astar = planner
    cc.initialize(dim)
    ctx.upload(_abi.GRID_PATCH, patch)
    want = oracle.process_safety_patch(patch, 100, 130, cc.disk_r_c_)
    for use_tma in (0, 1):
        ctx.check(ctx.lib.ga_set_safety_staging(ctx.h, use_tma))
        cc.processSafetyPatch()
        assert np.array_equal(cc.patch_safety_arr_, want), use_tma
        assert np.array_equal(astar.astar_grid_, oracle.pool_max(want)), use_tma


def test_empty_and_full_patches(planner, oracle):
    ctx, cc, astar = planner
    cc.initialize(256)
    cc.processSafetyPatch()  # fresh patch: all 127 -> all UNKNOWN
    assert np.all(cc.patch_safety_arr_ == 1) and np.all(astar.astar_grid_ == 1)
    ctx.upload(_abi.GRID_PATCH, np.zeros((256, 256), np.uint8))
    cc.processSafetyPatch()
    assert np.all(cc.patch_safety_arr_ == 0) and np.all(astar.astar_grid_ == 0)
    ctx.upload(_abi.GRID_PATCH, np.full((256, 256), 255, np.uint8))
    cc.processSafetyPatch()
    assert np.all(cc.patch_safety_arr_ == 2)


def test_safety_idempotent_and_monotone_4096(planner):
    """size-independent properties at the full config-2 size: dilation only grows values, a
    second build of the same patch gives the same bytes, pooled grid == block max."""
    dim = 4096
    patch = ga_b200.gen_map(dim, 20240)
    ctx, cc, astar = _setup(planner, dim, patch)
    s1 = cc.patch_safety_arr_
    cc.processSafetyPatch()
    assert np.array_equal(s1, cc.patch_safety_arr_)
    thr = np.where(patch > 130, 2, np.where(patch < 100, 0, 1)).astype(np.uint8)
    assert np.all(s1 >= thr)
    a = astar.astar_grid_
    assert np.array_equal(a, s1.reshape(1024, 4, 1024, 4).max(axis=(1, 3)))


# ---- K7: window insertion ------------------------------------------------------------------------

def test_patch_insert_semantics(planner, oracle):
    ctx, cc, astar = planner
    dim = 500
    cc.initialize(dim)
    want = np.full((dim, dim), 127, np.uint8)
    rng = np.random.default_rng(0)
    for (ox, oy, d) in [(10, 20, 100), (450, 10, 100), (-5, 10, 50), (100, -30, 80), (200, 460, 90), (0, 0, 500),
                        (300, 300, 1)]:
        win = rng.integers(0, 256, (d, d)).astype(np.uint8)
        cc.passLocalMap(win, (ox, oy), d)
        oracle.pass_local_map(want, win, ox, oy)
    assert np.array_equal(cc.patch_arr_, want)


# ---- K6: collision checks -- bit-exact ------------------------------------------------------------

def _random_poses(rng, n, patch_dim):
    ext = patch_dim * ga_b200.GM_RES
    x = rng.uniform(-2.0, ext + 2.0, n)
    y = rng.uniform(-2.0, ext + 2.0, n)
    yaw = rng.uniform(-3 * math.pi, 3 * math.pi, n)
    return x, y, yaw


@pytest.mark.parametrize("dim,seed,variant", [(333, 1, 0), (1492, 2, 0), (1492, 3, 1), (2048, 4, 0)])
def test_check_poses_bit_exact(planner, oracle, dim, seed, variant):
    patch = ga_b200.gen_map(dim, seed)
    ctx, cc, astar = _setup(planner, dim, patch)
    safety = cc.patch_safety_arr_
    rng = np.random.default_rng(seed)
    n = 200000
    x, y, yaw = _random_poses(rng, n, dim)
    # ragged segments, including empty ones
    cuts = np.sort(rng.integers(0, n + 1, 5000))
    seg = np.concatenate([[0], cuts, [n]]).astype(np.int32)
    got = cc.checkSegments(x, y, yaw, seg, variant)
    want = oracle.check_segments(safety, cc.disk_centers_, 3, x, y, yaw, seg, variant)
    assert np.array_equal(got, want)
    assert (want >= 0).any() and (want < 0).any()


def test_reference_known_answers(planner):
    """LIB/test/cpp/collision_checker_lib_test/test_collision_checking.cpp:9-30: vehicle
    (0.6, 2.7, 3.0, 1.5, 2.0), patch 1000, all-UNKNOWN grid."""
    ctx, cc, astar = planner
    cc.vehicle = ga_b200.Vehicle(0.6, 2.7, 3.0, 1.5, 2.0, False)
    cc.initialize(1000)
    assert cc.checkGrid((500, 500, 0)) is True
    assert cc.checkGrid((0, 0, 0)) is False
    assert cc.checkGrid((999, 999, 0)) is False


def test_single_pose_api(planner, oracle):
    patch = ga_b200.gen_map(333, 9)
    ctx, cc, astar = _setup(planner, 333, patch)
    safety = cc.patch_safety_arr_
    rng = np.random.default_rng(9)
    x, y, yaw = _random_poses(rng, 64, 333)
    for i in range(64):
        assert cc.checkPose((x[i], y[i], yaw[i])) == bool(oracle.check_pose(safety, cc.disk_centers_, 3, x[i], y[i], yaw[i], 0))
        assert cc.checkPoseVariant((x[i], y[i], yaw[i])) == bool(oracle.check_pose(safety, cc.disk_centers_, 3, x[i], y[i], yaw[i], 1))
    assert cc.getPathCollisionIndex(x, y, yaw) == int(oracle.check_segments(safety, cc.disk_centers_, 3, x, y, yaw, [0, 64])[0])
    assert cc.getPathCollisionIndex([], [], []) == -1


@pytest.mark.parametrize("ushift", [False, True])
def test_expand_matches_oracle(planner, oracle, ushift):
    dim = 1492
    patch = ga_b200.gen_map(dim, 11)
    ctx, cc, astar = _setup(planner, dim, patch)
    cc.setExpandParams()
    safety = cc.patch_safety_arr_
    rng = np.random.default_rng(12)
    n = 3000
    nodes = np.zeros(n, _abi.NODE_DTYPE)
    ext = dim * ga_b200.GM_RES
    nodes["x"] = rng.uniform(1.0, ext - 1.0, n)
    nodes["y"] = rng.uniform(1.0, ext - 1.0, n)
    nodes["yaw"] = rng.uniform(-math.pi, math.pi, n)
    nodes["motion_res"] = rng.choice([0.1, 0.25, 0.5, 0.9375, 1.0], n)
    nodes["with_rear_axis"] = 1 if ushift else 0
    got = cc.expandNodes(nodes)
    import ga_oracle_py as go
    ep = go.ExpandParams(cc.vehicle.max_steer_, cc.vehicle.w_b_, 0.625 * 1.5, 10.0, 13, int(ushift), 90.0, dim // 4)
    stride = got.shape[1]
    assert stride == 26 + 6
    n_flagged = 0
    for i in range(n):
        v, disc, end, npos = oracle.expand_node(safety, cc.disk_centers_, 3, ep, float(nodes["x"][i]), float(nodes["y"][i]),
                                                float(nodes["yaw"][i]), float(nodes["motion_res"][i]))
        k = len(v)
        g = got[i]
        assert np.array_equal(g["n_poses"][:k], npos)
        assert np.all(g["n_poses"][k:] == 0)
        # continuous end poses: device libm vs glibc, a few ulp
        assert np.allclose(np.stack([g["x"][:k], g["y"][:k], g["yaw"][:k]], 1), end, rtol=0, atol=1e-12)
        # every primitive, bit for bit: the ones the device flags near_cell_edge were rolled out again on the host
        # with the reference's arithmetic (CollisionChecker.expandNodes(exact=True))
        assert np.all(g["near_cell_edge"][:k] == 0)
        assert np.array_equal(g["valid"][:k], v)
        gd = np.stack([g["xi"][:k], g["yi"][:k], g["yawi"][:k]], 1)
        assert np.array_equal(gd, disc)
    raw = cc.expandNodes(nodes, exact=False)
    n_flagged = int((raw["near_cell_edge"] != 0).sum())
    assert n_flagged < 0.001 * n * stride
    # nodes ON cell edges (coordinates that are multiples of the cell size, axis-aligned yaws): many flagged primitives
    m = 256
    edge = np.zeros(m, _abi.NODE_DTYPE)
    edge["x"] = rng.integers(40, int(ext / 0.625) - 40, m) * 0.625
    edge["y"] = rng.integers(40, int(ext / 0.625) - 40, m) * 0.625
    edge["yaw"] = rng.choice([0.0, math.pi / 2, -math.pi / 2, math.pi / 6], m)
    edge["motion_res"] = rng.choice([0.1, 0.3125, 0.625], m)
    edge["with_rear_axis"] = 1 if ushift else 0
    got = cc.expandNodes(edge)
    assert int((cc.expandNodes(edge, exact=False)["near_cell_edge"] != 0).sum()) > 0
    for i in range(m):
        v, disc, end, npos = oracle.expand_node(safety, cc.disk_centers_, 3, ep, float(edge["x"][i]), float(edge["y"][i]),
                                                float(edge["yaw"][i]), float(edge["motion_res"][i]))
        k = len(v)
        g = got[i]
        assert np.array_equal(g["valid"][:k], v)
        assert np.array_equal(np.stack([g["xi"][:k], g["yi"][:k], g["yawi"][:k]], 1), disc)


# ---- K5: heuristic grid ---------------------------------------------------------------------------

def _heur_inputs(astar, ctx, rng, with_prox):
    dim = astar.astar_dim_
    grid = astar.astar_grid_
    mc = np.full((dim, dim), 5.0)
    lanes = rng.random((dim, dim)) < 0.05
    mc[lanes] = 3.0
    prox = np.zeros((dim, dim))
    if with_prox:
        prox = rng.random((dim, dim)) * (rng.random((dim, dim)) < 0.3)
    fence = (rng.random((dim, dim)) < 0.002).astype(np.uint8)
    ctx.upload(_abi.GRID_MOVEMENT_COST, mc)
    ctx.upload(_abi.GRID_H_PROX, prox)
    ctx.upload(_abi.GRID_GEOFENCE, fence)
    return grid, mc, prox, fence


@pytest.mark.parametrize("dim,seed,with_prox,unk_w", [(333, 1, False, 0.0), (1000, 2, True, 0.0), (1492, 3, True, 2.5),
                                                      (2048, 4, True, 0.0), (4096, 20240, True, 0.0)])
def test_heuristic_fixed_point(planner, oracle, dim, seed, with_prox, unk_w):
    patch = ga_b200.gen_map(dim, seed)
    ctx, cc, astar = planner
    astar.cfg.ASTAR_UNKNOWN_COST = unk_w
    astar.set_params()
    cc.initialize(dim)
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    rng = np.random.default_rng(seed)
    grid, mc, prox, fence = _heur_inputs(astar, ctx, rng, with_prox)
    goal = ga_b200.sample_free_cells(np.where(fence == 1, 2, grid), 1, seed)[0]
    n = astar.calcDistanceHeuristic(goal, None, for_path=False)
    import ga_oracle_py as go
    hp = go.HeurParams(astar.astar_dim_, 5.0, 3.0, unk_w, 0, 0, 0)
    want = oracle.heuristic_fixedpoint(hp, grid, fence, mc, prox, goal)
    cs = astar.closed_set_guidance_
    assert n == want["n_closed"] == cs.size()
    assert np.array_equal(cs.flags, want["flags"])
    assert np.array_equal(cs.cost, want["cost"])          # bit-exact (tolerance allowed: 1e-5 rel)
    assert np.array_equal(cs.parent, want["parent"])
    m = want["flags"] & 1 > 0
    assert np.array_equal(cs.dist[m], want["dist"][m])
    # path walk
    far = np.argwhere(m)[-1]
    xs, ys = astar.getAstarPath(int(far[1]), int(far[0]))
    oxs, oys = oracle.astar_path(want["parent"], want["flags"], int(far[1]), int(far[0]))
    assert np.array_equal(xs, oxs) and np.array_equal(ys, oys)
    assert (xs[-1], ys[-1]) == (goal[0], goal[1])


def test_heuristic_batch_matches_oracle(oracle):
    """ga_build_heuristic_batch: independent grids of different sizes relaxed by one kernel chain are
    bit-identical to the oracle's fixed point of each grid (and so to the one-by-one builds)."""
    import ga_oracle_py as go
    cases = [(512, 11, False, 0.0), (1000, 12, True, 0.0), (333, 13, True, 2.5), (2048, 14, False, 0.0), (640, 15, True, 0.0)]
    ctxs, astars, goals, wants = [], [], [], []
    try:
        for dim, seed, with_prox, unk_w in cases:
            ctx = ga_b200.GridContext(0)
            ctxs.append(ctx)
            cfg = ga_b200.LibConfig()
            cfg.ASTAR_UNKNOWN_COST = unk_w
            astar = ga_b200.AStar(ctx, cfg)
            cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
            cc.initialize(dim)
            ctx.upload(_abi.GRID_PATCH, ga_b200.gen_map(dim, seed))
            cc.processSafetyPatch()
            rng = np.random.default_rng(seed)
            grid, mc, prox, fence = _heur_inputs(astar, ctx, rng, with_prox)
            goal = ga_b200.sample_free_cells(np.where(fence == 1, 2, grid), 1, seed)[0]
            hp = go.HeurParams(astar.astar_dim_, 5.0, 3.0, unk_w, 0, 0, 0)
            wants.append(oracle.heuristic_fixedpoint(hp, grid, fence, mc, prox, goal))
            astars.append(astar)
            goals.append(goal)
        for rep in range(2):  # twice: the epoch-stamped tile flags and the ring must be reusable
            n = ga_b200.AStar.calcDistanceHeuristicBatch(astars, goals, for_path=bool(rep))
            for i, (astar, want) in enumerate(zip(astars, wants)):
                cs = astar.closed_set_path_ if rep else astar.closed_set_guidance_
                assert n[i] == want["n_closed"] == cs.size()
                assert np.array_equal(cs.cost, want["cost"])
                assert np.array_equal(cs.parent, want["parent"])
                assert np.array_equal(cs.flags, want["flags"])
                m = want["flags"] & 1 > 0
                assert np.array_equal(cs.dist[m], want["dist"][m])
        # a batch of three runs in the latency mode (<= 4 grids: four warps per tile, every changing sweep published)
        n3 = ga_b200.AStar.calcDistanceHeuristicBatch(astars[1:4], goals[1:4], for_path=False)
        for i, (astar, want) in enumerate(zip(astars[1:4], wants[1:4])):
            cs = astar.closed_set_guidance_
            assert n3[i] == want["n_closed"]
            assert np.array_equal(cs.cost, want["cost"]) and np.array_equal(cs.parent, want["parent"])
        # a single build on one of the contexts afterwards still works (own ring, own counters)
        assert astars[1].calcDistanceHeuristic(goals[1], None, for_path=False) == wants[1]["n_closed"]
        assert np.array_equal(astars[1].closed_set_guidance_.cost, wants[1]["cost"])
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("with_prox", [False, True])
def test_heuristic_vs_reference_order(planner, oracle, with_prox):
    """against the reference's own pop order (early exit off): the fixed point closes every cell the
    reference closes, is never worse, and agrees within 1e-5 relative wherever the reference itself
    is optimal.  The reference is an A* without re-opening whose Euclidean term is taken at the
    parent (a_star.cpp:260-263), so some of ITS cells are suboptimal: counted here."""
    dim = 1000
    patch = ga_b200.gen_map(dim, 77)
    ctx, cc, astar = _setup(planner, dim, patch)
    rng = np.random.default_rng(77)
    grid, mc, prox, fence = _heur_inputs(astar, ctx, rng, with_prox)
    if not with_prox:
        mc[:] = 5.0
        ctx.upload(_abi.GRID_MOVEMENT_COST, mc)
    goal = ga_b200.sample_free_cells(np.where(fence == 1, 2, grid), 1, 5)[0]
    start = ga_b200.sample_free_cells(np.where(fence == 1, 2, grid), 1, 6)[0]
    astar.calcDistanceHeuristic(goal, start, for_path=True, reference_order=False)
    cs = astar.closed_set_path_
    import ga_oracle_py as go
    hp = go.HeurParams(astar.astar_dim_, 5.0, 3.0, 0.0, 0, 10000, 0)
    ref = oracle.heuristic_reforder(hp, grid, fence, mc, prox, goal, start)
    m = ref["flags"] & 1 > 0
    assert np.all((cs.flags[m] & 1) > 0)
    assert np.all(cs.cost[m] <= ref["cost"][m])
    rel = (ref["cost"][m] - cs.cost[m]) / np.maximum(cs.cost[m], 1e-300)
    rel[cs.cost[m] == 0] = 0
    n_sub = int((rel > 1e-5).sum())
    print("reference-suboptimal cells (prox=%s): %d of %d, max rel %.3g" % (with_prox, n_sub, int(m.sum()), rel.max()))
    if not with_prox:
        assert n_sub < 0.02 * m.sum()
    # where the reference is optimal the two agree to the last bit or within 1e-5 relative
    assert np.all(rel[rel <= 1e-5] >= 0)


def test_heuristic_slots_independent_and_rebuild(planner, oracle):
    dim = 500
    patch = ga_b200.gen_map(dim, 8)
    ctx, cc, astar = _setup(planner, dim, patch)
    grid = astar.astar_grid_
    g0, g1 = ga_b200.sample_free_cells(grid, 2, 3)
    astar.calcDistanceHeuristic(g0, None, for_path=False)
    astar.calcDistanceHeuristic(g1, None, for_path=True)
    c0 = astar.closed_set_guidance_.cost.copy()
    c1 = astar.closed_set_path_.cost.copy()
    assert c0[g0[1], g0[0]] == 0 and c1[g1[1], g1[0]] == 0
    astar.calcDistanceHeuristic(g0, None, for_path=True)  # rebuild slot 1 with goal 0
    assert np.array_equal(astar.closed_set_path_.cost, c0)


# ---- K3 + K4: distance transforms + potential field ---------------------------------------------

@pytest.mark.parametrize("variant", [_abi.DT_FIXEDPOINT, _abi.DT_IPPFLOAT])
@pytest.mark.parametrize("ego", [(100, 100), (60, 250), (273, 273), (150, 180)])
def test_potential_field(planner, oracle, variant, ego):
    dim = 1492
    ctx, cc, astar = planner
    astar.cfg.DT_VARIANT = variant
    astar.set_params()
    cc.initialize(dim)
    patch = ga_b200.gen_map(dim, 21)
    ctx.upload(_abi.GRID_PATCH, patch)
    cc.processSafetyPatch()
    grid = astar.astar_grid_
    adim = astar.astar_dim_
    import ga_oracle_py as go
    pp = go.PotParams(100.0, 7.0, 3.0, 0.1, 1.0, 3.0, 10.0, 200, variant)
    o_od = np.zeros((200, 200), np.float32)
    o_vd = np.zeros((200, 200), np.float32)
    o_prox = np.zeros((adim, adim))
    o_mres = np.full((adim, adim), 1.0)
    rng = np.random.default_rng(ego[0])
    for rep in range(2):  # second round exercises the persistent voronoi_dist_ window
        sites = astar.potentialSites(ego)
        ox, oy = max(ego[0] - 100, 0), max(ego[1] - 100, 0)
        wy, wx = np.nonzero(grid[oy:oy + 200, ox:ox + 200] == 2)
        assert np.array_equal(sites, np.stack([wx, wy], 1))
        vor = rng.integers(0, 200, (300, 2)).astype(np.int32)
        had = astar.calcVoronoiPotentialField(ego, vor)
        want_had = oracle.voronoi_potential_field(pp, grid, ego, vor, o_od, o_vd, o_prox, o_mres)
        assert had == bool(want_had)
        assert np.array_equal(astar.obstacle_dist_, o_od)
        assert np.array_equal(astar.voronoi_dist_, o_vd)
        assert np.array_equal(astar.h_prox_arr_, o_prox, equal_nan=True)
        assert np.array_equal(astar.motion_res_map_, o_mres)


def test_potential_no_obstacle_early_return(planner, oracle):
    ctx, cc, astar = planner
    cc.initialize(1000)
    ctx.upload(_abi.GRID_PATCH, np.zeros((1000, 1000), np.uint8))
    cc.processSafetyPatch()
    before = astar.h_prox_arr_
    had = astar.calcVoronoiPotentialField((125, 125), np.zeros((0, 2), np.int32))
    assert had is False
    assert np.array_equal(astar.h_prox_arr_, before)
    assert np.all(astar.voronoi_dist_ == 0)


# ---- ABI behaviour --------------------------------------------------------------------------------

def test_error_conventions(planner):
    ctx, cc, astar = planner
    lib = ctx.lib
    assert lib.ga_build_safety(ctx.h) == -3          # GA_ERR_STATE: no patch yet
    assert b"no patch" in lib.ga_last_error(ctx.h)
    cc.initialize(128)
    assert lib.ga_build_heuristic(ctx.h, 0, 999, 0, None) == -1
    assert lib.ga_build_heuristic(ctx.h, 2, 0, 0, None) == -1
    with pytest.raises(ga_b200.GaError):
        ctx.upload(_abi.GRID_PATCH, np.zeros((64, 64), np.uint8))
    assert lib.ga_launch_count(ctx.h) >= 0


def test_two_contexts_are_independent(oracle):
    a = ga_b200.GridContext(0)
    b = ga_b200.GridContext(0)
    veh = ga_b200.Vehicle.from_config(ga_b200.VehConfig())
    out = []
    for ctx, seed, dim in ((a, 1, 400), (b, 2, 600)):
        ga_b200.AStar(ctx)
        cc = ga_b200.CollisionChecker(ctx, veh)
        cc.initialize(dim)
        ctx.upload(_abi.GRID_PATCH, ga_b200.gen_map(dim, seed))
        cc.processSafetyPatch()
        out.append(cc)
    for cc, seed, dim in ((out[0], 1, 400), (out[1], 2, 600)):
        assert np.array_equal(cc.patch_safety_arr_, oracle.process_safety_patch(ga_b200.gen_map(dim, seed), 100, 130, 10))
    a.close()
    b.close()


def _sensor_window(rng, dim):
    w = np.full((dim, dim), 127, np.uint8)
    m = rng.random((dim, dim))
    w[m < 0.45] = 255
    w[m > 0.9] = 0
    w[rng.integers(0, dim, 40), rng.integers(0, dim, 40)] = rng.integers(1, 255, 40)  # values cartograph must ignore
    return w


def test_cartographing_matches_oracle(planner, oracle):
    """ga_cart_* (Cartographing::resetPatch / cartograph / passLocalMap / loadPrevPatch) against the
    oracle, which tests/test_ref_pins_oracle.py pins on the reference's own cartographing.cpp:
    accumulated map and collision-checker patch bit for bit, windows overhanging every edge, patch
    moves in every direction and with changing patch sizes."""
    ctx, cc, astar = planner
    rng = np.random.default_rng(77)
    dim = 600
    cc.initialize(dim)
    cart_api = ga_b200.Cartographing(ctx)
    cart_api.resetPatch(dim)
    cart = np.full((dim, dim), 127, np.uint8)
    cc_patch = ctx.download(_abi.GRID_PATCH)
    assert np.all(cc_patch == 127)
    for ox, oy, wd in [(100, 120, 201), (150, 90, 201), (-30, 200, 101), (520, 300, 101), (250, -40, 151), (260, 500, 151),
                       (0, 0, 600), (399, 399, 201), (-700, 0, 50), (3, 5, 1)]:
        win = _sensor_window(rng, wd)
        cart_api.cartograph(win, (ox, oy))
        oracle.cartograph(cart, win, ox, oy)
        assert np.array_equal(cart_api.getMap(), cart)
    for ox, oy, wd in [(100, 120, 201), (0, 0, 600), (399, 399, 201), (10, 300, 300)]:
        cart_api.passLocalMap((ox, oy), wd)
        oracle.pass_local_map(cc_patch, cart[oy:oy + wd, ox:ox + wd], ox, oy)
        assert np.array_equal(ctx.download(_abi.GRID_PATCH), cc_patch)
    with pytest.raises(ga_b200.GaError):
        cart_api.passLocalMap((500, 0), 200)  # leaves the map: the reference's out-of-bounds read
    # the safety pipeline runs on what the cartographed map handed over
    cc.processSafetyPatch()
    assert np.array_equal(cc.patch_safety_arr_, oracle.process_safety_patch(cc_patch, 100, 130, cc.disk_r_c_))
    prev_utm = np.array([1000.0, 2000.0])
    c2g = 6.4
    for k, (shift, ndim) in enumerate([((12.3, -7.9), 600), ((-20.0, 15.5), 600), ((5.1, 5.1), 520), ((-3.3, 30.2), 700),
                                       ((200.0, 0.0), 600)]):
        utm = prev_utm + np.array(shift)
        cc.initialize(ndim)           # CollisionChecker::resetPatch + AStar::reinit of create_patch
        cart_api.resetPatch(ndim)
        cart_api.loadPrevPatch(prev_utm, utm)
        nxt = np.full((ndim, ndim), 127, np.uint8)
        oracle.copy_patch_to_patch(cart, [ga_b200.Cartographing._to_int(v * c2g) for v in prev_utm], nxt,
                                   [ga_b200.Cartographing._to_int(v * c2g) for v in utm])
        assert np.array_equal(cart_api.getMap(), nxt), k
        assert np.array_equal(ctx.download(_abi.GRID_PATCH), nxt), k
        cart, prev_utm = nxt, utm
        win = _sensor_window(rng, 201)
        cart_api.cartograph(win, (50 + 10 * k, 60))
        oracle.cartograph(cart, win, 50 + 10 * k, 60)
        assert np.array_equal(cart_api.getMap(), cart)


def test_raytracing_matches_oracle(planner, oracle):
    """ga_raytrace (GridMapSim::gmSimulation) against the oracle pinned on the reference's
    gridmap_sim.cpp: bit-exact windows for empty, sparse, walled, dense and full ground truths."""
    ctx, cc, astar = planner
    sim = ga_b200.GridMapSim(ctx)
    rng = np.random.default_rng(5)
    for density, walls in [(0.0, False), (0.0005, True), (0.01, True), (0.2, False), (1.0, False)]:
        gt = np.where(rng.random((801, 801)) < density, 255, 0).astype(np.uint8)
        if walls:
            gt[rng.integers(0, 801), 100:700] = 255
            gt[150:650, rng.integers(0, 801)] = 255
        got = sim.gmSimulation(gt)
        assert np.array_equal(got, oracle.raytrace(gt))
        assert np.array_equal(ctx.download(_abi.GRID_PMAP), got)
    gt = np.zeros((801, 801), np.uint8)
    gt[400, 400] = 255
    assert np.array_equal(sim.gmSimulation(gt), oracle.raytrace(gt))
    with pytest.raises(ValueError):
        sim.gmSimulation(np.zeros((800, 800), np.uint8))


def test_unknown_environment_cycle(planner, oracle):
    """BASELINE config 3, one vehicle driving through an unknown map: per cycle ray-cast the 801^2
    ground-truth window around the ego cell, merge it into the accumulated map, hand the window to the
    collision checker, rebuild safety + pooled grid (simulation.py:262-277, planning.py:595-603).
    Device-resident chain (ga_raytrace -> ga_cartograph_raytraced -> ga_cart_pass_local_map ->
    ga_build_safety) against the oracle's chain, every cycle, every grid, bit for bit."""
    ctx, cc, astar = planner
    dim = 1492
    world = ga_b200.gen_map(2600, 31)
    world = np.where(world == 255, 255, 0).astype(np.uint8)  # ground truth: occupied or free
    cc.initialize(dim)
    cart_api, sim = ga_b200.Cartographing(ctx), ga_b200.GridMapSim(ctx)
    cart_api.resetPatch(dim)
    cart = np.full((dim, dim), 127, np.uint8)
    patch = np.full((dim, dim), 127, np.uint8)
    patch_origin = np.array([500, 400])      # patch origin inside the world, gm cells
    for cyc in range(6):
        ego = patch_origin + np.array([420 + 35 * cyc, 430 + 20 * cyc])   # ego cell in the world
        gt = world[ego[1] - 400:ego[1] + 401, ego[0] - 400:ego[0] + 401]
        win_o = ego - patch_origin - 400      # window origin inside the patch
        sim.gmSimulation(gt, fetch=False)
        cart_api.cartographRaytraced(win_o)
        cart_api.passLocalMap(win_o, 801)
        cc.processSafetyPatch()
        pm = oracle.raytrace(gt)
        oracle.cartograph(cart, pm, win_o[0], win_o[1])
        oracle.pass_local_map(patch, cart[win_o[1]:win_o[1] + 801, win_o[0]:win_o[0] + 801], win_o[0], win_o[1])
        want = oracle.process_safety_patch(patch, 100, 130, cc.disk_r_c_)
        assert np.array_equal(cart_api.getMap(), cart), cyc
        assert np.array_equal(ctx.download(_abi.GRID_PATCH), patch), cyc
        assert np.array_equal(cc.patch_safety_arr_, want), cyc
        assert np.array_equal(astar.astar_grid_, oracle.pool_max(want)), cyc
    assert (cart != 127).mean() > 0.05


def test_prox_scoring_matches_oracle(planner, oracle):
    """ga_prox_of_corners / ga_path_prox_costs (HybridAStar::getProxOfCorners and the prox-cost loops of
    getRSPathCosts / getRAPathCosts) against the oracle pinned on the reference's util::getBilinInterp:
    a prox field built by the potential-field kernel's own format (zeros, NaN cells), poses on cell
    centres (no-interpolation branches), poses whose corners leave the grid (clamped on both sides),
    ragged paths with empty ones, an empty batch.  1e-12 relative: device sin / cos."""
    ctx, cc, astar = planner
    dim = 1492
    cc.initialize(dim)
    ad = astar.astar_dim_
    rng = np.random.default_rng(99)
    grid = rng.random((ad, ad)) * (rng.random((ad, ad)) < 0.6)
    grid[rng.integers(0, ad, 20), rng.integers(0, ad, 20)] = np.nan
    ctx.upload(_abi.GRID_H_PROX, grid)
    hy = ga_b200.HybridAStar(ctx, cc.vehicle, 5.0, 0.1)
    params = (cc.vehicle.lf_, cc.vehicle.lb_, cc.vehicle.width_, 1 / ga_b200.ASTAR_RES, 5.0, 0.1)
    n = 50000
    span = ad * ga_b200.ASTAR_RES
    poses = np.stack([(rng.random(n) * 1.1 - 0.05) * span, (rng.random(n) * 1.1 - 0.05) * span,
                      (rng.random(n) - 0.5) * 4 * np.pi], 1)
    poses[:2000, :2] = np.round(poses[:2000, :2] / ga_b200.ASTAR_RES) * ga_b200.ASTAR_RES
    poses[2000:2100, 2] = 0.0
    got = hy.getProxOfCornersBatch(poses)
    want = oracle.prox_of_corners(grid, params, poses)
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-14)
    assert hy.getProxOfCorners(poses[7]) == got[7]
    cuts = np.sort(rng.integers(0, n + 1, 999))
    cuts[100:110] = cuts[100]
    offsets = np.concatenate([[0], np.sort(cuts), [n]]).astype(np.int32)
    np.testing.assert_allclose(hy.getPathProxCosts(poses, offsets), oracle.path_prox_costs(grid, params, poses, offsets),
                               rtol=1e-12, atol=1e-13)
    assert len(hy.getProxOfCornersBatch(np.zeros((0, 3)))) == 0
    assert np.array_equal(hy.getPathProxCosts(np.zeros((0, 3)), [0, 0, 0]), [0.0, 0.0])
    with pytest.raises(ValueError):
        hy.getPathProxCosts(poses, [0, 5])
    with pytest.raises(ga_b200.GaError):
        ctx.check(ctx.lib.ga_path_prox_costs(ctx.h, _abi.ptr(poses), _abi.ptr(np.array([0, 9, 4], np.int32)), 2,
                                             _abi.ptr(np.zeros(2))))


def test_valid_close_pose_matches_oracle(planner, oracle):
    """ga_valid_close_pose (step 3 of HybridAStar::getValidClosePose) against the oracle pinned on the
    reference's parts: centres in the open, next to obstacles and at the patch border, with a zero prox
    field (every cost ties eight ways: the LAST candidate in loop order must win), a random one, an empty
    neighbourhood (-> the ego pose) and a neighbourhood that is entirely in collision."""
    ctx, cc, astar = planner
    dim = 1200
    cc.initialize(dim)
    cc.passLocalMap(ga_b200.gen_map(dim, 321), (0, 0), dim)
    cc.processSafetyPatch()
    grid, safety = astar.astar_grid_, cc.patch_safety_arr_
    ad = grid.shape[0]
    hy = ga_b200.HybridAStar(ctx, cc.vehicle, 5.0, 0.1)
    params = (cc.vehicle.lf_, cc.vehicle.lb_, cc.vehicle.width_, 1 / ga_b200.ASTAR_RES, 5.0, 0.1)
    rng = np.random.default_rng(11)
    ys, xs = np.mgrid[0:ad, 0:ad]
    free = np.argwhere(grid == 0)
    picks = [free[i] for i in rng.integers(0, len(free), 5)] + [np.array([1, 2]), np.array([ad - 2, ad - 3])]
    n_found = 0
    for prox in (np.zeros((ad, ad)), rng.random((ad, ad)) * (rng.random((ad, ad)) < 0.5)):
        ctx.upload(_abi.GRID_H_PROX, prox)
        for cy, cx in picks:
            near = (np.sqrt((xs - cx) ** 2.0 + (ys - cy) ** 2.0) < 10.0) & (grid != 2)
            cells = (ys[near] * ad + xs[near]).astype(np.int32)
            center = (cx * 0.625 + 0.3, cy * 0.625 + 0.05, float(rng.uniform(-3, 3)))
            want = oracle.valid_close_pose(safety, cc.disk_centers_, 3, prox, params, cells, center)
            got = hy.getValidClosePose(center, cells)
            assert (want is None) == (got is None)
            if want is not None:
                n_found += 1
                assert np.array_equal(got, want), (center, got, want)
    assert n_found >= 6
    ego = np.array([1.0, 2.0, 3.0])
    assert hy.getValidClosePose((50.0, 50.0, 0.0), np.zeros(0, np.int32), ego) is ego
    blocked = np.argwhere(grid == 2)
    cy, cx = blocked[len(blocked) // 2]
    cells = np.array([cy * ad + cx], np.int32)
    center = (cx * 0.625 + 0.3, cy * 0.625 + 0.3, 0.0)
    want = oracle.valid_close_pose(safety, cc.disk_centers_, 3, prox, params, cells, center)
    got = hy.getValidClosePose(center, cells)
    assert (want is None) == (got is None) and (want is None or np.array_equal(want, got))


def test_download_async_and_interleaved_contexts(oracle):
    """ga_download_async + ga_sync return what ga_download returns, also when three contexts (one stream
    each) have their uploads, builds and downloads in flight together (bench.py's pipelined e2e leg)."""
    import torch
    envs = []
    for k in range(3):
        ctx = ga_b200.GridContext(0)
        cfg = ga_b200.LibConfig()
        astar = ga_b200.AStar(ctx, cfg)
        cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
        cc.initialize(1024)
        patch = torch.from_numpy(ga_b200.gen_map(1024, 900 + k)).pin_memory()
        ctx.upload(_abi.GRID_PATCH, patch.numpy())
        cc.processSafetyPatch()
        goal = ga_b200.sample_free_cells(astar.astar_grid_, 1, 5 + k)[0]
        outs = {gid: torch.empty(256 * 256, dtype=dt).pin_memory()
                for gid, dt in ((_abi.GRID_HEUR_COST, torch.float64), (_abi.GRID_HEUR_PARENT, torch.int32))}
        envs.append((ctx, astar, patch, goal, outs))
    for rep in range(3):
        for ctx, astar, patch, goal, outs in envs:
            ctx.check(ctx.lib.ga_upload(ctx.h, _abi.GRID_PATCH, patch.data_ptr(), patch.numel()))
            ctx.check(ctx.lib.ga_build_safety(ctx.h))
            ctx.check(ctx.lib.ga_build_heuristic(ctx.h, 0, int(goal[0]), int(goal[1]), None))
            for gid, t in outs.items():
                ctx.check(ctx.lib.ga_download_async(ctx.h, gid, t.data_ptr(), t.numel() * t.element_size()))
    for ctx, astar, patch, goal, outs in envs:
        ctx.sync()
        for gid, t in outs.items():
            assert np.array_equal(t.numpy().reshape(256, 256), ctx.download(gid), equal_nan=True)
        fp = oracle.heuristic_fixedpoint(oracle.HeurParams(256, 5.0, 3.0, 0.0, 0, 0, 0), astar.astar_grid_,
                                         np.zeros((256, 256), np.uint8), np.full((256, 256), 5.0), np.zeros((256, 256)),
                                         (int(goal[0]), int(goal[1])))
        cost = outs[_abi.GRID_HEUR_COST].numpy().reshape(256, 256)
        closed = (fp["flags"] & 1) != 0
        assert np.array_equal(cost[closed], fp["cost"][closed])
        ctx.close()


# ga_set_heur_knobs: { warps per SM, idle back-off ns, early publications, lock-step rounds, halo refresh, linger ns, stage-ahead (+ 4: one warp per tile, + 8: two; default four in single builds) }
@pytest.mark.parametrize("knobs", [
    (8, 20, 0, 0, 0, 0, 0),            # plain: publish when converged
    (8, 20, 1, 0, 1, 0, 0),
    (8, 20, 1, 1, 1, 0, 0),            # throughput mode
    (8, 20, 255, 1, 2, 8000, 1),
    (8, 20, 255, 0, 2, 8000, 1),       # latency mode
    (1, 20, 255, 0, 2, 100000, 1),     # one worker per SM
    (2, 20, 3, 0, 2, 0, 1),
    (8, 20, 255, 0, 2, 8000, 5),       # latency mode without helper warps (stage-ahead knob + 4)
    (6, 20, 255, 0, 1, 0, 8),          # three leader / helper pairs
    (8, 20, 255, 0, 2, 8000, 9),       # leader / helper pairs instead of four warps per tile (stage-ahead knob + 8)
])
def test_heuristic_schedules_agree(oracle, knobs):
    """the sweep kernel's scheduling knobs (sweep order, when a tile publishes, in-place halo refresh, waiting,
    staging ahead, worker count) only change WHEN cells are relaxed: every combination must give the oracle's
    fixed point bit for bit, with prox and unknown-cell costs on (all four sweep_step variants)."""
    dim, seed = 1492, 3
    ctx = ga_b200.GridContext(0)
    k7 = np.array(knobs, np.int32)
    ctx.check(ctx.lib.ga_set_heur_knobs(ctx.h, _abi.ptr(k7)))
    cfg = ga_b200.LibConfig(ASTAR_UNKNOWN_COST=2.5)
    astar = ga_b200.AStar(ctx, cfg)
    cc = ga_b200.CollisionChecker(ctx, ga_b200.Vehicle.from_config(ga_b200.VehConfig()), cfg)
    cc.initialize(dim)
    ctx.upload(_abi.GRID_PATCH, ga_b200.gen_map(dim, seed))
    cc.processSafetyPatch()
    rng = np.random.default_rng(seed)
    grid, mc, prox, fence = _heur_inputs(astar, ctx, rng, True)
    goal = ga_b200.sample_free_cells(np.where(fence == 1, 2, grid), 1, seed)[0]
    hp = oracle.HeurParams(astar.astar_dim_, 5.0, 3.0, 2.5, 0, 0, 0)
    want = oracle.heuristic_fixedpoint(hp, grid, fence, mc, prox, goal)
    for rep in range(3):
        n = astar.calcDistanceHeuristic(goal, None, for_path=False)
        cs = astar.closed_set_guidance_
        assert n == want["n_closed"]
        assert np.array_equal(cs.flags, want["flags"])
        closed = (want["flags"] & 1) != 0
        assert np.array_equal(cs.cost[closed], want["cost"][closed])
        assert np.array_equal(cs.parent[closed], want["parent"][closed])
        assert np.array_equal(cs.dist[closed], want["dist"][closed])
    ctx.close()
