"""Round-2 entry points through the C ABI on the B200: the reference-order heuristic (GA_HEUR_REFORDER) on the
device-resident inputs, nodes_near_goal_, the packed plane, window downloads, the successor epilogue of ga_expand_ex,
the asynchronous expansion, and the parameter validation the advisor asked for."""
import ctypes as C
import math
import os

import numpy as np
import pytest

import ga_b200
from ga_b200 import _abi

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
R = np.load(os.path.join(HERE, "golden", "ref_golden.npz"))


def _env(planner, dim, seed, **cfg):
    ctx, cc, astar = planner
    for k, v in cfg.items():
        setattr(astar.cfg, k, v)
    astar.set_params()
    cc.initialize(dim)
    ctx.upload(_abi.GRID_PATCH, ga_b200.gen_map(dim, seed))
    cc.processSafetyPatch()
    return ctx, cc, astar


@pytest.mark.parametrize("dim,seed,early,extra,unk,prox,only_near", [
    (512, 7, True, 10000, 0.0, False, False),
    (512, 8, False, 10000, 0.0, True, False),
    (768, 9, True, 200, 2.5, True, False),
    (333, 10, True, 0, 0.0, False, False),
    (512, 12, True, 50, 0.0, False, True),
    (2048, 13, True, 10000, 0.0, True, False),
])
def test_heuristic_reference_order_on_device(planner, oracle, dim, seed, early, extra, unk, prox, only_near):
    """ga_build_heuristic_ex(GA_HEUR_REFORDER ...) == gao_heuristic_reforder (pinned bit for bit on the reference's
    a_star.cpp in tests/test_ref_pins_oracle.py, same five configurations): cost, cost_dist, parent, flags, the closing
    order and nodes_near_goal_, with the inputs taken from the device (pooled grid of K1, uploaded costs)."""
    ctx, cc, astar = _env(planner, dim, seed, ASTAR_UNKNOWN_COST=unk, HEURISTIC_EARLY_EXIT=early, MAX_EXTRA_NODES_ASTAR=extra)
    grid = astar.astar_grid_
    ad = grid.shape[0]
    rng = np.random.default_rng(seed + 100)
    free = np.argwhere(grid != 2)
    gy, gx = free[rng.integers(0, len(free))]
    sy, sx = free[rng.integers(0, len(free))]
    mc, px = np.full((ad, ad), 5.0), np.zeros((ad, ad))
    if prox:
        px = rng.uniform(0, 1, (ad, ad)) * (rng.uniform(0, 1, (ad, ad)) < 0.3)
        mc = np.where(rng.uniform(0, 1, (ad, ad)) < 0.1, 3.0, 5.0)
    ctx.upload(_abi.GRID_H_PROX, px)
    ctx.upload(_abi.GRID_MOVEMENT_COST, mc)
    n = astar.calcDistanceHeuristic((gx, gy), (sx, sy), for_path=True, get_only_near=only_near)
    import ga_oracle_py as go
    hp = go.HeurParams(ad, 5.0, 3.0, unk, int(early), extra, int(only_near))
    want = oracle.heuristic_reforder(hp, grid, np.zeros((ad, ad), np.uint8), mc, px, (gx, gy), (sx, sy))
    cs = astar.closed_set_path_
    assert n == want["n_closed"]
    for k in ("cost", "dist", "parent", "flags"):
        assert np.array_equal(getattr(cs, k), want[k]), k
    assert np.array_equal(astar.closing_order(), want["order"])
    near = astar.nodes_near_goal_
    assert [i for _, i in near] == want["near_ids"].tolist()
    assert [d for d, _ in near] == want["near_dist"].tolist()
    # the packed plane of the same slot
    packed = ctx.download(_abi.GRID_HEUR_PACKED + 1)
    closed = (want["flags"] & 1) != 0
    assert np.array_equal((packed & 16) != 0, closed) and np.array_equal((packed & 32) != 0, (want["flags"] & 2) != 0)
    mdx = np.array([1, 0, -1, 0, -1, -1, 1, 1]), np.array([0, 1, 0, -1, -1, 1, -1, 1])
    ys, xs = np.nonzero(closed & (want["parent"] >= 0))
    code = packed[ys, xs] & 15
    assert np.all(code < 8)
    assert np.array_equal((ys - mdx[1][code]) * ad + (xs - mdx[0][code]), want["parent"][ys, xs])
    assert (packed[gy, gx] & 15) == 8 and np.all((packed[~closed] & 15) == 15)


def test_heuristic_early_exit_equals_reference_golden(planner):
    """the committed outputs of the reference itself (tests/golden/ref_golden.npz, A_heur_early_*: HEURISTIC_EARLY_EXIT
    on, 500 extra nodes) reproduced bit for bit by the CUDA-side entry point"""
    dim, seed = (int(v) for v in R["A_dim_seed"])
    ctx, cc, astar = _env(planner, dim, seed, HEURISTIC_EARLY_EXIT=True, MAX_EXTRA_NODES_ASTAR=500)
    gx, gy, sx, sy = (int(v) for v in R["A_goal_start"])
    astar.calcDistanceHeuristic((gx, gy), (sx, sy), for_path=False)
    cs = astar.closed_set_guidance_
    for k in ("cost", "dist", "parent", "flags"):
        assert np.array_equal(getattr(cs, k), R["A_heur_early_" + k]), k
    xs, ys = astar.getAstarPath(sx, sy)
    assert np.array_equal(np.stack([xs, ys], 1), R["A_heur_early_path"])


def test_near_nodes_after_fixed_point(planner):
    """get_only_near with the GPU fixed point: the list is derived from the slot's planes (row-major order)"""
    ctx, cc, astar = _env(planner, 512, 5)
    grid = astar.astar_grid_
    free = ga_b200.sample_free_cells(grid, 2, 5)
    goal, start = free[0], free[1]
    astar.calcDistanceHeuristic(goal, start, for_path=False, get_only_near=True, reference_order=False)
    cs = astar.closed_set_guidance_
    ys, xs = np.nonzero((cs.flags & 1) != 0)
    d = np.sqrt((xs - start[0]) ** 2.0 + (ys - start[1]) ** 2.0)
    want = sorted((int(y) * 128 + int(x)) for x, y, dd in zip(xs, ys, d) if dd < 10)
    assert [i for _, i in astar.nodes_near_goal_] == want


@pytest.mark.parametrize("dim,seed,unk,slot", [(1000, 21, 0.0, 0), (1492, 22, 2.5, 1), (4096, 23, 0.0, 0)])
def test_heur_records_equal_the_four_planes(planner, dim, seed, unk, slot):
    """GA_GRID_HEUR_REC (12 B per cell) unpacks to GA_GRID_HEUR_COST / _DIST / _PARENT / _FLAGS bit for bit after a device
    build (random proximity field, with and without the unknown cost, both slots); after a reference-order replay of the
    slot the record is refused (its planes are not the device's)."""
    ctx, cc, astar = _env(planner, dim, seed, ASTAR_UNKNOWN_COST=unk)
    ad = astar.astar_dim_
    rng = np.random.default_rng(seed)
    ctx.upload(_abi.GRID_H_PROX, rng.random((ad, ad)) * (rng.random((ad, ad)) < 0.3))
    free = ga_b200.sample_free_cells(astar.astar_grid_, 2, seed)
    goal, start = free[0], free[1]
    astar.calcDistanceHeuristic(goal, None, for_path=bool(slot), reference_order=False)
    planes = [ctx.download(g + slot) for g in (_abi.GRID_HEUR_COST, _abi.GRID_HEUR_DIST, _abi.GRID_HEUR_PARENT, _abi.GRID_HEUR_FLAGS)]
    rec = ctx.download(_abi.GRID_HEUR_REC + slot)
    assert rec.nbytes == 12 * ad * ad
    got = ga_b200.unpack_heur_records(rec)
    assert int((planes[3] & 1).sum()) > ad * ad // 4
    for name, a, b in zip(("cost", "dist", "parent", "flags"), planes, got):
        assert np.array_equal(a, b), name
    # asynchronous download into caller memory + ga_sync gives the same bytes
    rec2 = np.zeros_like(rec)
    ctx.check(ctx.lib.ga_download_async(ctx.h, _abi.GRID_HEUR_REC + slot, _abi.ptr(rec2), rec2.nbytes))
    ctx.sync()
    assert np.array_equal(rec.view(np.uint32), rec2.view(np.uint32))
    if dim <= 1500:
        astar.calcDistanceHeuristic(goal, start, for_path=bool(slot), reference_order=True)
        with pytest.raises(ga_b200.GaError, match="GA_GRID_HEUR_REC"):
            ctx.download(_abi.GRID_HEUR_REC + slot)


def test_download_window(planner):
    ctx, cc, astar = _env(planner, 1024, 3)
    ad = 256
    prox = np.random.default_rng(0).random((ad, ad))
    ctx.upload(_abi.GRID_H_PROX, prox)
    win = np.zeros((40, 56))
    ctx.check(ctx.lib.ga_download_window(ctx.h, _abi.GRID_H_PROX, 17, 33, 56, 40, _abi.ptr(win)))
    assert np.array_equal(win, prox[33:73, 17:73])
    assert ctx.lib.ga_download_window(ctx.h, _abi.GRID_H_PROX, 250, 0, 56, 40, _abi.ptr(win)) == -1  # outside the grid


def test_expand_epilogue_and_async(planner):
    """ga_expand_ex: per-successor heuristic cost / dist, movement cost, motion resolution and the parent's corner
    proximity equal host lookups in the downloaded planes; ga_expand_submit / collect equals ga_expand."""
    ctx, cc, astar = _env(planner, 1492, 11)
    ad = astar.astar_dim_
    rng = np.random.default_rng(4)
    prox = rng.random((ad, ad)) * (rng.random((ad, ad)) < 0.4)
    mres = rng.uniform(0.1, 1.0, (ad, ad))
    mc = np.where(rng.random((ad, ad)) < 0.1, 3.0, 5.0)
    ctx.upload(_abi.GRID_H_PROX, prox)
    ctx.upload(_abi.GRID_MOTION_RES, mres)
    ctx.upload(_abi.GRID_MOVEMENT_COST, mc)
    goal = ga_b200.sample_free_cells(astar.astar_grid_, 1, 9)[0]
    astar.calcDistanceHeuristic(goal, (ad // 2, ad // 2), for_path=True)  # early exit: part of the grid is NOT closed
    cs = astar.closed_set_path_
    hy = ga_b200.HybridAStar(ctx, cc.vehicle)
    cc.setExpandParams()
    n = 500
    nodes = np.zeros(n, _abi.NODE_DTYPE)
    ext = 1492 * ga_b200.GM_RES
    nodes["x"], nodes["y"] = rng.uniform(2, ext - 2, n), rng.uniform(2, ext - 2, n)
    nodes["yaw"], nodes["motion_res"], nodes["with_rear_axis"] = rng.uniform(-math.pi, math.pi, n), 0.3125, 1
    succ, cost, pprox = cc.expandNodesWithCosts(nodes, for_path=True)
    plain = cc.expandNodes(nodes, exact=False)
    assert np.array_equal(succ, plain)
    on = (succ["n_poses"] > 0) & (succ["xi"] >= 0) & (succ["xi"] < ad) & (succ["yi"] >= 0) & (succ["yi"] < ad)
    yi, xi = succ["yi"][on], succ["xi"][on]
    closed = (cs.flags[yi, xi] & 1) != 0
    assert np.array_equal(cost["heur_cost"][on], np.where(closed, cs.cost[yi, xi], _abi.OUT_OF_HEURISTIC))
    assert np.array_equal(cost["heur_dist"][on], np.where(closed, cs.dist[yi, xi], _abi.OUT_OF_HEURISTIC))
    assert np.array_equal(cost["movement_cost"][on], mc[yi, xi]) and np.array_equal(cost["motion_res"][on], mres[yi, xi])
    assert np.all(cost["heur_cost"][~on] == _abi.OUT_OF_HEURISTIC)
    assert (~closed).any() and closed.any()
    want = hy.getProxOfCornersBatch(np.stack([nodes["x"], nodes["y"], nodes["yaw"]], 1))
    assert np.array_equal(pprox, want)
    # asynchronous form, 26 nodes per submission, several in flight
    tickets = []
    for k in range(4):
        t = C.c_int()
        ctx.check(ctx.lib.ga_expand_submit(ctx.h, _abi.ptr(nodes[26 * k:26 * k + 26]), 26, C.byref(t)))
        tickets.append(t.value)
    for k, t in enumerate(tickets):
        out = np.zeros((32, cc.expand_stride_), _abi.SUCC_DTYPE)
        nn = C.c_int()
        ctx.check(ctx.lib.ga_expand_collect(ctx.h, t, _abi.ptr(out), C.byref(nn)))
        assert nn.value == 26 and np.array_equal(out[:26], plain[26 * k:26 * k + 26])


def test_cost_validation(planner):
    """zero or negative edge weights would let parent pointers form cycles (advisor, round 1): rejected at the boundary"""
    ctx, cc, astar = _env(planner, 512, 2)
    p = astar.astar_params()
    p.astar_movement_cost = 0.0
    assert ctx.lib.ga_set_astar_params(ctx.h, C.byref(p)) == -1
    p.astar_movement_cost = 5.0
    p.astar_prox_cost = -1.0
    assert ctx.lib.ga_set_astar_params(ctx.h, C.byref(p)) == -1
    mc = np.full((128, 128), 5.0)
    mc[7, 9] = 0.0
    assert ctx.lib.ga_upload(ctx.h, _abi.GRID_MOVEMENT_COST, _abi.ptr(mc), mc.nbytes) == -1
    mc[7, 9] = float("nan")
    assert ctx.lib.ga_upload(ctx.h, _abi.GRID_MOVEMENT_COST, _abi.ptr(mc), mc.nbytes) == -1
    px = np.zeros((128, 128))
    px[3, 3] = -0.5
    assert ctx.lib.ga_upload(ctx.h, _abi.GRID_H_PROX, _abi.ptr(px), px.nbytes) == -1
    px[3, 3] = float("nan")  # the reference's 0/0 of setCostmapValue is kept
    assert ctx.lib.ga_upload(ctx.h, _abi.GRID_H_PROX, _abi.ptr(px), px.nbytes) == 0
    # the smallest legal weights still give a finite, acyclic result
    tiny = np.full((128, 128), 1e-300)
    ctx.upload(_abi.GRID_MOVEMENT_COST, tiny)
    ctx.upload(_abi.GRID_H_PROX, np.zeros((128, 128)))
    goal = ga_b200.sample_free_cells(astar.astar_grid_, 1, 1)[0]
    n = astar.calcDistanceHeuristic(goal, None, for_path=False)
    cs = astar.closed_set_guidance_
    assert n == cs.size() > 1000
    far = np.argwhere((cs.flags & 1) != 0)[-1]
    xs, ys = astar.getAstarPath(int(far[1]), int(far[0]))
    assert (xs[-1], ys[-1]) == (goal[0], goal[1]) and len(xs) < 128 * 128
