"""Plan-level parity (north_star: "drops into the serial Hybrid A* search unchanged ... expanded-node and path
sequences bit-exact"): the reference's own hybrid_a_star_lib.cpp is run twice on the same query,
  arm A  on the reference's AStar / CollisionChecker / Vehicle            (oracle/_ref/libga_ref.so),
  arm B  on the drop-in static classes over libgridaccel.so, CUDA on cuda:0 (oracle/_ref/libga_planb.so),
each in a fresh process (one planner per process).  Compared: the sequence of expanded nodes, a hash
over every collision query and its answer, the closed set of the search (keys, parents, costs) and the final path.
Both libraries are built where /root/reference is mounted and travel to the GPU box prebuilt."""
import os

import numpy as np
import pytest

import ga_plan_py as gp

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not (os.path.exists(gp.LIBS["ref"]) and os.path.exists(gp.LIBS["b200"])),
                                 reason="oracle/_ref plan libraries not built (need /root/reference at build time)")]


def _same_plan(a, b, prefix=""):
    k = lambda s: prefix + s
    assert int(a[k("found")]) == int(b[k("found")]) == 1
    assert int(a[k("n_expanded")]) == int(b[k("n_expanded")])
    assert np.array_equal(a[k("expanded")], b[k("expanded")]), "expanded-node sequence differs"
    assert int(a[k("n_checks")]) == int(b[k("n_checks")]) and int(a[k("n_check_poses")]) == int(b[k("n_check_poses")])
    assert int(a[k("check_hash")]) == int(b[k("check_hash")]), "a collision query or its answer differs"
    assert np.array_equal(a[k("closed_keys")], b[k("closed_keys")])
    assert np.array_equal(a[k("closed_cost")], b[k("closed_cost")])
    for f in ("path_x", "path_y", "path_yaw", "path_delta", "path_dir", "path_type"):
        assert np.array_equal(a[k(f)], b[k(f)]), f
    assert float(a[k("path_cost")]) == float(b[k("path_cost")])


def test_config1_valet_parking_sequences(tmp_path):
    """BASELINE config 1: default SCENARIO_NAME map, ALL_VISIBLE = true, patch 1492^2, shipped config."""
    a = gp.run_query_subprocess("ref", str(tmp_path / "a.npz"), scenario="Valet_parking_015", extra=["--heuristic"])
    b = gp.run_query_subprocess("b200", str(tmp_path / "b.npz"), scenario="Valet_parking_015", mode=0, extra=["--heuristic"])
    assert int(a["patch_dim"]) == int(b["patch_dim"]) == 1492
    # the environment the search runs on: pooled grid, potential field, both heuristic sets, bit for bit
    assert np.array_equal(a["astar_grid"], b["astar_grid"])
    assert np.array_equal(a["h_prox"], b["h_prox"], equal_nan=True)
    assert np.array_equal(a["motion_res"], b["motion_res"])
    for s in ("hg_", "hp_"):
        for f in ("cost", "dist", "parent", "flags"):
            assert np.array_equal(a[s + f], b[s + f]), s + f
    _same_plan(a, b)
    assert int(a["n_expanded"]) > 1000


def test_config1_fixed_point_heuristic_reports_difference(tmp_path):
    """the same query guided by the GPU fixed point (every reachable cell closed, optimal costs): the search still
    finds a path; how far its sequence departs from the reference's is reported, and bounded loosely."""
    a = gp.run_query_subprocess("ref", str(tmp_path / "a.npz"), scenario="Valet_parking_015", extra=["--heuristic"])
    b = gp.run_query_subprocess("b200", str(tmp_path / "b.npz"), scenario="Valet_parking_015", mode=1, extra=["--heuristic"])
    assert int(b["found"]) == 1
    # the drop-in's closed sets read the device's 12-byte records in place (GA_GRID_HEUR_REC): what their accessors
    # return equals the device's four planes bit for bit (unclosed cells: the dump leaves its zero defaults)
    for s in ("hg_", "hp_"):
        m = (b["dev_" + s + "flags"] & 1) != 0
        assert np.array_equal((b[s + "flags"] & 1) != 0, m) and m.sum() > 1000
        for f in ("cost", "dist", "parent", "flags"):
            assert np.array_equal(b[s + f][m], b["dev_" + s + f][m]), s + f
    closed = (a["hp_flags"] & 1) != 0
    assert np.all(((b["hp_flags"] & 1) != 0)[closed])             # the fixed point closes a superset
    ra, rb = a["hp_cost"][closed], b["hp_cost"][closed]
    assert np.all(rb <= ra)                                          # never above the reference
    n_diff = int(np.count_nonzero(ra != rb))
    rel = float(np.max((ra - rb) / np.maximum(ra, 1e-300)))
    print("fixed point vs reference on the real Voronoi field: %d / %d cells differ, max rel %.3g; expanded %d vs %d"
          % (n_diff, int(closed.sum()), rel, int(b["n_expanded"]), int(a["n_expanded"])))
    assert n_diff <= 0.05 * closed.sum() and rel < 0.02
    assert abs(int(b["n_expanded"]) - int(a["n_expanded"])) <= 0.25 * int(a["n_expanded"])


def test_config4_ushift_sequences(tmp_path):
    """BASELINE config 4: U-Shift extended motion model (rear-axis rotation primitives) on the config-1 map."""
    a = gp.run_query_subprocess("ref", str(tmp_path / "a.npz"), scenario="Valet_parking_015", ushift=True)
    b = gp.run_query_subprocess("b200", str(tmp_path / "b.npz"), scenario="Valet_parking_015", ushift=True, mode=0)
    _same_plan(a, b)


def test_config5_synthetic_queries_sequences(tmp_path):
    """BASELINE config 5's unit at a size the CPU arm finishes in seconds: three queries on one synthetic 1024^2 map,
    run back to back in one process per arm (the reference's search state leaks from plan to plan, H5 -- in both arms)."""
    a = gp.run_query_subprocess("ref", str(tmp_path / "a.npz"), synthetic=1024, seed=50000, queries=3)
    b = gp.run_query_subprocess("b200", str(tmp_path / "b.npz"), synthetic=1024, seed=50000, queries=3, mode=0)
    assert int(a["n_queries"]) == int(b["n_queries"]) == 3
    for i in range(3):
        _same_plan(a, b, "q%d_" % i)


def test_patch_move_carries_the_potential_field(tmp_path):
    """AStar::reinit (a_star.cpp:70-99): h_prox_arr_ and motion_res_map_ survive a patch move (util::saveTemp +
    copyPatch2Patch).  Both arms build the environment, move the patch with the ego, and build again: the carried
    maps -- on the host AND in the device copy the kernels read -- and the guidance heuristic computed on them are equal."""
    a = gp.run_query_subprocess("ref", str(tmp_path / "a.npz"), scenario="Valet_parking_015", extra=["--patch-move"])
    b = gp.run_query_subprocess("b200", str(tmp_path / "b.npz"), scenario="Valet_parking_015", mode=0, extra=["--patch-move"])
    assert int(a["patch_dim"]) == int(b["patch_dim"]) and np.array_equal(a["origin"], b["origin"])
    assert np.count_nonzero(a["carried_h_prox"]) > 1000            # the field really is carried, not zeroed
    for k in ("carried_h_prox", "carried_motion_res"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k
    assert np.array_equal(a["carried_h_prox"], b["device_h_prox"], equal_nan=True)
    assert np.array_equal(a["carried_motion_res"], b["device_motion_res"])
    for k in ("h_prox", "motion_res", "hg_cost", "hg_dist", "hg_parent", "hg_flags"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k
