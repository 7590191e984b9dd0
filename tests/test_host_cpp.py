"""The C++ host mirror (guided-extended-hybrid-astar_b200/host/gridaccel_host.hpp): builds on the CPU,
runs its self-test (the reference's gtest known answers + grid-builder properties) on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "guided-extended-hybrid-astar_b200", "host")
BIN = os.path.join(ROOT, "guided-extended-hybrid-astar_b200", "lib", "host_selftest")


def test_host_mirror_builds():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "guided-extended-hybrid-astar_b200", "csrc")])
    subprocess.check_call(["make", "-s", "-C", HOST])
    assert os.path.exists(BIN)


def test_host_mirror_names_match_reference_interface():
    """the members hybrid_a_star_lib.cpp / planning.py call on the two classes"""
    src = open(os.path.join(HOST, "gridaccel_host.hpp")).read()
    for name in ("initialize", "calculateDisks", "resetPatch", "processSafetyPatch", "passLocalMap", "passLocalMapData",
                 "getYawIdx", "checkGrid", "checkGridVariant", "checkPose", "checkPoseVariant", "getPathCollisionIndex",
                 "checkPathCollision", "returnDiskPositions", "disk_r_c_", "patch_arr_", "patch_safety_arr_",
                 "calcDistanceHeuristic", "calcVoronoiPotentialField", "calcAstarGridCuda", "getAstarPath",
                 "findValidNeighborIndex", "calcIndex", "reinit", "resetMovementMap", "closed_set_path_",
                 "closed_set_guidance_", "astar_grid_", "movement_cost_map_", "restr_geofence_grid_", "h_prox_arr_",
                 "motion_res_map_", "obstacle_dist_", "voronoi_dist_", "astar_dim_", "alpha_", "do_max_", "do_min_"):
        assert name in src, name


@pytest.mark.gpu
def test_host_mirror_selftest_on_gpu():
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-s", "-C", HOST])
    out = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "host selftest ok" in out.stdout
