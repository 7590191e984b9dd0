"""CPU: the C-ABI library loads and exports every declared symbol; host-side logic of the mirror."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import ga_b200
from ga_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gridaccel.h")).read()
    declared = set(re.findall(r"^GA_API [\w \*]+?\b(ga_\w+)\(", hdr, re.M))
    assert len(declared) >= 28
    lib = C.CDLL(_abi.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(_abi.SYMBOLS), declared ^ set(_abi.SYMBOLS)
    assert _abi.load().ga_abi_version() == 1


def test_struct_layouts_match_header():
    assert C.sizeof(_abi.CollisionParams) == 24
    assert C.sizeof(_abi.ExpandParams) == 48
    assert C.sizeof(_abi.AstarParams) == 96
    assert _abi.NODE_DTYPE.itemsize == 40 and _abi.SUCC_DTYPE.itemsize == 48


def test_no_cuda_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ga_b200.GaError):
        ga_b200.GridContext(0)


def test_null_context_calls_return_errors():
    lib = _abi.load()
    assert lib.ga_sync(None) == -1
    assert lib.ga_build_safety(None) == -1
    assert lib.ga_destroy(None) == -1
    assert lib.ga_launch_count(None) == 0
    assert lib.ga_last_error(None) == b"null context"


def test_calc_disks_matches_oracle(oracle):
    for veh, cfg in [(ga_b200.Vehicle.from_config(ga_b200.VehConfig()), ga_b200.LibConfig()),
                     (ga_b200.Vehicle(0.6, 2.7, 3.0, 1.5, 2.0, False), ga_b200.LibConfig()),
                     (ga_b200.Vehicle(0.5, 2.0, 2.5, 0.8, 1.8, True), ga_b200.LibConfig(DOUBLE_DISK_ROWS=True, YAW_RES_COLL=5))]:
        nb, r, rc, cen = ga_b200.calc_disks(veh, cfg)
        info, ocen = oracle.calc_disks(veh.length_, veh.width_, veh.lb_, cfg.LEN_PER_DISK, cfg.DOUBLE_DISK_ROWS,
                                       cfg.SAFETY_DISTANCE_M, cfg.YAW_RES_COLL)
        assert (nb, rc) == (info.nb_disks, info.disk_r_c) and r == info.disk_r
        assert np.array_equal(cen, ocen)


def test_gen_map_is_seeded_and_walled():
    a, b = ga_b200.gen_map(512, 5), ga_b200.gen_map(512, 5)
    assert np.array_equal(a, b) and not np.array_equal(a, ga_b200.gen_map(512, 6))
    assert np.all(a[:2] == 255) and np.all(a[:, -2:] == 255)
    assert set(np.unique(a)) <= {0, 127, 255}


def test_closed_set_read_api():
    from ga_b200.planner import ClosedSet
    cost = np.full((4, 4), np.inf)
    flags = np.zeros((4, 4), np.uint8)
    parent = np.full((4, 4), -2, np.int32)
    cost[1, 2], flags[1, 2], parent[1, 2] = 0.0, 1, -1
    cost[1, 3], flags[1, 3], parent[1, 3] = 3.125, 3, 6
    cs = ClosedSet(cost, cost.copy(), parent, flags)
    assert cs.contains(6) and 7 in cs and not cs.contains(0) and not cs.contains(99)
    n = cs.at(7)
    assert n.pos == (3, 1) and n.cost_ == 3.125 and n.parent_index_ == 6 and n.is_unknown_
    assert cs.size() == 2 and list(cs.keys()) == [6, 7]
    with pytest.raises(KeyError):
        cs.at(0)
