"""Host-side mirror of the reference's grid interfaces on top of libgridaccel.so (B200, sm_100a).

Python twin of the C++ mirror in ../host/: the class and method names are those of the
reference's pybind module `_hybridastar_planning_lib_api`
(LIB/src/hybridastar_planning_lib/wrapper_hybrid_a_star_lib.cpp:176-199, 327-346)."""
from ._abi import GaError, LIB_PATH, load  # noqa: F401
from .planner import (AStar, Cartographing, CollisionChecker, GridContext, GridMapSim, HybridAStar, Vehicle, VehConfig, LibConfig,  # noqa: F401
                      calc_disks, unpack_heur_records, GM_RES, ASTAR_RES)
from .synth import gen_map, sample_free_cells  # noqa: F401
