import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "guided-extended-hybrid-astar_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import ga_oracle_py
    ga_oracle_py.lib()
    return ga_oracle_py


@pytest.fixture()
def planner():
    """(ctx, CollisionChecker, AStar) on cuda:0 with the shipped config and the default vehicle."""
    import ga_b200
    ctx = ga_b200.GridContext(0)
    veh = ga_b200.Vehicle.from_config(ga_b200.VehConfig())
    cfg = ga_b200.LibConfig()
    astar = ga_b200.AStar(ctx, cfg)
    cc = ga_b200.CollisionChecker(ctx, veh, cfg)
    yield ctx, cc, astar
    ctx.close()
