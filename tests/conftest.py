import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as g
    return g.load_package()


@pytest.fixture(scope="session")
def small_world():
    import flivo_scenes as scenes
    return scenes.make_world(seed=7, n_points=60000, ds=0.5)


@pytest.fixture(scope="session")
def small_scan(small_world):
    import flivo_scenes as scenes
    return scenes.make_scan(small_world, seed=8, n_points=4000)
