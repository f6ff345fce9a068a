import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("OMP_WAIT_POLICY", "passive")  # the oracle's OpenMP regions are short

GOLDEN = os.path.join(ROOT, "tests", "golden")
CACHE = os.path.join(ROOT, "scenes", "_cache")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def teapot():
    from fluctus_b200 import scene as fs
    return fs.load_scene(os.path.join(GOLDEN, "teapot_sbvh.npz"))


@pytest.fixture(scope="session")
def night_env():
    from fluctus_b200 import scene as fs
    return fs.load_envmap(os.path.join(GOLDEN, "night_env.npz"))


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


def cached_scene(name):
    """Scenes generated from the reference's assets by tools/make_fixtures.py (git-ignored cache)."""
    from fluctus_b200 import scene as fs
    path = os.path.join(CACHE, name + "_sbvh.npz")
    if not os.path.exists(path):
        pytest.skip("scene cache %s missing (tools/make_fixtures.py needs /root/reference)" % name)
    return fs.load_scene(path)


@pytest.fixture(scope="session")
def egyptcat():
    return cached_scene("egyptcat")


@pytest.fixture(scope="session")
def psor_cube():
    return cached_scene("psor_cube")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def pytest_terminal_summary(terminalreporter):
    """Largest out-of-tolerance fractions the stage comparisons saw (the limits in test_stages_gpu.py are set from these)."""
    import sys
    mod = sys.modules.get("test_stages_gpu")
    if mod is not None and mod.OBSERVED["checks"]:
        o = mod.OBSERVED
        terminalreporter.write_line("stage parity: %d state comparisons; largest fraction of paths beyond 1e-4 or with a differing "
                                    "integer field %.5f (limit %.5f); largest fraction between 1e-5 and 1e-4 %.5f (limit %.5f)"
                                    % (o["checks"], o["flip"], mod.MAX_FLIP_FRACTION, o["amplified"], mod.MAX_AMPLIFIED_FRACTION))
