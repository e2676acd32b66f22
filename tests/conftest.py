import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure(path, target, cwd=ROOT):
    if not os.path.exists(path):
        subprocess.run(["make", target], cwd=cwd, check=True, capture_output=True)


@pytest.fixture(scope="session", autouse=True)
def built_artifacts():
    """CPU-side artefacts the suite needs (normally produced by __graft_entry__.build())."""
    _ensure(os.path.join(ROOT, "oracle", "liboracle.so"), "restate", cwd=os.path.join(ROOT, "oracle"))
    _ensure(os.path.join(ROOT, "global-illumination_b200", "bin", "gi_scene_tool"), "tools")
    _ensure(os.path.join(ROOT, "tests", "host_sim", "libisect_sim.so"), "host_sim")
    yield


@pytest.fixture(scope="session")
def oracle():
    from gi_testlib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def gpu_ctx():
    import gi_b200
    ctx = gi_b200.Context(0)      # raises if the library is missing or there is no CUDA device
    yield ctx
    ctx.close()
