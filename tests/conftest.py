import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")
    if not config.pluginmanager.hasplugin("timeout"):
        config.addinivalue_line("markers", "timeout: per-test time limit (pytest-timeout; inert without the plugin)")


def pytest_collection_modifyitems(config, items):
    """A GPU test that stalls (a kernel that never returns blocks the ctypes call for good) must end the run with a
    failure instead of hanging it: every gpu-marked test without a timeout of its own gets 900 s (pytest-timeout's
    thread method ends the process; the marker is inert when the plugin is absent)."""
    if not config.pluginmanager.hasplugin("timeout"):
        return
    for item in items:
        if item.get_closest_marker("gpu") and not item.get_closest_marker("timeout"):
            item.add_marker(pytest.mark.timeout(900, method="thread"))


@pytest.fixture(scope="session")
def crt():
    """The product binding (loads libcrt_b200.so; fails loudly when it is not built)."""
    mod = importlib.import_module("raytracer-cuda_b200")
    mod.load_library()
    return mod


@pytest.fixture(scope="session")
def gpu_available():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
