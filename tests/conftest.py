import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


HAS_GPU = _has_gpu()


def pytest_collection_modifyitems(config, items):
    if HAS_GPU:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def lib():
    import __graft_entry__ as g
    g.build()
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
    return _lib.load()


@pytest.fixture(scope="session")
def gpu_engine(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine
    return Engine(0)
