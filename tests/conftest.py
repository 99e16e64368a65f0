import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device; run with -m gpu on the GPU box")


def pytest_collection_modifyitems(config, items):
    """tests marked `gpu` are skipped (not errored) on a machine without a B200: a plain `pytest tests` then reports the CPU-side
    results cleanly"""
    try:
        import torch
        ok = torch.cuda.is_available() and torch.cuda.get_device_capability(0) == (10, 0)
    except Exception:
        ok = False
    if ok:
        return
    skip = pytest.mark.skip(reason="needs a B200 (compute capability 10.0)")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built_lib():
    """libb200iq.so, (re)built in-tree with nvcc when sources changed (cross-compiles without a GPU)"""
    from blt_vqg_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
