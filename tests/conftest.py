import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def native():
    """The native library, built in-tree; GPU tests must never run on a silent fallback."""
    from gpu_mpi_tests_b200.ops import _lib
    from gpu_mpi_tests_b200._native.build import ensure_built
    ensure_built()     # also rebuilds after an edit of csrc/ or include/: never test a stale library
    _lib.require_native()
    return _lib.lib


@pytest.fixture(scope="session")
def dev(native):
    import torch
    torch.cuda.set_device(0)
    return torch.device("cuda", 0)
