import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib():
    """The in-tree CUDA backend library (host-layer helpers work without a GPU)."""
    from heatray_b200 import capi
    from heatray_b200.build import LIB, build_cuda
    if not LIB.exists():
        build_cuda()
    return capi.load_library()


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import loader
    return loader


@pytest.fixture()
def oracle_backend(oracle_mod, lib):
    b = oracle_mod.backend()
    yield b
    b.close()


@pytest.fixture()
def gpu_backend(lib):
    from heatray_b200 import capi
    b = capi.Backend(lib, "hrc", device=0, collect_stats=True)
    yield b
    b.close()
