import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-object-detection_b200")
for p in (ROOT, PKG, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """The tests exercise the built library.  The .so is not in history, so on a fresh checkout it is compiled
    here first, the way __graft_entry__.build() does; the product itself never builds or falls back
    (ypp._lib.lib() raises when the library is missing)."""
    import shutil
    from ypp import _lib
    if not os.path.exists(_lib.LIB_PATH) and shutil.which("nvcc"):
        _lib.build()
    yield


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
