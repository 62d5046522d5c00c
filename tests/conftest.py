import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: test needs >= 2 CUDA devices")


@pytest.fixture(scope="session", autouse=True)
def _native_library():
    """The C++/CUDA library is git-ignored: build it in-tree (nvcc cross-compiles without a GPU) if this is a
    fresh checkout, so that no test depends on another one having built it."""
    from rocm_mpi_b200.ops import native
    native.load(build_if_missing=True)
    yield


@pytest.fixture(autouse=True)
def _fresh_grid():
    """Every test starts and ends without a global grid singleton."""
    import rocm_mpi_b200 as igg
    if igg.grid_is_initialized():
        igg.finalize_global_grid(finalize_dist=False)
    yield
    if igg.grid_is_initialized():
        igg.finalize_global_grid(finalize_dist=False)
