import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: test needs >= 2 CUDA devices")


@pytest.fixture(autouse=True)
def _fresh_grid():
    """Every test starts and ends without a global grid singleton."""
    import rocm_mpi_b200 as igg
    if igg.grid_is_initialized():
        igg.finalize_global_grid(finalize_dist=False)
    yield
    if igg.grid_is_initialized():
        igg.finalize_global_grid(finalize_dist=False)
