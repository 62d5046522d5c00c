"""Multi-GPU tests (one process per GPU).  Skipped when fewer than 2 GPUs are visible."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from mp_util import run_workers  # noqa: E402

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("nprocs,args", [(2, (66, 260, 0, 0, 12)), (2, (530, 260, 0, 1, 6)), (2, (66, 260, 1, 1, 8)), (4, (40, 132, 0, 1, 6)),
                                         (8, (34, 130, 0, 0, 6))])
def test_peer_fabric_and_solver_multi_gpu(nprocs, args):
    if _ngpu() < nprocs:
        pytest.skip(f"needs {nprocs} GPUs")
    outs = run_workers("multi_gpu.py", nprocs, args, timeout=600)
    assert all("ok" in o for o in outs)


@pytest.mark.parametrize("nprocs", [2, 4, 8])
def test_apps_multi_gpu(nprocs, tmp_path):
    if _ngpu() < nprocs:
        pytest.skip(f"needs {nprocs} GPUs")
    outs = run_workers("apps_multi_gpu.py", nprocs, (str(tmp_path), nprocs), timeout=900)
    assert all("ok" in o for o in outs)


@pytest.mark.parametrize("nprocs,args", [(2, (34, 66, 0, 0, 600)), (2, (20, 34, 1, 1, 400)), (8, (18, 34, 1, 0, 400))])
def test_flag_protocol_stress_small_grids(nprocs, args):
    if _ngpu() < nprocs:
        pytest.skip(f"needs {nprocs} GPUs")
    outs = run_workers("stress_multi_gpu.py", nprocs, args, timeout=600)
    assert all("ok" in o for o in outs)
