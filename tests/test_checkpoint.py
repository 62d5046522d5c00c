import torch
import pytest

import rocm_mpi_b200 as igg
from rocm_mpi_b200.utils.checkpoint import latest_checkpoint, load_checkpoint, save_checkpoint


def test_checkpoint_roundtrip_and_layout_check(tmp_path):
    igg.init_global_grid(12, 10, 1, device_type="cpu", quiet=True)
    T = torch.rand(12, 10, dtype=torch.float64)
    Cp = torch.rand(12, 10, dtype=torch.float64)
    save_checkpoint(str(tmp_path), 5, T=T, Cp=Cp)
    p = save_checkpoint(str(tmp_path), 40, T=T + 1, Cp=Cp)
    assert latest_checkpoint(str(tmp_path)) == p
    step, f = load_checkpoint(p)
    assert step == 40 and torch.equal(f["T"], T + 1) and torch.equal(f["Cp"], Cp)
    igg.finalize_global_grid()
    igg.init_global_grid(14, 10, 1, device_type="cpu", quiet=True)
    with pytest.raises(ValueError):
        load_checkpoint(p)
