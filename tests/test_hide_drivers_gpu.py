"""GPU tests of the drivers composed from the frame / interior kernels (``gpu`` tier; they first passed on a B200 in
round 2, ``gpurun_out/t2_pytest.log``):

* ``Diffusion3D(mode="hide")``: frame kernel + push on a high-priority stream, interior kernel on a low-priority one;
* ``Diffusion2D(mode="hide_exchange")``: the reference's commented variant (3), ``scripts/diffusion_2D_perf_hide.jl:94-101``
  (frame kernel, ``update_halo(T2)`` over a transport while the interior kernel runs).
"""
import pytest
import torch

import rocm_mpi_b200 as igg

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")]
TOL = 1e-12


@pytest.mark.parametrize("periods", [(1, 1, 1), (0, 1, 0), (0, 0, 0)])
def test_3d_hide_mode_equals_fused_mode(periods):
    from rocm_mpi_b200.models.diffusion3d import Diffusion3D
    out = {}
    for mode in ("fused", "hide"):
        igg.init_global_grid(40, 36, 66, periodx=periods[0], periody=periods[1], periodz=periods[2], quiet=True,
                             init_dist=False)
        s = Diffusion3D(40, 36, 66, mode=mode, b_width=(4, 4, 8))
        s.begin()
        s.step(5)
        s.finish()
        out[mode] = s.T.clone()
        igg.finalize_global_grid(finalize_dist=False)
    assert float((out["fused"] - out["hide"]).abs().max()) < TOL


@pytest.mark.parametrize("transport", ["peer", "host"])
@pytest.mark.parametrize("periods", [(1, 1), (1, 0), (0, 0)])
def test_2d_hide_exchange_equals_plain(transport, periods):
    from rocm_mpi_b200.models.solver import Diffusion2D
    out = {}
    for mode in ("plain", "hide_exchange"):
        igg.init_global_grid(200, 260, 1, periodx=periods[0], periody=periods[1], quiet=True, init_dist=False)
        s = Diffusion2D(200, 260, mode=mode, variant="tma", transport=transport, b_width=(16, 4))
        s.begin()
        s.step(6)
        s.finish()
        torch.cuda.synchronize()
        out[mode] = s.T.clone()
        igg.finalize_global_grid(finalize_dist=False)
    assert float((out["plain"] - out["hide_exchange"]).abs().max()) < TOL
