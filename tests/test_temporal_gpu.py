"""GPU tests of the two-steps-per-pass kernel (``csrc/temporal2d.cu``).  Part of the ``gpu`` tier since the kernel
passed its first run on a B200 (``gpurun_out/t2_check.log``: bit-identical to two one-step launches without
neighbours, <= 1.2e-16 through the solver with a self-neighbour).

Each test compares against two launches of the GPU-tested one-step kernel (same arithmetic, same FMA
contraction), which itself is tested against the PyTorch fp32/fp64 statement in ``tests/test_kernels_gpu.py``.
"""
import pytest
import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.ops import native
from rocm_mpi_b200.ops import reference as ref

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")]
TOL = 1e-12
PH = (1.3, 1e-3, 2.0, 3.0)


def _fields(nx, ny, seed=1):
    g = torch.Generator().manual_seed(seed)
    T = torch.rand(nx, ny, generator=g, dtype=torch.float64)
    Cp = torch.rand(nx, ny, generator=g, dtype=torch.float64) + 0.5
    return T, Cp


@pytest.mark.parametrize("shape", [(3, 4), (16, 16), (17, 130), (40, 126), (64, 250), (130, 258), (300, 380),
                                   (1030, 2050)])
def test_two_step_kernel_matches_reference(shape):
    T, Cp = _fields(*shape)
    a = T.clone()
    ref.perf_step(a, T, Cp, *PH)
    want = a.clone()
    ref.perf_step(want, a, Cp, *PH)
    Td, Cpd = T.cuda(), Cp.cuda()
    T2d = Td.clone()
    n0 = native.launch_count()
    native.diffusion2d_step2(T2d, Td, Cpd, *PH)
    assert native.launch_count() == n0 + 1
    got = T2d.cpu()
    assert float((got - want).abs().max()) < TOL
    assert torch.equal(got[0], T[0]) and torch.equal(got[:, -1], T[:, -1])      # boundary never written


def test_two_step_kernel_diagnostics_and_per_step_parameters():
    T, Cp = _fields(300, 380, seed=3)
    ph1, ph2 = (1.3, 1e-3, 2.0, 3.0), (0.9, 7e-4, 2.0, 3.0)
    a = T.clone()
    ref.perf_step(a, T, Cp, *ph1)
    b = a.clone()
    ref.perf_step(b, a, Cp, *ph2)
    Td, Cpd = T.cuda(), Cp.cuda()
    T2d = Td.clone()
    params = torch.tensor(ph1 + ph2, dtype=torch.float64).cuda()
    diag = torch.zeros(4, dtype=torch.float64, device="cuda")
    for _ in range(2):                                  # twice: the scratch must be left ready for the next launch
        native.diffusion2d_step2(T2d, Td, Cpd, 0.0, 0.0, 0.0, 0.0, dev_params=params, diag_out=diag)
        d = diag.cpu()
        assert float((T2d.cpu() - b).abs().max()) < TOL
        ia, ib = a[1:-1, 1:-1], b[1:-1, 1:-1]
        assert abs(float(d[0]) - float(ia.max())) < TOL and abs(float(d[2]) - float(ib.max())) < TOL
        assert abs(float(d[1]) - float(ia.sum())) < 1e-9 * ia.numel()
        assert abs(float(d[3]) - float(ib.sum())) < 1e-9 * ib.numel()


@pytest.mark.parametrize("periods,steps", [((1, 1), 4), ((1, 0), 5), ((0, 1), 6), ((0, 0), 7)])
def test_temporal_mode_equals_fused_mode_with_self_neighbour(periods, steps):
    """Single process: periodic dimensions make the rank its own neighbour, so the mid-line kernel, the mid
    flags and the pushes of the pass all run (against local pointers)."""
    from rocm_mpi_b200.models.solver import Diffusion2D
    out = {}
    for mode in ("fused", "temporal"):
        igg.init_global_grid(300, 380, 1, periodx=periods[0], periody=periods[1], quiet=True, init_dist=False)
        s = Diffusion2D(300, 380, mode=mode)
        s.begin()
        s.step(steps)
        s.finish()
        out[mode] = s.T.clone()
        igg.finalize_global_grid(finalize_dist=False)
    assert float((out["fused"] - out["temporal"]).abs().max()) < TOL


def test_controlled_pairs_match_controlled_steps():
    from rocm_mpi_b200.models.solver import Diffusion2D
    igg.init_global_grid(130, 258, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(130, 258, mode="fused")
    want = [s.step_controlled() for _ in range(6)]
    s.finish()
    Tf = s.T.clone()
    igg.finalize_global_grid(finalize_dist=False)
    igg.init_global_grid(130, 258, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(130, 258, mode="temporal")
    got = []
    for _ in range(3):
        got += list(s.step_controlled_pair())
    s.finish()
    assert float((s.T - Tf).abs().max()) < TOL
    for g, w in zip(got, want):
        assert abs(g[0] - w[0]) < TOL and abs(g[1] - w[1]) < 1e-9 * 130 * 258
    igg.finalize_global_grid(finalize_dist=False)


def test_pipelined_pairs_with_an_odd_tail_match_blocking_steps():
    """Pipelined control loop in mode 'temporal': pairs, then ONE single step (an odd step count), every diagnostic
    seen exactly once and equal to the blocking one-step loop's."""
    from rocm_mpi_b200.models.solver import Diffusion2D
    igg.init_global_grid(130, 258, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(130, 258, mode="fused")
    want = [s.step_controlled() for _ in range(7)]
    s.finish()
    Tf = s.T.clone()
    igg.finalize_global_grid(finalize_dist=False)
    igg.init_global_grid(130, 258, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(130, 258, mode="temporal")
    got = []
    for _ in range(3):
        r = s.step_controlled_pair(pipelined=True)
        if r is not None:
            got += list(r)
    r = s.step_controlled(pipelined=True)
    if r is not None:
        got += list(r) if isinstance(r[0], tuple) else [r]
    for r in s.drain_controlled():
        got += list(r) if isinstance(r[0], tuple) else [r]
    s.finish()
    assert len(got) == 7
    assert float((s.T - Tf).abs().max()) < TOL
    for g_, w in zip(got, want):
        assert abs(g_[0] - w[0]) < TOL and abs(g_[1] - w[1]) < 1e-9 * 130 * 258
    igg.finalize_global_grid(finalize_dist=False)
