"""Peer-memory protocol on ONE GPU: a periodic single-rank grid is its own neighbour, so the same
kernels (one-launch update_halo, fused step + push + epoch flags) run against local pointers."""
import pytest
import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.ops import native, reference as ref
from rocm_mpi_b200.parallel.fabric import PushPlan, fabric_zeros, get_fabric

pytestmark = pytest.mark.gpu


def test_fabric_tensor_aliases_raw_allocation():
    igg.init_global_grid(16, 16, 1, periodx=1, quiet=True, init_dist=False)
    t = fabric_zeros((16, 16))
    assert t.is_cuda and t.dtype == torch.float64 and float(t.abs().sum()) == 0.0
    t.fill_(3.0)
    u = t + 1
    assert float(u.sum()) == 4.0 * 256
    fab = get_fabric()
    assert fab.is_registered(t) and fab.peer_ptr(t, 0) == t.data_ptr()


@pytest.mark.parametrize("shape,periods", [((12, 10), (1, 1)), ((12, 10), (1, 0)), ((8, 9, 10), (1, 1, 1)),
                                           ((8, 9, 10), (0, 1, 0))])
def test_update_halo_peer_periodic_self(shape, periods):
    n = list(shape) + [1] * (3 - len(shape))
    igg.init_global_grid(*n, periodx=periods[0], periody=periods[1], periodz=periods[2] if len(periods) > 2 else 0,
                         quiet=True, init_dist=False)
    vg = igg.VirtualGrid(shape, (1,) * len(shape), periods)
    g = torch.Generator().manual_seed(3)
    stag = tuple(s + (1 if d == 0 else 0) for d, s in enumerate(shape))
    A, B = torch.rand(shape, generator=g, dtype=torch.float64), torch.rand(stag, generator=g, dtype=torch.float64)
    Ad, Bd = A.cuda(), B.cuda()
    for _ in range(3):                       # repeated calls: epochs advance
        igg.update_halo(Ad, Bd, transport="peer")
        vg.update_halo([A], [B])
        assert torch.equal(Ad.cpu(), A) and torch.equal(Bd.cpu(), B)
        A += 1.0
        Ad += 1.0
    get_fabric().check_errors(0)


@pytest.mark.parametrize("variant", ["tma", "direct"])
@pytest.mark.parametrize("mode", ["fused", "hide", "plain", "persistent", "resident"])
def test_solver_periodic_self_neighbor(variant, mode):
    from rocm_mpi_b200.models.solver import Diffusion2D
    nx, ny = 96, 260
    igg.init_global_grid(nx, ny, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(nx, ny, mode=mode, variant=variant, b_width=(16, 4), transport="peer")
    T = s.T.cpu().clone()
    s.begin()
    s.step(3)
    s.step(2)
    s.finish()
    vg = igg.VirtualGrid((nx, ny), (1, 1), (1, 1))
    T2 = T.clone()
    Cp = torch.ones_like(T)
    vg.update_halo([T])
    for _ in range(5):
        ref.perf_step(T2, T, Cp, s.lam, s.dt, s._dx, s._dy)
        vg.update_halo([T2])
        T, T2 = T2, T
    assert float((s.T.cpu() - T).abs().max()) < 1e-13


def test_update_halo_peer_other_dtypes():
    igg.init_global_grid(12, 10, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    vg = igg.VirtualGrid((12, 10), (1, 1), (1, 1))
    g = torch.Generator().manual_seed(9)
    for dt in (torch.float32, torch.int64, torch.int16, torch.uint8):
        A = (torch.rand(12, 10, generator=g) * 100).to(dt)
        Ad = A.cuda()
        igg.update_halo(Ad, transport="peer")
        vg.update_halo([A])
        assert torch.equal(Ad.cpu(), A), dt


def test_strip_kernel_periodic_self_neighbor():
    """Large enough for the strip-marching kernel: its in-kernel push / flags against itself."""
    from rocm_mpi_b200.models.solver import Diffusion2D
    nx, ny = 600, 260
    igg.init_global_grid(nx, ny, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(nx, ny, mode="fused", variant="tma")
    T = s.T.cpu().clone()
    s.begin()
    s.step(1)
    s.step(4)
    s.finish()
    vg = igg.VirtualGrid((nx, ny), (1, 1), (1, 1))
    T2 = T.clone()
    Cp = torch.ones_like(T)
    vg.update_halo([T])
    for _ in range(5):
        ref.perf_step(T2, T, Cp, s.lam, s.dt, s._dx, s._dy)
        vg.update_halo([T2])
        T, T2 = T2, T
    assert float((s.T.cpu() - T).abs().max()) < 1e-13


def test_push_plan_3d_periodic_self():
    shape = (20, 18, 66)
    igg.init_global_grid(*shape, periodx=1, periody=1, periodz=1, quiet=True, init_dist=False)
    T, T2 = fabric_zeros(shape), fabric_zeros(shape)
    g = torch.Generator().manual_seed(5)
    T.copy_(torch.rand(shape, generator=g, dtype=torch.float64))
    T2.copy_(T)
    Cp = torch.ones(shape, dtype=torch.float64, device="cuda")
    plan = PushPlan(T, T2)
    plan.begin()
    Th = T.cpu().clone()                      # after begin(): halos consistent
    for variant in ("tma", "direct"):
        native.diffusion3d_step(T2, T, Cp, 1.0, 1e-3, 2.0, 3.0, 4.0, variant=variant, push=plan.next(T2))
        T, T2 = T2, T
    plan.finish()
    vg = igg.VirtualGrid(shape, (1, 1, 1), (1, 1, 1))
    Th2 = Th.clone()
    Cph = torch.ones_like(Th)
    for _ in range(2):
        ref.perf_step_3d(Th2, Th, Cph, 1.0, 1e-3, 2.0, 3.0, 4.0)
        vg.update_halo([Th2])
        Th, Th2 = Th2, Th
    assert float((T.cpu() - Th).abs().max()) < 1e-13


def test_peer_copy_and_wait_value():
    igg.init_global_grid(16, 16, 1, periodx=1, quiet=True, init_dist=False)
    import ctypes as C
    fab = get_fabric()
    src = torch.rand(100003, dtype=torch.float64, device="cuda")
    dst = torch.zeros_like(src)
    flag = torch.zeros(2, dtype=torch.int64, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    lib = native.require()
    native._check(lib.igg_peer_copy(dst.data_ptr(), src.data_ptr(), src.numel() * 8, flag.data_ptr(), 7,
                                    cnt.data_ptr(), native._stream()))
    native._check(lib.igg_wait_value(flag.data_ptr(), 7, flag[1:].data_ptr(), native._stream()))
    torch.cuda.synchronize()
    assert torch.equal(dst, src) and int(flag[0]) == 7 and int(flag[1]) == 0
