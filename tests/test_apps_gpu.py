"""Applications on ONE GPU (single process; periodic layouts exercise the peer protocol against itself)."""
import os

import pytest
import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.ops import native, reference as ref
from rocm_mpi_b200.utils.config import DiffusionConfig

pytestmark = pytest.mark.gpu


def test_kp_app_single_gpu_matches_oracle(tmp_path):
    from rocm_mpi_b200.models.diffusion2d_kp import diffusion2D_kp
    cfg = DiffusionConfig(nx=128, ny=128, nt=100, outdir=str(tmp_path), quiet=True)
    res = diffusion2D_kp(cfg)
    T_ref, _ = ref.oracle_2d(128, 128, 100, "kp")
    assert float((res.T_v - T_ref[1:-1, 1:-1]).abs().max()) < 1e-12
    assert res.launches == 300 and os.path.exists(res.png)


def test_kp_app_periodic_fused_push(tmp_path):
    from rocm_mpi_b200.models.diffusion2d_kp import diffusion2D_kp
    nx, ny, nt = 66, 40, 30
    cfg = DiffusionConfig(nx=nx, ny=ny, nt=nt, periodx=1, periody=1, do_vis=False, quiet=True, transport="peer")
    res = diffusion2D_kp(cfg, finalize=False)
    T = res.T_local.cpu()
    g = igg.global_grid()
    dx, dy = 10.0 / igg.nx_g(), 10.0 / igg.ny_g()
    probe = torch.empty(nx, ny, device="meta")
    Th = ref.gaussian_ic(igg.coords_g(0, dx, probe), igg.coords_g(1, dy, probe), dx, dy, 10.0, 10.0)
    igg.finalize_global_grid(finalize_dist=False)
    vg = igg.VirtualGrid((nx, ny), (1, 1), (1, 1))
    Cp = torch.ones_like(Th)
    qx, qy, dd = torch.zeros(nx - 1, ny - 2, dtype=torch.float64), torch.zeros(nx - 2, ny - 1, dtype=torch.float64), \
        torch.zeros(nx - 2, ny - 2, dtype=torch.float64)
    dt = ref.time_step(dx, dy)
    vg.update_halo([Th])
    for _ in range(nt):
        ref.kp_step(Th, Cp, qx, qy, dd, 1.0, dt, 1.0 / dx, 1.0 / dy)
        vg.update_halo([Th])
    assert float((T - Th).abs().max()) < 1e-12


@pytest.mark.parametrize("path", ["resident", "persistent"])
def test_kp_app_single_launch_paths_match_oracle_and_three_kernels(tmp_path, path):
    """``kp_path``: all nt steps in one launch with the kp arithmetic (reference scripts/diffusion_2D_kp.jl:20-51)."""
    from rocm_mpi_b200.models.diffusion2d_kp import diffusion2D_kp
    T_ref, _ = ref.oracle_2d(128, 128, 100, "kp")
    three = diffusion2D_kp(DiffusionConfig(nx=128, ny=128, nt=100, outdir=str(tmp_path), quiet=True))
    res = diffusion2D_kp(DiffusionConfig(nx=128, ny=128, nt=100, outdir=str(tmp_path), quiet=True, kp_path=path))
    assert float((res.T_v - T_ref[1:-1, 1:-1]).abs().max()) < 1e-12
    assert float((res.T_v - three.T_v).abs().max()) < 1e-13
    assert res.launches <= 3          # IC + ONE stepping launch (+ interior extraction), not 300


@pytest.mark.parametrize("mode", ["resident", "persistent"])
def test_kp_form_divides_by_cp(mode):
    """A non-constant Cp separates the kp form (1/Cp) from the perf form (Cp *); periodic = self-neighbour pushes."""
    from rocm_mpi_b200.models.solver import Diffusion2D
    nx, ny, nt = 66, 128, 12
    igg.init_global_grid(nx, ny, 1, periodx=1, periody=1, quiet=True, init_dist=False)
    s = Diffusion2D(nx, ny, mode=mode, form="kp")
    g = torch.Generator().manual_seed(5)
    Cp = torch.rand(nx, ny, generator=g, dtype=torch.float64) + 0.5
    s.Cp.copy_(Cp)
    Th = s.T.cpu().clone()
    s.begin()
    s.step(nt)
    s.finish()
    got = s.T.cpu()
    dx, dy, dt = s.dx, s.dy, s.dt
    igg.finalize_global_grid(finalize_dist=False)
    vg = igg.VirtualGrid((nx, ny), (1, 1), (1, 1))
    qx, qy, dd = torch.zeros(nx - 1, ny - 2, dtype=torch.float64), torch.zeros(nx - 2, ny - 1, dtype=torch.float64), \
        torch.zeros(nx - 2, ny - 2, dtype=torch.float64)
    vg.update_halo([Th])
    for _ in range(nt):
        ref.kp_step(Th, Cp, qx, qy, dd, 1.0, dt, 1.0 / dx, 1.0 / dy)
        vg.update_halo([Th])
    assert float((got - Th).abs().max()) < 1e-12


def test_perf_and_hide_apps_small(tmp_path):
    from rocm_mpi_b200.models.diffusion2d_perf import diffusion2D_perf
    from rocm_mpi_b200.models.diffusion2d_perf_hide import diffusion2D_perf_hide
    T_ref, _ = ref.oracle_2d(254, 254, 60, "perf")
    for fn, kw in ((diffusion2D_perf, dict(fused=True)), (diffusion2D_perf, dict(fused=False)),
                   (diffusion2D_perf_hide, dict(fused=False, b_width=(16, 4, 4)))):
        cfg = DiffusionConfig(nx=254, ny=254, nt=60, warmup=10, outdir=str(tmp_path), quiet=True, **kw)
        res = fn(cfg)
        assert res.t_eff > 0 and res.wtime > 0
        assert float((res.T_v - T_ref[1:-1, 1:-1]).abs().max()) < 1e-12
        assert abs(res.max_T - float(T_ref[1:-1, 1:-1].max())) < 1e-12


def test_perf_hide_prof_writes_profile(tmp_path):
    from rocm_mpi_b200.models.diffusion2d_perf_hide_prof import diffusion2D_perf_hide_prof
    cfg = DiffusionConfig(nx=512, ny=512, nt=40, warmup=10, do_vis=False, quiet=True, b_width=(32, 8, 8), fused=False)
    prof = str(tmp_path / "prof.txt")
    res = diffusion2D_perf_hide_prof(cfg, prof)
    assert res.t_eff > 0
    text = open(prof).read()
    assert "cumulative" in text and "step" in text


def test_selectdevice_single_rank():
    from rocm_mpi_b200.models.selectdevice import selectdevice_test
    out = selectdevice_test(4, quiet=True)
    assert out["peer"].tolist() == [0.0] * 4 and out["nccl"].tolist() == [0.0] * 4


@pytest.mark.parametrize("variant", ["tma", "direct"])
def test_3d_app_matches_oracle(tmp_path, variant):
    from rocm_mpi_b200.models.diffusion3d import diffusion3D
    cfg = DiffusionConfig(nx=40, ny=36, nz=66, nt=12, warmup=2, kernel=variant, outdir=str(tmp_path), quiet=True,
                          do_vis=True, b_width=(4, 4, 8))
    res = diffusion3D(cfg)
    T_ref, _ = ref.oracle_3d(40, 36, 66, 12)
    assert float((res.T_v - T_ref[1:-1, 1:-1, 1:-1]).abs().max()) < 1e-12
    assert os.path.exists(res.png) and os.path.exists(res.raw)


def test_hide_communication_generic_gpu():
    igg.init_global_grid(200, 180, 1, periodx=1, quiet=True, init_dist=False)
    g = torch.Generator().manual_seed(8)
    T = torch.rand(200, 180, generator=g, dtype=torch.float64).cuda()
    Cp = (torch.rand(200, 180, generator=g, dtype=torch.float64) + 0.5).cuda()
    T2 = T.clone()

    def step(region):
        T2[region] = ref._perf_update(T, Cp, 1.0, 1e-3, 2.0, 3.0, region[0], region[1])

    igg.hide_communication((16, 4), step, T2)
    torch.cuda.synchronize()
    want = T.cpu().clone()
    ref.perf_step(want, T.cpu(), Cp.cpu(), 1.0, 1e-3, 2.0, 3.0)
    igg.VirtualGrid((200, 180), (1, 1), (1, 0)).update_halo([want])
    assert float((T2.cpu() - want).abs().max()) < 1e-13


def test_persistent_multistep_matches_oracle_and_counts_one_launch():
    from rocm_mpi_b200.models.solver import Diffusion2D
    igg.init_global_grid(128, 128, 1, quiet=True, init_dist=False)
    s = Diffusion2D(128, 128, mode="persistent")
    s.begin()
    n0 = native.launch_count()
    s.step(101)
    assert native.launch_count() - n0 == 1            # 101 time steps, one kernel launch
    s.step(99)
    s.finish()
    T_ref, _ = ref.oracle_2d(128, 128, 200, "perf")
    assert float((s.T.cpu() - T_ref).abs().max()) < 1e-13


@pytest.mark.parametrize("mode", ["fused", "plain", "hide"])
def test_step_controlled_end_to_end_values(mode):
    from rocm_mpi_b200.models.solver import Diffusion2D
    igg.init_global_grid(200, 260, 1, periodx=1, quiet=True, init_dist=False)
    s = Diffusion2D(200, 260, mode=mode, variant="tma", transport="peer")
    ref_s = Diffusion2D(200, 260, mode="plain", variant="direct", transport="peer")
    s.begin()
    ref_s.begin()
    for k in range(4):
        mx, sm = s.step_controlled(dt=s.dt * (1.0 + 0.1 * k))       # host-chosen parameters take effect
        ref_s.dt_backup = ref_s.dt
        ref_s.dt = ref_s.dt_backup * (1.0 + 0.1 * k)
        ref_s.step(1)
        ref_s.dt = ref_s.dt_backup
        inner = ref_s.T[1:-1, 1:-1]
        assert abs(mx - float(inner.max())) < 1e-13
        assert abs(sm - float(inner.sum())) < 1e-9 * abs(float(inner.sum())) + 1e-9
    s.finish()
    ref_s.finish()
    assert float((s.T - ref_s.T).abs().max()) < 1e-13


@pytest.mark.parametrize("shape", [(128, 128), (65, 128), (254, 254), (33, 70), (17, 40)])
def test_cluster_resident_kernel_matches_oracle(shape):
    from rocm_mpi_b200.models.solver import Diffusion2D
    nx, ny = shape
    igg.init_global_grid(nx, ny, 1, quiet=True, init_dist=False)
    s = Diffusion2D(nx, ny, mode="resident")
    s.begin()
    n0 = native.launch_count()
    s.step(61)
    assert native.launch_count() - n0 == 1
    s.step(40)
    s.finish()
    T_ref, _ = ref.oracle_2d(nx, ny, 101, "perf")
    assert float((s.T.cpu() - T_ref).abs().max()) < 1e-13


def test_cluster_resident_kernel_rejects_large_fields():
    from rocm_mpi_b200.models.solver import Diffusion2D
    igg.init_global_grid(1024, 1024, 1, quiet=True, init_dist=False)
    s = Diffusion2D(1024, 1024, mode="resident")
    s.begin()
    with pytest.raises(native.NativeError):
        s.step(2)
