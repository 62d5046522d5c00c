"""GPU numerics: every native kernel against a plain PyTorch fp64 statement of the same op."""
import pytest
import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.ops import native, reference as ref

pytestmark = pytest.mark.gpu

TOL = 1e-13


@pytest.fixture(scope="module", autouse=True)
def _need_native():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    native.require()


def _fields(nx, ny, seed=0):
    g = torch.Generator().manual_seed(seed)
    T = torch.rand(nx, ny, generator=g, dtype=torch.float64)
    Cp = torch.rand(nx, ny, generator=g, dtype=torch.float64) + 0.5
    return T, Cp


@pytest.mark.parametrize("variant", ["direct", "tma"])
@pytest.mark.parametrize("shape", [(64, 64), (130, 258), (33, 130), (257, 512), (3, 4), (40, 1026)])
def test_fused_step_matches_reference(variant, shape):
    nx, ny = shape
    T, Cp = _fields(nx, ny)
    want = T.clone()
    ref.perf_step(want, T, Cp, 1.3, 1e-3, 2.0, 3.0)
    Td, Cpd = T.cuda(), Cp.cuda()
    T2d = Td.clone()
    n0 = native.launch_count()
    native.diffusion2d_step(T2d, Td, Cpd, 1.3, 1e-3, 2.0, 3.0, variant=variant)
    assert native.launch_count() == n0 + 1
    assert float((T2d.cpu() - want).abs().max()) < TOL
    # boundary cells of T2 are never written
    assert torch.equal(T2d.cpu()[0], T[0]) and torch.equal(T2d.cpu()[:, -1], T[:, -1])


def test_direct_variant_handles_odd_ny_and_tma_rejects_it():
    T, Cp = _fields(50, 131)
    want = T.clone()
    ref.perf_step(want, T, Cp, 1.0, 1e-3, 2.0, 3.0)
    T2d = T.cuda().clone()
    native.diffusion2d_step(T2d, T.cuda(), Cp.cuda(), 1.0, 1e-3, 2.0, 3.0, variant="direct")
    assert float((T2d.cpu() - want).abs().max()) < TOL
    with pytest.raises(native.NativeError):
        native.diffusion2d_step(T2d, T.cuda(), Cp.cuda(), 1.0, 1e-3, 2.0, 3.0, variant="tma")


@pytest.mark.parametrize("variant", ["direct", "tma"])
@pytest.mark.parametrize("shape,bw", [((200, 520), (16, 4)), ((70, 300), (32, 8)), ((40, 40), (30, 30)),
                                      ((300, 260), (1, 1)), ((600, 300), (16, 4)), ((1100, 260), (40, 130))])
def test_split_step_frame_then_interior(variant, shape, bw):
    nx, ny = shape
    T, Cp = _fields(nx, ny, 3)
    full = T.clone()
    ref.perf_step(full, T, Cp, 1.0, 2e-3, 1.5, 2.5)
    m = ref.frame_mask(T.shape, bw)
    Td, Cpd = T.cuda(), Cp.cuda()
    T2d = Td.clone()
    native.diffusion2d_step(T2d, Td, Cpd, 1.0, 2e-3, 1.5, 2.5, mode=1, b_width=bw, variant=variant)
    got = T2d.cpu()
    assert float((got[m] - full[m]).abs().max()) < TOL          # frame computed
    assert torch.equal(got[~m], T[~m])                           # everything else untouched
    native.diffusion2d_step(T2d, Td, Cpd, 1.0, 2e-3, 1.5, 2.5, mode=2, b_width=bw, variant=variant)
    assert float((T2d.cpu() - full).abs().max()) < TOL          # frame + interior == full step


def test_dev_params_override_scalars():
    T, Cp = _fields(66, 130, 5)
    want = T.clone()
    ref.perf_step(want, T, Cp, 0.7, 3e-3, 2.0, 3.0)
    params = torch.tensor([0.7, 3e-3, 2.0, 3.0], dtype=torch.float64, device="cuda")
    for variant in ("direct", "tma"):
        T2d = T.cuda().clone()
        native.diffusion2d_step(T2d, T.cuda(), Cp.cuda(), 9.0, 9.0, 9.0, 9.0, variant=variant, dev_params=params)
        assert float((T2d.cpu() - want).abs().max()) < TOL


def test_kp_kernels_match_reference():
    nx, ny = 77, 101
    T, Cp = _fields(nx, ny, 7)
    lam, dt, _dx, _dy = 1.1, 1e-3, 4.0, 5.0
    qx = torch.zeros(nx - 1, ny - 2, dtype=torch.float64)
    qy = torch.zeros(nx - 2, ny - 1, dtype=torch.float64)
    dTdt = torch.zeros(nx - 2, ny - 2, dtype=torch.float64)
    Tr = T.clone()
    ref.kp_step(Tr, Cp, qx, qy, dTdt, lam, dt, _dx, _dy)
    Td, Cpd = T.cuda(), Cp.cuda()
    qxd, qyd, dd = torch.zeros_like(qx).cuda(), torch.zeros_like(qy).cuda(), torch.zeros_like(dTdt).cuda()
    native.flux2d(qxd, qyd, Td, lam, _dx, _dy)
    native.residual2d(dd, qxd, qyd, Cpd, _dx, _dy)
    native.update2d(Td, dd, dt)
    assert float((qxd.cpu() - qx).abs().max()) < TOL
    assert float((qyd.cpu() - qy).abs().max()) < TOL
    assert float((dd.cpu() - dTdt).abs().max()) < 1e-11   # 1/Cp amplifies rounding differences
    assert float((Td.cpu() - Tr).abs().max()) < TOL


@pytest.mark.parametrize("variant", ["direct", "tma"])
@pytest.mark.parametrize("shape", [(20, 18, 66), (9, 17, 130), (34, 10, 64), (3, 3, 4)])
def test_3d_step_matches_reference(variant, shape):
    g = torch.Generator().manual_seed(11)
    T = torch.rand(shape, generator=g, dtype=torch.float64)
    Cp = torch.rand(shape, generator=g, dtype=torch.float64) + 0.5
    want = T.clone()
    ref.perf_step_3d(want, T, Cp, 1.2, 1e-3, 2.0, 3.0, 4.0)
    T2d = T.cuda().clone()
    native.diffusion3d_step(T2d, T.cuda(), Cp.cuda(), 1.2, 1e-3, 2.0, 3.0, 4.0, variant=variant)
    assert float((T2d.cpu() - want).abs().max()) < TOL


@pytest.mark.parametrize("variant", ["direct", "tma"])
def test_3d_split_step(variant):
    shape, bw = (40, 36, 140), (4, 3, 8)
    g = torch.Generator().manual_seed(12)
    T = torch.rand(shape, generator=g, dtype=torch.float64)
    Cp = torch.rand(shape, generator=g, dtype=torch.float64) + 0.5
    full = T.clone()
    ref.perf_step_3d(full, T, Cp, 1.0, 1e-3, 2.0, 3.0, 4.0)
    m = ref.frame_mask(shape, bw)
    T2d = T.cuda().clone()
    native.diffusion3d_step(T2d, T.cuda(), Cp.cuda(), 1.0, 1e-3, 2.0, 3.0, 4.0, mode=1, b_width=bw, variant=variant)
    got = T2d.cpu()
    assert float((got[m] - full[m]).abs().max()) < TOL
    assert torch.equal(got[~m], T[~m])
    native.diffusion3d_step(T2d, T.cuda(), Cp.cuda(), 1.0, 1e-3, 2.0, 3.0, 4.0, mode=2, b_width=bw, variant=variant)
    assert float((T2d.cpu() - full).abs().max()) < TOL


def test_gaussian_ic_matches_host_comprehension():
    nx, ny, dx, dy = 70, 50, 10.0 / 254, 10.0 / 254
    T = torch.empty(nx, ny, dtype=torch.float64, device="cuda")
    native.gaussian_ic_2d(T, 3 * dx, 5 * dy, dx, dy, 10.0, 10.0)
    xs = (torch.arange(nx, dtype=torch.float64) + 3) * dx
    ys = (torch.arange(ny, dtype=torch.float64) + 5) * dy
    want = ref.gaussian_ic(xs, ys, dx, dy, 10.0, 10.0)
    assert float((T.cpu() - want).abs().max()) < 1e-15
    T3 = torch.empty(12, 10, 14, dtype=torch.float64, device="cuda")
    native.gaussian_ic_3d(T3, 0.0, dy, 0.0, dx, dy, dx, 10.0, 10.0, 10.0)
    want3 = ref.gaussian_ic_3d(torch.arange(12, dtype=torch.float64) * dx,
                               (torch.arange(10, dtype=torch.float64) + 1) * dy,
                               torch.arange(14, dtype=torch.float64) * dx, dx, dy, dx, 10.0, 10.0, 10.0)
    assert float((T3.cpu() - want3).abs().max()) < 1e-15


@pytest.mark.parametrize("shape", [(9, 14), (6, 7, 10)])
def test_pack_unpack_extract(shape):
    A = torch.arange(torch.Size(shape).numel(), dtype=torch.float64).reshape(shape)
    Ad = A.cuda()
    for dim in range(len(shape)):
        for idx in (0, 1, shape[dim] - 1):
            plane = A.select(dim, idx)
            buf = torch.empty(plane.numel(), dtype=torch.float64, device="cuda")
            native.pack_plane(Ad, dim, idx, buf)
            assert torch.equal(buf.cpu(), plane.reshape(-1))
            B = torch.zeros_like(Ad)
            native.unpack_plane(B, dim, idx, buf)
            assert torch.equal(B.cpu().select(dim, idx), plane)
            assert float(B.sum()) == float(plane.sum())
    inner = native.extract_inner(Ad)
    assert torch.equal(inner.cpu(), igg.inner(A))


def test_reduce_max_sum_repeated_calls():
    for n in (1, 1000, 1 << 20, (1 << 22) + 3):
        x = torch.rand(n, dtype=torch.float64, device="cuda") - 0.25
        for _ in range(2):   # scratch is reused: second call must be right too
            out = native.reduce_max_sum(x).cpu()
            assert float(out[0]) == float(x.max())
            assert abs(float(out[1]) - float(x.sum())) <= 1e-9 * max(1.0, abs(float(x.sum())))


def test_colormap_matches_numpy_lut():
    from rocm_mpi_b200.utils import output
    f = torch.rand(37, 53, dtype=torch.float64)
    want = output.colorize(f, 0.0, 1.0)
    got = output.colorize(f.cuda(), 0.0, 1.0)
    assert (want == got).all()


def test_smoke_entry_point():
    import __graft_entry__ as ge
    ge.smoke()


def test_in_kernel_diagnostic_matches_torch():
    for (nx, ny) in ((66, 130), (300, 1026), (1100, 2050)):
        T, Cp = _fields(nx, ny, 21)
        T = T - 0.3
        want = T.clone()
        ref.perf_step(want, T, Cp, 1.0, 1e-3, 2.0, 3.0)
        inner = want[1:-1, 1:-1]
        T2d = T.cuda().clone()
        out = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
        for rep in range(2):          # the scratch is reused between calls
            native.diffusion2d_step(T2d, T.cuda(), Cp.cuda(), 1.0, 1e-3, 2.0, 3.0, variant="tma", diag_out=out)
            got = out.cpu()
            assert abs(float(got[0]) - float(inner.max())) < 1e-14
            assert abs(float(got[1]) - float(inner.sum())) <= 1e-10 * abs(float(inner.sum())) + 1e-9
    with pytest.raises(native.NativeError):
        native.diffusion2d_step(T2d, T.cuda(), Cp.cuda(), 1.0, 1e-3, 2.0, 3.0, variant="direct", diag_out=out)


@pytest.mark.parametrize("shape", [(512, 130), (520, 260), (777, 130), (1030, 1026), (2048, 384)])
def test_strip_marching_kernel_matches_reference(shape):
    """nx >= 512 selects the strip-marching TMA kernel for whole-interior steps (partial chunks / blocks / strips)."""
    nx, ny = shape
    T, Cp = _fields(nx, ny, 31)
    want = T.clone()
    ref.perf_step(want, T, Cp, 1.3, 1e-3, 2.0, 3.0)
    T2d = T.cuda().clone()
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    native.diffusion2d_step(T2d, T.cuda(), Cp.cuda(), 1.3, 1e-3, 2.0, 3.0, variant="tma", diag_out=out)
    got = T2d.cpu()
    assert float((got - want).abs().max()) < TOL
    assert torch.equal(got[0], T[0]) and torch.equal(got[-1], T[-1]) and torch.equal(got[:, 0], T[:, 0])
    inner = want[1:-1, 1:-1]
    assert abs(float(out[0]) - float(inner.max())) < 1e-14
    assert abs(float(out[1]) - float(inner.sum())) <= 1e-10 * abs(float(inner.sum())) + 1e-9


def test_strip_kernel_agrees_with_direct_kernel_at_bench_like_size():
    """GPU-vs-GPU: the strip-marching TMA kernel and the direct-load kernel at a multi-chunk, many-strip size."""
    nx, ny = 4100, 8200
    g = torch.Generator(device="cuda").manual_seed(5)
    T = torch.rand(nx, ny, generator=g, dtype=torch.float64, device="cuda")
    Cp = torch.rand(nx, ny, generator=g, dtype=torch.float64, device="cuda") + 0.5
    A, B = T.clone(), T.clone()
    native.diffusion2d_step(A, T, Cp, 1.0, 1e-3, 2.0, 3.0, variant="tma")
    native.diffusion2d_step(B, T, Cp, 1.0, 1e-3, 2.0, 3.0, variant="direct")
    assert float((A - B).abs().max()) < 1e-13
    assert torch.equal(A[0], T[0]) and torch.equal(A[:, -1], T[:, -1])


@pytest.mark.parametrize("shape", [(5, 6), (66, 131), (300, 258), (17, 20, 23), (40, 36, 66)])
def test_extract_inner_and_device_gather_single_rank(shape):
    """Row-wise interior copy (16-byte stores, 8-byte aligned source): dense extraction and the gather kernel that
    stores the tile into the (here: own) root buffer -- reference scripts/diffusion_2D_ap.jl:45-46."""
    import rocm_mpi_b200 as igg
    from rocm_mpi_b200.parallel.gather import gather_interior_device, inner
    A = torch.rand(*shape, dtype=torch.float64, device="cuda")
    want = inner(A).contiguous()
    assert torch.equal(native.extract_inner(A), want)
    n = list(shape) + [1] * (3 - len(shape))
    igg.init_global_grid(n[0], n[1], n[2], quiet=True, init_dist=False)
    G = gather_interior_device(A)
    assert tuple(G.shape) == tuple(want.shape) and torch.equal(G, want)
    G2 = gather_interior_device(A * 2.0)             # cached buffer, second use
    assert torch.equal(G2, want * 2.0)
    igg.finalize_global_grid(finalize_dist=False)


def test_bench_size_kernels_against_cpu_oracle():
    """16384 x 16384 (the size bench.py times): one step of the strip-marching kernel and one two-step pass against
    the OpenMP C++ restatement of reference scripts/diffusion_2D_perf.jl:8-10 (csrc/cpu_oracle.cpp)."""
    import psutil
    if psutil.virtual_memory().available < 40 << 30:
        pytest.skip("needs ~14 GB of host memory for the CPU oracle fields")
    nx = ny = 16384
    g = torch.Generator().manual_seed(11)
    T = torch.rand(nx, ny, generator=g, dtype=torch.float64)
    Cp = torch.rand(nx, ny, generator=g, dtype=torch.float64) + 0.5
    ph = (1.3, 1e-3, 2.0, 3.0)
    a = T.clone()
    native.cpu_diffusion2d_step(a, T, Cp, *ph)
    b = a.clone()
    native.cpu_diffusion2d_step(b, a, Cp, *ph)
    Td, Cpd = T.cuda(), Cp.cuda()
    T2d = Td.clone()
    native.diffusion2d_step(T2d, Td, Cpd, *ph, variant="tma")            # nx >= 512: the strip kernel
    assert float((T2d.cpu() - a).abs().max()) < 1e-12
    T2d.copy_(Td)
    native.diffusion2d_step2(T2d, Td, Cpd, *ph)
    assert float((T2d.cpu() - b).abs().max()) < 1e-12


@pytest.mark.parametrize("n,offset", [(1, 0), (2, 0), (7, 0), (1000, 1), (100003, 0), (262144 + 5, 1)])
def test_reduce_max_sum_odd_lengths_and_unaligned_views(n, offset):
    """The reduction reads 16-byte pairs when the input is 16-byte aligned and single values otherwise (a view that
    starts at an odd element); odd lengths have a one-element tail."""
    g = torch.Generator().manual_seed(n)
    base = torch.rand(n + offset, generator=g, dtype=torch.float64).cuda()
    x = base[offset:]
    out = native.reduce_max_sum(x).cpu()
    assert float(out[0]) == float(x.max())
    assert abs(float(out[1]) - float(x.sum())) < 1e-9 * max(1.0, n)
    y = x.clone()
    y[n // 2] = float("nan")
    out = native.reduce_max_sum(y).cpu()
    assert out[0] != out[0] and out[1] != out[1]          # NaN shows up in the max as well as in the sum


def test_in_kernel_diagnostics_report_nan():
    """A blown-up field must show in the max, not only in the sum (the reference's maximum(T_v) propagates NaN,
    scripts/diffusion_2D_perf_hide.jl:115): one-step kernels (tile and strip) and the two-step kernel."""
    for shape in [(300, 380), (1100, 512)]:
        T = torch.rand(*shape, dtype=torch.float64, device="cuda")
        T[shape[0] // 2, shape[1] // 2] = float("nan")
        Cp = torch.ones_like(T)
        T2 = T.clone()
        d = torch.zeros(2, dtype=torch.float64, device="cuda")
        native.diffusion2d_step(T2, T, Cp, 1.0, 1e-4, 2.0, 3.0, variant="tma", diag_out=d)
        v = d.cpu()
        assert v[0] != v[0] and v[1] != v[1], (shape, v)
        d4 = torch.zeros(4, dtype=torch.float64, device="cuda")
        T2.copy_(T)
        native.diffusion2d_step2(T2, T, Cp, 1.0, 1e-4, 2.0, 3.0, diag_out=d4)
        v = d4.cpu()
        assert all(bool(x != x) for x in v), (shape, v)
