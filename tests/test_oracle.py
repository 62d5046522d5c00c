"""Golden values of SURVEY.md 0.3 / BASELINE.md and equivalence of decomposed and single-domain runs."""
import pytest
import torch

from rocm_mpi_b200.ops import reference as ref
from rocm_mpi_b200.parallel.virtual import VirtualGrid

GOLD_DT = 0.0003780495365868781
GOLD_MAX = 0.3978647302669094
GOLD_SUM = 2026.7584320460926


@pytest.fixture(scope="module")
def oracle_ap():
    return ref.oracle_2d(254, 254, 1000, "ap")


def test_golden_values(oracle_ap):
    T, dt = oracle_ap
    assert dt == GOLD_DT
    T_v = T[1:-1, 1:-1]
    assert T_v.shape == (252, 252)
    assert abs(float(T_v.max()) - GOLD_MAX) < 1e-15
    assert abs(float(T_v.sum()) - GOLD_SUM) < 1e-9
    assert abs(float(T[0, 0]) - 2.857e-22) < 1e-24   # global corner never updated


def test_forms_agree(oracle_ap):
    T_ap, _ = oracle_ap
    T_perf, _ = ref.oracle_2d(254, 254, 1000, "perf")
    T_kp, _ = ref.oracle_2d(254, 254, 1000, "kp")
    assert float((T_ap - T_perf).abs().max()) < 1e-15
    assert float((T_ap - T_kp).abs().max()) < 1e-15


@pytest.mark.parametrize("nxy,dims", [((128, 128), (2, 2)), ((65, 128), (4, 2)), ((128, 65), (2, 4))])
def test_decomposed_equals_single_domain(nxy, dims):
    nt = 60
    vg = VirtualGrid(nxy, dims)
    assert vg.nxyz_g[:2] == (254, 254)
    dx = dy = 10.0 / 254
    dt = ref.time_step(dx, dy)
    xs = torch.arange(254, dtype=torch.float64) * dx
    T0 = ref.gaussian_ic(xs, xs, dx, dy, 10.0, 10.0)
    locs = vg.scatter(T0)
    Cp = torch.ones(nxy, dtype=torch.float64)
    for _ in range(nt):
        for T in locs:
            ref.ap_step(T, Cp, 1.0, dt, dx, dy)
        vg.update_halo(locs)
    T_v = vg.gather([T[1:-1, 1:-1] for T in locs])
    T_ref, _ = ref.oracle_2d(254, 254, nt, "ap")
    assert torch.equal(T_v, T_ref[1:-1, 1:-1])   # bit-identical


def test_split_step_covers_domain():
    torch.manual_seed(0)
    T = torch.rand(40, 36, dtype=torch.float64)
    Cp = torch.rand(40, 36, dtype=torch.float64) + 0.5
    full = T.clone()
    ref.perf_step(full, T, Cp, 1.0, 1e-3, 3.0, 2.0)
    split = T.clone()
    ref.perf_step_region(split, T, Cp, 1.0, 1e-3, 3.0, 2.0, (5, 3), 1)
    m = ref.frame_mask(T.shape, (5, 3))
    assert torch.equal(split[m], full[m])
    assert torch.equal(split[~m], T[~m])
    ref.perf_step_region(split, T, Cp, 1.0, 1e-3, 3.0, 2.0, (5, 3), 2)
    assert torch.equal(split, full)


def test_oracle_3d_runs():
    T, dt = ref.oracle_3d(20, 18, 16, 5)
    assert T.shape == (20, 18, 16)
    assert 0 < float(T.max()) < 1


def test_frame_regions_tile_the_frame_exactly():
    from rocm_mpi_b200.parallel.hide import frame_regions, interior_region
    for shape, bw in (((40, 36), (5, 3)), ((12, 50), (8, 2)), ((20, 18, 16), (3, 2, 4))):
        cover = torch.zeros(shape, dtype=torch.int32)
        for r in frame_regions(shape, bw):
            cover[r] += 1
        m = ref.frame_mask(shape, bw)
        assert torch.equal(cover == 1, m) and int(cover.max()) <= 1
        inner = interior_region(shape, bw)
        if inner is not None:
            assert int(cover[inner].sum()) == 0


def test_hide_communication_cpu_single_rank_equals_full_step():
    import rocm_mpi_b200 as igg
    igg.init_global_grid(40, 36, 1, device_type="cpu", quiet=True)
    torch.manual_seed(4)
    T = torch.rand(40, 36, dtype=torch.float64)
    Cp = torch.rand(40, 36, dtype=torch.float64) + 0.5
    want = T.clone()
    ref.perf_step(want, T, Cp, 1.0, 1e-3, 2.0, 3.0)
    T2 = T.clone()

    def step(region):
        T2[region] = ref._perf_update(T, Cp, 1.0, 1e-3, 2.0, 3.0, region[0], region[1])

    igg.hide_communication((5, 3), step, T2)
    assert torch.equal(T2, want)
