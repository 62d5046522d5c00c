"""update_halo semantics on the virtual-rank model (no processes, no GPU)."""
import torch

from rocm_mpi_b200.parallel.halo import halo_planes
from rocm_mpi_b200.parallel.virtual import VirtualGrid


def test_halo_planes_indices():
    specs = halo_planes((8, 6, 1), (2, 2, 2), (8, 6))
    assert specs[0].send == (1, 6) and specs[0].recv == (0, 7)
    assert specs[1].send == (1, 4) and specs[1].recv == (0, 5)
    assert specs[2] is None
    # staggered (n+1): overlap 3 -> planes 2 and size-3
    specs = halo_planes((8, 6, 1), (2, 2, 2), (9, 6))
    assert specs[0].send == (2, 6) and specs[0].recv == (0, 8)
    # staggered (n-1): overlap 1 -> skipped in that dim
    specs = halo_planes((8, 6, 1), (2, 2, 2), (7, 6))
    assert specs[0] is None and specs[1] is not None


def test_corner_propagation_2x2():
    vg = VirtualGrid((6, 5), (2, 2))
    locs = [torch.full((6, 5), float(r)) for r in range(4)]
    for r in range(4):
        locs[r][1:-1, 1:-1] = 10.0 * (r + 1)
    vg.update_halo(locs)
    # rank 0 at coords (0,0): high-x neighbour is rank 2, high-y neighbour rank 1, diagonal rank 3
    assert torch.all(locs[0][5, 1:4] == 30.0)
    assert torch.all(locs[0][1:5, 4] == 20.0)
    assert locs[0][5, 4] == 40.0             # corner value came from the diagonal neighbour
    assert locs[0][0, 0] == 0.0              # global boundary untouched
    assert locs[3][0, 0] == 10.0


def test_periodic_self_neighbor():
    vg = VirtualGrid((6, 5), (1, 1), periods=(1, 0))
    A = torch.arange(30, dtype=torch.float64).reshape(6, 5)
    ref = A.clone()
    vg.update_halo([A])
    assert torch.equal(A[0], ref[4]) and torch.equal(A[5], ref[1])
    assert torch.equal(A[1:5], ref[1:5])


def test_3d_global_consistency():
    # after one update_halo every overlapping cell pair holds the same value
    vg = VirtualGrid((6, 5, 4), (2, 2, 2))
    G = torch.arange(10 * 8 * 6, dtype=torch.float64).reshape(10, 8, 6)
    locs = vg.scatter(G)
    for r, L in enumerate(locs):      # scramble halos, keep interiors
        keep = L[1:-1, 1:-1, 1:-1].clone()
        L.fill_(-1.0)
        L[1:-1, 1:-1, 1:-1] = keep
    vg.update_halo(locs)
    for r, L in enumerate(locs):
        o = vg.origin(r)
        want = G[o[0]:o[0] + 6, o[1]:o[1] + 5, o[2]:o[2] + 4]
        interior_of_global = torch.zeros_like(G, dtype=torch.bool)
        interior_of_global[1:-1, 1:-1, 1:-1] = True
        m = interior_of_global[o[0]:o[0] + 6, o[1]:o[1] + 5, o[2]:o[2] + 4]
        assert torch.equal(L[m], want[m])


import pytest


@pytest.mark.parametrize("nxyz,dims,periods", [
    ((7, 6), (2, 2), (0, 0)), ((7, 6), (3, 2), (0, 0)), ((7, 6), (1, 1), (1, 1)), ((7, 6), (2, 1), (1, 0)),
    ((7, 6), (2, 2), (1, 1)), ((7, 6), (3, 1), (0, 1)), ((6, 5, 5), (2, 2, 2), (0, 0, 0)),
    ((6, 5, 5), (2, 1, 3), (1, 0, 1)), ((6, 5, 5), (1, 1, 1), (1, 1, 1)), ((6, 5, 5), (3, 2, 1), (0, 1, 0)),
])
def test_box_form_equals_sequential_form(nxyz, dims, periods):
    """The one-phase box copies used by the peer-memory kernel reproduce the x->y->z exchange exactly,
    for arbitrary (inconsistent) field contents, staggered fields included."""
    vg = VirtualGrid(nxyz, dims, periods)
    gen = torch.Generator().manual_seed(7)
    nd = len(nxyz)
    shapes = [tuple(nxyz), tuple(n + (1 if d == 0 else 0) for d, n in enumerate(nxyz)),
              tuple(n - (1 if d == nd - 1 else 0) for d, n in enumerate(nxyz)),
              tuple(n + 1 for n in nxyz)]
    for shp in shapes:
        a = [torch.rand(shp, generator=gen, dtype=torch.float64) for _ in range(vg.nprocs)]
        b = [t.clone() for t in a]
        vg.update_halo(a)
        vg.update_halo_boxes(b)
        for r in range(vg.nprocs):
            assert torch.equal(a[r], b[r]), (shp, r)


@pytest.mark.parametrize("overlaps", [(3, 3, 3), (2, 3, 2), (4, 2, 2)])
def test_wider_overlaps_box_form_and_global_size(overlaps):
    """overlaps > 2 (upstream keyword): plane indices shift, the box form still equals the sequential form."""
    vg = VirtualGrid((9, 8), (3, 2), (0, 1), overlaps)
    assert vg.nxyz_g[0] == 3 * (9 - overlaps[0]) + overlaps[0]
    assert vg.nxyz_g[1] == 2 * (8 - overlaps[1])            # periodic: overlap not counted
    specs = halo_planes((9, 8, 1), overlaps, (9, 8))
    assert specs[0].send == (overlaps[0] - 1, 9 - overlaps[0]) and specs[0].recv == (0, 8)
    gen = torch.Generator().manual_seed(11)
    a = [torch.rand(9, 8, generator=gen, dtype=torch.float64) for _ in range(vg.nprocs)]
    b = [t.clone() for t in a]
    vg.update_halo(a)
    vg.update_halo_boxes(b)
    for r in range(vg.nprocs):
        assert torch.equal(a[r], b[r])


def test_per_dimension_transport_switches(monkeypatch):
    """IGG_B200_TRANSPORT_DIMX/Y/Z: the analogue of upstream's IGG_ROCMAWARE_MPI_DIMX/_DIMY/_DIMZ."""
    from rocm_mpi_b200.parallel.halo import per_dim_staging
    for d in "XYZ":
        monkeypatch.delenv(f"IGG_B200_TRANSPORT_DIM{d}", raising=False)
    assert per_dim_staging("nccl") == (False, False, False)
    assert per_dim_staging("host") == (True, True, True)
    monkeypatch.setenv("IGG_B200_TRANSPORT_DIMY", "host")
    assert per_dim_staging("nccl") == (False, True, False)
    monkeypatch.setenv("IGG_B200_TRANSPORT_DIMX", "nccl")
    assert per_dim_staging("host") == (False, True, True)
    monkeypatch.setenv("IGG_B200_TRANSPORT_DIMZ", "mpi")
    with pytest.raises(ValueError):
        per_dim_staging("nccl")


def test_default_transport_leaves_the_node_over_nccl(monkeypatch):
    """Peer mappings are CUDA IPC (one node); a job that spans nodes defaults to the NCCL transport."""
    import types
    from rocm_mpi_b200.parallel.comm import spans_nodes
    from rocm_mpi_b200.parallel.halo import default_transport
    dev_field, cpu_field = types.SimpleNamespace(is_cuda=True), types.SimpleNamespace(is_cuda=False)
    monkeypatch.delenv("IGG_B200_TRANSPORT", raising=False)
    monkeypatch.setenv("WORLD_SIZE", "8")
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    assert not spans_nodes() and default_transport(dev_field) == "peer"
    monkeypatch.setenv("WORLD_SIZE", "16")
    assert spans_nodes() and default_transport(dev_field) == "nccl"
    assert default_transport(cpu_field) == "gloo"
    monkeypatch.setenv("IGG_B200_TRANSPORT", "host")
    assert default_transport(dev_field) == "host"
    monkeypatch.delenv("LOCAL_WORLD_SIZE")
    assert not spans_nodes()                         # no launcher information: assume one node
