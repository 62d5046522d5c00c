import pytest
import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.parallel.topology import (PROC_NULL, CartTopology, cart_coords, cart_rank, cart_shift,
                                              dims_create)


def test_dims_create_matches_mpi_conventions():
    assert dims_create(4, [0, 0, 1]) == [2, 2, 1]
    assert dims_create(8, [0, 0, 1]) == [4, 2, 1]
    assert dims_create(8, [0, 0, 0]) == [2, 2, 2]
    assert dims_create(6, [0, 0, 1]) == [3, 2, 1]
    assert dims_create(12, [0, 0, 0]) == [3, 2, 2]
    assert dims_create(8, [2, 0, 1]) == [2, 4, 1]
    assert dims_create(7, [0, 0, 1]) == [7, 1, 1]
    assert dims_create(1, [0, 0, 0]) == [1, 1, 1]
    with pytest.raises(ValueError):
        dims_create(8, [3, 0, 1])


def test_cart_coords_roundtrip_and_shift():
    dims = (4, 2, 1)
    for r in range(8):
        c = cart_coords(r, dims)
        assert cart_rank(c, dims, (0, 0, 0)) == r
    # row-major: last dim fastest
    assert cart_coords(1, dims) == [0, 1, 0]
    assert cart_coords(2, dims) == [1, 0, 0]
    assert cart_shift(0, dims, (0, 0, 0), 0) == (PROC_NULL, 2)
    assert cart_shift(0, dims, (1, 0, 0), 0) == (6, 2)
    assert cart_shift(7, dims, (0, 0, 0), 1) == (6, PROC_NULL)
    t = CartTopology.create(8, 3, dims, (0, 0, 0))
    assert t.coords == (1, 1, 0)
    assert t.neighbors == ((1, 2, PROC_NULL), (5, PROC_NULL, PROC_NULL))
    assert t.rank_at_offset((1, -1, 0)) == 4
    assert t.rank_at_offset((-1, 1, 0)) == PROC_NULL


def test_init_global_grid_single_process():
    me, dims, nprocs, coords, comm_cart = igg.init_global_grid(128, 128, 1, device_type="cpu", quiet=True)
    assert (me, dims, nprocs, coords) == (0, (1, 1, 1), 1, (0, 0, 0))
    assert comm_cart.size == 1
    assert (igg.nx_g(), igg.ny_g(), igg.nz_g()) == (128, 128, 1)
    with pytest.raises(RuntimeError):
        igg.init_global_grid(4, 4, 1, device_type="cpu", quiet=True)
    igg.finalize_global_grid()
    with pytest.raises(RuntimeError):
        igg.nx_g()
    # periodic: the overlap is not counted in the global size
    igg.init_global_grid(128, 64, 1, periodx=1, device_type="cpu", quiet=True)
    assert (igg.nx_g(), igg.ny_g()) == (126, 64)


def test_init_argument_checks():
    with pytest.raises(ValueError):
        igg.init_global_grid(1, 8, 1, device_type="cpu", quiet=True)
    with pytest.raises(ValueError):
        igg.init_global_grid(8, 1, 8, device_type="cpu", quiet=True)
    with pytest.raises(ValueError):
        igg.init_global_grid(8, 8, 1, dimx=-1, device_type="cpu", quiet=True)
    with pytest.raises(ValueError):
        igg.init_global_grid(2, 8, 1, periodx=1, device_type="cpu", quiet=True)


def test_global_coordinates():
    igg.init_global_grid(8, 6, 1, device_type="cpu", quiet=True)
    dx = 0.5
    T = torch.zeros(8, 6)
    assert igg.x_g(0, dx, T) == 0.0
    assert igg.x_g(7, dx, T) == 3.5
    # staggered arrays are shifted by half a cell per missing/extra point
    qx = torch.zeros(7, 6)
    assert igg.x_g(0, dx, qx) == 0.25
    Vx = torch.zeros(9, 6)
    assert igg.x_g(0, dx, Vx) == -0.25
    xs = igg.coords_g(0, dx, T)
    assert torch.equal(xs, torch.arange(8, dtype=torch.float64) * dx)
    assert igg.ol(0, T) == 2 and igg.ol(0, qx) == 1 and igg.ol(0, Vx) == 3
    assert igg.nx_g(T) == 8 and igg.nx_g(qx) == 7 and igg.nx_g(Vx) == 9 and igg.ny_g(qx) == 6


def test_global_coordinates_periodic():
    igg.init_global_grid(8, 6, 1, periodx=1, device_type="cpu", quiet=True)
    dx = 1.0
    T = torch.zeros(8, 6)
    assert igg.nx_g() == 6
    # first cell is a ghost cell -> wraps to the far end
    assert igg.x_g(0, dx, T) == 5.0
    assert igg.x_g(1, dx, T) == 0.0
    assert igg.x_g(7, dx, T) == 0.0
    xs = igg.coords_g(0, dx, T)
    assert xs.tolist() == [5.0, 0.0, 1.0, 2.0, 3.0, 4.0, 5.0, 0.0]
