"""The implicit global grid: ``init_global_grid`` and friends.

The reference scripts call (never implement) this surface -- e.g.
``me, dims, nprocs, coords, comm_cart = init_global_grid(nx, ny, 1)`` at reference
``scripts/diffusion_2D_ap.jl:17``, ``nx_g()/ny_g()`` at ``:19``, ``x_g/y_g`` at ``:28`` and
``finalize_global_grid()`` at ``:48``.  The semantics implemented here are the ones written
down in SURVEY.md section 2.2 (G1-G3, G7, G8):

* the arguments are **local** array sizes including the overlap;
* ``dims`` comes from a ``Dims_create``-style balanced factorisation, with ``dims[k] = 1``
  forced where ``n[k] == 1``;
* the global size is ``n_g = dims*(n - overlap) + overlap*(period == 0)``;
* the device is chosen from the node-local rank (reference
  ``scripts/rocmaware_test_selectdevice.jl:7-9``).

Index convention of this framework (decided once, everywhere): API index order
``(ix, iy, iz)`` is tensor dimension order ``(0, 1, 2)`` of a row-major ``torch`` tensor and
all indices are **0-based**.  The contiguous tensor dimension is therefore the *last*
spatial dimension.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Optional, Sequence, Tuple

import torch

from .comm import Comm, local_rank, make_comm
from .topology import NDIMS, PROC_NULL, CartTopology, dims_create

__all__ = [
    "GlobalGrid", "init_global_grid", "finalize_global_grid", "global_grid", "grid_is_initialized",
    "select_device", "nx_g", "ny_g", "nz_g", "x_g", "y_g", "z_g", "coords_g", "ol", "CartComm",
]


@dataclass
class CartComm:
    """What the reference receives as ``comm_cart``: the communicator + its topology."""

    comm: Comm
    topo: CartTopology

    @property
    def rank(self) -> int:
        return self.topo.me

    @property
    def size(self) -> int:
        return self.topo.nprocs


@dataclass
class GlobalGrid:
    nxyz: Tuple[int, int, int]
    nxyz_g: Tuple[int, int, int]
    dims: Tuple[int, int, int]
    overlaps: Tuple[int, int, int]
    periods: Tuple[int, int, int]
    nprocs: int
    me: int
    coords: Tuple[int, int, int]
    neighbors: Tuple[Tuple[int, int, int], Tuple[int, int, int]]
    disp: int
    reorder: int
    comm: Comm
    topo: CartTopology
    device_type: str
    device: torch.device
    quiet: bool = False
    # lazily created helpers (halo engines, peer fabric, timers) hang off the singleton
    cache: dict = field(default_factory=dict)

    @property
    def ndims_active(self) -> int:
        return sum(1 for n in self.nxyz if n > 1)

    def cart_comm(self) -> CartComm:
        return CartComm(self.comm, self.topo)


_GRID: Optional[GlobalGrid] = None


def grid_is_initialized() -> bool:
    return _GRID is not None


def global_grid() -> GlobalGrid:
    if _GRID is None:
        raise RuntimeError("No global grid: call init_global_grid() first "
                           "(or it has already been finalized).")
    return _GRID


def select_device(device_type: str = "auto") -> torch.device:
    """Bind this process to the GPU with the index of its node-local rank.

    Restates reference ``scripts/rocmaware_test_selectdevice.jl:7-9`` (``device!(rank_l+1)``,
    1-based there, 0-based here).  More ranks than GPUs on the node is an error, as it is
    upstream.
    """
    if device_type == "auto":
        device_type = "cuda" if torch.cuda.is_available() else "cpu"
    if device_type == "cpu":
        return torch.device("cpu")
    if device_type != "cuda":
        raise ValueError(f"device_type must be 'auto', 'cuda' or 'cpu', got {device_type!r}")
    if not torch.cuda.is_available():
        raise RuntimeError("device_type='cuda' requested but no CUDA device is available")
    ndev = torch.cuda.device_count()
    lr = local_rank()
    if lr >= ndev:
        raise RuntimeError(f"node-local rank {lr} has no GPU: only {ndev} device(s) visible")
    torch.cuda.set_device(lr)
    return torch.device("cuda", lr)


def init_global_grid(nx: int, ny: int, nz: int = 1, *, dimx: int = 0, dimy: int = 0, dimz: int = 0,
                     periodx: int = 0, periody: int = 0, periodz: int = 0,
                     overlaps: Sequence[int] = (2, 2, 2), disp: int = 1, reorder: int = 1,
                     comm: Optional[Comm] = None, init_dist: bool = True, device_type: str = "auto",
                     select_device: bool = True, quiet: bool = False):
    """Create the (singleton) implicit global grid from *local* sizes.

    Returns ``(me, dims, nprocs, coords, comm_cart)`` exactly as the reference unpacks it
    (``scripts/diffusion_2D_ap.jl:17``).  Keyword names mirror the upstream API
    (``init_MPI`` is spelled ``init_dist`` because the launcher is ``torchrun``).
    """
    global _GRID
    if _GRID is not None:
        raise RuntimeError("The global grid has already been initialized.")
    nxyz = (int(nx), int(ny), int(nz))
    overlaps = tuple(int(o) for o in overlaps)
    periods = (int(bool(periodx)), int(bool(periody)), int(bool(periodz)))
    if len(overlaps) != NDIMS:
        raise ValueError("overlaps must have 3 entries")
    if any(n < 1 for n in nxyz):
        raise ValueError("Invalid arguments: nx, ny, and nz cannot be less than 1.")
    if any(d < 0 for d in (dimx, dimy, dimz)):
        raise ValueError("Invalid arguments: dimx, dimy, and dimz cannot be negative.")
    if nxyz[0] == 1:
        raise ValueError("Invalid arguments: nx can never be 1.")
    if nxyz[1] == 1 and nxyz[2] > 1:
        raise ValueError("Invalid arguments: ny cannot be 1 if nz is greater than 1.")
    for n, p, what in zip(nxyz, periods, "xyz"):
        if n == 1 and p:
            raise ValueError(f"Invalid arguments: n{what} == 1 cannot be periodic.")
    for n, o, p, what in zip(nxyz, overlaps, periods, "xyz"):
        if p and n < 2 * o - 1:
            raise ValueError(f"Incoherent arguments: periodic n{what}={n} must be >= 2*overlap-1.")

    if device_type == "auto":
        device_type = "cuda" if torch.cuda.is_available() else "cpu"
    device = torch.device("cpu")
    if device_type == "cuda":
        if select_device:
            device = globals()["select_device"]("cuda")
        else:
            device = torch.device("cuda", torch.cuda.current_device())

    if comm is None:
        comm = make_comm(device_type, init_dist=init_dist)
    nprocs, me = comm.size, comm.rank

    dims = [int(dimx), int(dimy), int(dimz)]
    for k in range(NDIMS):
        if nxyz[k] == 1 and dims[k] == 0:
            dims[k] = 1
    dims = tuple(dims_create(nprocs, dims))
    topo = CartTopology.create(nprocs, me, dims, periods, disp)
    nxyz_g = tuple(dims[k] * (nxyz[k] - overlaps[k]) + overlaps[k] * (periods[k] == 0)
                   for k in range(NDIMS))
    # a dimension of local size 1 is not decomposed and has no overlap: keep it at 1
    nxyz_g = tuple(1 if nxyz[k] == 1 else nxyz_g[k] for k in range(NDIMS))

    _GRID = GlobalGrid(nxyz=nxyz, nxyz_g=nxyz_g, dims=dims, overlaps=overlaps, periods=periods,
                       nprocs=nprocs, me=me, coords=topo.coords, neighbors=topo.neighbors, disp=disp,
                       reorder=reorder, comm=comm, topo=topo, device_type=device_type, device=device,
                       quiet=quiet)
    if not quiet and me == 0:
        print("Global grid: %dx%dx%d (nprocs: %d, dims: %dx%dx%d)"
              % (nxyz_g + (nprocs,) + dims), flush=True)
    return me, dims, nprocs, topo.coords, _GRID.cart_comm()


def finalize_global_grid(*, finalize_dist: bool = True) -> None:
    """Release halo engines / peer mappings and (optionally) the process group.

    Counterpart of ``finalize_global_grid()`` at reference ``scripts/diffusion_2D_ap.jl:48``.
    """
    global _GRID
    g = global_grid()
    fab = g.cache.get("fabric")
    pending = None
    if fab is not None:
        try:
            fab.check_all_errors()
        except RuntimeError as e:           # still release everything, then report
            pending = e
    for key, obj in list(g.cache.items()):
        close = getattr(obj, "close", None)
        if callable(close):
            close()
    g.cache.clear()
    if finalize_dist:
        g.comm.finalize()
    _GRID = None
    if pending is not None:
        raise pending


def _n_g(dim: int, A=None) -> int:
    g = global_grid()
    if A is None:
        return g.nxyz_g[dim]
    size = A.shape[dim] if dim < A.dim() else 1
    return g.nxyz_g[dim] + (size - g.nxyz[dim])       # staggered arrays: same shift as locally


def nx_g(A=None) -> int:
    """Global number of grid points in x (reference ``scripts/diffusion_2D_ap.jl:19``); with a field
    argument, the global size of that (possibly staggered) field, as upstream ``nx_g(A)``."""
    return _n_g(0, A)


def ny_g(A=None) -> int:
    return _n_g(1, A)


def nz_g(A=None) -> int:
    return _n_g(2, A)


def ol(dim: int, A=None) -> int:
    """Overlap of field ``A`` along ``dim``: ``overlap[dim] + (size(A,dim) - n[dim])``."""
    g = global_grid()
    if A is None:
        return g.overlaps[dim]
    size = A.shape[dim] if dim < A.dim() else 1
    return g.overlaps[dim] + (size - g.nxyz[dim])


def _coord_g(dim: int, idx, d: float, A):
    g = global_grid()
    n = g.nxyz[dim]
    size = (A.shape[dim] if dim < A.dim() else 1) if hasattr(A, "shape") else int(A)
    x0 = 0.5 * (n - size) * d
    x = (g.coords[dim] * (n - g.overlaps[dim]) + idx) * d + x0
    if g.periods[dim]:
        ng = g.nxyz_g[dim]
        x = x - d  # first cell of a periodic global problem is a ghost cell
        if torch.is_tensor(x):
            x = torch.where(x > (ng - 1) * d, x - ng * d, x)
            x = torch.where(x < 0, x + ng * d, x)
        else:
            if x > (ng - 1) * d:
                x -= ng * d
            if x < 0:
                x += ng * d
    return x


def x_g(ix, dx: float, A):
    """Global x-coordinate of local index ``ix`` (0-based) of field ``A``.

    Same formula the reference uses through IGG at ``scripts/diffusion_2D_ap.jl:28``:
    ``x0 = 0.5*(nx - size(A,1))*dx`` (staggered-array shift) and
    ``x = (coords[0]*(nx - overlap) + ix)*dx + x0``, with periodic wrap.  ``ix`` may be an int
    or an integer/floating tensor (vectorised).
    """
    return _coord_g(0, ix, dx, A)


def y_g(iy, dy: float, A):
    return _coord_g(1, iy, dy, A)


def z_g(iz, dz: float, A):
    return _coord_g(2, iz, dz, A)


def coords_g(dim: int, d: float, A, device=None, dtype=torch.float64) -> torch.Tensor:
    """All global coordinates of ``A`` along ``dim`` as a 1-D tensor."""
    size = A.shape[dim] if dim < A.dim() else 1
    idx = torch.arange(size, dtype=dtype, device=device)
    return _coord_g(dim, idx, d, A)
