"""Cartesian process topology: the pure-arithmetic part of ``init_global_grid``.

The reference never implements this itself; it calls ImplicitGlobalGrid's
``init_global_grid(nx, ny, 1)`` (reference ``scripts/diffusion_2D_ap.jl:17``) which
in turn leans on ``MPI.Dims_create`` / ``Cart_create`` / ``Cart_coords`` /
``Cart_shift``.  Here the same *behaviour* is restated without MPI: ranks are laid out
row-major over ``dims`` (last dimension varies fastest, which is what MPI Cartesian
communicators do), and a missing neighbour is ``PROC_NULL``.

Everything in this file is plain Python and runs without a GPU.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

PROC_NULL = -1
NDIMS = 3


def _factorizations(n: int, k: int, max_factor: int | None = None):
    """Yield all non-increasing k-tuples of positive ints whose product is n."""
    if max_factor is None:
        max_factor = n
    if k == 1:
        if n <= max_factor:
            yield (n,)
        return
    f = min(n, max_factor)
    while f >= 1:
        if n % f == 0:
            for rest in _factorizations(n // f, k - 1, f):
                yield (f,) + rest
        f -= 1


def dims_create(nprocs: int, dims: Sequence[int]) -> List[int]:
    """Balanced factorisation of ``nprocs`` over the zero entries of ``dims``.

    Same contract as ``MPI_Dims_create``: non-zero entries are kept, the free entries
    are filled with factors that are as close to each other as possible and placed in
    non-increasing order (4 -> (2,2), 8 -> (4,2) in 2-D, (2,2,2) in 3-D, 6 -> (3,2)).
    """
    dims = [int(d) for d in dims]
    if nprocs < 1:
        raise ValueError("nprocs must be >= 1")
    if any(d < 0 for d in dims):
        raise ValueError("dims entries must be >= 0")
    fixed = 1
    for d in dims:
        if d > 0:
            fixed *= d
    free = [i for i, d in enumerate(dims) if d == 0]
    if nprocs % fixed != 0:
        raise ValueError(
            f"nprocs={nprocs} is not divisible by the product of the fixed dims {dims}")
    rest = nprocs // fixed
    if not free:
        if rest != 1:
            raise ValueError(f"fixed dims {dims} do not multiply to nprocs={nprocs}")
        return dims
    best = None
    for cand in _factorizations(rest, len(free)):
        key = (cand[0] - cand[-1], cand)  # smallest spread; ties -> lexicographically smallest
        if best is None or key < best[0]:
            best = (key, cand)
    for i, f in zip(free, best[1]):
        dims[i] = f
    return dims


def cart_coords(rank: int, dims: Sequence[int]) -> List[int]:
    """Row-major rank -> coordinates (last dim fastest), as MPI_Cart_coords."""
    coords = [0] * len(dims)
    r = rank
    for d in range(len(dims) - 1, -1, -1):
        coords[d] = r % dims[d]
        r //= dims[d]
    return coords


def cart_rank(coords: Sequence[int], dims: Sequence[int], periods: Sequence[int]) -> int:
    """Coordinates -> rank; out-of-range coordinates wrap if periodic, else PROC_NULL."""
    r = 0
    for d in range(len(dims)):
        c = coords[d]
        if c < 0 or c >= dims[d]:
            if periods[d]:
                c %= dims[d]
            else:
                return PROC_NULL
        r = r * dims[d] + c
    return r


def cart_shift(rank: int, dims: Sequence[int], periods: Sequence[int], dim: int,
               disp: int = 1) -> Tuple[int, int]:
    """(source, dest) ranks for a shift by ``disp`` along ``dim`` (MPI_Cart_shift).

    ``source`` is the neighbour on the low side, ``dest`` the neighbour on the high side.
    """
    c = cart_coords(rank, dims)
    lo = list(c)
    hi = list(c)
    lo[dim] -= disp
    hi[dim] += disp
    return cart_rank(lo, dims, periods), cart_rank(hi, dims, periods)


@dataclass(frozen=True)
class CartTopology:
    """Immutable description of one rank's place in the Cartesian process grid."""

    nprocs: int
    me: int
    dims: Tuple[int, int, int]
    periods: Tuple[int, int, int]
    coords: Tuple[int, int, int]
    # neighbors[side][dim]; side 0 = low, side 1 = high
    neighbors: Tuple[Tuple[int, int, int], Tuple[int, int, int]]
    disp: int = 1

    @staticmethod
    def create(nprocs: int, me: int, dims: Sequence[int], periods: Sequence[int],
               disp: int = 1) -> "CartTopology":
        dims = tuple(int(d) for d in dims)
        periods = tuple(int(bool(p)) for p in periods)
        prod = 1
        for d in dims:
            prod *= d
        if prod != nprocs:
            raise ValueError(f"dims {dims} do not multiply to nprocs={nprocs}")
        coords = tuple(cart_coords(me, dims))
        lo, hi = [], []
        for d in range(len(dims)):
            s, t = cart_shift(me, dims, periods, d, disp)
            lo.append(s)
            hi.append(t)
        return CartTopology(nprocs, me, dims, periods, coords, (tuple(lo), tuple(hi)), disp)

    def neighbor(self, side: int, dim: int) -> int:
        return self.neighbors[side][dim]

    def rank_at_offset(self, offset: Sequence[int]) -> int:
        """Rank whose coordinates are ``coords + offset`` (PROC_NULL if outside)."""
        c = [self.coords[d] + int(offset[d]) for d in range(len(self.dims))]
        return cart_rank(c, self.dims, self.periods)

    def has_neighbor(self, side: int, dim: int) -> bool:
        return self.neighbors[side][dim] != PROC_NULL
