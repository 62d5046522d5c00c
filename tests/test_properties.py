"""Property-based checks (hypothesis) of the decomposition arithmetic and of the two halo formulations.

The fixed cases live in ``test_topology.py`` / ``test_halo_virtual.py``; here shapes, rank grids, periodicity,
overlaps and staggering are drawn at random.  (The reference has no tests at all; its
decomposition arithmetic is ImplicitGlobalGrid's: ``dims_create``, ``Cart_*``, ``n_g = dims*(n-ol) + ol*(1-period)``.)
"""
import torch
from hypothesis import given, settings, strategies as st

from rocm_mpi_b200.parallel.topology import PROC_NULL, CartTopology, cart_coords, cart_rank, dims_create
from rocm_mpi_b200.parallel.virtual import VirtualGrid

SET = dict(max_examples=60, deadline=None)


@settings(**SET)
@given(st.integers(1, 64), st.tuples(st.integers(0, 4), st.integers(0, 4), st.integers(0, 4)))
def test_dims_create_properties(nprocs, fixed):
    prod_fixed = 1
    for f in fixed:
        prod_fixed *= f if f else 1
    if nprocs % prod_fixed or (all(fixed) and prod_fixed != nprocs):
        return                                        # no valid factorisation: covered by the error tests
    dims = dims_create(nprocs, list(fixed))
    p = 1
    for d in dims:
        p *= d
    assert p == nprocs
    assert all(d == f for d, f in zip(dims, fixed) if f)                       # fixed entries are kept
    free = [d for d, f in zip(dims, fixed) if not f]
    assert free == sorted(free, reverse=True)                                  # MPI_Dims_create: non-increasing
    if len(free) > 1:                                                          # ... and as balanced as possible
        n_free = nprocs // prod_fixed
        best = min(max(c) - min(c) for c in _factorisations(n_free, len(free)))
        assert max(free) - min(free) == best


def _factorisations(n, k):
    if k == 1:
        yield (n,)
        return
    for a in range(1, n + 1):
        if n % a == 0:
            for rest in _factorisations(n // a, k - 1):
                yield (a,) + rest


@settings(**SET)
@given(st.tuples(st.integers(1, 4), st.integers(1, 4), st.integers(1, 3)),
       st.tuples(st.booleans(), st.booleans(), st.booleans()))
def test_cartesian_topology_is_consistent(dims, periods):
    n = dims[0] * dims[1] * dims[2]
    per = tuple(int(p) for p in periods)
    for r in range(n):
        c = cart_coords(r, dims)
        assert cart_rank(c, dims, per) == r
        topo = CartTopology.create(n, r, dims, per)
        for d in range(3):
            lo, hi = topo.neighbors[0][d], topo.neighbors[1][d]
            if hi != PROC_NULL:                       # my high neighbour's low neighbour is me
                assert CartTopology.create(n, hi, dims, per).neighbors[0][d] == r
            else:
                assert not per[d] and c[d] == dims[d] - 1
            if lo != PROC_NULL:
                assert CartTopology.create(n, lo, dims, per).neighbors[1][d] == r
            else:
                assert not per[d] and c[d] == 0


@settings(max_examples=40, deadline=None)
@given(st.data())
def test_box_exchange_equals_sequential_exchange(data):
    """``update_halo`` as independent box copies (what the one-launch peer kernel does) == the x, y, z sequential
    definition, for random layouts, overlaps, periodicity and staggered field sizes, arbitrary field contents."""
    nd = data.draw(st.integers(2, 3))
    dims = tuple(data.draw(st.integers(1, 3)) for _ in range(nd))
    overlaps = tuple(data.draw(st.integers(2, 3)) for _ in range(3))
    periods = tuple(int(data.draw(st.booleans())) for _ in range(nd))
    n = tuple(data.draw(st.integers(2 * overlaps[d] + 1, 2 * overlaps[d] + 4)) for d in range(nd))
    vg = VirtualGrid(n, dims, periods, overlaps)
    stag = tuple(data.draw(st.integers(-1, 1)) for _ in range(nd))
    shape = tuple(n[d] + stag[d] for d in range(nd))
    gen = torch.Generator().manual_seed(data.draw(st.integers(0, 1000)))
    a = [torch.rand(shape, generator=gen, dtype=torch.float64) for _ in range(vg.nprocs)]
    b = [t.clone() for t in a]
    vg.update_halo(a)
    vg.update_halo_boxes(b)
    for r in range(vg.nprocs):
        assert torch.equal(a[r], b[r]), (n, dims, periods, overlaps, shape, r)


@settings(max_examples=40, deadline=None)
@given(st.tuples(st.integers(1, 3), st.integers(1, 3)), st.tuples(st.integers(5, 9), st.integers(5, 9)),
       st.integers(0, 1000))
def test_scatter_update_gather_roundtrip(dims, n, seed):
    """Cutting a global field into overlapping local arrays, exchanging halos (a no-op on consistent data) and
    gathering the halo-free interiors gives back the global field minus its outermost ring."""
    vg = VirtualGrid(n, dims)
    gen = torch.Generator().manual_seed(seed)
    G = torch.rand(vg.nxyz_g[0], vg.nxyz_g[1], generator=gen, dtype=torch.float64)
    loc = vg.scatter(G)
    before = [t.clone() for t in loc]
    vg.update_halo(loc)
    assert all(torch.equal(x, y) for x, y in zip(loc, before))
    out = vg.gather([t[1:-1, 1:-1].contiguous() for t in loc])
    assert torch.equal(out, G[1:-1, 1:-1])


@settings(max_examples=40, deadline=None)
@given(st.integers(3, 70), st.integers(2, 160), st.sampled_from([(8, 4), (8, 12), (8, 28), (16, 12), (16, 28)]),
       st.sampled_from([2, 4]), st.integers(1, 9), st.integers(0, 1000))
def test_two_step_kernel_index_code_on_random_shapes(nx, half_ny, srb_scx, rpt, workers, seed):
    """The index code of the two-steps-per-pass CUDA kernel (run on the host, thread by thread) equals two plain
    steps bit for bit for random field shapes, block heights, item heights and CTA counts."""
    from rocm_mpi_b200.ops import native, reference as ref
    ny = 2 * half_ny
    srb, scx = srb_scx
    gen = torch.Generator().manual_seed(seed)
    T = torch.rand(nx, ny, generator=gen, dtype=torch.float64)
    Cp = torch.rand(nx, ny, generator=gen, dtype=torch.float64) + 0.5
    a = T.clone()
    ref.perf_step(a, T, Cp, 1.1, 2e-3, 1.5, 2.5)
    b = a.clone()
    ref.perf_step(b, a, Cp, 1.1, 2e-3, 1.5, 2.5)
    T2 = T.clone()
    native.cpu_temporal2d_emulate(T2, T, Cp, 1.1, 2e-3, 1.5, 2.5, scx=scx, rpt=rpt, workers=workers, srb=srb)
    assert torch.equal(T2, b), (nx, ny, srb, scx, rpt, workers)
