"""``gather``: assemble equally-shaped local blocks into one global array on the root rank.

Reference call sites: ``gather!(T_nh, T_v)`` at ``scripts/diffusion_2D_ap.jl:46`` (also
``kp.jl:95``, ``perf.jl:61``, ``perf_hide.jl:114``), always with the halo-free interior
``T[2:end-1,2:end-1]`` so that blocks abut exactly and the global array has
``(n-2)*dims`` points per dimension (``ap.jl:31-33``).  The block of the rank with Cartesian
coordinates ``(cx, cy, cz)`` lands at ``[cx*lx:(cx+1)*lx, cy*ly:..., cz*lz:...]``.
"""
from __future__ import annotations

from typing import Optional

import torch

from .grid import global_grid
from .topology import cart_coords

__all__ = ["gather", "inner", "inner_to_host", "gather_interior", "gather_interior_device", "device_gather_ok"]


def inner(A: torch.Tensor) -> torch.Tensor:
    """Halo-free interior view: ``A[1:-1, 1:-1(, 1:-1)]`` over every dimension of size > 2."""
    sl = tuple(slice(1, -1) if s > 2 else slice(None) for s in A.shape)
    return A[sl]


def gather(A_local: torch.Tensor, A_global: Optional[torch.Tensor], root: int = 0) -> None:
    """Gather ``A_local`` from every rank into ``A_global`` on ``root``.

    ``A_global`` only needs to exist on the root (pass ``None`` elsewhere).  Device tensors
    are accepted; they are brought to the host first, as the reference does with
    ``Array(T[2:end-1,2:end-1])`` (``scripts/diffusion_2D_ap.jl:45``).
    """
    g = global_grid()
    fab = g.cache.get("fabric")
    if fab is not None:
        fab.check_all_errors()              # never gather a field whose halo exchange timed out
    block = A_local.detach()
    if block.is_cuda:
        block = block.contiguous().cpu()
    blocks = g.comm.gather_blocks(block.contiguous(), root)
    if g.me != root:
        return
    if A_global is None:
        raise ValueError("A_global must be provided on the root rank")
    lshape = tuple(block.shape)
    nd = len(lshape)
    want = tuple(lshape[d] * g.dims[d] for d in range(nd))
    if tuple(A_global.shape) != want:
        raise ValueError(f"A_global has shape {tuple(A_global.shape)}, expected {want}")
    for r, blk in enumerate(blocks):
        c = cart_coords(r, g.dims)
        sl = tuple(slice(c[d] * lshape[d], (c[d] + 1) * lshape[d]) for d in range(nd))
        A_global[sl].copy_(blk)


def inner_to_host(A: torch.Tensor) -> torch.Tensor:
    """Halo-free interior of ``A`` as a contiguous host tensor.

    Restates ``Array(T[2:end-1,2:end-1])`` (reference ``scripts/diffusion_2D_ap.jl:45``): on CUDA the strided
    interior is compacted by the native ``extract_inner`` kernel and copied into pinned host memory."""
    if A.is_cuda and A.dtype == torch.float64 and A.is_contiguous():
        from ..ops import native
        if native.available():
            dev = native.extract_inner(A)
            host = torch.empty(dev.shape, dtype=dev.dtype, pin_memory=True)
            host.copy_(dev, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return host
    return inner(A).contiguous().cpu()


def gather_interior(A: torch.Tensor, root: int = 0) -> Optional[torch.Tensor]:
    """Convenience: strip the halo, gather, and return the global array (a HOST tensor) on the root.

    CUDA fields on one node take the device path (tiles stored into the root's memory over NVLink, then one
    device-to-host copy on the root); everything else is staged through the host as the reference does."""
    g = global_grid()
    if device_gather_ok(A) and all(g.nxyz[d] > 2 or g.dims[d] == 1 for d in range(A.dim())):
        G = gather_interior_device(A, root)
        return G.cpu() if G is not None else None
    blk = inner_to_host(A)
    out = None
    if g.me == root:
        out = torch.zeros(tuple(blk.shape[d] * g.dims[d] for d in range(blk.dim())),
                          dtype=A.dtype)
    gather(blk, out, root)
    return out


def device_gather_ok(A: torch.Tensor) -> bool:
    """Can ``gather_interior_device`` serve this field?  (CUDA fp64, contiguous, native library, one node.)"""
    if not (A.is_cuda and A.dtype == torch.float64 and A.is_contiguous()):
        return False
    from ..ops import native
    if not native.available():
        return False
    from .comm import spans_nodes
    return not spans_nodes()


def gather_interior_device(A: torch.Tensor, root: int = 0, sync: bool = True) -> Optional[torch.Tensor]:
    """``gather!`` of the halo-free interior WITHOUT host staging: returns the global array as a device tensor
    on ``root`` (``None`` elsewhere).

    The reference copies every rank's interior to the host and lets MPI collect the blocks
    (``Array(T[2:end-1,2:end-1])`` then ``gather!``, ``scripts/diffusion_2D_ap.jl:45-46``; 1.2 GB per rank in
    the perf run, ``scripts/diffusion_2D_perf.jl:59-63``).  Here the global array exists once, in the root's
    memory; every rank runs ONE kernel that stores its tile into it at its Cartesian position -- through a
    peer-mapped pointer over NVLink for the other ranks (``igg_gather_inner``).  Collective.  The returned tensor
    is a cached buffer of the fabric: the next ``gather_interior_device`` of the same shape overwrites it."""
    from ..ops import native
    from .fabric import get_fabric
    g = global_grid()
    fab = get_fabric(g)
    fab.check_all_errors()                  # never gather a field whose halo exchange timed out
    nd = A.dim()
    m = tuple(s - 2 if s > 2 else s for s in A.shape)
    gshape = tuple(m[d] * g.dims[d] for d in range(nd))
    G, ptr = fab.root_buffer(gshape, A.dtype, root)
    native.gather_inner(A, ptr, g.dims[:nd], g.coords[:nd])
    if sync:
        torch.cuda.current_stream().synchronize()
        g.comm.barrier()                    # every tile has landed in the root's buffer
    return G if g.me == root else None
