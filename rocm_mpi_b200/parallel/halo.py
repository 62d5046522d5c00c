"""``update_halo``: the dimension-sequential halo exchange.

The reference calls ``update_halo!(T)`` once per time step (``scripts/diffusion_2D_ap.jl:42``,
``scripts/diffusion_2D_kp.jl:91``, ``scripts/diffusion_2D_perf.jl:51``) and leaves the body
to ImplicitGlobalGrid.  The semantics restated here:

* dimensions are processed **in order x, y, z, one after the other** -- that ordering is
  what carries corner values to diagonal neighbours;
* for a field ``A`` with overlap ``ol = overlap[dim] + (size(A,dim) - n[dim])`` the plane
  sent to the low neighbour is index ``ol-1`` and to the high neighbour ``size-ol``
  (0-based); it spans the **whole extent of the other dimensions**, halo cells included;
* received planes land in index ``0`` (from the low neighbour) and ``size-1`` (from the high
  neighbour); fields with ``ol < 2`` in a dimension are skipped in that dimension;
* sides without a neighbour are left untouched (Dirichlet: the global boundary keeps its
  value).

Transports (``IGG_B200_TRANSPORT`` or the ``transport=`` argument; the analogue of the
reference's ``IGG_ROCMAWARE_MPI`` switch, ``scripts/setenv.sh:11-18``):

========  ==============================================================================
``peer``  product path: one kernel stores the planes straight into the neighbours' halo
          cells over NVLink peer memory and signals with flags (``parallel/fabric.py``)
``nccl``  baseline: pack kernel -> grouped NCCL send/recv -> unpack kernel
``host``  device buffers staged through pinned host memory and sent with gloo (the
          ``IGG_ROCMAWARE_MPI=0`` analogue, reference ``README.md:25-35``)
``gloo``  CPU tensors (plumbing / tests)
========  ==============================================================================

Upstream ImplicitGlobalGrid also has per-dimension switches (``IGG_ROCMAWARE_MPI_DIMX/_DIMY/_DIMZ``, SURVEY.md
G9).  Their analogue: ``IGG_B200_TRANSPORT_DIMX`` / ``_DIMY`` / ``_DIMZ`` = ``nccl`` or ``host`` choose, per
dimension, between device-direct messages and pinned-host staging inside the pack/exchange/unpack engine
(transports ``nccl`` and ``host``; the one-launch ``peer`` engine has no per-dimension stages to switch).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .grid import GlobalGrid, global_grid
from .topology import NDIMS, PROC_NULL

__all__ = ["update_halo", "halo_planes", "halo_boxes", "HaloSpec", "default_transport", "ExchangeHalo",
           "per_dim_staging"]


@dataclass(frozen=True)
class HaloSpec:
    """Send/receive plane indices of one field along one dimension (0-based)."""

    dim: int
    send: Tuple[int, int]   # (to low neighbour, to high neighbour)
    recv: Tuple[int, int]   # (from low neighbour, from high neighbour)


def halo_planes(nxyz: Sequence[int], overlaps: Sequence[int], shape: Sequence[int]) -> List[Optional[HaloSpec]]:
    """Per-dimension :class:`HaloSpec` of a field of ``shape`` (``None`` = skip that dim)."""
    specs: List[Optional[HaloSpec]] = []
    for dim in range(NDIMS):
        size = shape[dim] if dim < len(shape) else 1
        o = overlaps[dim] + (size - nxyz[dim])
        if size < 2 or o < 2 or nxyz[dim] == 1:
            specs.append(None)
            continue
        specs.append(HaloSpec(dim=dim, send=(o - 1, size - o), recv=(0, size - 1)))
    return specs


def halo_boxes(nxyz: Sequence[int], overlaps: Sequence[int], shape: Sequence[int], has_neighbor,
               offsets: Sequence[Sequence[int]]):
    """The dimension-sequential exchange written as independent box copies, one per neighbour offset.

    Exchanging x, then y, then z with full-extent planes delivers to every halo cell the *original*
    value of exactly one remote cell: shift by -1/+1 in each dimension where the cell is a halo cell
    with a neighbour.  Such source cells are never themselves received, so all boxes can be copied
    concurrently.  For offset ``o``: in dimensions with ``o[d] != 0`` the box is the single send plane
    (landing in the neighbour's receive plane); in dimensions with ``o[d] == 0`` it spans every index
    that is not a receive plane with a neighbour.  ``has_neighbor(side, dim) -> bool``.

    Yields ``(offset, src_start, dst_start, extent)``; this is what the one-launch peer-memory
    ``update_halo`` executes (``csrc/fabric.cu``), and ``tests/test_halo_virtual.py`` checks it against
    the sequential definition.
    """
    specs = halo_planes(nxyz, overlaps, shape)
    shp = list(shape) + [1] * (NDIMS - len(shape))
    for o in offsets:
        if all(x == 0 for x in o):
            continue
        src, dst, ext = [0, 0, 0], [0, 0, 0], [1, 1, 1]
        ok = True
        for d in range(NDIMS):
            sp = specs[d]
            if o[d] != 0:
                side = 0 if o[d] < 0 else 1
                if sp is None or not has_neighbor(side, d):
                    ok = False
                    break
                src[d], dst[d], ext[d] = sp.send[side], sp.recv[1 - side], 1
            else:
                lo, hi = 0, shp[d]
                if sp is not None:
                    if has_neighbor(0, d):
                        lo = 1
                    if has_neighbor(1, d):
                        hi = shp[d] - 1
                src[d] = dst[d] = lo
                ext[d] = hi - lo
        if ok and all(e > 0 for e in ext):
            yield tuple(o), tuple(src), tuple(dst), tuple(ext)


def default_transport(field: torch.Tensor) -> str:
    env = os.environ.get("IGG_B200_TRANSPORT", "").strip().lower()
    if not field.is_cuda:
        return "gloo"
    if env:
        if env not in ("peer", "nccl", "host"):
            raise ValueError(f"IGG_B200_TRANSPORT must be peer|nccl|host, got {env!r}")
        return env
    from .comm import spans_nodes
    return "nccl" if spans_nodes() else "peer"      # CUDA IPC peer mappings end at the node boundary


def per_dim_staging(transport: str) -> Tuple[bool, bool, bool]:
    """Per dimension: are messages staged through pinned host memory?  (``transport`` is ``nccl`` or ``host``;
    ``IGG_B200_TRANSPORT_DIMX/Y/Z`` override it for single dimensions.)"""
    out = []
    for d in "XYZ":
        env = os.environ.get(f"IGG_B200_TRANSPORT_DIM{d}", "").strip().lower()
        if env and env not in ("nccl", "host"):
            raise ValueError(f"IGG_B200_TRANSPORT_DIM{d} must be nccl|host, got {env!r}")
        out.append((env or transport) == "host")
    return tuple(out)


def _check_fields(g: GlobalGrid, fields: Sequence[torch.Tensor]) -> None:
    if not fields:
        raise ValueError("update_halo needs at least one field")
    seen = set()
    for A in fields:
        if not torch.is_tensor(A):
            raise TypeError("update_halo fields must be torch tensors")
        if A.dim() > NDIMS or A.dim() < 1:
            raise ValueError("fields must have 1 to 3 dimensions")
        if not A.is_contiguous():
            raise ValueError("fields must be contiguous (row-major)")
        if A.data_ptr() in seen:
            raise ValueError("the same field was passed twice to update_halo")
        seen.add(A.data_ptr())
        for dim in range(A.dim()):
            if g.nxyz[dim] > 1 and abs(A.shape[dim] - g.nxyz[dim]) > 1:
                raise ValueError(f"field size {tuple(A.shape)} is incompatible with the local grid "
                                 f"{g.nxyz}: staggered arrays may differ by at most 1")


class ExchangeHalo:
    """Pack -> point-to-point exchange -> unpack engine (transports ``gloo``/``nccl``/``host``)."""

    def __init__(self, grid: GlobalGrid, transport: str) -> None:
        self.g = grid
        self.transport = transport
        self._bufs: Dict[tuple, torch.Tensor] = {}
        self.launches = 0  # native kernels launched (pack/unpack), for accounting
        self.host_dims = per_dim_staging(transport) if transport in ("nccl", "host") else (False,) * NDIMS

    # -- buffers -----------------------------------------------------------------------
    def _buf(self, key: tuple, numel: int, like: torch.Tensor, pinned_host: bool = False) -> torch.Tensor:
        buf = self._bufs.get(key)
        if buf is None or buf.numel() != numel or buf.dtype != like.dtype:
            if pinned_host:
                buf = torch.empty(numel, dtype=like.dtype,
                                  pin_memory=torch.cuda.is_available())
            else:
                buf = torch.empty(numel, dtype=like.dtype, device=like.device)
            self._bufs[key] = buf
        return buf

    # -- pack / unpack -----------------------------------------------------------------
    def _pack(self, A: torch.Tensor, dim: int, idx: int, out: torch.Tensor) -> None:
        if A.is_cuda and A.dtype == torch.float64:
            from ..ops import native
            if native.available():
                native.pack_plane(A, dim, idx, out)
                self.launches += 1
                return
        out.view(A.select(dim, idx).shape).copy_(A.select(dim, idx))

    def _unpack(self, A: torch.Tensor, dim: int, idx: int, src: torch.Tensor) -> None:
        if A.is_cuda and A.dtype == torch.float64:
            from ..ops import native
            if native.available():
                native.unpack_plane(A, dim, idx, src)
                self.launches += 1
                return
        A.select(dim, idx).copy_(src.view(A.select(dim, idx).shape))

    # -- the exchange ------------------------------------------------------------------
    def update(self, fields: Sequence[torch.Tensor]) -> None:
        g = self.g
        specs = [halo_planes(g.nxyz, g.overlaps, A.shape) for A in fields]
        for dim in range(NDIMS):
            host = self.host_dims[dim]
            sends, recvs, unpacks = [], [], []
            for f, A in enumerate(fields):
                spec = specs[f][dim]
                if spec is None:
                    continue
                plane_numel = A.numel() // A.shape[dim]
                for side in (0, 1):
                    nb = g.neighbors[side][dim]
                    if nb == PROC_NULL:
                        continue
                    # message sent towards `side` arrives on the neighbour's opposite side
                    sbuf = self._buf(("s", f, dim, side), plane_numel, A)
                    self._pack(A, dim, spec.send[side], sbuf)
                    rbuf = self._buf(("r", f, dim, side), plane_numel, A)
                    if host:
                        hs = self._buf(("hs", f, dim, side), plane_numel, A, pinned_host=True)
                        hr = self._buf(("hr", f, dim, side), plane_numel, A, pinned_host=True)
                        hs.copy_(sbuf, non_blocking=True)
                        sends.append((hs, nb, 2 * f + (1 - side)))
                        recvs.append((hr, nb, 2 * f + side))
                        unpacks.append((A, spec.recv[side], rbuf, hr))
                    else:
                        sends.append((sbuf, nb, 2 * f + (1 - side)))
                        recvs.append((rbuf, nb, 2 * f + side))
                        unpacks.append((A, spec.recv[side], rbuf, None))
            if not sends and not recvs:
                continue
            if host and torch.cuda.is_available():
                torch.cuda.current_stream().synchronize()
            g.comm.exchange(sends, recvs)
            for (A, idx, rbuf, hr) in unpacks:
                if hr is not None:
                    rbuf.copy_(hr, non_blocking=True)
                self._unpack(A, dim, idx, rbuf)

    def close(self) -> None:
        self._bufs.clear()


def _engine(g: GlobalGrid, transport: str):
    key = ("halo", transport)
    eng = g.cache.get(key)
    if eng is None:
        if transport == "peer":
            from .fabric import PeerHalo
            eng = PeerHalo(g)
        else:
            eng = ExchangeHalo(g, transport)
        g.cache[key] = eng
    return eng


def update_halo(*fields: torch.Tensor, transport: Optional[str] = None) -> None:
    """Update the halo cells of one or more fields (variadic like upstream ``update_halo!``)."""
    g = global_grid()
    _check_fields(g, fields)
    if transport is None:
        transport = default_transport(fields[0])
    if fields[0].is_cuda and transport == "gloo":
        raise ValueError("transport 'gloo' is for CPU tensors; use peer|nccl|host on CUDA")
    if not fields[0].is_cuda and transport != "gloo":
        transport = "gloo"
    _engine(g, transport).update(fields)
