"""Single-process "virtual ranks" model of the global grid.

The reference has no fake backend (its only multi-rank testing is
``srun -n 4`` on a GPU server, ``README.md:16-18``).  This module holds every rank's local
array in one process and applies the ``update_halo`` / ``gather`` semantics of
``parallel/halo.py`` / ``parallel/gather.py`` with plain copies.  It is the executable
specification the distributed transports (gloo, NCCL, peer fabric) are tested against, and
the way a multi-rank layout is emulated when fewer GPUs than ranks are available.
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from .halo import halo_boxes, halo_planes
from .topology import NDIMS, PROC_NULL, CartTopology, cart_coords

__all__ = ["VirtualGrid"]


class VirtualGrid:
    def __init__(self, nxyz: Sequence[int], dims: Sequence[int], periods: Sequence[int] = (0, 0, 0),
                 overlaps: Sequence[int] = (2, 2, 2)) -> None:
        self.nxyz = tuple(int(n) for n in nxyz) + (1,) * (NDIMS - len(nxyz))
        self.dims = tuple(int(d) for d in dims) + (1,) * (NDIMS - len(dims))
        self.periods = tuple(int(p) for p in periods) + (0,) * (NDIMS - len(periods))
        self.overlaps = tuple(int(o) for o in overlaps)
        self.nprocs = self.dims[0] * self.dims[1] * self.dims[2]
        self.topos = [CartTopology.create(self.nprocs, r, self.dims, self.periods)
                      for r in range(self.nprocs)]
        self.nxyz_g = tuple(
            1 if self.nxyz[k] == 1 else
            self.dims[k] * (self.nxyz[k] - self.overlaps[k]) + self.overlaps[k] * (self.periods[k] == 0)
            for k in range(NDIMS))

    # -- layout ------------------------------------------------------------------------
    def origin(self, rank: int) -> List[int]:
        """Global index of local index 0 on ``rank`` (non-periodic numbering)."""
        c = cart_coords(rank, self.dims)
        return [c[k] * (self.nxyz[k] - self.overlaps[k]) for k in range(NDIMS)]

    def scatter(self, A_global: torch.Tensor) -> List[torch.Tensor]:
        """Cut the overlapping local arrays out of a global array (non-periodic grids)."""
        nd = A_global.dim()
        out = []
        for r in range(self.nprocs):
            o = self.origin(r)
            sl = tuple(slice(o[k], o[k] + self.nxyz[k]) for k in range(nd))
            out.append(A_global[sl].clone())
        return out

    # -- update_halo -------------------------------------------------------------------
    def update_halo(self, *fields: List[torch.Tensor]) -> None:
        """``fields`` are lists (one entry per rank) of local arrays; updated in place."""
        for dim in range(NDIMS):
            for locs in fields:
                spec = halo_planes(self.nxyz, self.overlaps, locs[0].shape)[dim]
                if spec is None:
                    continue
                # take all send planes first (a real exchange sends before it receives)
                planes = {}
                for r in range(self.nprocs):
                    for side in (0, 1):
                        nb = self.topos[r].neighbors[side][dim]
                        if nb != PROC_NULL:
                            planes[(nb, 1 - side)] = locs[r].select(dim, spec.send[side]).clone()
                for (r, side), p in planes.items():
                    locs[r].select(dim, spec.recv[side]).copy_(p)

    def update_halo_boxes(self, *fields: List[torch.Tensor]) -> None:
        """Same result as :meth:`update_halo`, computed the way the peer-memory kernel does it:
        all boxes read from the pre-exchange state, one box per neighbour offset."""
        import itertools
        for locs in fields:
            before = [L.clone() for L in locs]
            for r in range(self.nprocs):
                topo = self.topos[r]
                rng = [(-1, 0, 1) if self.nxyz[d] > 1 else (0,) for d in range(NDIMS)]
                offs = [o for o in itertools.product(*rng) if topo.rank_at_offset(o) != PROC_NULL]
                has_nb = lambda side, d, t=topo: t.neighbors[side][d] != PROC_NULL
                for (o, src, dst, ext) in halo_boxes(self.nxyz, self.overlaps, locs[r].shape, has_nb, offs):
                    nb = topo.rank_at_offset(o)
                    nd = locs[r].dim()
                    ssl = tuple(slice(src[d], src[d] + ext[d]) for d in range(nd))
                    dsl = tuple(slice(dst[d], dst[d] + ext[d]) for d in range(nd))
                    locs[nb][dsl] = before[r][ssl]

    # -- gather ------------------------------------------------------------------------
    def gather(self, blocks: List[torch.Tensor]) -> torch.Tensor:
        lshape = tuple(blocks[0].shape)
        nd = len(lshape)
        out = torch.zeros(tuple(lshape[d] * self.dims[d] for d in range(nd)), dtype=blocks[0].dtype)
        for r, blk in enumerate(blocks):
            c = cart_coords(r, self.dims)
            sl = tuple(slice(c[d] * lshape[d], (c[d] + 1) * lshape[d]) for d in range(nd))
            out[sl].copy_(blk)
        return out
