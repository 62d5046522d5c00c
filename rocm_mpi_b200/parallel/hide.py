"""``hide_communication``: overlap the halo exchange with the interior computation.

The reference emulates ParallelStencil's ``@hide_communication b_width begin ... end`` by hand
(``scripts/diffusion_2D_perf_hide.jl:15-29, 63-71, 84-101``): (1) compute the boundary frame of width
``b_width`` on a high-priority queue, (2) compute the interior on a low-priority queue, (3) as soon as
(1) is done exchange the halo of the *new* field while (2) is still running, (4) join, swap.

This module offers that structure for **any** user step written against regions:

    def step(region):            # region = tuple of slices into the local arrays
        T2[region] = ...         # anything that only WRITES inside `region`

    hide_communication((16, 4), step, T2)          # boundary, update_halo(T2) || interior

For the built-in diffusion kernels the same structure is available natively and faster
(``Diffusion2D(mode="hide")`` -- the frame kernel pushes its edges itself, no separate exchange).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch

from .grid import global_grid
from .halo import update_halo

__all__ = ["hide_communication", "frame_regions", "interior_region"]

Region = Tuple[slice, ...]


def interior_region(shape: Sequence[int], b_width: Sequence[int]) -> Optional[Region]:
    """Cells strictly inside the boundary frame (``None`` if the frame covers everything)."""
    sl = tuple(slice(1 + b_width[d], shape[d] - 1 - b_width[d]) for d in range(len(shape)))
    if any(s.stop <= s.start for s in sl):
        return None
    return sl


def frame_regions(shape: Sequence[int], b_width: Sequence[int]) -> List[Region]:
    """Disjoint boxes that tile the boundary frame of width ``b_width`` inside the interior ``[1, n-2]``.

    The frame contains every cell that is sent to a neighbour (index 1 and n-2), which is the
    constraint ParallelStencil states for ``b_width``.
    """
    nd = len(shape)
    if any(b < 1 for b in b_width[:nd]):
        raise ValueError("b_width entries must be >= 1")
    inner = interior_region(shape, b_width)
    full = [slice(1, shape[d] - 1) for d in range(nd)]
    if inner is None:
        return [tuple(full)]
    regions: List[Region] = []
    cur = list(full)
    for d in range(nd):
        lo = list(cur)
        lo[d] = slice(cur[d].start, inner[d].start)
        hi = list(cur)
        hi[d] = slice(inner[d].stop, cur[d].stop)
        regions += [tuple(lo), tuple(hi)]
        cur[d] = inner[d]
    return [r for r in regions if all(s.stop > s.start for s in r)]


_streams = {}


def _get_streams(device: torch.device):
    key = device.index
    if key not in _streams:
        _streams[key] = (torch.cuda.Stream(device=device, priority=-1), torch.cuda.Stream(device=device, priority=0))
    return _streams[key]


def hide_communication(b_width: Sequence[int], step: Callable[[Region], None], *halo_fields: torch.Tensor,
                       transport: Optional[str] = None) -> None:
    """Run ``step`` on the boundary frame, then exchange ``halo_fields`` while ``step`` runs on the interior."""
    if not halo_fields:
        raise ValueError("hide_communication needs the field(s) whose halo is to be updated")
    g = global_grid()
    shape = tuple(halo_fields[0].shape)
    frame = frame_regions(shape, b_width)
    inner = interior_region(shape, b_width)
    if halo_fields[0].is_cuda:
        main = torch.cuda.current_stream()
        s_hi, s_lo = _get_streams(halo_fields[0].device)
        ev = torch.cuda.Event()
        ev.record(main)
        s_hi.wait_event(ev)
        s_lo.wait_event(ev)
        with torch.cuda.stream(s_hi):                 # (1) boundary first, high priority ...
            for r in frame:
                step(r)
            if g.nprocs > 1 or any(g.periods):
                update_halo(*halo_fields, transport=transport)      # (3) ... then its exchange
        with torch.cuda.stream(s_lo):                 # (2) interior, concurrently
            if inner is not None:
                step(inner)
        main.wait_stream(s_hi)                        # (4) join
        main.wait_stream(s_lo)
    else:
        for r in frame:
            step(r)
        if inner is not None:
            step(inner)
        if g.nprocs > 1 or any(g.periods):
            update_halo(*halo_fields)
