"""``tic`` / ``toc`` and the ``T_eff`` metric.

The reference brackets its time loop with IGG's ``tic()`` (at iteration 11, i.e. after ten
warm-up steps) and ``toc()`` (``scripts/diffusion_2D_perf.jl:48,53``): barrier + host wall
clock.  Here the clock is the **device**: a CUDA event pair on the current stream, and
``toc`` returns the **maximum over ranks** (all multi-GPU numbers are max-over-ranks device
times).  On CPU it falls back to ``time.perf_counter`` with the same barrier semantics.

``t_eff`` restates reference ``scripts/diffusion_2D_perf.jl:55-57``:
``A_eff = (2+1)/1e9*nx*ny*sizeof(Float64)``, ``t_it = wtime/(nt-10)``, ``T_eff = A_eff/t_it``.
"""
from __future__ import annotations

import time
from typing import Optional

import torch

from .grid import global_grid, grid_is_initialized

__all__ = ["tic", "toc", "t_eff", "a_eff", "DeviceTimer"]


class DeviceTimer:
    def __init__(self) -> None:
        self._t0: Optional[float] = None
        self._e0 = None

    def _cuda(self) -> bool:
        return grid_is_initialized() and global_grid().device_type == "cuda"

    def start(self) -> None:
        if grid_is_initialized():
            global_grid().comm.barrier()
        if self._cuda():
            torch.cuda.synchronize()
            self._e0 = torch.cuda.Event(enable_timing=True)
            self._e0.record()
        self._t0 = time.perf_counter()

    def stop(self) -> float:
        if self._t0 is None:
            raise RuntimeError("toc() called before tic()")
        if self._cuda():
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            e1.synchronize()
            dt = self._e0.elapsed_time(e1) * 1e-3
        else:
            dt = time.perf_counter() - self._t0
        if grid_is_initialized():
            g = global_grid()
            fab = g.cache.get("fabric")
            if fab is not None:
                fab.check_all_errors()      # a timed-out flag wait must not go unnoticed past a timing point
            dt = g.comm.allreduce_max(dt)
        return dt


_TIMER = DeviceTimer()


def tic() -> None:
    """Barrier, then start the clock (device event on CUDA)."""
    _TIMER.start()


def toc() -> float:
    """Seconds since ``tic()``, measured on the device, max over ranks."""
    return _TIMER.stop()


def a_eff(shape, itemsize: int = 8, n_rw: int = 2, n_ro: int = 1) -> float:
    """Effective memory access per iteration in GB (T read + T2 write + Cp read)."""
    n = 1
    for s in shape:
        n *= int(s)
    return (n_rw + n_ro) / 1e9 * n * itemsize


def t_eff(shape, wtime: float, nt_timed: int, itemsize: int = 8) -> float:
    """Effective memory throughput in GB/s for ``nt_timed`` timed iterations."""
    return a_eff(shape, itemsize) / (wtime / nt_timed)
