"""Profiling hooks.

The reference's only profiling is Julia's host-side sampling profiler around the time loop:
``Profile.@profile`` (``scripts/diffusion_2D_perf_hide_prof.jl:34``), a warm-up call followed by
``Profile.clear()`` (``:110-112``) and a ``Profile.print(...; maxdepth=30)`` dump to ``prof.txt`` on rank 0
(``:119-121``), with the GC switched off around the loop (``scripts/diffusion_2D_perf_hide.jl:73,106``).

Here: NVTX ranges (visible in Nsight Systems / Compute), a host profile with ``cProfile`` dumped to
``prof.txt`` in the same spirit, and ``scripts/profile_kernel.sh`` for the per-kernel ``ncu`` capture that
is the real performance evidence on a GPU.
"""
from __future__ import annotations

import contextlib
import cProfile
import io
import pstats

import torch

__all__ = ["nvtx_range", "HostProfile", "cuda_profiler_range"]


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


@contextlib.contextmanager
def cuda_profiler_range():
    """``cudaProfilerStart/Stop`` bracket, for ``ncu --profile-from-start off`` / ``nsys -c cudaProfilerApi``."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    try:
        yield
    finally:
        if on:
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()


class HostProfile:
    """``with HostProfile() as hp: ...; hp.dump("prof.txt")`` -- the ``Profile.@profile`` analogue."""

    def __init__(self) -> None:
        self._p = cProfile.Profile()

    def __enter__(self) -> "HostProfile":
        self._p.enable()
        return self

    def __exit__(self, *exc) -> None:
        self._p.disable()

    def clear(self) -> None:
        """Discard what has been collected so far (``Profile.clear()``)."""
        self._p = cProfile.Profile()

    def text(self, limit: int = 30) -> str:
        s = io.StringIO()
        pstats.Stats(self._p, stream=s).strip_dirs().sort_stats("cumulative").print_stats(limit)
        return s.getvalue()

    def dump(self, path: str, limit: int = 30) -> str:
        with open(path, "w") as f:
            f.write(self.text(limit))
        return path
