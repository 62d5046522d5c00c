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

__all__ = ["nvtx_range", "HostProfile", "cuda_profiler_range", "Tracer"]


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


class Tracer:
    """Timeline of named spans, written as a Chrome / Perfetto trace (``chrome://tracing``, ``ui.perfetto.dev``).

    Nsight Systems is not part of every image this runs in; this records the spans that matter for a time loop --
    warm-up, the timed loop, optionally every step -- with CUDA events on the current stream (device time) or
    ``perf_counter`` on the CPU, one process track per rank.  The reference has no timeline at all (its only
    profile is the host sampling dump of ``scripts/diffusion_2D_perf_hide_prof.jl:119-121``)."""

    def __init__(self, enabled: bool = True, rank: int = 0) -> None:
        self.enabled, self.rank = enabled, rank
        self._cuda = torch.cuda.is_available()
        self._spans = []           # (name, start, end, track) -- CUDA events or floats
        self._t0 = None

    def _now(self):
        if self._cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            return e
        import time
        return time.perf_counter()

    @contextlib.contextmanager
    def span(self, name: str, track: str = "main"):
        if not self.enabled:
            yield
            return
        if self._t0 is None:
            self._t0 = self._now()
        a = self._now()
        try:
            yield
        finally:
            self._spans.append((name, a, self._now(), track))

    def events(self):
        """``[(name, start_us, duration_us, track)]`` relative to the first span."""
        if self._cuda:
            torch.cuda.synchronize()
            rel = lambda e: self._t0.elapsed_time(e) * 1e3
        else:
            rel = lambda t: (t - self._t0) * 1e6
        return [(n, rel(a), max(rel(b) - rel(a), 0.0), tr) for (n, a, b, tr) in self._spans]

    def dump(self, path: str) -> str:
        import json
        import os
        tracks = {}
        ev = []
        for (n, ts, dur, tr) in self.events():
            tid = tracks.setdefault(tr, len(tracks))
            ev.append({"name": n, "ph": "X", "ts": ts, "dur": dur, "pid": self.rank, "tid": tid})
        meta = [{"name": "process_name", "ph": "M", "pid": self.rank, "args": {"name": f"rank {self.rank}"}}]
        meta += [{"name": "thread_name", "ph": "M", "pid": self.rank, "tid": tid, "args": {"name": tr}}
                 for tr, tid in tracks.items()]
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with open(path, "w") as f:
            json.dump({"traceEvents": meta + ev, "displayTimeUnit": "ms"}, f)
        return path
