"""2-D heat diffusion, performance variant (``perf``): one fused double-buffered kernel per step.

Behavioural counterpart of reference ``scripts/diffusion_2D_perf.jl``: fused ``diffusion_step!`` (``:3-13``),
pointer swap (``:50``), ``update_halo!(T)`` (``:51``), ``tic()`` after the warm-up iterations (``:48``), ``toc()``
and the ``T_eff`` report (``:53-58``), optional gather + heat map (``:59-63``).

``--fused true`` (default): the halo exchange happens inside the stencil kernel (peer stores + epoch
flags), so a step is ONE launch.  ``--fused false``: the reference's serial structure -- kernel, swap,
``update_halo(T)`` over ``--transport`` (``nccl`` is the same-box baseline of BASELINE.md).

Run:  ``torchrun --nproc-per-node 8 -m rocm_mpi_b200.models.diffusion2d_perf --nx 16384 --ny 16384 --nt 110``
"""
from __future__ import annotations

import argparse
import gc

import torch

from ..ops import native
from ..parallel import a_eff, finalize_global_grid, global_grid, tic, toc
from ..utils.config import DiffusionConfig, add_config_args, config_from_args
from ..utils.profiling import nvtx_range
from .common import detach_local, RunResult, log0, setup_2d, visualise_2d
from .solver import Diffusion2D
from .solver_host import Diffusion2DHost

__all__ = ["diffusion2D_perf", "run_timed", "make_solver", "main"]


def make_solver(cfg: DiffusionConfig, mode: str):
    """The GPU stepper, or its CPU stand-in when the grid lives on the host (``--device-type cpu``): there a
    ``fused`` request degrades to the serial structure, ``hide`` keeps its frame / exchange / interior order."""
    if global_grid().device_type != "cuda":
        return Diffusion2DHost(cfg.nx, cfg.ny, lx=cfg.lx, ly=cfg.ly, lam=cfg.lam, Cp0=cfg.Cp0,
                               mode="hide" if mode == "hide" else "plain", b_width=cfg.b_width[:2])
    from ..parallel.comm import spans_nodes
    if mode == "hide" and (cfg.transport in ("nccl", "host") or spans_nodes()):
        mode = "hide_exchange"      # overlap with a library exchange instead of in-kernel peer stores
    if spans_nodes() and mode not in ("plain", "hide_exchange"):
        log0(f"ranks span several nodes: mode '{mode}' needs peer memory, falling back to plain + NCCL")
        mode = "plain"
    return Diffusion2D(cfg.nx, cfg.ny, lx=cfg.lx, ly=cfg.ly, lam=cfg.lam, Cp0=cfg.Cp0, mode=mode, variant=cfg.kernel,
                       b_width=cfg.b_width[:2], transport=cfg.transport)


def _steps(s, n: int, tracer, per_step: bool) -> None:
    if tracer is None or not per_step:
        s.step(n)
        return
    for _ in range(n):                 # one span per step (slower than the native loop: tracing only)
        with tracer.span(f"step {s.steps_done}", "steps"):
            s.step(1)


def run_timed(s, nt: int, warmup: int, tracer=None, trace_steps: bool = False) -> float:
    """The reference's time loop: ``tic()`` once ``warmup`` iterations are done, ``toc()`` at the end."""
    import contextlib
    span = tracer.span if tracer is not None else (lambda *a, **k: contextlib.nullcontext())
    s.begin()
    gc_was = gc.isenabled()
    gc.disable()                       # reference scripts/diffusion_2D_perf_hide.jl:73 (GC.enable(false))
    try:
        with nvtx_range("warmup"), span("warmup"):
            _steps(s, warmup, tracer, trace_steps)
        tic()
        with nvtx_range("timed_loop"), span("timed_loop"):
            _steps(s, nt - warmup, tracer, trace_steps)
        wtime = toc()
    finally:
        if gc_was:
            gc.enable()
    s.finish()
    return wtime


def _run_chunked(s, nt: int, warmup: int, every: int, directory: str, tracer=None, trace_steps: bool = False):
    """Time loop cut at checkpoint steps: chunks end where ``steps_done`` is a multiple of ``every``; only the
    first chunk has untimed warm-up steps.  Returns ``(timed seconds, untimed steps)``."""
    wtime, untimed, first = 0.0, 0, True
    while s.steps_done < nt:
        n = nt - s.steps_done
        if every > 0:
            n = min(n, every - s.steps_done % every)
        w = min(warmup, max(n - 1, 0)) if first else 0
        wtime += run_timed(s, n, w, tracer, trace_steps)
        untimed += w
        first = False
        if every > 0 and s.steps_done % every == 0:
            s.save(directory)
    return wtime, untimed


def diffusion2D_perf(cfg: DiffusionConfig = None, mode: str = None, variant_name: str = "perf",
                     finalize: bool = True) -> RunResult:
    cfg = cfg or DiffusionConfig(nx=16384, ny=16384, nt=110, do_vis=False)
    su = setup_2d(cfg)
    if mode is None:
        mode = "fused" if cfg.fused else "plain"
        if (mode == "fused" and int(cfg.steps_per_pass) == 2 and cfg.kernel == "tma" and cfg.ny % 2 == 0
                and global_grid().device_type == "cuda"):
            from ..parallel.comm import spans_nodes
            if not spans_nodes():
                mode = "temporal"      # same field, half the trips through memory
    n0 = native.launch_count()
    s = make_solver(cfg, mode)
    nt, warm = int(cfg.nt), min(int(cfg.warmup), max(int(cfg.nt) - 1, 0))
    if cfg.resume:
        from ..utils.checkpoint import latest_checkpoint
        path = latest_checkpoint(cfg.ckpt_dir)
        if path is not None:
            step = s.load(path)                       # collective
            log0(f"Resuming from {path} (step {step})")
    start_step = s.steps_done
    tracer = None
    if cfg.trace:
        from ..utils.profiling import Tracer
        tracer = Tracer(rank=su.me)
    log0("Starting the time loop ...", end="")
    if cfg.ckpt_every > 0 or s.steps_done > 0:
        wtime, warm = _run_chunked(s, nt, warm, int(cfg.ckpt_every), cfg.ckpt_dir, tracer, cfg.trace_steps)
    else:
        wtime = run_timed(s, nt, warm, tracer, cfg.trace_steps)
    log0("done")
    if tracer is not None:
        tracer.dump(f"{cfg.trace}.rank{su.me}.json")
    steps_here = nt - start_step                     # fewer than nt after a resume
    A = a_eff((cfg.nx, cfg.ny))                      # reference scripts/diffusion_2D_perf.jl:55
    wtime_it = wtime / max(steps_here - warm, 1)     # :56
    T_eff = A / wtime_it                             # :57
    log0("Executed %d steps in = %1.3e sec (@ T_eff = %1.2f GB/s) " % (nt, wtime, float("%.3g" % T_eff)))
    res = RunResult(variant_name, su.me, su.nprocs, su.dims, nt, wtime=wtime, t_eff=T_eff, T_local=s.T,
                    launches=native.launch_count() - n0)
    if cfg.do_vis:
        visualise_2d(su, s.T, variant_name, res)
        log0("maximum(T_v) = %r" % res.max_T)        # reference scripts/diffusion_2D_perf_hide.jl:115
    if finalize:
        detach_local(res)
        finalize_global_grid()
    return res


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    add_config_args(p, DiffusionConfig(nx=16384, ny=16384, nt=110, do_vis=False))
    diffusion2D_perf(config_from_args(p.parse_args(argv)))


if __name__ == "__main__":
    main()
