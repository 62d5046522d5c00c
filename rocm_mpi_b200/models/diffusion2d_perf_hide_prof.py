"""``perf_hide`` under a profiler (``perf_hide_prof``).

Behavioural counterpart of reference ``scripts/diffusion_2D_perf_hide_prof.jl``: the time loop lives in a
function (``compute_step``, ``:32-63``) that is called once as warm-up (``:110``), the profile is cleared
(``:112``), the measured call follows (``:114``) and rank 0 dumps the profile to ``prof.txt`` (``:119-121``).
Defaults follow that script: ``8*1024`` cells per side, ``b_width = (32, 8)``, ``nt = 300`` (``:71-77``).

Host time is collected with ``cProfile`` (what the Julia sampling profiler shows: launch overhead, waits);
the device side is annotated with NVTX ranges and bracketed by ``cudaProfilerStart/Stop`` so that
``ncu --profile-from-start off`` / ``nsys -c cudaProfilerApi`` capture exactly the measured call.
"""
from __future__ import annotations

import argparse
import os

from ..ops import native
from ..parallel import a_eff, finalize_global_grid, global_grid
from ..utils.config import DiffusionConfig, add_config_args, config_from_args
from ..utils.profiling import HostProfile, cuda_profiler_range
from .common import detach_local, RunResult, log0, setup_2d
from .diffusion2d_perf import make_solver, run_timed

__all__ = ["diffusion2D_perf_hide_prof", "main"]

DEFAULTS = dict(nx=8192, ny=8192, nt=300, do_vis=False, b_width=(32, 8, 8), fused=False)


def diffusion2D_perf_hide_prof(cfg: DiffusionConfig = None, prof_path: str = "prof.txt",
                               finalize: bool = True) -> RunResult:
    cfg = cfg or DiffusionConfig(**DEFAULTS)
    su = setup_2d(cfg)
    g = global_grid()
    n0 = native.launch_count()
    s = make_solver(cfg, "fused" if cfg.fused else "hide")
    nt, warm = int(cfg.nt), min(int(cfg.warmup), max(int(cfg.nt) - 1, 0))
    hp = HostProfile()
    with hp:
        run_timed(s, min(12, nt), min(warm, 2))          # warm-up call (reference :110, nt = 12)
    hp.clear()                                           # reference :112
    log0("Starting the time loop ...", end="")
    with cuda_profiler_range(), hp:
        wtime = run_timed(s, nt, warm)                   # measured call (reference :114)
    log0("done")
    T_eff = a_eff((cfg.nx, cfg.ny)) / (wtime / max(nt - warm, 1))
    log0("Executed %d steps in = %1.3e sec (@ T_eff = %1.2f GB/s) " % (nt, wtime, float("%.3g" % T_eff)))
    res = RunResult("hide_prof", su.me, su.nprocs, su.dims, nt, wtime=wtime, t_eff=T_eff, T_local=s.T,
                    launches=native.launch_count() - n0)
    if g.me == 0:
        os.makedirs(os.path.dirname(os.path.abspath(prof_path)), exist_ok=True)
        hp.dump(prof_path)                               # reference :119-121
    if finalize:
        detach_local(res)
        finalize_global_grid()
    return res


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    add_config_args(p, DiffusionConfig(**DEFAULTS))
    p.add_argument("--prof-path", default="prof.txt")
    ns = p.parse_args(argv)
    diffusion2D_perf_hide_prof(config_from_args(ns), ns.prof_path)


if __name__ == "__main__":
    main()
