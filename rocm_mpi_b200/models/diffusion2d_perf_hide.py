"""2-D heat diffusion with communication hiding (``perf_hide``).

Behavioural counterpart of reference ``scripts/diffusion_2D_perf_hide.jl``.  The reference splits the
kernel with ``(b_width, istep)`` predicates (``:15-29``), creates a HIGH- and a LOW-priority queue
(``:63-68``) and launches both parts per step (``:84-92``); its real overlap variant -- wait for the
boundary, ``update_halo!(T2)`` while the interior runs, join, swap -- is commented "not ready yet"
(``:94-101``) and the live variant neither exchanges halos nor covers the whole domain.
This application delivers the *intended* semantics: the boundary frame of width ``b_width`` is computed
first on a high-priority stream and pushes its edge cells to the neighbours as it goes; the interior
runs concurrently on a low-priority stream; both join before the swap.

``--fused true`` selects the single-launch form instead (frame tiles first inside one kernel).

Run:  ``torchrun --nproc-per-node 8 -m rocm_mpi_b200.models.diffusion2d_perf_hide --b-width 16 4``
"""
from __future__ import annotations

import argparse

from ..utils.config import DiffusionConfig, add_config_args, config_from_args
from .common import RunResult
from .diffusion2d_perf import diffusion2D_perf

__all__ = ["diffusion2D_perf_hide", "main"]

DEFAULTS = dict(nx=16384, ny=16384, nt=100, do_vis=False, b_width=(16, 4, 4), fused=False)


def diffusion2D_perf_hide(cfg: DiffusionConfig = None, finalize: bool = True) -> RunResult:
    cfg = cfg or DiffusionConfig(**DEFAULTS)
    return diffusion2D_perf(cfg, mode="fused" if cfg.fused else "hide", variant_name="hide", finalize=finalize)


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    add_config_args(p, DiffusionConfig(**DEFAULTS))
    diffusion2D_perf_hide(config_from_args(p.parse_args(argv)))


if __name__ == "__main__":
    main()
