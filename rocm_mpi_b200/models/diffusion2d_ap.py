"""2-D heat diffusion, array-programming variant (``ap``).

Behavioural counterpart of reference ``scripts/diffusion_2D_ap.jl``: per step four broadcast
assignments (``qx``, ``qy``, ``dTdt``, interior update -- ``:38-41``) followed by
``update_halo!(T)`` (``:42``), then gather + heat map (``:45-47``).  The broadcasts are plain
``torch`` expressions, so the same code runs on CPU tensors over gloo (BASELINE config 1) and
on CUDA tensors with any of the device transports.

Run:  ``torchrun --nproc-per-node 4 -m rocm_mpi_b200.models.diffusion2d_ap``
"""
from __future__ import annotations

import argparse

import torch

from ..ops import reference as ref
from ..parallel import finalize_global_grid, update_halo
from ..utils.config import DiffusionConfig, add_config_args, config_from_args
from .common import RunResult, initial_condition_2d, log0, setup_2d, visualise_2d

__all__ = ["diffusion2D_ap", "main"]


def diffusion2D_ap(cfg: DiffusionConfig = None, finalize: bool = True) -> RunResult:
    cfg = cfg or DiffusionConfig()
    s = setup_2d(cfg)
    dev, f64 = s.device, torch.float64
    nx, ny = cfg.nx, cfg.ny
    # Array initializations (reference scripts/diffusion_2D_ap.jl:22-28)
    dTdt = torch.zeros(nx - 2, ny - 2, dtype=f64, device=dev)
    qx = torch.zeros(nx - 1, ny - 2, dtype=f64, device=dev)
    qy = torch.zeros(nx - 2, ny - 1, dtype=f64, device=dev)
    Cp = cfg.Cp0 * torch.ones(nx, ny, dtype=f64, device=dev)
    T = initial_condition_2d(s)
    log0("Starting the time loop ...", end="")
    for _ in range(int(cfg.nt)):
        ref.ap_step(T, Cp, cfg.lam, s.dt, s.dx, s.dy, qx, qy, dTdt)
        update_halo(T, transport=cfg.transport)
    log0("done")
    res = RunResult("ap", s.me, s.nprocs, s.dims, int(cfg.nt), T_local=T)
    if cfg.do_vis:
        visualise_2d(s, T, "ap", res)
    if finalize:
        finalize_global_grid()
    return res


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    add_config_args(p)
    diffusion2D_ap(config_from_args(p.parse_args(argv)))


if __name__ == "__main__":
    main()
