"""Pieces shared by every diffusion application: set-up, initial condition, output."""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from ..ops import reference as ref
from ..parallel import (coords_g, finalize_global_grid, gather, global_grid, init_global_grid, inner,
                        nx_g, ny_g, nz_g)
from ..parallel.gather import inner_to_host
from ..utils.config import DiffusionConfig
from ..utils.output import output_name, save_heatmap, save_raw

__all__ = ["Setup", "setup_2d", "initial_condition_2d", "visualise_2d", "RunResult", "log0", "detach_local"]


def log0(*a, **k) -> None:
    g = global_grid()
    if g.me == 0 and not g.quiet:
        print(*a, **k, flush=True)


@dataclass
class Setup:
    cfg: DiffusionConfig
    me: int
    dims: tuple
    nprocs: int
    coords: tuple
    device: torch.device
    dx: float
    dy: float
    dt: float


@dataclass
class RunResult:
    variant: str
    me: int
    nprocs: int
    dims: tuple
    nt: int
    wtime: Optional[float] = None      # seconds for nt - warmup iterations (device time, max over ranks)
    t_eff: Optional[float] = None      # GB/s per GPU, reference scripts/diffusion_2D_perf.jl:55-57
    T_v: Optional[torch.Tensor] = None  # gathered halo-free field (root only)
    T_local: Optional[torch.Tensor] = None
    png: Optional[str] = None
    raw: Optional[str] = None
    max_T: Optional[float] = None
    launches: int = 0


def detach_local(res: "RunResult", limit_bytes: int = 64 << 20) -> None:
    """Before ``finalize_global_grid()`` frees peer-addressable fields: keep a private copy of a small local
    field in the result, drop a large one (it only aliases fabric memory)."""
    T = res.T_local
    if T is None or not T.is_cuda:
        return
    res.T_local = T.clone() if T.numel() * T.element_size() <= limit_bytes else None


def setup_2d(cfg: DiffusionConfig) -> Setup:
    """``init_global_grid`` + derived numerics (reference ``scripts/diffusion_2D_ap.jl:17-20``)."""
    me, dims, nprocs, coords, _ = init_global_grid(
        cfg.nx, cfg.ny, 1, dimx=cfg.dimx, dimy=cfg.dimy, periodx=cfg.periodx, periody=cfg.periody,
        device_type=cfg.device_type, quiet=cfg.quiet)
    g = global_grid()
    if not cfg.quiet:
        # reference scripts/diffusion_2D_ap.jl:18
        print(f"Process {me} selecting device {g.device}", flush=True)
    dx, dy = cfg.lx / nx_g(), cfg.ly / ny_g()
    dt = ref.time_step(dx, dy, cfg.Cp0, cfg.lam)
    return Setup(cfg, me, dims, nprocs, coords, g.device, dx, dy, dt)


def initial_condition_2d(s: Setup, dtype=torch.float64, on_host: bool = False) -> torch.Tensor:
    """Gaussian IC from the global coordinates (reference ``scripts/diffusion_2D_ap.jl:28``).

    The reference builds it in a host comprehension and uploads; on CUDA this runs the native
    ``gaussian_ic`` kernel directly on the device unless ``on_host`` is set.
    """
    cfg = s.cfg
    g = global_grid()
    shape = (cfg.nx, cfg.ny)
    probe = torch.empty(shape, device="meta")
    if s.device.type == "cuda" and not on_host and dtype == torch.float64:
        from ..ops import native
        if native.available():
            T = torch.empty(shape, dtype=dtype, device=s.device)
            x0 = float(coords_g(0, s.dx, probe)[0])
            y0 = float(coords_g(1, s.dy, probe)[0])
            if not (g.periods[0] or g.periods[1]):
                native.gaussian_ic_2d(T, x0, y0, s.dx, s.dy, cfg.lx, cfg.ly)
                return T
    xs = coords_g(0, s.dx, probe)
    ys = coords_g(1, s.dy, probe)
    T = ref.gaussian_ic(xs, ys, s.dx, s.dy, cfg.lx, cfg.ly).to(dtype)
    return T.to(s.device)


def visualise_2d(s: Setup, T: torch.Tensor, variant: str, res: RunResult) -> None:
    """Strip halo, gather to rank 0, write PNG + raw dump (reference ``scripts/diffusion_2D_ap.jl:45-47``)."""
    cfg = s.cfg
    g = global_grid()
    from ..parallel.gather import device_gather_ok, gather_interior_device
    T_dev = None
    if device_gather_ok(T):
        # tiles go straight into the root GPU's memory; the heat map is rasterised there, and only the raw dump
        # needs the field on the host
        T_dev = gather_interior_device(T)
        T_v = T_dev.cpu() if g.me == 0 else None
    else:
        T_nh = inner_to_host(T)
        T_v = None
        if g.me == 0:
            T_v = torch.zeros((cfg.nx - 2) * s.dims[0], (cfg.ny - 2) * s.dims[1], dtype=T.dtype)
        gather(T_nh, T_v)
    if g.me == 0:
        res.T_v = T_v
        res.max_T = float(T_v.max())
        base = os.path.join(cfg.outdir, output_name(variant, s.nprocs, nx_g(), ny_g()))
        res.png = save_heatmap(base + ".png", T_dev if T_dev is not None else T_v,
                               title=f"{variant}: T after {cfg.nt} steps, {s.nprocs} ranks {s.dims[0]}x{s.dims[1]}")
        res.raw, _ = save_raw(base, T_v, {"variant": variant, "nprocs": s.nprocs, "dims": list(s.dims),
                                           "nx_g": nx_g(), "ny_g": ny_g(), "nt": cfg.nt})
