"""3-D heat diffusion on the implicit global grid (7-point stencil, fused kernel + in-kernel halo push).

The reference is 2-D only (``init_global_grid(nx, ny, 1)``, ``scripts/diffusion_2D_ap.jl:17``); 3-D is the
ImplicitGlobalGrid capability the scripts' API exposes (``nz``, ``dimz``, ``nz_g()``, ``z_g()``) and the north
star asks for.  Same structure as the 2-D perf application (``scripts/diffusion_2D_perf.jl:47-58``):
double-buffered fused step, halo exchange, ``tic/toc`` after warm-up, ``T_eff`` -- here with up to 26
neighbours per rank, all served by peer stores from inside the stencil kernel.

Run:  ``torchrun --nproc-per-node 8 -m rocm_mpi_b200.models.diffusion3d --nx 512 --ny 512 --nz 512 --nt 60``
"""
from __future__ import annotations

import argparse
import os
from typing import Optional, Sequence

import torch

from ..ops import native
from ..ops import reference as ref
from ..parallel import (a_eff, coords_g, finalize_global_grid, gather, global_grid, init_global_grid, inner,
                        nx_g, ny_g, nz_g, tic, toc, update_halo)
from ..parallel.fabric import PushPlan, fabric_zeros
from ..parallel.halo import default_transport
from ..utils.config import DiffusionConfig, add_config_args, config_from_args
from ..utils.output import output_name, save_heatmap, save_raw
from .common import detach_local, RunResult, log0

__all__ = ["Diffusion3D", "diffusion3D", "main"]


class Diffusion3D:
    """3-D time stepper: ``fused`` (one launch per step, edges pushed in-kernel), ``plain`` (kernel, swap,
    ``update_halo`` over a transport) or ``hide`` (the two-stream structure of reference
    ``scripts/diffusion_2D_perf_hide.jl:63-68, 94-101`` carried to 3-D: boundary frame of width ``b_width`` +
    edge push on a high-priority stream, interior on a low-priority stream, join, swap; its two kernels are
    GPU-tested individually, the two-stream driver itself has not been run on hardware yet)."""

    def __init__(self, nx: int, ny: int, nz: int, *, lx=10.0, ly=10.0, lz=10.0, lam=1.0, Cp0=1.0,
                 mode: str = "fused", variant: str = "tma", b_width: Sequence[int] = (4, 4, 8),
                 transport: Optional[str] = None) -> None:
        g = global_grid()
        if g.device_type != "cuda":
            raise RuntimeError("Diffusion3D runs on CUDA devices")
        if (nx, ny, nz) != tuple(g.nxyz):
            raise ValueError("nx, ny, nz must be the local sizes given to init_global_grid")
        if mode not in ("fused", "plain", "hide"):
            raise ValueError("mode must be fused | plain | hide")
        if variant == "tma" and nz % 2:
            raise ValueError("variant 'tma' needs an even nz; use variant='direct'")
        native.require()
        self.g, self.shape = g, (nx, ny, nz)
        self.lam = lam
        self.dx, self.dy, self.dz = lx / nx_g(), ly / ny_g(), lz / nz_g()
        self._d = (1.0 / self.dx, 1.0 / self.dy, 1.0 / self.dz)
        self.dt = ref.time_step_3d(self.dx, self.dy, self.dz, Cp0, lam)
        self.mode, self.variant, self.b_width = mode, variant, tuple(int(b) for b in b_width)
        dev = g.device
        self.transport = transport or default_transport(torch.empty(0, device=dev))
        peer = mode in ("fused", "hide") or self.transport == "peer"
        mk = fabric_zeros if peer else (lambda s: torch.zeros(s, dtype=torch.float64, device=dev))
        self.T, self.T2 = mk(self.shape), mk(self.shape)
        self.Cp = torch.full(self.shape, float(Cp0), dtype=torch.float64, device=dev)
        probe = torch.empty(self.shape, device="meta")
        if any(g.periods):
            xs, ys, zs = (coords_g(d, h, probe) for d, h in enumerate((self.dx, self.dy, self.dz)))
            self.T.copy_(ref.gaussian_ic_3d(xs, ys, zs, self.dx, self.dy, self.dz, lx, ly, lz))
        else:
            o = [float(coords_g(d, h, probe)[0]) for d, h in enumerate((self.dx, self.dy, self.dz))]
            native.gaussian_ic_3d(self.T, o[0], o[1], o[2], self.dx, self.dy, self.dz, lx, ly, lz)
        self.T2.copy_(self.T)
        self.plan = PushPlan(self.T, self.T2) if mode in ("fused", "hide") else None
        self._running = False
        self._s_hi = self._s_lo = None
        if mode == "hide":
            self._s_hi = torch.cuda.Stream(device=dev, priority=-1)
            self._s_lo = torch.cuda.Stream(device=dev, priority=0)
            self._ev_main, self._ev_hi, self._ev_lo = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()

    def begin(self) -> None:
        if self.plan is not None:
            self.plan.begin()
        elif self.g.nprocs > 1 or any(self.g.periods):
            update_halo(self.T, self.T2, transport=self.transport)
        self._running = True

    def finish(self) -> None:
        if self._s_hi is not None:
            torch.cuda.current_stream().wait_stream(self._s_hi)
            torch.cuda.current_stream().wait_stream(self._s_lo)
        if self.plan is not None and self._running:
            self.plan.finish()
        self._running = False

    def _kernel(self, mode: int, push) -> None:
        native.diffusion3d_step(self.T2, self.T, self.Cp, self.lam, self.dt, *self._d, mode=mode,
                                b_width=self.b_width, variant=self.variant, push=push)

    def step(self, n: int = 1) -> None:
        if not self._running:
            self.begin()
        for _ in range(n):
            push = self.plan.next(self.T2) if self.plan is not None else None
            if self.mode == "hide":
                main = torch.cuda.current_stream()
                self._ev_main.record(main)
                self._s_hi.wait_event(self._ev_main)
                self._s_lo.wait_event(self._ev_main)
                with torch.cuda.stream(self._s_hi):
                    self._kernel(1, push)                 # frame + edge push
                    self._ev_hi.record(self._s_hi)
                with torch.cuda.stream(self._s_lo):
                    self._kernel(2, None)                 # interior
                    self._ev_lo.record(self._s_lo)
                main.wait_event(self._ev_hi)
                main.wait_event(self._ev_lo)
            else:
                self._kernel(0, push)
            self.T, self.T2 = self.T2, self.T
            if self.plan is None and (self.g.nprocs > 1 or any(self.g.periods)):
                update_halo(self.T, transport=self.transport)


def diffusion3D(cfg: DiffusionConfig = None, finalize: bool = True, mode: Optional[str] = None) -> RunResult:
    cfg = cfg or DiffusionConfig(nx=256, ny=256, nz=256, nt=60, do_vis=False)
    me, dims, nprocs, coords, _ = init_global_grid(
        cfg.nx, cfg.ny, cfg.nz, dimx=cfg.dimx, dimy=cfg.dimy, dimz=cfg.dimz, periodx=cfg.periodx,
        periody=cfg.periody, periodz=cfg.periodz, device_type=cfg.device_type, quiet=cfg.quiet)
    g = global_grid()
    if not cfg.quiet:
        print(f"Process {me} selecting device {g.device}", flush=True)
    if g.device_type != "cuda":
        return _diffusion3D_cpu(cfg, me, dims, nprocs, finalize)
    n0 = native.launch_count()
    s = Diffusion3D(cfg.nx, cfg.ny, cfg.nz, lx=cfg.lx, ly=cfg.ly, lz=cfg.lz, lam=cfg.lam, Cp0=cfg.Cp0,
                    mode=mode or ("fused" if cfg.fused else "plain"), variant=cfg.kernel, b_width=cfg.b_width,
                    transport=cfg.transport)
    nt, warm = int(cfg.nt), min(int(cfg.warmup), max(int(cfg.nt) - 1, 0))
    log0("Starting the time loop ...", end="")
    s.begin()
    s.step(warm)
    tic()
    s.step(nt - warm)
    wtime = toc()
    s.finish()
    log0("done")
    T_eff = a_eff(s.shape) / (wtime / max(nt - warm, 1))
    log0("Executed %d steps in = %1.3e sec (@ T_eff = %1.2f GB/s) " % (nt, wtime, float("%.3g" % T_eff)))
    res = RunResult("3D", me, nprocs, dims, nt, wtime=wtime, t_eff=T_eff, T_local=s.T,
                    launches=native.launch_count() - n0)
    if cfg.do_vis:
        from ..parallel.gather import device_gather_ok, gather_interior_device, inner_to_host
        if device_gather_ok(s.T):          # tiles stored into the root GPU's memory (3-D blocks, 16-byte rows)
            G = gather_interior_device(s.T)
            T_v = G.cpu() if me == 0 else None
        else:
            T_nh = inner_to_host(s.T)
            T_v = None
            if me == 0:
                T_v = torch.zeros(tuple((n - 2) * d for n, d in zip(s.shape, dims)), dtype=torch.float64)
            gather(T_nh, T_v)
        if me == 0:
            res.T_v, res.max_T = T_v, float(T_v.max())
            base = os.path.join(cfg.outdir, output_name("3D", nprocs, nx_g(), ny_g(), nz_g()))
            res.png = save_heatmap(base + ".png", T_v[:, :, T_v.shape[2] // 2], title="3D: mid-plane z slice")
            res.raw, _ = save_raw(base, T_v, {"variant": "3D", "nprocs": nprocs, "dims": list(dims), "nt": nt})
    if finalize:
        detach_local(res)
        finalize_global_grid()
    return res


def _diffusion3D_cpu(cfg: DiffusionConfig, me: int, dims, nprocs: int, finalize: bool) -> RunResult:
    """Array-programming 3-D run on CPU tensors over gloo (plumbing / tests; same structure as the 2-D
    ``ap`` application, reference ``scripts/diffusion_2D_ap.jl:37-43``)."""
    shape = (cfg.nx, cfg.ny, cfg.nz)
    dx, dy, dz = cfg.lx / nx_g(), cfg.ly / ny_g(), cfg.lz / nz_g()
    dt = ref.time_step_3d(dx, dy, dz, cfg.Cp0, cfg.lam)
    probe = torch.empty(shape, device="meta")
    xs, ys, zs = (coords_g(d, h, probe) for d, h in enumerate((dx, dy, dz)))
    T = ref.gaussian_ic_3d(xs, ys, zs, dx, dy, dz, cfg.lx, cfg.ly, cfg.lz)
    Cp = cfg.Cp0 * torch.ones(shape, dtype=torch.float64)
    for _ in range(int(cfg.nt)):
        ref.ap_step_3d(T, Cp, cfg.lam, dt, dx, dy, dz)
        update_halo(T)
    res = RunResult("3D", me, nprocs, dims, int(cfg.nt), T_local=T)
    if cfg.do_vis:
        T_v = torch.zeros(tuple((n - 2) * d for n, d in zip(shape, dims)), dtype=torch.float64) if me == 0 else None
        gather(inner(T).contiguous(), T_v)
        if me == 0:
            res.T_v, res.max_T = T_v, float(T_v.max())
            base = os.path.join(cfg.outdir, output_name("3D", nprocs, nx_g(), ny_g(), nz_g()))
            res.png = save_heatmap(base + ".png", T_v[:, :, T_v.shape[2] // 2], title="3D: mid-plane z slice")
            res.raw, _ = save_raw(base, T_v, {"variant": "3D", "nprocs": nprocs, "dims": list(dims), "nt": int(cfg.nt)})
    if finalize:
        finalize_global_grid()
    return res


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    add_config_args(p, DiffusionConfig(nx=256, ny=256, nz=256, nt=60, do_vis=False, b_width=(4, 4, 8)))
    p.add_argument("--mode", default=None, choices=["fused", "plain", "hide"],
                   help="time-stepper structure (default: fused, or plain with --fused false)")
    ns = p.parse_args(argv)
    diffusion3D(config_from_args(ns), mode=ns.mode)


if __name__ == "__main__":
    main()
