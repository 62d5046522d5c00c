"""2-D heat diffusion, kernel-programming variant (``kp``): three explicit kernels per step.

Behavioural counterpart of reference ``scripts/diffusion_2D_kp.jl``: ``Flux!`` (``:16-26``), ``Residual!``
(``:33-40``), ``Update!`` (``:47-54``), then ``update_halo!(T)`` (``:91``), gather and heat map (``:93-96``).
The reference blocks the host after every launch (``wait(@roc ...)``, ``:88-90``); here the kernels are
queued on one stream with no host waits.

Halo handling (``--transport``):

* ``peer`` (default on several GPUs): ``Update!`` itself stores the edge cells into the neighbours' halo
  cells and releases the step's epoch.  Because the update is in place, two tiny flag kernels bracket the
  step: wait for the neighbours' previous epoch before ``Flux!`` reads the halo, and a neighbour barrier
  before ``Update!`` overwrites what the neighbours' ``Flux!`` is still reading.
* ``nccl`` / ``host``: the three kernels, then ``update_halo(T)`` with pack -> send/recv -> unpack.

Fast path (``--kp-path auto|resident|persistent``, CUDA only): the reference runs kp at sizes where a step is pure
launch latency (128 x 128, ``scripts/diffusion_2D_kp.jl:61``; three launches, three host waits and an exchange per
step, ``:88-91``).  The single-launch kernels behind ``Diffusion2D(mode='resident' | 'persistent')`` therefore
also come with the kp arithmetic (``form='kp'``: rounded flux products, division by ``Cp``, ``T + dt*dTdt``): all
``nt`` steps are ONE launch, halo cells travel as in-kernel peer stores, and ``qx``, ``qy``, ``dTdt`` never exist in
memory.  ``three`` (default) keeps the literal three kernels.

Run:  ``torchrun --nproc-per-node 8 -m rocm_mpi_b200.models.diffusion2d_kp --nx 65 --ny 128 --dimx 4 --dimy 2``
"""
from __future__ import annotations

import argparse
import ctypes as C

import torch

from ..ops import native
from ..ops import reference as ref
from ..parallel import finalize_global_grid, global_grid, update_halo
from ..parallel.fabric import PushPlan, fabric_zeros, get_fabric
from ..parallel.halo import default_transport
from ..utils.config import DiffusionConfig, add_config_args, config_from_args
from .common import detach_local, RunResult, initial_condition_2d, log0, setup_2d, visualise_2d

__all__ = ["diffusion2D_kp", "KPThree", "main"]


class KPThree:
    """The literal kp time loop on one rank: ``Flux!`` / ``Residual!`` / ``Update!`` launches per step
    (reference ``scripts/diffusion_2D_kp.jl:87-92``) with the halo handled by ``transport`` (see module docstring).
    ``step(n)`` queues ``n`` steps; ``finish()`` waits for the last pushes.  Used by the application below and by
    ``benchmarks/small_grid_latency.py``."""

    def __init__(self, nx: int, ny: int, dx: float, dy: float, dt: float, lam: float, Cp0: float, T0: torch.Tensor,
                 transport: str, fused: bool = True) -> None:
        g = global_grid()
        dev, f64 = T0.device, torch.float64
        self.g, self.nx, self.ny, self.dt, self.lam = g, nx, ny, dt, lam
        self._dx, self._dy = 1.0 / dx, 1.0 / dy
        self.on_gpu = dev.type == "cuda"
        if self.on_gpu:
            native.require()
        self.transport = transport
        self.comm_needed = g.nprocs > 1 or any(g.periods)
        self.fused_push = self.on_gpu and transport == "peer" and fused and self.comm_needed
        # Array initializations (reference scripts/diffusion_2D_kp.jl:73-78)
        self.dTdt = torch.zeros(nx - 2, ny - 2, dtype=f64, device=dev)
        self.qx = torch.zeros(nx - 1, ny - 2, dtype=f64, device=dev)
        self.qy = torch.zeros(nx - 2, ny - 1, dtype=f64, device=dev)
        self.Cp = Cp0 * torch.ones(nx, ny, dtype=f64, device=dev)
        if self.fused_push:
            self.T = fabric_zeros((nx, ny))
            self.T.copy_(T0)
        else:
            self.T = T0
        self.cpu_native = (not self.on_gpu) and native.loaded() and self.T.is_contiguous()
        self.plan = None
        if self.fused_push:
            self.fab = get_fabric()
            self.plan = PushPlan(self.T)
            self.bar_bank = self.fab.new_bank()
            self.bar_epoch = 0
            self.plan.begin()

    def step(self, n: int = 1) -> None:
        T, Cp, qx, qy, dTdt = self.T, self.Cp, self.qx, self.qy, self.dTdt
        for _ in range(int(n)):
            if self.on_gpu:
                if self.fused_push:
                    # neighbours' Update! of the previous step has landed in my halo cells
                    sync = self.fab.sync_struct(self.plan.bank, self.plan.epoch)
                    native._check(self.fab.lib.igg_wait_epoch(C.byref(sync), native._stream()))
                native.flux2d(qx, qy, T, self.lam, self._dx, self._dy)
                native.residual2d(dTdt, qx, qy, Cp, self._dx, self._dy)
                if self.fused_push:
                    # every neighbour has finished reading its halo cells (its Flux!) before I overwrite them
                    self.bar_epoch += 1
                    bsync = self.fab.sync_struct(self.bar_bank, self.bar_epoch)
                    native._check(self.fab.lib.igg_neighbor_barrier(C.byref(bsync), native._stream()))
                    native.update2d(T, dTdt, self.dt, push=self.plan.next(T))
                else:
                    native.update2d(T, dTdt, self.dt)
                    if self.comm_needed:
                        update_halo(T, transport=self.transport)
            else:
                if self.cpu_native:            # the three kernels as OpenMP loop nests (csrc/cpu_oracle.cpp)
                    native.cpu_kp_step(T, Cp, qx, qy, dTdt, self.lam, self.dt, self._dx, self._dy)
                else:
                    ref.kp_step(T, Cp, qx, qy, dTdt, self.lam, self.dt, self._dx, self._dy)
                if self.comm_needed:
                    update_halo(T)

    def finish(self) -> None:
        if self.plan is not None:
            self.plan.finish()

    def close(self) -> None:
        """Collective: give the flag banks back (a sweep that builds several steppers in one grid)."""
        if self.plan is not None:
            self.plan.close()
            self.fab.release_bank(self.bar_bank)
            self.plan = None


def diffusion2D_kp(cfg: DiffusionConfig = None, finalize: bool = True) -> RunResult:
    cfg = cfg or DiffusionConfig()
    s = setup_2d(cfg)
    g = global_grid()
    dev = s.device
    on_gpu = dev.type == "cuda"
    transport = cfg.transport or default_transport(torch.empty(0, device=dev))
    if cfg.kp_path not in ("three", "persistent", "resident", "auto"):
        raise ValueError("kp_path must be three | persistent | resident | auto")
    if on_gpu and cfg.kp_path != "three":
        return _single_launch(cfg, s, finalize)
    T0 = initial_condition_2d(s)
    n0 = native.launch_count()            # launches of the time loop (the reference's 3 per step), not of the set-up
    kp = KPThree(cfg.nx, cfg.ny, s.dx, s.dy, s.dt, cfg.lam, cfg.Cp0, T0, transport, cfg.fused)
    log0("Starting the time loop ...", end="")
    kp.step(int(cfg.nt))
    kp.finish()
    log0("done")
    T = kp.T
    res = RunResult("kp", s.me, s.nprocs, s.dims, int(cfg.nt), T_local=T, launches=native.launch_count() - n0)
    if cfg.do_vis:
        visualise_2d(s, T, "kp", res)
    if finalize:
        detach_local(res)
        finalize_global_grid()
    return res


def _single_launch(cfg: DiffusionConfig, s, finalize: bool) -> RunResult:
    """All ``nt`` steps of the kp arithmetic in one launch (``Diffusion2D(form='kp')``)."""
    from .solver import Diffusion2D
    n0 = native.launch_count()
    modes = {"resident": ["resident"], "persistent": ["persistent"], "auto": ["resident", "persistent"]}[cfg.kp_path]
    sol = None
    log0("Starting the time loop ...", end="")
    for k, mode in enumerate(modes):
        sol = Diffusion2D(cfg.nx, cfg.ny, lx=cfg.lx, ly=cfg.ly, lam=cfg.lam, Cp0=cfg.Cp0, mode=mode, form="kp")
        try:
            sol.begin()
            sol.step(int(cfg.nt))
            sol.finish()
            break
        except native.NativeError as e:       # field too large for one cluster: every rank takes the same branch
            if k + 1 == len(modes) or "does not fit" not in str(e):
                raise
            sol.close()
    log0("done")
    T = sol.T
    res = RunResult("kp", s.me, s.nprocs, s.dims, int(cfg.nt), T_local=T, launches=native.launch_count() - n0)
    if cfg.do_vis:
        visualise_2d(s, T, "kp", res)
    if finalize:
        detach_local(res)
        finalize_global_grid()
    return res


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    add_config_args(p)
    diffusion2D_kp(config_from_args(p.parse_args(argv)))


if __name__ == "__main__":
    main()
