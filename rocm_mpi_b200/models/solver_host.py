"""``Diffusion2DHost``: the CPU stand-in for :class:`~rocm_mpi_b200.models.solver.Diffusion2D`.

The reference has no CPU path at all (every array is a ``ROCArray``).  This class lets the perf / perf_hide /
perf_hide_prof applications -- argument handling, time loop, ``tic``/``toc``, ``T_eff`` report, gather, heat
map -- run on machines without a GPU (laptops, CI, the timed CPU baseline) with exactly the arithmetic of the
fused kernel (``scripts/diffusion_2D_perf.jl:8-10``):

``plain``  native OpenMP step (``csrc/cpu_oracle.cpp``; PyTorch statement if the library is missing), swap,
           ``update_halo(T)`` over gloo -- the reference's serial structure (``scripts/diffusion_2D_perf.jl:49-51``);
``hide``   the intended ``hide_communication`` order (``scripts/diffusion_2D_perf_hide.jl:94-101``): boundary
           frame of width ``b_width``, halo exchange of the new field, interior.  On a CPU nothing overlaps;
           the point is that the result is the same field.

Both are bit-identical to the single-domain oracle for every rank layout.
"""
from __future__ import annotations

from typing import Sequence, Tuple

import torch

from ..ops import native
from ..ops import reference as ref
from ..parallel import coords_g, global_grid, hide_communication, nx_g, ny_g, update_halo

__all__ = ["Diffusion2DHost"]


class Diffusion2DHost:
    def __init__(self, nx: int, ny: int, *, lx: float = 10.0, ly: float = 10.0, lam: float = 1.0, Cp0: float = 1.0,
                 mode: str = "plain", b_width: Sequence[int] = (16, 4), use_native: bool = True) -> None:
        g = global_grid()
        if (nx, ny) != tuple(g.nxyz[:2]):
            raise ValueError("nx, ny must be the local sizes given to init_global_grid")
        if mode not in ("plain", "hide"):
            raise ValueError("on the CPU mode must be plain | hide")
        self.g = g
        self.nx, self.ny = nx, ny
        self.lx, self.ly, self.lam, self.Cp0 = lx, ly, lam, Cp0
        self.dx, self.dy = lx / nx_g(), ly / ny_g()
        self._dx, self._dy = 1.0 / self.dx, 1.0 / self.dy
        self.dt = ref.time_step(self.dx, self.dy, Cp0, lam)
        self.mode = mode
        self.b_width = (int(b_width[0]), int(b_width[1]))
        self.native = bool(use_native) and native.loaded()
        probe = torch.empty((nx, ny), device="meta")
        self.T = ref.gaussian_ic(coords_g(0, self.dx, probe), coords_g(1, self.dy, probe), self.dx, self.dy, lx, ly)
        self.T = self.T.contiguous()
        self.T2 = self.T.clone()                   # reference scripts/diffusion_2D_perf.jl:36
        self.Cp = torch.full((nx, ny), float(Cp0), dtype=torch.float64)
        self.steps_done = 0
        self._halo = g.nprocs > 1 or any(g.periods)

    def begin(self) -> None:
        if self._halo:
            update_halo(self.T, self.T2)

    def finish(self) -> None:
        pass

    def _region_step(self, region) -> None:
        self.T2[region] = ref._perf_update(self.T, self.Cp, self.lam, self.dt, self._dx, self._dy, region[0], region[1])

    def step(self, n: int = 1) -> None:
        for _ in range(n):
            if self.mode == "hide":
                hide_communication(self.b_width, self._region_step, self.T2)
                self.T, self.T2 = self.T2, self.T
            else:
                if self.native:
                    native.cpu_diffusion2d_step(self.T2, self.T, self.Cp, self.lam, self.dt, self._dx, self._dy)
                else:
                    ref.perf_step(self.T2, self.T, self.Cp, self.lam, self.dt, self._dx, self._dy)
                self.T, self.T2 = self.T2, self.T
                if self._halo:
                    update_halo(self.T)
            self.steps_done += 1

    # -- checkpoint / resume (the reference persists only the final PNG) -------------------
    def save(self, directory: str) -> str:
        """Collective: write ``T`` and ``Cp`` (halo included) of every rank; returns the checkpoint path."""
        from ..utils.checkpoint import save_checkpoint
        return save_checkpoint(directory, self.steps_done, T=self.T, Cp=self.Cp)

    def load(self, path: str) -> int:
        """Collective: restore the fields and the step counter from ``path``; returns the step."""
        from ..utils.checkpoint import load_checkpoint
        step, f = load_checkpoint(path)
        self.T.copy_(f["T"])
        self.T2.copy_(f["T"])
        self.Cp.copy_(f["Cp"])
        self.steps_done = step
        return step

    def max_sum(self) -> Tuple[float, float]:
        inner = self.T[1:-1, 1:-1]
        return float(inner.max()), float(inner.sum())
