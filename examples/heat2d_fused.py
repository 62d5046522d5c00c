#!/usr/bin/env python
"""Heat diffusion on GPUs with the fused kernel driven by hand: fields in peer-addressable memory, a
PushPlan that lets the stencil kernel store its edge cells straight into the neighbours' halo cells, and the
explicit begin / step / finish protocol that `Diffusion2D` wraps.

    torchrun --nproc-per-node 8 examples/heat2d_fused.py --size 4096 --nt 200
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.ops import native  # noqa: E402
from rocm_mpi_b200.parallel.fabric import PushPlan  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=4096)
ap.add_argument("--nt", type=int, default=200)
a = ap.parse_args()

lx = ly = 10.0
me, dims, nprocs, coords, _ = igg.init_global_grid(a.n, a.n, 1)
dx, dy = lx / igg.nx_g(), ly / igg.ny_g()
dt = min(dx, dy) ** 2 / 4.1

T, T2 = igg.zeros((a.n, a.n)), igg.zeros((a.n, a.n))       # collective: peer-addressable fields
Cp = torch.ones(a.n, a.n, dtype=torch.float64, device=T.device)
x0 = float(igg.x_g(0, dx, T)), float(igg.y_g(0, dy, T))
native.gaussian_ic_2d(T, x0[0], x0[1], dx, dy, lx, ly)     # initial condition computed on the device
T2.copy_(T)

plan = PushPlan(T, T2)                                      # collective: flag bank + neighbour pointers
plan.begin()                                                # halos consistent, neighbours in step
igg.tic()
for it in range(a.nt):
    # ONE launch: TMA stencil + edge cells pushed into the neighbours' T2 + epoch flag released
    native.diffusion2d_step(T2, T, Cp, 1.0, dt, 1 / dx, 1 / dy, push=plan.next(T2))
    T, T2 = T2, T
wtime = igg.toc()
plan.finish()                                               # every neighbour's last push has landed

T_eff = igg.t_eff((a.n, a.n), wtime, a.nt)
mx = float(native.reduce_max_sum(T.view(-1))[0])
if me == 0:
    print(f"{nprocs} GPUs {dims[0]}x{dims[1]}: {wtime / a.nt * 1e3:.4f} ms/step, T_eff = {T_eff:.0f} GB/s per GPU, "
          f"local max(T) = {mx:.6f}")
igg.finalize_global_grid()
