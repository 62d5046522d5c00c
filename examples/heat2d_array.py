#!/usr/bin/env python
"""Heat diffusion with nothing but the global-grid API and torch expressions (CPU or GPU).

The smallest complete program: decompose, initialise from global coordinates, step with array
expressions, exchange halos, gather, write the picture.  Works on CPU tensors over gloo
(`torchrun --nproc-per-node 4 examples/heat2d_array.py --cpu`) and on GPUs.
(What the reference's array-programming script does, scripts/diffusion_2D_ap.jl, as a user of THIS API
would write it.)
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.utils.output import save_heatmap  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cpu", action="store_true")
ap.add_argument("--size", dest="n", type=int, default=128, help="local grid points per dimension")
ap.add_argument("--nt", type=int, default=500)
ap.add_argument("--out", default="output/heat2d_array.png")
a = ap.parse_args()

lx = ly = 10.0
me, dims, nprocs, coords, _ = igg.init_global_grid(a.n, a.n, 1, device_type="cpu" if a.cpu else "auto")
dev = igg.global_grid().device
dx, dy = lx / igg.nx_g(), ly / igg.ny_g()
dt = min(dx, dy) ** 2 / 4.1

T = torch.zeros(a.n, a.n, dtype=torch.float64, device=dev)
x = igg.coords_g(0, dx, T, device=dev)                      # global coordinates of my rows / columns
y = igg.coords_g(1, dy, T, device=dev)
T.copy_(torch.exp(-(x[:, None] + dx / 2 - lx / 2) ** 2 - (y[None, :] + dy / 2 - ly / 2) ** 2))

for it in range(a.nt):
    qx = -(T[1:, 1:-1] - T[:-1, 1:-1]) / dx                 # Fourier's law
    qy = -(T[1:-1, 1:] - T[1:-1, :-1]) / dy
    T[1:-1, 1:-1] -= dt * ((qx[1:] - qx[:-1]) / dx + (qy[:, 1:] - qy[:, :-1]) / dy)   # conservation of energy
    igg.update_halo(T)

T_v = igg.gather_interior(T)                                # halo-free global field on rank 0
if me == 0:
    print(f"{nprocs} ranks {dims[0]}x{dims[1]}, global {igg.nx_g()}x{igg.ny_g()}: max(T) = {float(T_v.max()):.6f}")
    save_heatmap(a.out, T_v, title=f"T after {a.nt} steps")
igg.finalize_global_grid()
