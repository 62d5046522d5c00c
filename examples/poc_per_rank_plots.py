#!/usr/bin/env python
"""Proof of concept: per-rank pictures before and after, to see the heat cross the rank seam.

The reference documents its halo exchange with a screenshot of four interactive plot windows -- two ranks,
each before and after the run, hot spot straddling the seam (``README.md:5-7``, ``docs/poc_rocmaware.png``).
Interactive windows are unavailable on this box (no graphics stack); this writes the same four pictures as PNG
files, plus the seam check as numbers: after ``update_halo`` the halo column of one rank equals the first inner
column of its neighbour.

    torchrun --nproc-per-node 2 examples/poc_per_rank_plots.py [--cpu] [--size 128] [--nt 300]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.ops import reference as ref  # noqa: E402
from rocm_mpi_b200.utils.output import save_heatmap  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cpu", action="store_true")
ap.add_argument("--size", dest="n", type=int, default=128)
ap.add_argument("--nt", type=int, default=300)
ap.add_argument("--outdir", default="output")
a = ap.parse_args()

lx = ly = 10.0
me, dims, nprocs, coords, _ = igg.init_global_grid(a.n, a.n, 1, device_type="cpu" if a.cpu else "auto", quiet=True)
dev = igg.global_grid().device
dx, dy = lx / igg.nx_g(), ly / igg.ny_g()
dt = ref.time_step(dx, dy, 1.0, 1.0)
T = torch.zeros(a.n, a.n, dtype=torch.float64, device=dev)
x, y = igg.coords_g(0, dx, T, device=dev), igg.coords_g(1, dy, T, device=dev)
T.copy_(ref.gaussian_ic(x, y, dx, dy, lx, ly))
Cp = torch.ones_like(T)
os.makedirs(a.outdir, exist_ok=True)
vmax = 1.0
before = save_heatmap(os.path.join(a.outdir, f"poc_rank{me}_before.png"), T, vmin=0.0, vmax=vmax,
                      title=f"rank {me} {tuple(coords[:2])}: initial")
for _ in range(a.nt):
    ref.ap_step(T, Cp, 1.0, dt, dx, dy)
    igg.update_halo(T)
after = save_heatmap(os.path.join(a.outdir, f"poc_rank{me}_after.png"), T, vmin=0.0, vmax=vmax,
                     title=f"rank {me} {tuple(coords[:2])}: after {a.nt} steps")

# the seam in numbers: my high-x halo row == the x-high neighbour's first inner row (and so on); gather the edge
# rows on rank 0 through the public gather and compare
g = igg.global_grid()
edges = torch.stack([T[1], T[-2], T[0], T[-1]]).cpu()                    # send-low, send-high, halo-low, halo-high
allv = g.comm.allgather_object(edges)
worst = 0.0
if g.neighbors[1][0] >= 0:
    nb = allv[g.neighbors[1][0]]
    worst = max(worst, float((edges[3] - nb[0]).abs().max()))              # my high halo == neighbour's row 1
if g.neighbors[0][0] >= 0:
    nb = allv[g.neighbors[0][0]]
    worst = max(worst, float((edges[2] - nb[1]).abs().max()))              # my low halo == neighbour's row n-2
worst = g.comm.allreduce_max(worst)
if me == 0:
    print(f"{nprocs} ranks {dims[0]}x{dims[1]}: wrote poc_rank*_before.png / _after.png in {a.outdir}; "
          f"max |halo - neighbour's edge| across x seams = {worst:.1e}")
igg.finalize_global_grid()
