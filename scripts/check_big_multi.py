"""Consistency check at benchmark-like sizes: fused strip-marching kernel with in-kernel push versus the
direct-load kernel with NCCL update_halo, same global problem, field compared after several steps.
usage: torchrun --nproc-per-node N scripts/check_big_multi.py [nx ny steps]"""
import sys

import torch

sys.path.insert(0, ".")
import rocm_mpi_b200 as igg
from rocm_mpi_b200.models.solver import Diffusion2D

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
ny = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 12
me, dims, nprocs, _, _ = igg.init_global_grid(nx, ny, 1, quiet=True)
a = Diffusion2D(nx, ny, mode="fused", variant="tma")
b = Diffusion2D(nx, ny, mode="plain", variant="direct", transport="nccl" if nprocs > 1 else "peer")
for s in (a, b):
    s.begin()
    s.step(steps)
    s.finish()
torch.cuda.synchronize()
err = float((a.T - b.T).abs().max())
errh = max(float((a.T[0] - b.T[0]).abs().max()), float((a.T[-1] - b.T[-1]).abs().max()),
           float((a.T[:, 0] - b.T[:, 0]).abs().max()), float((a.T[:, -1] - b.T[:, -1]).abs().max()))
mx = igg.global_grid().comm.allreduce_max(err)
if me == 0:
    print(f"{nprocs} ranks {dims[0]}x{dims[1]}, local {nx}x{ny}, {steps} steps: max|fused - baseline| = {mx:.3e} "
          f"(halo/boundary cells {errh:.3e})", flush=True)
assert mx < 1e-13, mx
igg.finalize_global_grid()
