"""Per-step device times of the fused solver on every rank (debug aid)."""
import os, sys
import torch
sys.path.insert(0, ".")
import rocm_mpi_b200 as igg
from rocm_mpi_b200.models.solver import Diffusion2D

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
mode = sys.argv[2] if len(sys.argv) > 2 else "fused"
me, dims, nprocs, _, _ = igg.init_global_grid(n, n, 1, quiet=True)
s = Diffusion2D(n, n, mode=mode, variant="tma")
s.begin(); s.step(10)
igg.global_grid().comm.barrier(); torch.cuda.synchronize()
K = 24
ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
ev[0].record()
for k in range(K):
    s.step(1)
    ev[k + 1].record()
torch.cuda.synchronize()
s.finish()
d = [ev[k].elapsed_time(ev[k + 1]) for k in range(K)]
print(f"[rank {me}] " + " ".join(f"{x:.3f}" for x in d) + f" | total {ev[0].elapsed_time(ev[K]) / K:.4f}", flush=True)
igg.finalize_global_grid()
