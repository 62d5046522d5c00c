"""Worker (one per GPU): long runs on SMALL local grids, where the ranks are tightly coupled and a race in
the flag protocol would show up as a wrong field.  (compute-sanitizer's racecheck cannot see cross-GPU
traffic; this is the protocol's own stress test.)"""
import sys

import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.models.solver import Diffusion2D
from rocm_mpi_b200.ops import reference as ref

nx, ny, px, py, nt = map(int, sys.argv[1:6])
me, dims, nprocs, coords, comm = igg.init_global_grid(nx, ny, 1, periodx=px, periody=py, quiet=True)
vg = igg.VirtualGrid((nx, ny), dims, (px, py))
res = {}
for mode in ("fused", "temporal", "persistent", "resident", "hide"):
    if mode == "temporal" and ny % 2:
        continue
    s = Diffusion2D(nx, ny, mode=mode, variant="tma" if mode in ("fused", "hide", "temporal") else "direct", b_width=(4, 2))
    T0 = s.T.cpu().clone()
    s.begin()
    done = 0
    for chunk in (1, 7, 64, nt - 72):      # uneven chunks: epochs / barrier bases carry over between calls
        s.step(chunk)
        done += chunk
    s.finish()
    torch.cuda.synchronize()
    res[mode] = s.T.cpu().clone()
locs = igg.global_grid().comm.allgather_object(T0)
Cp = torch.ones(nx, ny, dtype=torch.float64)
vg.update_halo(locs)
l2 = [t.clone() for t in locs]
dx, dy = 10.0 / igg.nx_g(), 10.0 / igg.ny_g()
dt = ref.time_step(dx, dy)
for _ in range(nt):
    for a, b in zip(locs, l2):
        ref.perf_step(b, a, Cp, 1.0, dt, 1.0 / dx, 1.0 / dy)
    vg.update_halo(l2)
    locs, l2 = l2, locs
for mode, T in res.items():
    err = float((T - locs[me]).abs().max())
    assert err < 1e-12, (mode, me, err)
igg.finalize_global_grid()
print(f"rank {me} ok")
