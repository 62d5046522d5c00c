"""Worker (one per GPU): peer fabric, one-launch update_halo, fused solver and NCCL baseline."""
import sys

import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.models.solver import Diffusion2D
from rocm_mpi_b200.ops import native, reference as ref
from rocm_mpi_b200.parallel.fabric import get_fabric

nx, ny, px, py, nt = map(int, sys.argv[1:6])
me, dims, nprocs, coords, comm = igg.init_global_grid(nx, ny, 1, periodx=px, periody=py, quiet=True)
dev = igg.global_grid().device
vg = igg.VirtualGrid((nx, ny), dims, (px, py))

# 1) update_halo: peer / nccl / host transports against the virtual-rank model
gen = torch.Generator().manual_seed(99)
shapes = [(nx, ny), (nx + 1, ny), (nx, ny - 1)]
allf = [[torch.rand(s, generator=gen, dtype=torch.float64) for _ in range(nprocs)] for s in shapes]
want = [[t.clone() for t in f] for f in allf]
vg.update_halo(*want)
for transport in ("peer", "nccl", "host"):
    mine = [f[me].to(dev) for f in allf]
    for rep in range(2):
        igg.update_halo(*mine, transport=transport)
    torch.cuda.synchronize()
    for k in range(len(shapes)):
        assert torch.equal(mine[k].cpu(), want[k][me]), (transport, k, me)
get_fabric().check_errors(0)

# 2) solver modes against the single-domain / virtual oracle
gx, gy = igg.nx_g(), igg.ny_g()
results = {}
for mode, variant, transport in (("fused", "tma", "peer"), ("temporal", "tma", "peer"), ("fused", "direct", "peer"),
                                 ("hide", "tma", "peer"),
                                 ("plain", "tma", "nccl"), ("plain", "tma", "peer"), ("persistent", "direct", "peer"), ("resident", "direct", "peer")):
    s = Diffusion2D(nx, ny, mode=mode, variant=variant, transport=transport, b_width=(8, 4))
    T0 = s.T.cpu().clone()
    s.begin()
    s.step(nt)
    s.finish()
    torch.cuda.synchronize()
    results[(mode, variant, transport)] = s.T.cpu().clone()
# oracle by virtual ranks (handles periodic layouts too)
locs = igg.global_grid().comm.allgather_object(T0)
Cp = torch.ones(nx, ny, dtype=torch.float64)
vg.update_halo(locs)
l2 = [t.clone() for t in locs]
dx, dy = 10.0 / gx, 10.0 / gy
dt = ref.time_step(dx, dy)
for _ in range(nt):
    for a, b in zip(locs, l2):
        ref.perf_step(b, a, Cp, 1.0, dt, 1.0 / dx, 1.0 / dy)
    vg.update_halo(l2)
    locs, l2 = l2, locs
for key, T in results.items():
    err = float((T - locs[me]).abs().max())
    assert err < 1e-13, (key, me, err)
igg.finalize_global_grid()
print(f"rank {me} ok")
