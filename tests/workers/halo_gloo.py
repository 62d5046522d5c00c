"""Worker: update_halo/gather over gloo must equal the virtual-rank model bit for bit."""
import sys

import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.parallel.virtual import VirtualGrid

nx, ny, nz, dimx, dimy, dimz, px, py, pz = map(int, sys.argv[1:10])
me, dims, nprocs, coords, comm = igg.init_global_grid(nx, ny, nz, dimx=dimx, dimy=dimy, dimz=dimz,
                                                      periodx=px, periody=py, periodz=pz,
                                                      device_type="cpu", quiet=True)
vg = VirtualGrid((nx, ny, nz), dims, (px, py, pz))
assert vg.nxyz_g == (igg.nx_g(), igg.ny_g(), igg.nz_g())
nd = 3 if nz > 1 else 2
base = (nx, ny, nz)[:nd]
shapes = [base, tuple(b + (1 if d == 0 else 0) for d, b in enumerate(base)),
          tuple(b - (1 if d == 1 else 0) for d, b in enumerate(base))]
gen = torch.Generator().manual_seed(1234)
all_fields = [[torch.rand(s, generator=gen, dtype=torch.float64) for _ in range(nprocs)] for s in shapes]
mine = [f[me].clone() for f in all_fields]
igg.update_halo(*mine)
vg.update_halo(*all_fields)
for k in range(len(shapes)):
    assert torch.equal(mine[k], all_fields[k][me]), f"rank {me} field {k} mismatch"
# gather
blk = igg.inner(mine[0])
out = igg.gather_interior(mine[0])
if me == 0:
    want = vg.gather([igg.inner(f) for f in all_fields[0]])
    assert torch.equal(out, want)
# tic/toc plumbing
igg.tic()
t = igg.toc()
assert t >= 0
igg.finalize_global_grid()
print(f"rank {me} ok")
