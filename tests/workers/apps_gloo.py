"""Worker: kp and 3-D applications plus hide_communication on CPU tensors over gloo, against the oracles."""
import sys

import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.ops import reference as ref
from rocm_mpi_b200.utils.config import DiffusionConfig

outdir = sys.argv[1]
world = int(sys.argv[2])

# kp app (three explicit kernels -> here their PyTorch statements), 4 or 8 ranks, bit-exact vs the kp oracle
from rocm_mpi_b200.models.diffusion2d_kp import diffusion2D_kp
nx, ny, dimx, dimy = {4: (34, 30, 2, 2), 8: (18, 30, 4, 2)}[world]
res = diffusion2D_kp(DiffusionConfig(nx=nx, ny=ny, nt=40, dimx=dimx, dimy=dimy, device_type="cpu", outdir=outdir,
                                     quiet=True))
if res.me == 0:
    gx, gy = dimx * (nx - 2) + 2, dimy * (ny - 2) + 2
    T_ref, _ = ref.oracle_2d(gx, gy, 40, "kp")
    assert torch.equal(res.T_v, T_ref[1:-1, 1:-1])

# 3-D array-programming run, 2x2x1 or 2x2x2 ranks, bit-exact vs the ap-form 3-D oracle
from rocm_mpi_b200.models.diffusion3d import diffusion3D
dims3 = {4: (2, 2, 1), 8: (2, 2, 2)}[world]
loc = (10, 9, 8)
res = diffusion3D(DiffusionConfig(nx=loc[0], ny=loc[1], nz=loc[2], nt=12, dimx=dims3[0], dimy=dims3[1], dimz=dims3[2],
                                  device_type="cpu", outdir=outdir, quiet=True, do_vis=True))
if res.me == 0:
    g3 = tuple(d * (n - 2) + 2 for d, n in zip(dims3, loc))
    T3, _ = ref.oracle_3d(g3[0], g3[1], g3[2], 12, form="ap")
    assert torch.equal(res.T_v, T3[1:-1, 1:-1, 1:-1])

# hide_communication == step + update_halo
me, dims, nprocs, coords, _ = igg.init_global_grid(20, 18, 1, device_type="cpu", quiet=True)
g = torch.Generator().manual_seed(100 + me)
T = torch.rand(20, 18, generator=g, dtype=torch.float64)
Cp = torch.rand(20, 18, generator=g, dtype=torch.float64) + 0.5
igg.update_halo(T)
want = T.clone()
ref.perf_step(want, T, Cp, 1.0, 1e-3, 2.0, 3.0)
igg.update_halo(want)
T2 = T.clone()


def step(region):
    T2[region] = ref._perf_update(T, Cp, 1.0, 1e-3, 2.0, 3.0, region[0], region[1])


igg.hide_communication((4, 2), step, T2)
assert torch.equal(T2, want)
igg.finalize_global_grid()
print(f"rank {me} ok")
