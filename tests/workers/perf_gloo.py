"""Worker: perf / perf_hide / perf_hide_prof applications on CPU tensors over gloo (the CPU stand-in of the GPU
stepper), bit-exact against the single-domain oracle."""
import os
import sys

import torch

from rocm_mpi_b200.models.diffusion2d_perf import diffusion2D_perf
from rocm_mpi_b200.models.diffusion2d_perf_hide import diffusion2D_perf_hide
from rocm_mpi_b200.models.diffusion2d_perf_hide_prof import diffusion2D_perf_hide_prof
from rocm_mpi_b200.ops import reference as ref
from rocm_mpi_b200.utils.config import DiffusionConfig

outdir = sys.argv[1]
world = int(sys.argv[2])
nx, ny, dimx, dimy = {4: (34, 30, 2, 2), 8: (18, 30, 4, 2)}[world]
gx, gy = dimx * (nx - 2) + 2, dimy * (ny - 2) + 2
nt = 30


def cfg(**kw):
    return DiffusionConfig(nx=nx, ny=ny, nt=nt, dimx=dimx, dimy=dimy, device_type="cpu", outdir=outdir, quiet=True,
                           do_vis=True, warmup=5, fused=False, **kw)


T_ref, _ = ref.oracle_2d(gx, gy, nt, "perf")
res = diffusion2D_perf(cfg())
assert res.wtime > 0 and res.t_eff > 0
if res.me == 0:
    assert torch.equal(res.T_v, T_ref[1:-1, 1:-1])
    assert os.path.exists(res.png) and abs(res.max_T - float(T_ref[1:-1, 1:-1].max())) == 0.0

res = diffusion2D_perf_hide(cfg(b_width=(4, 3, 1)))
if res.me == 0:
    assert torch.equal(res.T_v, T_ref[1:-1, 1:-1])

# the profiled app runs a 12-step warm-up call first (reference scripts/diffusion_2D_perf_hide_prof.jl:110)
prof = os.path.join(outdir, "prof.txt")
res = diffusion2D_perf_hide_prof(cfg(b_width=(4, 3, 1)), prof_path=prof)
T_ref2, _ = ref.oracle_2d(gx, gy, nt + 12, "perf")
from rocm_mpi_b200.parallel.virtual import VirtualGrid                         # noqa: E402
loc = VirtualGrid((nx, ny), (dimx, dimy)).scatter(T_ref2)[res.me]
assert torch.equal(res.T_local[1:-1, 1:-1], loc[1:-1, 1:-1])
if res.me == 0:
    assert os.path.getsize(prof) > 0
print(f"rank {res.me} ok")
