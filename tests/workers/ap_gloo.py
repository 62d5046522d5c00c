"""Worker: BASELINE config 1 -- diffusion_2D_ap on CPU/gloo, compared bit-exactly with the oracle."""
import os
import sys

import torch

from rocm_mpi_b200.models.diffusion2d_ap import diffusion2D_ap
from rocm_mpi_b200.ops import reference as ref
from rocm_mpi_b200.utils.config import DiffusionConfig
from rocm_mpi_b200.utils.output import read_png_size

nx, ny, dimx, dimy, nt = map(int, sys.argv[1:6])
outdir = sys.argv[6]
cfg = DiffusionConfig(nx=nx, ny=ny, nt=nt, dimx=dimx, dimy=dimy, device_type="cpu", outdir=outdir, quiet=True)
res = diffusion2D_ap(cfg)
if res.me == 0:
    gx = res.dims[0] * (nx - 2) + 2
    gy = res.dims[1] * (ny - 2) + 2
    T_ref, _ = ref.oracle_2d(gx, gy, nt, "ap")
    assert torch.equal(res.T_v, T_ref[1:-1, 1:-1]), float((res.T_v - T_ref[1:-1, 1:-1]).abs().max())
    assert os.path.exists(res.png) and read_png_size(res.png) == (600, 400)
    assert os.path.basename(res.png) == f"Temp_ap_{res.nprocs}_{gx}_{gy}.png"
    assert os.path.getsize(res.raw) == res.T_v.numel() * 8
    print("MAX", repr(res.max_T))
print(f"rank {res.me} ok")
