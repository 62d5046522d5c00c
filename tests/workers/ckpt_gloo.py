"""Worker: checkpoint every N steps and resume, through the perf application on CPU ranks."""
import os
import sys

import torch

from rocm_mpi_b200.models.diffusion2d_perf import diffusion2D_perf
from rocm_mpi_b200.utils.config import DiffusionConfig

outdir = sys.argv[1]
ck = os.path.join(outdir, "ckpt")


def cfg(nt, **kw):
    return DiffusionConfig(nx=26, ny=22, nt=nt, dimx=2, dimy=2, device_type="cpu", outdir=outdir, quiet=True,
                           do_vis=False, warmup=3, fused=False, ckpt_dir=ck, **kw)


straight = diffusion2D_perf(cfg(40)).T_local.clone()                      # 40 steps in one go
part = diffusion2D_perf(cfg(30, ckpt_every=10))                           # 30 steps, checkpoints at 10, 20, 30
assert sorted(os.listdir(ck)) == ["ckpt_00000010", "ckpt_00000020", "ckpt_00000030"]
assert part.t_eff > 0
resumed = diffusion2D_perf(cfg(40, resume=True, ckpt_every=10))           # picks up at 30, runs 10 more
assert torch.equal(resumed.T_local, straight), float((resumed.T_local - straight).abs().max())
assert "ckpt_00000040" in os.listdir(ck)
print(f"rank {part.me} ok")
