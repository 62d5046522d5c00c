"""Worker (one per GPU): applications on several GPUs against the single-domain oracle."""
import sys

import torch

import rocm_mpi_b200 as igg
from rocm_mpi_b200.ops import reference as ref
from rocm_mpi_b200.utils.config import DiffusionConfig

outdir = sys.argv[1]
world = int(sys.argv[2])

# 1) ring test (reference rocmaware_test_selectdevice.jl)
from rocm_mpi_b200.models.selectdevice import selectdevice_test
out = selectdevice_test(4, quiet=True)
me = int(__import__("os").environ["RANK"])
assert out["peer"].tolist() == [float((me - 1) % world)] * 4
assert out["nccl"].tolist() == [float((me - 1) % world)] * 4

# 2) kp app: fused push and NCCL transport, 254 x 254 global problem of the reference image
from rocm_mpi_b200.models.diffusion2d_kp import diffusion2D_kp
layouts = {2: (128, 254, 2, 1), 4: (128, 128, 2, 2), 8: (65, 128, 4, 2)}
nx, ny, dimx, dimy = layouts[world]
T_ref, _ = ref.oracle_2d(254, 254, 120, "kp")
for transport in ("peer", "nccl"):
    cfg = DiffusionConfig(nx=nx, ny=ny, nt=120, dimx=dimx, dimy=dimy, outdir=outdir, quiet=True, transport=transport)
    res = diffusion2D_kp(cfg)
    if res.me == 0:
        err = float((res.T_v - T_ref[1:-1, 1:-1]).abs().max())
        assert err < 1e-12, (transport, err)

# 3) perf / hide apps
from rocm_mpi_b200.models.diffusion2d_perf import diffusion2D_perf
from rocm_mpi_b200.models.diffusion2d_perf_hide import diffusion2D_perf_hide
T_refp, _ = ref.oracle_2d(254, 254, 80, "perf")
ny2 = ny if ny % 2 == 0 else ny      # TMA needs even ny; all layouts above have even ny
for fn, kw in ((diffusion2D_perf, dict(fused=True)), (diffusion2D_perf_hide, dict(fused=False, b_width=(8, 4, 4)))):
    cfg = DiffusionConfig(nx=nx, ny=ny2, nt=80, warmup=10, dimx=dimx, dimy=dimy, outdir=outdir, quiet=True, **kw)
    res = fn(cfg)
    if res.me == 0:
        err = float((res.T_v - T_refp[1:-1, 1:-1]).abs().max())
        assert err < 1e-12, (fn.__name__, err)
        assert abs(res.max_T - float(T_refp[1:-1, 1:-1].max())) < 1e-12

# 4) 3-D app split along z (and x when enough ranks)
from rocm_mpi_b200.models.diffusion3d import diffusion3D
dims3 = {2: (1, 1, 2), 4: (2, 1, 2), 8: (2, 2, 2)}[world]
loc = (24, 22, 34)
g3 = tuple(d * (n - 2) + 2 for d, n in zip(dims3, loc))
cfg = DiffusionConfig(nx=loc[0], ny=loc[1], nz=loc[2], nt=10, warmup=2, dimx=dims3[0], dimy=dims3[1], dimz=dims3[2],
                      outdir=outdir, quiet=True, do_vis=True, b_width=(4, 4, 8))
res = diffusion3D(cfg)
if res.me == 0:
    T3, _ = ref.oracle_3d(g3[0], g3[1], g3[2], 10)
    err = float((res.T_v - T3[1:-1, 1:-1, 1:-1]).abs().max())
    assert err < 1e-12, err
print(f"rank {me} ok")
