"""Launch every kernel of the native library once at a representative size (single GPU), for one ncu capture
per kernel.  Multi-rank paths run against the rank itself (periodic self-neighbour), which exercises the same
peer-store / flag code."""
import sys

import torch

sys.path.insert(0, ".")
import rocm_mpi_b200 as igg
from rocm_mpi_b200.models.solver import Diffusion2D
from rocm_mpi_b200.models.diffusion3d import Diffusion3D
from rocm_mpi_b200.ops import native
from rocm_mpi_b200.utils.output import inferno_lut

dev = "cuda"
# --- 2-D kernels at 8192 x 8192, periodic so that the push / flag paths are live
igg.init_global_grid(8192, 8192, 1, periodx=1, periody=1, quiet=True, init_dist=False)
for mode in ("fused", "hide", "plain", "temporal"):
    s = Diffusion2D(8192, 8192, mode=mode, variant="tma", transport="peer")
    s.begin(); s.step(2); s.finish()
    if mode == "temporal":
        s.step_controlled_pair()                     # the variant with the in-kernel diagnostics
        s.finish()
    s.close()
s = Diffusion2D(8192, 8192, mode="fused", variant="direct")
s.begin(); s.step(1); s.finish()
T = s.T
nx, ny = T.shape
qx = torch.zeros(nx - 1, ny - 2, dtype=torch.float64, device=dev)
qy = torch.zeros(nx - 2, ny - 1, dtype=torch.float64, device=dev)
dT = torch.zeros(nx - 2, ny - 2, dtype=torch.float64, device=dev)
Cp = torch.ones(nx, ny, dtype=torch.float64, device=dev)
native.flux2d(qx, qy, T, 1.0, 10.0, 10.0)
native.residual2d(dT, qx, qy, Cp, 10.0, 10.0)
native.update2d(T, dT, 1e-6)
buf = torch.empty(nx, dtype=torch.float64, device=dev)
native.pack_plane(T, 1, 1, buf); native.unpack_plane(T, 1, 0, buf)
native.extract_inner(T)
from rocm_mpi_b200.parallel.gather import gather_interior_device
gather_interior_device(T)
native.reduce_max_sum(T.view(-1))
native.colormap(T, 0.0, 1.0, torch.from_numpy(inferno_lut()))
native.gaussian_ic_2d(T, 0.0, 0.0, 1e-3, 1e-3, 10.0, 10.0)
torch.cuda.synchronize()
igg.finalize_global_grid(finalize_dist=False)
# --- small-grid kernels
igg.init_global_grid(254, 254, 1, periodx=1, periody=1, quiet=True, init_dist=False)
for mode in ("persistent", "resident"):
    s = Diffusion2D(254, 254, mode=mode)
    s.begin(); s.step(200); s.finish()
torch.cuda.synchronize()
igg.finalize_global_grid(finalize_dist=False)
# --- 3-D
igg.init_global_grid(384, 384, 384, periodx=1, periody=1, periodz=1, quiet=True, init_dist=False)
for variant in ("tma", "direct"):
    s3 = Diffusion3D(384, 384, 384, mode="fused", variant=variant)
    s3.begin(); s3.step(2); s3.finish()
torch.cuda.synchronize()
igg.finalize_global_grid(finalize_dist=False)
print("all kernels launched; native launches:", native.launch_count())
