import sys, torch
sys.path.insert(0, '.')
from rocm_mpi_b200.ops import native, reference as ref
variant = sys.argv[1] if len(sys.argv) > 1 else "tma"
nx, ny = 64, 256
T = torch.rand(nx, ny, dtype=torch.float64, device="cuda")
Cp = torch.ones_like(T)
T2 = T.clone()
native.diffusion2d_step(T2, T, Cp, 1.0, 1e-3, 2.0, 3.0, variant=variant)
torch.cuda.synchronize()
want = T.cpu().clone()
ref.perf_step(want, T.cpu(), Cp.cpu(), 1.0, 1e-3, 2.0, 3.0)
print(variant, "ok, err", float((T2.cpu() - want).abs().max()))
