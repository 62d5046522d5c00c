"""Sustained loop of the two-step kernel WITH its in-kernel diagnostics but without any host copies (device parameters,
device diagnostics): separates the cost of the diagnostic arithmetic from the cost of the per-pass copies of the
end-to-end loop.  usage: python scripts/prof_diag_loop.py [passes]"""
import sys

import torch

sys.path.insert(0, ".")
from rocm_mpi_b200.ops import native  # noqa: E402

n = 16384
passes = int(sys.argv[1]) if len(sys.argv) > 1 else 30
T = torch.rand(n, n, dtype=torch.float64, device="cuda")
T2 = T.clone()
Cp = torch.ones_like(T)
params = torch.tensor([1.0, 1e-6, 100.0, 100.0] * 2, dtype=torch.float64, device="cuda")
diag = torch.zeros(4, dtype=torch.float64, device="cuda")
for with_diag in (False, True, False, True):
    for it in range(5):
        native.diffusion2d_step2(T2, T, Cp, 0, 0, 0, 0, dev_params=params, diag_out=diag if with_diag else None)
        T, T2 = T2, T
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(passes):
        native.diffusion2d_step2(T2, T, Cp, 0, 0, 0, 0, dev_params=params, diag_out=diag if with_diag else None)
        T, T2 = T2, T
    e1.record()
    torch.cuda.synchronize()
    print(f"diag={with_diag}: {e0.elapsed_time(e1) / passes / 2:.4f} ms/step")
