"""Short single-GPU program for ncu: a few steps of one stencil kernel at the headline size.

usage: python scripts/prof_step.py [tma|direct|temporal] [nx] [ny] [steps] [mode]
(Profiling counterpart of the reference's diffusion_2D_perf_hide_prof.jl, which wraps its loop in
Julia's host-side Profile.@profile -- here the evidence is an ncu capture of the kernel itself.)
"""
import sys

import torch

sys.path.insert(0, ".")
from rocm_mpi_b200.ops import native  # noqa: E402

variant = sys.argv[1] if len(sys.argv) > 1 else "tma"
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
ny = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 6
mode = int(sys.argv[5]) if len(sys.argv) > 5 else 0
T = torch.rand(nx, ny, dtype=torch.float64, device="cuda")
T2 = T.clone()
Cp = torch.ones_like(T)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(steps):
    if it == 2:
        e0.record()
    if variant == "temporal":       # one launch = two steps
        native.diffusion2d_step2(T2, T, Cp, 1.0, 1e-6, 100.0, 100.0)
    else:
        native.diffusion2d_step(T2, T, Cp, 1.0, 1e-6, 100.0, 100.0, mode=mode, b_width=(16, 4), variant=variant)
    T, T2 = T2, T
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / max(1, steps - 2) / (2 if variant == "temporal" else 1)
print(f"{variant} {nx}x{ny} mode={mode}: {ms:.4f} ms/step, T_eff = {3 * nx * ny * 8 / ms / 1e6:.1f} GB/s")
