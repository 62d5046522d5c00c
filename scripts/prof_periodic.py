"""1-GPU stand-in for neighbours: a periodic dimension makes the rank its own neighbour, so the mid-line kernel, the
flag waits and the edge pushes all run (against local memory).  usage: prof_periodic.py PX PY [mode] [steps]"""
import sys

import torch

sys.path.insert(0, ".")
import rocm_mpi_b200 as igg
from rocm_mpi_b200.models.solver import Diffusion2D

px, py = int(sys.argv[1]), int(sys.argv[2])
mode = sys.argv[3] if len(sys.argv) > 3 else "temporal"
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 40
n = 16384
igg.init_global_grid(n, n, 1, periodx=px, periody=py, quiet=True, init_dist=False)
s = Diffusion2D(n, n, mode=mode)
s.begin()
s.step(10)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
s.step(steps)
e1.record()
torch.cuda.synchronize()
s.finish()
print(f"periods ({px},{py}) mode={mode}: {e0.elapsed_time(e1) / steps:.4f} ms/step")
igg.finalize_global_grid(finalize_dist=False)
