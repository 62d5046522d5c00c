#!/usr/bin/env python
"""Timed CPU baseline: the C++ restatement of the reference's fused step (csrc/cpu_oracle.cpp) on this host.

The reference has no CPU code path; BASELINE.md asks for a timed CPU restatement of its maths as the
non-GPU comparison point.  Reports T_eff = 3*nx*ny*8 B / t_it like the GPU benchmark.

    OMP_NUM_THREADS=8 python benchmarks/cpu_baseline.py --n 4096 --nt 20
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rocm_mpi_b200.ops import native, reference as ref  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4096)
ap.add_argument("--nt", type=int, default=20)
a = ap.parse_args()
n = a.n
dx = dy = 10.0 / n
dt = ref.time_step(dx, dy)
xs = torch.arange(n, dtype=torch.float64) * dx
T = ref.gaussian_ic(xs, xs, dx, dy, 10.0, 10.0)
T2 = T.clone()
Cp = torch.ones_like(T)
native.cpu_diffusion2d_run(T, T2, Cp, 1.0, dt, 1 / dx, 1 / dy, 2)          # warm-up
t0 = time.perf_counter()
native.cpu_diffusion2d_run(T, T2, Cp, 1.0, dt, 1 / dx, 1 / dy, a.nt)
t = (time.perf_counter() - t0) / a.nt
print(json.dumps({"bench": "cpu_baseline", "n": n, "steps": a.nt, "threads": os.environ.get("OMP_NUM_THREADS", "default"),
                  "ms_per_step": round(t * 1e3, 3), "T_eff_GBps": round(3 * n * n * 8 / t / 1e9, 2),
                  "max_T": float(T.max())}))
