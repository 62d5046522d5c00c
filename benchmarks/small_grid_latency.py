#!/usr/bin/env python
"""Per-step latency on small local grids (BASELINE.json config 2: the kp / Temp_4_252_252 regime).

At 128 x 128 (or 65 x 128) cells per GPU a time step is launch + synchronisation latency, not bandwidth.
Compares, device-timed and max over ranks:

* ``kp-nccl``     three kernels + pack / NCCL p2p / unpack per step (the reference's structure, kp.jl:87-92)
* ``fused``       one fused kernel per step with in-kernel halo push
* ``persistent``  ONE cooperative launch for all steps (in-kernel grid barrier + halo push)
* ``resident``    ONE cluster launch, field resident in distributed shared memory, cluster barriers

Run:  ``torchrun --nproc-per-node 8 benchmarks/small_grid_latency.py --nx 65 --ny 128 --dimx 4 --dimy 2``
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.models.solver import Diffusion2D  # noqa: E402
from rocm_mpi_b200.ops import native  # noqa: E402


def timed(g, fn) -> float:
    g.comm.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    return g.comm.allreduce_max(e0.elapsed_time(e1))


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=128)
    ap.add_argument("--ny", type=int, default=128)
    ap.add_argument("--dimx", type=int, default=0)
    ap.add_argument("--dimy", type=int, default=0)
    ap.add_argument("--nt", type=int, default=1000)
    a = ap.parse_args()
    me, dims, nprocs, _, _ = igg.init_global_grid(a.nx, a.ny, 1, dimx=a.dimx, dimy=a.dimy, quiet=True)
    g = igg.global_grid()
    out = {"bench": "small_grid_latency", "n_gpus": nprocs, "local": [a.nx, a.ny], "dims": list(dims[:2]), "nt": a.nt}
    variant = "tma" if a.ny % 2 == 0 else "direct"
    for name, kw in (("plain-nccl", dict(mode="plain", variant=variant, transport="nccl")),
                     ("fused", dict(mode="fused", variant=variant)),
                     ("persistent", dict(mode="persistent")), ("resident", dict(mode="resident"))):
        s = Diffusion2D(a.nx, a.ny, **kw)
        s.begin()
        s.step(20)
        n0 = native.launch_count()
        ms = timed(g, lambda: s.step(a.nt))
        s.finish()
        mx, _ = s.max_sum()
        out[name] = {"us_per_step": round(ms * 1e3 / a.nt, 3), "native_launches": native.launch_count() - n0,
                     "max_T_local": mx}
    if me == 0:
        print(json.dumps(out), flush=True)
    igg.finalize_global_grid()
    return 0


if __name__ == "__main__":
    sys.exit(main())
