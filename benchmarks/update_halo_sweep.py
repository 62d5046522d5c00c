#!/usr/bin/env python
"""``update_halo`` API sweep: the public call, every transport, data verified on every rank.

The reference calls ``update_halo!(T)`` once per step (``scripts/diffusion_2D_ap.jl:42``, ``diffusion_2D_perf.jl:51``);
``benchmarks/halo_bw_sweep.py`` measures a contiguous ring copy, this one measures the API itself on a 2-D
field: x-planes are strided (one element per row), y-planes contiguous, corners travel to the diagonal
neighbours.  For local sizes 128^2 ... 16384^2 and transports ``peer`` (one kernel over NVLink peer memory),
``nccl`` (pack -> grouped NCCL send/recv -> unpack) and ``host`` (pinned staging + gloo, small sizes only):

* the field is ``f(global index)`` with every halo cell poisoned; after ONE ``update_halo`` every halo cell that has
  a neighbour must hold ``f`` of its global index and every other cell must be unchanged (``ok``);
* then ``iters`` back-to-back calls are timed on the device (CUDA events, max over ranks).

Run:  ``torchrun --nproc-per-node N benchmarks/update_halo_sweep.py [--out file.json]``
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.parallel.fabric import fabric_zeros  # noqa: E402
from rocm_mpi_b200.parallel.topology import PROC_NULL  # noqa: E402


def expected_field(g, n, dev):
    """f(global index) on the local block, and the same with the halo cells that have NO neighbour poisoned."""
    gx = g.coords[0] * (n - 2) + torch.arange(n, device=dev, dtype=torch.float64)
    gy = g.coords[1] * (n - 2) + torch.arange(n, device=dev, dtype=torch.float64)
    return gx[:, None] * 65536.0 + gy[None, :]


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", type=int, nargs="+", default=[128, 512, 2048, 8192, 16384])
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rows = []
    first = True
    for n in a.sizes:
        me, dims, nprocs, coords, _ = igg.init_global_grid(n, n, 1, quiet=True, init_dist=first)
        first = False
        g = igg.global_grid()
        dev = g.device
        want = expected_field(g, n, dev)
        has = [[g.neighbors[s][d] != PROC_NULL for d in range(2)] for s in range(2)]      # [side][dim]
        res = {"local": [n, n], "dims": list(dims[:2]), "halo_bytes_per_rank": 8 * 2 * (n * (has[0][0] + has[1][0]) +
                                                                                       n * (has[0][1] + has[1][1])) // 2}
        for tr in ("peer", "nccl", "host"):
            if tr == "host" and n > 2048:
                continue
            T = fabric_zeros((n, n)) if tr == "peer" else torch.zeros(n, n, dtype=torch.float64, device=dev)
            T.copy_(want)
            T[0, :] = -1.0
            T[-1, :] = -1.0
            T[:, 0] = -1.0
            T[:, -1] = -1.0
            ref = want.clone()                  # cells on a side without a neighbour keep the poison
            if not has[0][0]:
                ref[0, :] = -1.0
            if not has[1][0]:
                ref[-1, :] = -1.0
            if not has[0][1]:
                ref[:, 0] = -1.0
            if not has[1][1]:
                ref[:, -1] = -1.0
            igg.update_halo(T, transport=tr)
            torch.cuda.synchronize()
            ok = bool(torch.equal(T, ref))
            res[tr + "_ok"] = bool(g.comm.allreduce_max(0.0 if ok else 1.0) == 0.0)
            iters = 200 if n <= 2048 else 50
            if tr == "host":
                iters = 20
            for _ in range(5):
                igg.update_halo(T, transport=tr)
            g.comm.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                igg.update_halo(T, transport=tr)
            e1.record()
            torch.cuda.synchronize()
            res[tr + "_us"] = round(g.comm.allreduce_max(e0.elapsed_time(e1)) / iters * 1e3, 2)
            del T
        rows.append(res)
        if me == 0:
            print(json.dumps(res), flush=True)
        igg.finalize_global_grid(finalize_dist=False)
    if me == 0 and a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump({"bench": "update_halo_sweep", "n_gpus": nprocs, "rows": rows}, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
