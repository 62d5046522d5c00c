#!/usr/bin/env python
"""Output path at benchmark size: ``gather`` of the halo-free field to rank 0 and the heat map.

The reference's perf run gathers 12286 x 12286 cells (1.2 GB) per rank when ``do_vis`` is set
(``scripts/diffusion_2D_perf.jl:21-22, 59-63``): ``Array(T[2:end-1,2:end-1])`` to the host on every rank, ``gather!``
over MPI, ``heatmap`` on rank 0.  Here:

* ``device``  -- ``gather_interior_device``: every rank stores its tile into the root GPU's buffer over NVLink
  (one kernel per rank), timed on the device (max over ranks) and by the host clock including the barrier;
* ``host``    -- the reference's structure (device -> pinned host on every rank, gloo gather), once, for comparison;
* ``heatmap`` -- colour map of the gathered field on the root GPU + PNG encoding.

Run:  ``torchrun --nproc-per-node 8 benchmarks/gather_bench.py --size 12288 [--out file.json]``
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.parallel.gather import gather, gather_interior_device, inner_to_host  # noqa: E402
from rocm_mpi_b200.utils.output import save_heatmap  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", dest="n", type=int, default=12288, help="local grid is size x size")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--host-size", dest="host_n", type=int, default=4096, help="local size of the host-staged comparison (0 = skip)")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    res = {"bench": "gather", "local": [a.n, a.n]}
    me, dims, nprocs, coords, _ = igg.init_global_grid(a.n, a.n, 1, quiet=True)
    g = igg.global_grid()
    dev = g.device
    gx = coords[0] * (a.n - 2) + torch.arange(a.n, device=dev, dtype=torch.float64)
    gy = coords[1] * (a.n - 2) + torch.arange(a.n, device=dev, dtype=torch.float64)
    T = torch.exp(-((gx[:, None] / (dims[0] * a.n) - 0.5) ** 2) - (gy[None, :] / (dims[1] * a.n) - 0.5) ** 2)
    tile_bytes = (a.n - 2) ** 2 * 8
    G = gather_interior_device(T)                       # allocates the root buffer, warms up
    ms_dev, s_wall = [], []
    for _ in range(a.reps):
        g.comm.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        G = gather_interior_device(T, sync=False)
        e1.record()
        torch.cuda.synchronize()
        g.comm.barrier()
        s_wall.append(g.comm.allreduce_max(time.perf_counter() - t0))
        ms_dev.append(g.comm.allreduce_max(e0.elapsed_time(e1)))
    ok = True
    if me == 0:
        # every tile where it belongs: f(global index) is smooth, so compare against the closed form
        X = torch.arange(G.shape[0], device=dev, dtype=torch.float64)
        Y = torch.arange(G.shape[1], device=dev, dtype=torch.float64)
        gxx = X + 1                                        # global index of gathered row r is r + 1
        gyy = Y + 1
        want = torch.exp(-((gxx[:, None] / (dims[0] * a.n) - 0.5) ** 2) - (gyy[None, :] / (dims[1] * a.n) - 0.5) ** 2)
        ok = bool(torch.equal(G, want))
        del want
    ok = g.comm.allreduce_max(0.0 if ok else 1.0) == 0.0
    total = tile_bytes * nprocs
    remote = tile_bytes * (nprocs - 1)
    best = min(ms_dev)
    res.update({"n_gpus": nprocs, "dims": list(dims[:2]), "ok": bool(ok), "tile_bytes": tile_bytes,
                "device_ms": [round(x, 3) for x in ms_dev], "device_wall_ms": [round(x * 1e3, 3) for x in s_wall],
                "into_root_GBps": round(remote / (best * 1e-3) / 1e9, 1) if nprocs > 1 else None,
                "gathered_GBps": round(total / (best * 1e-3) / 1e9, 1)})
    # heat map from the device-resident gathered field
    if me == 0:
        t0 = time.perf_counter()
        save_heatmap(os.path.join("gpurun_out", f"gather_bench_{nprocs}.png"), G, title=f"gathered {tuple(G.shape)}")
        res["heatmap_s"] = round(time.perf_counter() - t0, 3)
        t0 = time.perf_counter()
        Gh = G.cpu()
        res["root_d2h_s"] = round(time.perf_counter() - t0, 3)
        del Gh
    igg.finalize_global_grid(finalize_dist=False)
    # host-staged comparison (the reference's structure) at a smaller size
    if a.host_n:
        n = a.host_n
        igg.init_global_grid(n, n, 1, quiet=True, init_dist=False)
        g = igg.global_grid()
        Th = torch.rand(n, n, dtype=torch.float64, device=g.device)
        out = torch.zeros((n - 2) * dims[0], (n - 2) * dims[1], dtype=torch.float64) if me == 0 else None
        g.comm.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        gather(inner_to_host(Th), out)
        g.comm.barrier()
        t_host = g.comm.allreduce_max(time.perf_counter() - t0)
        Gd = gather_interior_device(Th)
        g.comm.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        Gd = gather_interior_device(Th)
        t_devp = g.comm.allreduce_max(time.perf_counter() - t0)
        same = bool(torch.equal(Gd.cpu(), out)) if me == 0 else True
        res["host_staged"] = {"local": [n, n], "host_gloo_s": round(t_host, 4), "device_s": round(t_devp, 5),
                              "identical": bool(g.comm.allreduce_max(0.0 if same else 1.0) == 0.0),
                              "GBps_host": round((n - 2) ** 2 * 8 * nprocs / t_host / 1e9, 2)}
        igg.finalize_global_grid(finalize_dist=False)
    if me == 0:
        print(json.dumps(res), flush=True)
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                json.dump(res, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
