#!/usr/bin/env python
"""Headline benchmark: ``diffusion_2D_perf`` weak scaling, 16384 x 16384 fp64 cells per GPU.

Metric (reference ``scripts/diffusion_2D_perf.jl:55-58``): ``T_eff = (2+1)*nx*ny*8 B / t_it`` per GPU,
reported as the whole-job aggregate (sum over GPUs).  One rank per GPU; for N > 1 launch with
``python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N``.

Two timed regions, both bracketed by barrier + ``torch.cuda.synchronize()``, max over ranks:

* ``value``  -- K steps of the flagship fused kernel (stencil + in-kernel halo push), CUDA events;
* ``e2e``    -- K steps through the public ``Diffusion2D.step_controlled`` API: per step the parameters
  travel host(pinned) -> device, the step runs, ``[max, sum]`` of the new field travel device -> host.

``--impl reference``: the reference is Julia + AMDGPU.jl + ROCm-aware MPI; it cannot be installed or run
here, so that arm reports ``unavailable``.  ``--impl nccl`` runs the same-box baseline
BASELINE.md defines (separate stencil kernel + pack / NCCL send-recv / unpack).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl", "nccl_tma", "plain_peer", "hide", "direct", "temporal"])
    p.add_argument("--nx", type=int, default=16384)
    p.add_argument("--ny", type=int, default=16384)
    p.add_argument("--b-width", type=int, nargs=2, default=[16, 4])
    p.add_argument("--no-e2e", action="store_true")
    return p.parse_args()


class ClockSampler:
    """``nvidia-smi`` clocks / throttle reasons sampled during the timed region (rank 0 only)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
                pw.append(float(c[3]))
            except ValueError:
                continue
            for k, n in enumerate(names):
                if c[5 + k].lower().startswith("active"):
                    reasons.add(n)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def main() -> int:
    a = parse_args()
    if a.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:      # under torchrun: one line, from rank 0
            return 0
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is Julia scripts on AMDGPU.jl/ROCm-aware MPI (no Python package, no CUDA "
                          "path); julia/MPI/ROCm are not in this image and cannot be installed offline"}))
        return 0

    import torch
    import rocm_mpi_b200 as igg
    from rocm_mpi_b200.models.solver import Diffusion2D
    from rocm_mpi_b200.ops import native

    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device"}))
        return 1
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and a.gpus > 1:
        print(f"bench.py: --gpus {a.gpus} needs torchrun with {a.gpus} ranks", file=sys.stderr)
        return 2
    W = max(3, a.warmup)
    K = max(1, a.steps)
    nx, ny = a.nx, a.ny
    me, dims, nprocs, coords, _ = igg.init_global_grid(nx, ny, 1, quiet=True)
    g = igg.global_grid()
    mode, variant, transport = {"ours": ("fused", "tma", "peer"), "hide": ("hide", "tma", "peer"),
                                "direct": ("fused", "direct", "peer"),
                                "temporal": ("temporal", "tma", "peer"),   # experimental: two steps per pass
                                "plain_peer": ("plain", "tma", "peer"), "nccl_tma": ("plain", "tma", "nccl"),
                                "nccl": ("plain", "direct", "nccl")}[a.impl]
    s = Diffusion2D(nx, ny, mode=mode, variant=variant, transport=transport, b_width=a.b_width)
    A_eff = igg.a_eff((nx, ny))                      # GB per step per GPU

    # ---- device-timed region -------------------------------------------------------------------
    s.begin()
    s.step(W)
    g.comm.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(g.device.index) if me == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = native.launch_count()
    e0.record()
    s.step(K)
    e1.record()
    torch.cuda.synchronize()
    n1 = native.launch_count()
    g.comm.barrier()
    ms = g.comm.allreduce_max(e0.elapsed_time(e1)) / K
    s.finish()
    t_eff = A_eff / (ms * 1e-3)

    # ---- end-to-end region ---------------------------------------------------------------------
    e2e = None
    if not a.no_e2e:
        s.begin()
        pair = a.impl == "temporal"          # experimental: the host hands over / reads back two steps at a time
        for _ in range(W):
            s.step_controlled_pair() if pair else s.step_controlled()
        g.comm.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ee0.record()
        mx = 0.0
        if pair:
            for _ in range(K // 2):
                _, (mx, sm) = s.step_controlled_pair()
            for _ in range(K % 2):
                mx, sm = s.step_controlled()
        else:
            for _ in range(K):
                mx, sm = s.step_controlled()
        ee1.record()
        torch.cuda.synchronize()
        dt_wall = time.perf_counter() - t0
        g.comm.barrier()
        # device clock (CUDA events around the whole loop: includes every gap the host causes), max over ranks;
        # the host wall clock of the same region is reported next to it
        ms_e2e = g.comm.allreduce_max(ee0.elapsed_time(ee1)) / K
        ms_e2e_wall = g.comm.allreduce_max(dt_wall * 1e3) / K
        s.finish()
        e2e = {"value": round(nprocs * A_eff / (ms_e2e * 1e-3), 2), "unit": "GB/s", "ms_per_step": round(ms_e2e, 5),
               "ms_per_step_host_wall": round(ms_e2e_wall, 5),
               "h2d_bytes_per_step": Diffusion2D.h2d_bytes_per_step,
               "d2h_bytes_per_step": Diffusion2D.d2h_bytes_per_step, "last_max_T": mx,
               "api": "rocm_mpi_b200.models.solver.Diffusion2D." + ("step_controlled_pair" if pair else "step_controlled")}
    clocks = sampler.stop() if sampler is not None else None

    if me == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        out = {
            "metric": "T_eff effective memory throughput of diffusion_2D_perf (3*nx*ny*8 B / t_it), aggregate over GPUs",
            "value": round(nprocs * t_eff, 2), "unit": "GB/s", "n_gpus": nprocs, "steps": K, "warmup": W,
            "ms_per_step": round(ms, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp64", "data": "synthetic (Gaussian initial condition, Cp = 1; no datasets exist for this app)",
            "impl": a.impl,
            "config": {"model": "diffusion_2D_perf", "local_grid": [nx, ny], "global_grid": [igg.nx_g(), igg.ny_g()],
                       "global_batch": nprocs * nx * ny, "seq_len": None,
                       "parallelism": f"cartesian {dims[0]}x{dims[1]} domain decomposition, halo width 1",
                       "step": f"mode={mode} kernel={variant} transport={transport} b_width={list(a.b_width)}",
                       "l2": "inputs exceed L2 (3 x %.2f GiB per GPU per step vs 126 MB)" % (nx * ny * 8 / 2**30)},
            "per_gpu_T_eff": round(t_eff, 2),
            "roofline": {"hbm_gbs": hbm, "source": "MEASURED_PEAKS.json" if peaks else "fallback",
                         "fraction_of_hbm_peak": round(t_eff / hbm, 4)},
            "e2e": e2e, "gpu_launches": int(n1 - n0), "clocks": clocks,
        }
        print(json.dumps(out), flush=True)
    igg.finalize_global_grid()
    return 0


if __name__ == "__main__":
    sys.exit(main())
