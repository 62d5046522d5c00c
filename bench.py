#!/usr/bin/env python
"""Headline benchmark: ``diffusion_2D_perf`` weak scaling, 16384 x 16384 fp64 cells per GPU.

Metric (reference ``scripts/diffusion_2D_perf.jl:55-58``): ``T_eff = (2+1)*nx*ny*8 B / t_it`` per GPU,
reported as the whole-job aggregate (sum over GPUs).  One rank per GPU; for N > 1 launch with
``python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N``.

Two timed regions, both bracketed by barrier + ``torch.cuda.synchronize()``, max over ranks:

* ``value``  -- K steps of the flagship kernel (stencil + in-kernel halo push), CUDA events;
* ``e2e``    -- K steps through the public ``Diffusion2D.step_controlled`` API: per step the parameters
  travel host(pinned) -> device, the step runs, ``[max, sum]`` of the new field travel device -> host
  (pipelined: step k+1 is queued before the host waits for step k's diagnostic).

Outside both timed regions, on the same ranks (``--no-verify`` / ``--no-baseline`` skip them):

* ``verify``       -- the benchmarked mode against the serial baseline structure (direct-load kernel, swap,
  ``update_halo`` over NCCL) from the same initial condition at the benchmark size, every cell incl. halos;
  and a small 254 x 254 global problem (the reference's published field, ``docs/Temp_4_252_252.png``) run for
  200 steps on the same rank layout against the single-domain CPU oracle.  A mismatch makes the run fail.
* ``vs_baseline``  -- the same-box baseline BASELINE.json defines (separate stencil kernel + pack / NCCL
  send-recv / unpack), timed in this process with the same K/W: ``value / value_of(nccl arm)``.  It is NOT a
  ratio against the Julia reference (that arm is unavailable here).

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
    p.add_argument("--dims", type=int, nargs=2, default=[0, 0], help="force the process grid (default: balanced, 8 -> 4x2)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-verify", action="store_true")
    p.add_argument("--no-baseline", action="store_true")
    p.add_argument("--kernel", default=os.environ.get("IGG_BENCH_KERNEL", "auto"), choices=["auto", "fused", "temporal"],
                   help="what --impl ours runs: one step per pass (fused) or two steps per pass (temporal)")
    return p.parse_args()


class ClockSampler:
    """``nvidia-smi`` clocks / throttle reasons sampled during the timed region (rank 0 only)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        # nvidia-smi needs ~0.1-0.3 s to start (longer on an 8-GPU box), which can exceed a short timed region: an
        # NVML thread in this process samples the same quantities every 5 ms from the first moment and is used when
        # nvidia-smi delivered fewer than three samples
        self._nvml_rows, self._nvml_stop, self._nvml_thread = [], False, None
        try:
            import threading
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons

            def loop():
                while not self._nvml_stop:
                    try:
                        self._nvml_rows.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                                                pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM),
                                                pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
                                                get_reasons(h)))
                    except Exception:
                        pass
                    time.sleep(0.005)
            self._nvml = pynvml
            self._nvml_thread = threading.Thread(target=loop, daemon=True)
            self._nvml_thread.start()
        except Exception:
            self._nvml_thread = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self) -> dict:
        if self.p is None and not self._nvml_rows:
            self._nvml_stop = True
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        if self.p is None:
            self.p = subprocess.Popen(["true"])
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
        source = "nvidia-smi"
        self._nvml_stop = True
        if self._nvml_thread is not None:
            self._nvml_thread.join(timeout=1)
        if len(sm) < 3 and self._nvml_rows:
            source = "nvml"
            nv = self._nvml
            bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown,
                    "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown,
                    "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
            sm = [float(r[0]) for r in self._nvml_rows]
            mx = [float(r[1]) for r in self._nvml_rows]
            pw = [float(r[2]) for r in self._nvml_rows]
            reasons = {n for r in self._nvml_rows for n, b in bits.items() if r[3] & b}
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "source": source,
                "reasons": sorted(reasons)}


class NvlinkCounters:
    """NVLink payload counters of one GPU (NVML field values, all links summed; KiB resolution).  The bytes that
    cross NVLink during a timed region, next to the algorithmic halo bytes, are the NVLink side of the fused
    kernel's roofline; ``None`` where NVML does not expose the counters."""

    def __init__(self, gpu_index: int) -> None:
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            tx, rx = pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX
            # the all-links scope first; some drivers only answer per link
            for ids in ([(tx, 0xFFFFFFFF), (rx, 0xFFFFFFFF)],
                        [(tx, k) for k in range(18)] + [(rx, k) for k in range(18)]):
                self.ids = ids
                if self.read() is not None:
                    break
            else:
                self.h = None
        except Exception:
            self.h = None

    def read(self):
        if self.h is None:
            return None
        try:
            v = self.nv.nvmlDeviceGetFieldValues(self.h, self.ids)
            half = len(v) // 2
            out = [0, 0]
            for k, x in enumerate(v):
                if x.nvmlReturn != 0:
                    return None
                out[0 if k < half else 1] += int(x.value.ullVal) * 1024
            return out
        except Exception:
            return None


IMPLS = {"ours": ("fused", "tma", "peer"), "hide": ("hide", "tma", "peer"), "direct": ("fused", "direct", "peer"),
         "temporal": ("temporal", "tma", "peer"), "plain_peer": ("plain", "tma", "peer"),
         "nccl_tma": ("plain", "tma", "nccl"), "nccl": ("plain", "direct", "nccl")}
DEFAULT_OURS = "temporal"        # what `--kernel auto` resolves to (the faster of the two on B200, see profiles/)


def time_steps(s, g, W, K, native, torch, nvl=None):
    """W untimed + K timed steps of solver ``s``: (ms per step, max over ranks; native launches in the region;
    NVLink payload bytes [tx, rx] of this rank's GPU during the region, if ``nvl`` can read them)."""
    s.begin()
    s.step(W)
    g.comm.barrier()
    torch.cuda.synchronize()
    c0 = nvl.read() if nvl is not None else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = native.launch_count()
    e0.record()
    s.step(K)
    e1.record()
    torch.cuda.synchronize()
    n1 = native.launch_count()
    c1 = nvl.read() if nvl is not None else None
    g.comm.barrier()
    ms = g.comm.allreduce_max(e0.elapsed_time(e1)) / K
    s.finish()
    link = [c1[0] - c0[0], c1[1] - c0[1]] if (c0 is not None and c1 is not None) else None
    return ms, int(n1 - n0), link


def time_e2e(s, g, W, K, pair, torch):
    """K end-to-end steps through the public control API (pipelined), device clock + host wall clock."""
    s.begin()
    for _ in range(W):
        s.step_controlled_pair(pipelined=True) if pair else s.step_controlled(pipelined=True)
    s.drain_controlled()
    g.comm.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ee0.record()
    seen = 0
    last = None
    if pair:
        for _ in range(K // 2):
            r = s.step_controlled_pair(pipelined=True)
            if r is not None:
                last, seen = r[1], seen + 2
        for _ in range(K % 2):
            r = s.step_controlled(pipelined=True)
            if r is not None:
                last, seen = (r[1] if isinstance(r[0], tuple) else r), seen + (2 if isinstance(r[0], tuple) else 1)
    else:
        for _ in range(K):
            r = s.step_controlled(pipelined=True)
            if r is not None:
                last, seen = r, seen + 1
    for r in s.drain_controlled():                      # the last diagnostics: still inside the timed region
        if isinstance(r[0], tuple):
            last, seen = r[1], seen + 2
        else:
            last, seen = r, seen + 1
    ee1.record()
    torch.cuda.synchronize()
    dt_wall = time.perf_counter() - t0
    g.comm.barrier()
    ms = g.comm.allreduce_max(ee0.elapsed_time(ee1)) / K
    ms_wall = g.comm.allreduce_max(dt_wall * 1e3) / K
    s.finish()
    assert seen == K, (seen, K)
    return ms, ms_wall, last


def verify(igg, Diffusion2D, torch, nx, ny, mode, dims, nprocs):
    """Correctness of the benchmarked mode on THIS rank layout (outside the timed regions)."""
    from rocm_mpi_b200.ops import reference as ref
    out = {"flag_timeouts": 0}
    g = igg.global_grid()
    # (a) benchmark size: `mode` vs the reference's serial structure with the direct-load kernel and NCCL
    nsteps = 7                                  # odd: the two-step mode ends with a one-step launch
    a = Diffusion2D(nx, ny, mode=mode, variant="tma")
    b = Diffusion2D(nx, ny, mode="plain", variant="direct", transport="nccl" if nprocs > 1 else "peer")
    try:
        for s in (a, b):
            s.begin()
            s.step(nsteps)
            s.finish()
        torch.cuda.synchronize()
        err = float((a.T - b.T).abs().max())
    except RuntimeError as e:
        if "timed out" not in str(e):
            raise
        out["flag_timeouts"] += 1
        err = float("inf")
    out["max_abs_err_vs_nccl"] = g.comm.allreduce_max(err)
    out["vs_nccl"] = f"{mode} vs plain+direct+nccl, local {nx}x{ny}, {nsteps} steps, all cells incl. halos"
    a.close()
    b.close()
    igg.finalize_global_grid(finalize_dist=False)
    # (b) the reference's published problem (254 x 254 global, docs/Temp_4_252_252.png) on this layout vs the
    #     single-domain fp64 CPU oracle, 200 steps, gathered halo-free field
    if 252 % dims[0] or 252 % dims[1]:
        out["max_abs_err_vs_oracle"] = None
        out["vs_oracle"] = f"skipped: 252 is not divisible by dims {dims}"
    else:
        lx_, ly_ = 252 // dims[0] + 2, 252 // dims[1] + 2
        small_mode = mode if ly_ % 2 == 0 else "fused"
        igg.init_global_grid(lx_, ly_, 1, dimx=dims[0], dimy=dims[1], quiet=True, init_dist=False)
        g = igg.global_grid()
        s = Diffusion2D(lx_, ly_, mode=small_mode, variant="tma" if ly_ % 2 == 0 else "direct")
        err = 0.0
        try:
            s.begin()
            s.step(200)
            s.finish()
            Tv = igg.gather_interior(s.T)
            if g.me == 0:
                want, _ = ref.oracle_2d(igg.nx_g(), igg.ny_g(), 200, form="perf")
                err = float((Tv - want[1:-1, 1:-1]).abs().max())
                out["max_T_v"] = float(Tv.max())
        except RuntimeError as e:
            if "timed out" not in str(e):
                raise
            out["flag_timeouts"] += 1
            err = float("inf")
        out["max_abs_err_vs_oracle"] = g.comm.allreduce_max(err)
        out["vs_oracle"] = (f"{small_mode}, local {lx_}x{ly_} on {dims[0]}x{dims[1]} (global 254x254), 200 steps, "
                            "gathered field vs single-domain CPU oracle")
        s.close()
        igg.finalize_global_grid(finalize_dist=False)
    out["flag_timeouts"] = int(g.comm.allreduce_max(float(out["flag_timeouts"])))
    out["tolerance"] = 1e-12
    out["ok"] = bool(out["flag_timeouts"] == 0 and out["max_abs_err_vs_nccl"] <= 1e-12 and
                     (out["max_abs_err_vs_oracle"] is None or out["max_abs_err_vs_oracle"] <= 1e-12))
    return out


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
    me, dims, nprocs, coords, _ = igg.init_global_grid(nx, ny, 1, dimx=a.dims[0], dimy=a.dims[1], quiet=True)
    g = igg.global_grid()
    mode, variant, transport = IMPLS[a.impl]
    if a.impl == "ours":
        kern = DEFAULT_OURS if a.kernel == "auto" else a.kernel
        if kern == "temporal" and ny % 2 == 0:
            mode = "temporal"
    # ---- verification FIRST (own grids; leaves no grid behind): a kernel that does not reproduce the field is not
    #      timed.  If the two-step kernel should fail on this rank layout, the one-step fused kernel is verified and
    #      timed instead and the line says so (`fallback_from`).
    ver, fallback = None, None
    if not a.no_verify and mode in ("fused", "temporal", "hide"):
        try:
            ver = verify(igg, Diffusion2D, torch, nx, ny, mode, dims, nprocs)
        except Exception as ex:                       # a failing check must not cost the run its line
            if mode != "temporal":
                raise
            print(f"bench.py: verification of the two-step kernel raised {type(ex).__name__}: {ex}", file=sys.stderr)
            ver = {"ok": False, "exception": f"{type(ex).__name__}: {ex}"[:300]}
            if igg.grid_is_initialized():
                try:
                    igg.finalize_global_grid(finalize_dist=False)
                except Exception:
                    pass
        if not ver["ok"] and mode == "temporal":
            fallback = {"mode": "temporal", "verify": ver}
            mode = "fused"
            igg.init_global_grid(nx, ny, 1, dimx=dims[0], dimy=dims[1], quiet=True, init_dist=False)
            ver = verify(igg, Diffusion2D, torch, nx, ny, mode, dims, nprocs)
        igg.init_global_grid(nx, ny, 1, dimx=dims[0], dimy=dims[1], quiet=True, init_dist=False)
        g = igg.global_grid()
    steps_per_pass = 2 if mode == "temporal" else 1
    s = Diffusion2D(nx, ny, mode=mode, variant=variant, transport=transport, b_width=a.b_width)
    A_eff = igg.a_eff((nx, ny))                      # GB per step per GPU

    # ---- device-timed region -------------------------------------------------------------------
    sampler = ClockSampler(g.device.index) if me == 0 else None
    nvl = NvlinkCounters(g.device.index) if me == 0 else None
    ms, launches, link_bytes = time_steps(s, g, W, K, native, torch, nvl)
    t_eff = A_eff / (ms * 1e-3)

    # ---- end-to-end region ---------------------------------------------------------------------
    e2e = None
    if not a.no_e2e and mode in ("fused", "temporal") and variant == "tma":
        pair = mode == "temporal"
        ms_e2e, ms_e2e_wall, last = time_e2e(s, g, W, K, pair, torch)
        e2e = {"value": round(nprocs * A_eff / (ms_e2e * 1e-3), 2), "unit": "GB/s", "ms_per_step": round(ms_e2e, 5),
               "ms_per_step_host_wall": round(ms_e2e_wall, 5),
               "h2d_bytes_per_step": Diffusion2D.h2d_bytes_per_step,
               "d2h_bytes_per_step": Diffusion2D.d2h_bytes_per_step, "last_max_T": last[0] if last else None,
               "pipeline": "step k+1 is queued before the host waits for step k's diagnostic; every step's copies "
                           "are inside the timed region",
               "api": "rocm_mpi_b200.models.solver.Diffusion2D." +
                      ("step_controlled_pair(pipelined=True)" if pair else "step_controlled(pipelined=True)")}
    clocks = sampler.stop() if sampler is not None else None
    s.close()

    # ---- same-box baseline arms (BASELINE.json: separate stencil kernel + pack / NCCL p2p / unpack) ----------
    base = {}
    if not a.no_baseline and a.impl not in ("nccl", "nccl_tma"):
        # nccl / nccl_tma: the baseline BASELINE.json defines; plain_peer: the SAME serial structure (kernel, swap,
        # update_halo) with this framework's one-launch peer-memory update_halo -- the public API against NCCL
        for name in ("nccl", "nccl_tma", "plain_peer"):
            bm, bv, bt = IMPLS[name]
            sb = Diffusion2D(nx, ny, mode=bm, variant=bv, transport=bt if nprocs > 1 else "peer", b_width=a.b_width)
            bms, _, _ = time_steps(sb, g, W, K, native, torch)
            sb.close()
            base[name] = {"ms_per_step": round(bms, 5), "value": round(nprocs * A_eff / (bms * 1e-3), 2)}

    igg.finalize_global_grid(finalize_dist=False)

    rc = 0
    if me == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        nvl_gbs = 770.0                                # measured peer copy, GB/s per direction per GPU (profiling recipe)
        # bytes this rank sends per step: one edge line per existing face neighbour (+ 8 B per corner); the
        # two-step mode sends a mid line and an edge line per PAIR of steps -- the same per step
        # face neighbours of the busiest rank per dimension: none (dims 1), one (dims 2), two (dims > 2)
        n_face = [0 if dims[d] == 1 else (1 if dims[d] == 2 else 2) for d in range(2)]
        halo_bytes = 8 * (n_face[0] * ny + n_face[1] * nx)
        hbm_term = A_eff / steps_per_pass / hbm * 1e3 if steps_per_pass > 1 else A_eff / hbm * 1e3
        dram_bytes_step = A_eff / steps_per_pass if steps_per_pass > 1 else A_eff
        nvl_term = halo_bytes / 1e9 / nvl_gbs * 1e3
        roof_ms = max(hbm_term, nvl_term)
        value = round(nprocs * t_eff, 2)
        out = {
            "metric": "T_eff effective memory throughput of diffusion_2D_perf (3*nx*ny*8 B / t_it), aggregate over GPUs",
            "value": value, "unit": "GB/s", "n_gpus": nprocs, "steps": K, "warmup": W,
            "ms_per_step": round(ms, 5), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": round(value / base["nccl"]["value"], 4) if base else None,
            "dtype": "fp64", "data": "synthetic (Gaussian initial condition, Cp = 1; no datasets exist for this app)",
            "impl": a.impl,
            "config": {"model": "diffusion_2D_perf", "local_grid": [nx, ny], "global_grid": [dims[0] * (nx - 2) + 2,
                                                                                              dims[1] * (ny - 2) + 2],
                       "global_batch": nprocs * nx * ny, "seq_len": None,
                       "parallelism": f"cartesian {dims[0]}x{dims[1]} domain decomposition, halo width 1",
                       "step": f"mode={mode} kernel={variant} transport={transport} b_width={list(a.b_width)}",
                       "steps_per_pass": steps_per_pass,
                       "l2": "inputs exceed L2 (3 x %.2f GiB per GPU per step vs 126 MB)" % (nx * ny * 8 / 2**30)},
            "per_gpu_T_eff": round(t_eff, 2),
            "roofline": {"hbm_gbs": hbm, "nvlink_gbs_per_dir": nvl_gbs,
                         "source": "MEASURED_PEAKS.json" if peaks else "fallback",
                         "fraction_of_hbm_peak": round(t_eff / hbm, 4),
                         "dram_bytes_per_step_algorithmic": int(dram_bytes_step * 1e9),
                         "hbm_term_ms": round(hbm_term, 5), "halo_bytes_per_step": halo_bytes,
                         "nvlink_term_ms": round(nvl_term, 6), "bound": "hbm" if hbm_term >= nvl_term else "nvlink",
                         "nvlink_measured_bytes_per_step_rank0": ({"tx": link_bytes[0] // K, "rx": link_bytes[1] // K,
                                                                   "source": "NVML NVLINK_THROUGHPUT_DATA_TX/RX over the "
                                                                             "device-timed region (payload, KiB counters)"}
                                                                  if link_bytes is not None else None),
                         "fraction_of_roofline": round(roof_ms / ms, 4),
                         "note": ("T_eff keeps the reference's definition (3 arrays of the grid per STEP, "
                                  "scripts/diffusion_2D_perf.jl:55-57); a two-step pass moves those bytes once per TWO "
                                  "steps, so T_eff can exceed the copy bandwidth while the DRAM traffic does not")
                         if steps_per_pass > 1 else None},
            "vs_baseline_arms": ({"kind": "same-box baseline defined by BASELINE.json (separate stencil kernel + pack / "
                                          "NCCL p2p / unpack), timed in this process; NOT the Julia reference.  plain_peer = the same serial "
                                          "structure with this framework's one-launch peer-memory update_halo",
                                  **{k: {"ratio": round(value / v["value"], 4), **v} for k, v in base.items()}}
                                 if base else None),
            "verify": ver, "fallback_from": fallback,
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }
        print(json.dumps(out), flush=True)
        if ver is not None and not ver["ok"]:
            print("bench.py: VERIFICATION FAILED", ver, file=sys.stderr)
            rc = 3
    if ver is not None and not ver["ok"]:
        rc = 3
    g.comm.finalize()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()       # explicit: no "was not called before program exit" warning
    return rc


if __name__ == "__main__":
    sys.exit(main())
