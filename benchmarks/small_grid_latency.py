#!/usr/bin/env python
"""Per-step latency on small local grids (BASELINE.json config 2: the kp / Temp_4_252_252 regime).

At 128 x 128 (or 65 x 128) cells per GPU a time step is launch + synchronisation latency, not bandwidth.
Compares, device-timed and max over ranks:

* ``plain-nccl``  one fused kernel + pack / NCCL p2p / unpack per step (the reference's perf structure)
* ``fused``       one fused kernel per step with in-kernel halo push
* ``persistent``  ONE cooperative launch for all steps (in-kernel grid barrier + halo push)
* ``resident``    ONE cluster launch, field resident in distributed shared memory, cluster barriers

and, for the kp application itself (``Flux!`` / ``Residual!`` / ``Update!`` arithmetic, ``scripts/diffusion_2D_kp.jl:87-92``):

* ``kp-3kernel-nccl``  the three kernels + ``update_halo`` over NCCL per step (the reference's structure)
* ``kp-3kernel-peer``  the three kernels, ``Update!`` pushing its edge cells, two flag kernels around them
* ``kp-persistent`` / ``kp-resident``  the single-launch kernels with the kp arithmetic (``form='kp'``)

Every kp row is checked against the single-domain CPU oracle of the kp form when the global problem is small
(``max_abs_err_vs_oracle``, tolerance 1e-12).

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
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    me, dims, nprocs, _, _ = igg.init_global_grid(a.nx, a.ny, 1, dimx=a.dimx, dimy=a.dimy, quiet=True)
    g = igg.global_grid()
    out = {"bench": "small_grid_latency", "n_gpus": nprocs, "local": [a.nx, a.ny], "dims": list(dims[:2]), "nt": a.nt}
    variant = "tma" if a.ny % 2 == 0 else "direct"
    for name, kw in (("plain-nccl", dict(mode="plain", variant=variant, transport="nccl")),
                     ("fused", dict(mode="fused", variant=variant)),
                     ("persistent", dict(mode="persistent")), ("resident", dict(mode="resident"))):
        s = Diffusion2D(a.nx, a.ny, **kw)
        try:
            s.begin()
            s.step(20)
            n0 = native.launch_count()
            ms = timed(g, lambda: s.step(a.nt))
            s.finish()
            mx, _ = s.max_sum()
            out[name] = {"us_per_step": round(ms * 1e3 / a.nt, 3), "native_launches": native.launch_count() - n0,
                         "max_T_local": mx}
        except native.NativeError as e:            # e.g. 252 x 252 does not fit one cluster's shared memory
            out[name] = {"unsupported": str(e)[:120]}
        s.close()
    # ---- the kp application's arithmetic ------------------------------------------------------------------
    from rocm_mpi_b200.models.diffusion2d_kp import KPThree
    from rocm_mpi_b200.ops import reference as ref
    from rocm_mpi_b200.parallel.gather import gather_interior
    nt_total = 20 + a.nt
    want = None
    if me == 0 and igg.nx_g() * igg.ny_g() <= 1 << 20:
        want, _ = ref.oracle_2d(igg.nx_g(), igg.ny_g(), nt_total, "kp")

    def check(T):
        Tv = gather_interior(T)
        err = None
        if me == 0 and want is not None:
            err = float((Tv - want[1:-1, 1:-1]).abs().max())
        return err

    dx, dy = 10.0 / igg.nx_g(), 10.0 / igg.ny_g()
    dt = ref.time_step(dx, dy, 1.0, 1.0)
    for name, transport in (("kp-3kernel-nccl", "nccl"), ("kp-3kernel-peer", "peer")):
        probe = Diffusion2D(a.nx, a.ny, mode="plain", variant="direct", transport="nccl")      # for its initial condition
        T0 = probe.T.clone()
        probe.close()
        kp = KPThree(a.nx, a.ny, dx, dy, dt, 1.0, 1.0, T0, transport if nprocs > 1 else "peer")
        kp.step(20)
        n0 = native.launch_count()
        ms = timed(g, lambda: kp.step(a.nt))
        kp.finish()
        out[name] = {"us_per_step": round(ms * 1e3 / a.nt, 3), "native_launches": native.launch_count() - n0,
                     "max_abs_err_vs_oracle": check(kp.T)}
        kp.close()
    for name, mode in (("kp-persistent", "persistent"), ("kp-resident", "resident")):
        s = Diffusion2D(a.nx, a.ny, mode=mode, form="kp")
        try:
            s.begin()
            s.step(20)
            n0 = native.launch_count()
            ms = timed(g, lambda: s.step(a.nt))
            s.finish()
            out[name] = {"us_per_step": round(ms * 1e3 / a.nt, 3), "native_launches": native.launch_count() - n0,
                         "max_abs_err_vs_oracle": check(s.T)}
        except native.NativeError as e:
            out[name] = {"unsupported": str(e)[:120]}
        s.close()
    if me == 0:
        print(json.dumps(out), flush=True)
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                json.dump(out, f, indent=1)
    igg.finalize_global_grid()
    return 0


if __name__ == "__main__":
    sys.exit(main())
