#!/usr/bin/env python
"""First-contact check of the two-steps-per-pass kernel on a GPU (it has only met its host emulator so far).

    python scripts/check_temporal.py                       # 1 GPU: numerics, then timing at 16384^2
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_temporal.py

1. kernel alone (no neighbours) against two launches of the one-step kernel, several shapes;
2. through ``Diffusion2D(mode="temporal")`` against ``mode="fused"`` (periodic single rank = self-neighbour
   on 1 GPU; real neighbours under torchrun), even and odd step counts;
3. ms per step and T_eff of both modes at ``--nx/--ny``.
Run each stage under its own ``timeout``; a protocol bug shows up as a flag time-out error, not a hang.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rocm_mpi_b200 as igg                                     # noqa: E402
from rocm_mpi_b200.models.solver import Diffusion2D              # noqa: E402
from rocm_mpi_b200.ops import native                             # noqa: E402


def kernel_alone():
    worst = 0.0
    for (nx, ny) in [(3, 4), (16, 16), (17, 130), (40, 126), (64, 250), (130, 258), (300, 380), (1030, 2050)]:
        g = torch.Generator().manual_seed(nx * 7 + ny)
        T = torch.rand(nx, ny, generator=g, dtype=torch.float64).cuda()
        Cp = (torch.rand(nx, ny, generator=g, dtype=torch.float64) + 0.5).cuda()
        a, b, c = T.clone(), T.clone(), T.clone()
        native.diffusion2d_step(a, T, Cp, 1.3, 1e-3, 2.0, 3.0)
        native.diffusion2d_step(b, a, Cp, 1.3, 1e-3, 2.0, 3.0)
        native.diffusion2d_step2(c, T, Cp, 1.3, 1e-3, 2.0, 3.0)
        torch.cuda.synchronize()
        err = float((c - b).abs().max())
        worst = max(worst, err)
        print(f"  kernel alone {nx:5d} x {ny:5d}: max |two-step - 2 x one-step| = {err:.3e}", flush=True)
    return worst


DIMS = [0, 0]            # --dimx / --dimy: force the process grid (e.g. 1 x 2: neighbours in y only)


def through_solver(nx, ny, periods, nsteps):
    out = {}
    for mode in ("fused", "temporal"):
        igg.init_global_grid(nx, ny, 1, dimx=DIMS[0], dimy=DIMS[1], periodx=periods[0], periody=periods[1], quiet=True)
        s = Diffusion2D(nx, ny, mode=mode)
        s.begin()
        s.step(nsteps)
        s.finish()
        torch.cuda.synchronize()
        out[mode] = s.T.clone()
        igg.finalize_global_grid()
    return float((out["fused"] - out["temporal"]).abs().max())


def controlled_pairs(nx, ny, periods, npairs):
    """step_controlled_pair against single controlled steps of the fused mode: field and both diagnostics."""
    igg.init_global_grid(nx, ny, 1, periodx=periods[0], periody=periods[1], quiet=True)
    s = Diffusion2D(nx, ny, mode="fused")
    want = []
    for _ in range(2 * npairs):
        want.append(s.step_controlled())
    s.finish()
    Tf = s.T.clone()
    igg.finalize_global_grid()
    igg.init_global_grid(nx, ny, 1, periodx=periods[0], periody=periods[1], quiet=True)
    s = Diffusion2D(nx, ny, mode="temporal")
    got = []
    for _ in range(npairs):
        d1, d2 = s.step_controlled_pair()
        got += [d1, d2]
    s.finish()
    err = float((s.T - Tf).abs().max())
    igg.finalize_global_grid()
    derr = max(max(abs(g[0] - w[0]), abs(g[1] - w[1]) / max(1.0, abs(w[1]))) for g, w in zip(got, want))
    return err, derr


def timing(nx, ny, steps):
    res = {}
    for mode in ("fused", "temporal"):
        igg.init_global_grid(nx, ny, 1, quiet=True)
        g = igg.global_grid()
        s = Diffusion2D(nx, ny, mode=mode)
        s.begin()
        s.step(10)
        g.comm.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s.step(steps)
        e1.record()
        torch.cuda.synchronize()
        ms = g.comm.allreduce_max(e0.elapsed_time(e1)) / steps
        s.finish()
        res[mode] = ms
        if g.me == 0:
            print(f"  {mode:9s} {ms:.4f} ms/step   T_eff = {igg.a_eff((nx, ny)) / (ms * 1e-3):.0f} GB/s per GPU", flush=True)
        igg.finalize_global_grid()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=16384)
    ap.add_argument("--ny", type=int, default=16384)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--skip-timing", action="store_true")
    ap.add_argument("--timing-only", action="store_true")
    ap.add_argument("--dimx", type=int, default=0)
    ap.add_argument("--dimy", type=int, default=0)
    a = ap.parse_args()
    DIMS[0], DIMS[1] = a.dimx, a.dimy
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    ok = True
    if a.timing_only:
        timing(a.nx, a.ny, a.steps)
        return 0
    if world == 1:
        print("stage 1: kernel alone", flush=True)
        ok &= kernel_alone() < 1e-12
        print("stage 2: solver, periodic single rank (self-neighbour)", flush=True)
        for (nx, ny, per, n) in [(64, 256, (1, 1), 4), (300, 380, (1, 0), 5), (520, 1024, (0, 1), 6), (64, 256, (0, 0), 7)]:
            err = through_solver(nx, ny, per, n)
            print(f"  {nx} x {ny} periods {per} steps {n}: max |fused - temporal| = {err:.3e}", flush=True)
            ok &= err < 1e-12
        print("stage 2b: controlled pairs (per-step parameters in, both diagnostics out)", flush=True)
        for (nx, ny, per) in [(130, 258, (0, 0)), (300, 380, (1, 1))]:
            err, derr = controlled_pairs(nx, ny, per, 3)
            print(f"  {nx} x {ny} periods {per}: max |field diff| = {err:.3e}, max diagnostic diff = {derr:.3e}", flush=True)
            ok &= err < 1e-12 and derr < 1e-9
    else:
        if rank == 0:
            print(f"stage 2: solver on {world} ranks", flush=True)
        for (nx, ny, n) in [(64, 256, 4), (300, 380, 5), (1030, 2050, 6)]:
            err = through_solver(nx, ny, (0, 0), n)
            if rank == 0:
                print(f"  {nx} x {ny} steps {n}: max |fused - temporal| = {err:.3e} (rank 0)", flush=True)
            ok &= err < 1e-12
    if not a.skip_timing:
        if rank == 0:
            print(f"stage 3: timing at {a.nx} x {a.ny}", flush=True)
        timing(a.nx, a.ny, a.steps)
    if rank == 0:
        print("OK" if ok else "MISMATCH", flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
