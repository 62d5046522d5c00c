#!/usr/bin/env python
"""Heat diffusion with a host IN the loop and the GPU never waiting for it.

`Diffusion2D(mode="temporal")` advances two time steps per pass over memory.  The host chooses the time step of every
pair from the diagnostics it reads back (`max`, `sum` of the field after each of the two steps, computed inside the
stencil kernel) -- here a toy controller that shrinks `dt` while the peak still drops fast -- but it does so
PIPELINED: pass k+1 is queued before pass k's diagnostics are awaited, so every pass's 64 B of parameters travel
pinned -> device and its 32 B of diagnostics device -> pinned without ever leaving the GPU idle.  At the end the
halo-free field is gathered into the root GPU's memory (tiles stored over NVLink by every rank) and written as a heat
map + raw dump.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/heat2d_two_step_control.py --size 4096 --nt 400

(The reference's loop has no host control at all -- scripts/diffusion_2D_perf.jl:47-52 -- and prints
maximum(T_v) once at the end, scripts/diffusion_2D_perf_hide.jl:115; this is what that diagnostic looks like when an
application wants it every step.)
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.models.solver import Diffusion2D  # noqa: E402
from rocm_mpi_b200.parallel import gather_interior_device  # noqa: E402
from rocm_mpi_b200.utils.output import output_name, save_heatmap, save_raw  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=4096)
ap.add_argument("--nt", type=int, default=400)
ap.add_argument("--outdir", default="output")
a = ap.parse_args()

me, dims, nprocs, coords, _ = igg.init_global_grid(a.n, a.n, 1)
s = Diffusion2D(a.n, a.n, mode="temporal")          # fields in peer-addressable memory, flag bank, mid lines
dt_max = s.dt                                       # the stable explicit step; the controller only goes below it
dt = 0.5 * dt_max
history = []                                        # (step, max, sum) of this rank's part of the field
last_max = None


def control(diag, step):
    """Toy controller: relax dt towards dt_max once the local peak changes by < 0.1 % per step."""
    global dt, last_max
    for k, (mx, sm) in enumerate(diag):
        history.append((step + k, mx, sm))
        if last_max is not None and abs(mx - last_max) < 1e-3 * abs(last_max):
            dt = min(dt_max, 1.05 * dt)
        last_max = mx


igg.tic()
submitted = 0
for it in range(a.nt // 2):
    d = s.step_controlled_pair(dt=(dt, dt), pipelined=True)    # queues pass `it`, returns pass `it - 1`'s diagnostics
    submitted += 2
    if d is not None:
        control(d, submitted - 4)
for d in s.drain_controlled():                                 # the last pass
    control(d, submitted - 2)
s.finish()
wtime = igg.toc()

G = gather_interior_device(s.T)                                # device tensor on rank 0, None elsewhere
if me == 0:
    base = os.path.join(a.outdir, output_name("two_step", nprocs, igg.nx_g(), igg.ny_g()))
    save_heatmap(base + ".png", G, title=f"two steps per pass, {nprocs} ranks {dims[0]}x{dims[1]}, {submitted} steps")
    save_raw(base, G, {"variant": "two_step", "nprocs": nprocs, "dims": list(dims), "nt": submitted})
    print(f"{nprocs} GPUs {dims[0]}x{dims[1]}: {wtime / submitted * 1e3:.4f} ms/step end to end "
          f"(T_eff = {igg.t_eff((a.n, a.n), wtime, submitted):.0f} GB/s per GPU), dt ended at {dt / dt_max:.2f} dt_max, "
          f"max(T) {history[0][1]:.6f} -> {history[-1][1]:.6f}; wrote {base}.png")
s.close()
igg.finalize_global_grid()
