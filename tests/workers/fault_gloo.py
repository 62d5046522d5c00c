"""Worker: one rank dies after set-up; the survivors must get an error from the next halo exchange, not hang
(the reference has no failure handling: "an MPI failure aborts the job")."""
import os
import sys
import time

import torch

import rocm_mpi_b200 as igg

me, dims, nprocs, coords, _ = igg.init_global_grid(12, 10, 1, device_type="cpu", quiet=True)
T = torch.rand(12, 10, dtype=torch.float64)
igg.update_halo(T)                       # everybody alive: works
if me == nprocs - 1:
    os._exit(7)                          # simulated crash: no finalize, sockets torn down by the OS
t0 = time.time()
try:
    for _ in range(3):
        igg.update_halo(T)
    print(f"rank {me}: exchange with a dead neighbour did not fail", flush=True)
    sys.exit(1)
except Exception as e:                   # noqa: BLE001 - any error type from the backend is fine, a hang is not
    print(f"rank {me}: failure detected after {time.time() - t0:.1f} s ({type(e).__name__})", flush=True)
    os._exit(0)                          # skip the collective finalize: the group is broken
