#!/usr/bin/env python
"""Device-buffer point-to-point bandwidth sweep, 4 KB - 256 MB (BASELINE.json config 5).

Generalises the reference's 32-byte smoke test (``scripts/rocmaware_test_selectdevice.jl:15-22``) into a
measurement: every rank sends ``bytes`` to its ring successor and receives from its predecessor,

* ``peer``  -- ONE kernel: 16-byte stores through a peer-mapped pointer, release flag into the receiver's memory,
               wait for the own flag (``igg_peer_sendrecv``: what the halo fabric does);
* ``nccl``  -- grouped NCCL ``isend``/``irecv`` (the baseline).

Time is measured on the device (CUDA events around ``iters`` back-to-back exchanges, max over ranks) and
reported as GB/s per direction per GPU next to the NVLink references of the profiling guide (770 GB/s
measured peer copy, 900 GB/s nominal per direction).

Run:  ``torchrun --nproc-per-node 2 benchmarks/halo_bw_sweep.py [--max-mb 256]``
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import rocm_mpi_b200 as igg  # noqa: E402
from rocm_mpi_b200.ops import native  # noqa: E402
from rocm_mpi_b200.parallel.fabric import get_fabric  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-kb", type=int, default=4)
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--out", default=None)
    ap.add_argument("--label", default="", help="free text stored with the result (e.g. the NCCL_* settings of this run)")
    a = ap.parse_args()
    me, dims, nprocs, _, _ = igg.init_global_grid(4, 4, 1, dimy=1, periodx=1, quiet=True)
    g = igg.global_grid()
    fab = get_fabric()
    lib = fab.lib
    dst, src = (me + 1) % nprocs, (me - 1) % nprocs
    max_bytes = a.max_mb << 20
    recv = fab.zeros((max_bytes // 8,))
    ramp = (torch.arange(max_bytes // 8, device=g.device) % 4093).to(torch.float64)
    send = ramp + float(me)             # the receiver knows what its predecessor sends: ramp + src
    flag = fab.zeros((2,), torch.int64)
    cnt = torch.zeros(1, dtype=torch.int64, device=g.device)
    recv_n = torch.empty_like(send)
    dst_ptr = fab.peer_ptr(recv, dst) if nprocs > 1 else recv.data_ptr()
    flag_ptr = fab.peer_ptr(flag, dst) if nprocs > 1 else flag.data_ptr()
    rows = []
    epoch = 0
    size = a.min_kb << 10
    while size <= max_bytes:
        iters = 200 if size <= (1 << 20) else (50 if size <= (32 << 20) else 10)
        res = {"bytes": size}
        for tr in ("peer", "nccl"):
            def one():
                nonlocal epoch
                if tr == "peer":
                    epoch += 1
                    native._check(lib.igg_peer_sendrecv(dst_ptr, send.data_ptr(), size, flag_ptr, flag.data_ptr(), epoch,
                                                        cnt.data_ptr(), flag[1:].data_ptr(), native._stream()))
                elif nprocs > 1:
                    g.comm.exchange([(send[: size // 8], dst, 0)], [(recv_n[: size // 8], src, 0)])
                else:
                    recv_n[: size // 8].copy_(send[: size // 8])
            for _ in range(5):
                one()
            g.comm.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                one()
            e1.record()
            torch.cuda.synchronize()
            ms = g.comm.allreduce_max(e0.elapsed_time(e1)) / iters
            res[tr + "_us"] = round(ms * 1e3, 2)
            res[tr + "_GBps"] = round(size / (ms * 1e-3) / 1e9, 2)
        if tr == "peer" and int(flag[1]) != 0:
            raise RuntimeError("peer flag wait timed out")
        # both arms are checked on every rank against what the ring predecessor sent
        n = size // 8
        want = ramp[:n] + float(src)
        ok = bool(torch.equal(recv[:n], want)) and bool(torch.equal(recv_n[:n], want))
        res["ok"] = bool(g.comm.allreduce_max(0.0 if ok else 1.0) == 0.0)
        recv[:n].zero_()
        recv_n[:n].zero_()
        rows.append(res)
        if me == 0:
            print(json.dumps(res), flush=True)
        size *= 4
    if me == 0:
        summary = {"bench": "halo_bw_sweep", "n_gpus": nprocs, "label": a.label,
                   "nccl_env": {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}, "nvlink_ref_GBps": {"measured_peer_copy": 770, "nominal": 900},
                   "rows": rows}
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                json.dump(summary, f, indent=1)
    igg.finalize_global_grid()
    return 0


if __name__ == "__main__":
    sys.exit(main())
