"""Device selection + device-buffer point-to-point smoke test.

Behavioural counterpart of reference ``scripts/rocmaware_test_selectdevice.jl``: node-local rank -> GPU
(``:7-9``), ring neighbours ``dst = (r+1) % P``, ``src = (r-1) % P`` (``:11-13``), a 4-element fp64 **device**
buffer filled with the rank (``:15-18``), barrier, ``Sendrecv!`` straight on the device buffers (``:20-22``),
print what arrived (``:23``), barrier (``:24``).  Pass criterion: rank ``r`` receives ``[r-1, r-1, r-1, r-1]``.

Two transports are exercised, mirroring the reference's "GPU-aware or not" question:

* ``peer`` -- the receive buffer lives in fabric memory; the sender stores into it through a peer-mapped
  pointer with one kernel and raises a flag; the receiver waits on the flag (no library call);
* ``nccl`` -- grouped ``isend``/``irecv`` on the device tensors.

(The reference computes ``dst``/``src`` from the node-local rank but uses them as world ranks, ``:12-13,22``;
that is only right on one node, which is also the scope here.)

Run:  ``torchrun --nproc-per-node 8 -m rocm_mpi_b200.models.selectdevice``
"""
from __future__ import annotations

import argparse
from typing import Dict

import torch

from ..ops import native
from ..parallel import finalize_global_grid, global_grid, init_global_grid
from ..parallel.comm import local_rank
from ..parallel.fabric import get_fabric

__all__ = ["selectdevice_test", "main"]


def selectdevice_test(N: int = 4, transports=("peer", "nccl"), quiet: bool = False,
                      finalize: bool = True) -> Dict[str, torch.Tensor]:
    # the grid is only used for bring-up: a 1-D periodic ring of all ranks
    me, dims, nprocs, coords, _ = init_global_grid(4, 4, 1, dimy=1, periodx=1, quiet=True)
    g = global_grid()
    if g.device_type != "cuda":
        transports = ("gloo",)          # CPU plumbing run: same ring, host buffers over gloo
    rank_l = local_rank()
    size = nprocs
    dst, src = (me + 1) % size, (me - 1) % size
    if not quiet:
        print(f"rank={rank_l} (gpu_id={g.device.index}), size={size}, dst={dst}, src={src}", flush=True)
    out: Dict[str, torch.Tensor] = {}
    send = torch.full((N,), float(me), dtype=torch.float64, device=g.device)
    for tr in transports:
        g.comm.barrier()
        if me == 0 and not quiet:
            print(f"start sending ({tr})...", flush=True)
        if tr == "peer":
            fab = get_fabric()
            lib = fab.lib
            recv = fab.zeros((N,))                       # collective; addressable by the neighbours
            flag = fab.zeros((2,), torch.int64)
            cnt = torch.zeros(1, dtype=torch.int64, device=g.device)
            torch.cuda.synchronize()
            g.comm.barrier()                             # everybody's receive buffer exists and is zero
            if size > 1:
                dst_ptr, flag_ptr = fab.peer_ptr(recv, dst), fab.peer_ptr(flag, dst)
            else:
                dst_ptr, flag_ptr = recv.data_ptr(), flag.data_ptr()
            # send + receive in ONE launch: store into the successor's buffer, release its flag, wait for mine
            native._check(lib.igg_peer_sendrecv(dst_ptr, send.data_ptr(), N * 8, flag_ptr, flag.data_ptr(), 1,
                                                cnt.data_ptr(), flag[1:].data_ptr(), native._stream()))
            torch.cuda.synchronize()
            if int(flag[1]) != 0:
                raise RuntimeError("peer flag wait timed out")
        elif tr in ("nccl", "gloo"):
            recv = torch.empty(N, dtype=torch.float64, device=g.device)
            if size > 1:
                g.comm.exchange([(send, dst, 0)], [(recv, src, 0)])
            else:
                recv.copy_(send)
            if g.device_type == "cuda":
                torch.cuda.synchronize()
        else:
            raise ValueError(f"unknown transport {tr!r}")
        out[tr] = recv.cpu()
        if not quiet:
            print(f"recv_mesg on proc {rank_l} ({tr}): {out[tr].tolist()}", flush=True)
        expect = float(src)
        if not bool((out[tr] == expect).all()):
            raise AssertionError(f"rank {me}: expected {expect} from rank {src} over {tr}, got {out[tr].tolist()}")
        g.comm.barrier()
    if me == 0 and not quiet:
        print("done.", flush=True)
    if finalize:
        finalize_global_grid()
    return out


def main(argv=None) -> None:
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    p.add_argument("-N", type=int, default=4)
    p.add_argument("--transports", nargs="+", default=["peer", "nccl"])
    ns = p.parse_args(argv)
    selectdevice_test(ns.N, tuple(ns.transports))


if __name__ == "__main__":
    main()
