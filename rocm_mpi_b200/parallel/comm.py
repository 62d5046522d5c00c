"""Process-level communication plumbing (launch, rendezvous, barriers, object exchange).

This layer replaces what the reference gets from MPI.jl + PMIx
(``MPI.Init`` / ``Comm_rank`` / ``Comm_size`` / ``Comm_split_type`` / ``Barrier`` in
reference ``scripts/rocmaware_test_selectdevice.jl:3-24``): one process per GPU started by
``torchrun``; ``torch.distributed`` carries rendezvous, CPU-side (gloo) traffic and the
NCCL baseline.  The *hot* halo path does not go through here at all -- it is peer-memory
stores issued from inside the stencil kernels (see ``parallel/fabric.py``).

Two implementations share one small interface:

* :class:`TorchComm` -- ``torch.distributed`` process group (gloo on CPU, gloo+nccl on GPU).
* :class:`SelfComm`  -- a single process that is every one of its own neighbours
  (used when no launcher environment is present; still supports periodic grids).
"""
from __future__ import annotations

import os
from typing import Any, List, Optional, Sequence

import torch

__all__ = ["Comm", "SelfComm", "TorchComm", "local_rank", "launcher_env_present", "make_comm", "spans_nodes"]


def launcher_env_present() -> bool:
    return "RANK" in os.environ and "WORLD_SIZE" in os.environ


def local_rank() -> int:
    """Node-local rank: the analogue of ``Comm_rank(Comm_split_type(COMM_TYPE_SHARED))``.

    Reference ``scripts/rocmaware_test_selectdevice.jl:7-8`` derives it by splitting the
    world communicator by shared-memory node; ``torchrun`` exports it as ``LOCAL_RANK``.
    """
    for key in ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID", "MV2_COMM_WORLD_LOCAL_RANK"):
        if key in os.environ:
            return int(os.environ[key])
    return 0


def spans_nodes() -> bool:
    """True when the job has more ranks than this node hosts (``torchrun`` exports ``LOCAL_WORLD_SIZE``).

    The peer fabric maps memory with CUDA IPC, i.e. inside one NVSwitch domain; across nodes the halo exchange
    has to use the NCCL transport (the reference's multi-node story is likewise the launcher's:
    ``srun -n 4 --mpi=pmix``, ``README.md:16-18``)."""
    try:
        return int(os.environ.get("WORLD_SIZE", "1")) > int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
    except ValueError:
        return False


class Comm:
    rank: int = 0
    size: int = 1
    group = None

    def barrier(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def allgather_object(self, obj: Any) -> List[Any]:
        raise NotImplementedError

    def allreduce_max(self, value: float) -> float:
        raise NotImplementedError

    def allreduce_sum(self, value: float) -> float:
        raise NotImplementedError

    def exchange(self, sends: Sequence[tuple], recvs: Sequence[tuple]) -> None:
        """Complete a set of tagged point-to-point transfers.

        ``sends``/``recvs`` are sequences of ``(tensor, peer, tag)``.  Tensors must be
        contiguous.  Transfers whose peer is this rank are matched locally by tag.
        """
        raise NotImplementedError

    def gather_blocks(self, block: torch.Tensor, root: int) -> Optional[List[torch.Tensor]]:
        raise NotImplementedError

    def finalize(self) -> None:
        pass


def _match_local(sends, recvs, me):
    """Resolve self-addressed messages by tag (periodic grid with one rank in a dim)."""
    local_s = {tag: t for (t, peer, tag) in sends if peer == me}
    rest_r = []
    for (t, peer, tag) in recvs:
        if peer == me:
            t.copy_(local_s.pop(tag))
        else:
            rest_r.append((t, peer, tag))
    if local_s:
        raise RuntimeError("unmatched self-addressed halo messages: tags %s" % sorted(local_s))
    rest_s = [(t, peer, tag) for (t, peer, tag) in sends if peer != me]
    return rest_s, rest_r


class SelfComm(Comm):
    """One process, no launcher.  Every transfer is a local copy."""

    def __init__(self) -> None:
        self.rank = 0
        self.size = 1

    def barrier(self) -> None:
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def allgather_object(self, obj):
        return [obj]

    def allreduce_max(self, value):
        return float(value)

    def allreduce_sum(self, value):
        return float(value)

    def exchange(self, sends, recvs):
        s, r = _match_local(sends, recvs, 0)
        if s or r:
            raise RuntimeError("SelfComm cannot talk to other ranks")

    def gather_blocks(self, block, root):
        return [block]


class TorchComm(Comm):
    """``torch.distributed`` backed communicator.

    ``device_type == 'cuda'`` initialises the default group as ``cpu:gloo,cuda:nccl`` so
    that host objects/handles travel over gloo while device tensors handed to
    ``exchange`` use NCCL send/recv (the *baseline* transport; the product path is the
    peer fabric).
    """

    def __init__(self, device_type: str = "cpu", init: bool = True, timeout_s: Optional[int] = None) -> None:
        import datetime
        import torch.distributed as dist

        if timeout_s is None:      # how long a collective / p2p call waits for a rank that never arrives
            timeout_s = int(os.environ.get("IGG_B200_TIMEOUT_S", "600"))

        self._dist = dist
        self._owns_pg = False
        if not dist.is_initialized():
            if not init:
                raise RuntimeError("torch.distributed is not initialised and init_dist=False")
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            backend = "cpu:gloo,cuda:nccl" if device_type == "cuda" else "gloo"
            kwargs = {}
            if device_type == "cuda":
                kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
            dist.init_process_group(backend=backend,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kwargs)
            self._owns_pg = True
        self.rank = dist.get_rank()
        self.size = dist.get_world_size()
        self.device_type = device_type
        self._cpu_group = None

    # A gloo-only group for host objects when the default group is NCCL-only.
    def _host_group(self):
        dist = self._dist
        backend = dist.get_backend()
        if "gloo" in str(dist.get_backend_config()) or backend == "gloo":
            return None  # default group can do CPU tensors
        if self._cpu_group is None:
            self._cpu_group = dist.new_group(backend="gloo")
        return self._cpu_group

    def barrier(self) -> None:
        if torch.cuda.is_available() and self.device_type == "cuda":
            torch.cuda.synchronize()
        # a CPU-tensor all_reduce is a barrier that never touches NCCL
        t = torch.zeros(1)
        self._dist.all_reduce(t, group=self._host_group())

    def allgather_object(self, obj):
        out = [None] * self.size
        self._dist.all_gather_object(out, obj, group=self._host_group())
        return out

    def allreduce_max(self, value):
        t = torch.tensor([float(value)], dtype=torch.float64)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX, group=self._host_group())
        return float(t.item())

    def allreduce_sum(self, value):
        t = torch.tensor([float(value)], dtype=torch.float64)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._host_group())
        return float(t.item())

    def exchange(self, sends, recvs):
        dist = self._dist
        sends, recvs = _match_local(list(sends), list(recvs), self.rank)
        if not sends and not recvs:
            return
        on_cuda = any(t.is_cuda for (t, _, _) in list(sends) + list(recvs))
        if on_cuda:
            # NCCL matches by issue order per peer, not by tag: order both lists by
            # (peer, tag) so the two ends of every pair agree.
            ops = []
            for (t, peer, tag) in sorted(recvs, key=lambda x: (x[1], x[2])):
                ops.append(dist.P2POp(dist.irecv, t, peer))
            for (t, peer, tag) in sorted(sends, key=lambda x: (x[1], x[2])):
                ops.append(dist.P2POp(dist.isend, t, peer))
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        else:
            grp = self._host_group()
            reqs = [dist.irecv(t, src=peer, tag=tag, group=grp) for (t, peer, tag) in recvs]
            reqs += [dist.isend(t, dst=peer, tag=tag, group=grp) for (t, peer, tag) in sends]
            for r in reqs:
                r.wait()

    def gather_blocks(self, block, root):
        dist = self._dist
        block = block.contiguous()
        if block.is_cuda:
            block = block.cpu()
        out = [torch.empty_like(block) for _ in range(self.size)] if self.rank == root else None
        dist.gather(block, out, dst=root, group=self._host_group())
        return out

    def finalize(self) -> None:
        """Release the communicator.

        The process group itself is torn down once, at interpreter exit: an application may call
        ``finalize_global_grid()`` and ``init_global_grid()`` again in the same process (several runs in one
        script), and re-creating a process group on the same rendezvous port is neither cheap nor
        reliable.  (MPI cannot be re-initialised at all; this is the more forgiving behaviour.)
        """
        if self._owns_pg and self._dist.is_initialized():
            _destroy_pg_at_exit()


_atexit_registered = False


def _destroy_pg_at_exit() -> None:
    global _atexit_registered
    if _atexit_registered:
        return
    _atexit_registered = True
    import atexit

    def _destroy():
        import torch.distributed as dist
        try:
            if dist.is_available() and dist.is_initialized():
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                dist.destroy_process_group()
        except Exception:
            pass

    atexit.register(_destroy)


def make_comm(device_type: str, init_dist: bool = True) -> Comm:
    import torch.distributed as dist

    if dist.is_available() and (dist.is_initialized() or launcher_env_present()):
        return TorchComm(device_type=device_type, init=init_dist)
    return SelfComm()
