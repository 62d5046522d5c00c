"""Peer-memory fabric: symmetric device buffers, flag pads and the engines built on them.

Where the reference relies on a ROCm-aware MPI that accepts device pointers
(``MPI.Sendrecv!`` on ``ROCArray``s, ``scripts/rocmaware_test_selectdevice.jl:15-22``; the
``IGG_ROCMAWARE_MPI=1`` mode of ``update_halo!``, ``scripts/setenv.sh:11-13``), a B200 node lets
every GPU address every other GPU's memory through NVSwitch.  The fabric therefore

* allocates (or registers) fields so that every rank can obtain a **peer-mapped pointer** to the
  same field on each Cartesian neighbour (CUDA IPC handles travel once over gloo);
* owns a **flag pad** per rank -- ``FLAG_BANKS`` independent banks of epoch flags that neighbours
  write with system-scope release stores -- plus a private pad of arrival counters;
* provides :class:`PeerHalo` (``update_halo`` as ONE kernel launch: handshake, all boxes stored
  into the neighbours' arrays, epoch released, own epoch awaited) and :class:`PushPlan` (the
  descriptor that lets the stencil kernels push their edge cells themselves).

With a single process (no launcher) every neighbour of a periodic grid is the rank itself and the
same kernels run against local pointers, which is how the protocol is tested on one GPU.
"""
from __future__ import annotations

import ctypes as C
import itertools
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..ops import native
from .grid import GlobalGrid, global_grid
from .halo import halo_boxes, halo_planes
from .topology import NDIMS, PROC_NULL

__all__ = ["Fabric", "PeerHalo", "PushPlan", "get_fabric", "fabric_zeros", "offset_slot", "neighbor_offsets"]


def offset_slot(o: Sequence[int]) -> int:
    """Slot index of neighbour offset ``(ox, oy, oz)`` in a flag bank."""
    return ((o[0] + 1) * 3 + (o[1] + 1)) * 3 + (o[2] + 1)


class _DeviceBlock:
    """Exposes a raw device allocation to ``torch.as_tensor`` via ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, shape: Tuple[int, ...], typestr: str) -> None:
        self.ptr = ptr
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False),
                                         "version": 3, "strides": None}


_TYPESTR = {torch.float64: "<f8", torch.float32: "<f4", torch.float16: "<f2", torch.int64: "<i8",
            torch.int32: "<i4", torch.int16: "<i2", torch.uint8: "|u1"}


def neighbor_offsets(g: GlobalGrid) -> List[Tuple[Tuple[int, int, int], int]]:
    """All ``(offset, rank)`` pairs of existing Cartesian neighbours (up to 26)."""
    out = []
    rng = [(-1, 0, 1) if g.nxyz[d] > 1 else (0,) for d in range(NDIMS)]
    for o in itertools.product(*rng):
        if o == (0, 0, 0):
            continue
        r = g.topo.rank_at_offset(o)
        if r != PROC_NULL:
            out.append((o, r))
    return out


class Fabric:
    """Per-grid registry of peer-mapped memory."""

    def __init__(self, g: GlobalGrid) -> None:
        if g.device_type != "cuda":
            raise RuntimeError("the peer fabric needs CUDA devices")
        self.g = g
        self.lib = native.require()
        self._owned: Dict[int, int] = {}                # pointer -> bytes, from igg_fabric_alloc
        self._handles: Dict[int, Dict[int, bytes]] = {}  # local data_ptr -> {rank: IPC handle it was mapped from}
        self._opened: Dict[bytes, int] = {}             # IPC handle -> mapped base pointer
        self._peer_ptrs: Dict[int, Dict[int, int]] = {}  # local data_ptr -> {rank: peer pointer}
        self._keepalive: List[object] = []
        # strong references to every registered tensor: while a field is registered its storage cannot be freed,
        # so the caching allocator can never hand the same address to another tensor (a reused address would
        # look "already registered" on this rank only -> ranks disagree about entering the collective, or
        # the neighbours keep storing into freed memory)
        self._registered: Dict[int, torch.Tensor] = {}
        self._free_banks: List[int] = []
        self._root_bufs: Dict[tuple, tuple] = {}        # (shape, dtype, root) -> (tensor on root | None, my pointer)
        self.neighbors = neighbor_offsets(g)
        self.neighbor_ranks = sorted({r for _, r in self.neighbors})
        self._next_bank = 0
        pad_elems = native.FLAG_BANKS * native.BANK_STRIDE
        self.flag_pad = self.zeros((pad_elems,), torch.int64)       # collective
        self.count_pad = torch.zeros(pad_elems, dtype=torch.int64, device=g.device)
        self._check_peer_access()

    # -- allocation / registration --------------------------------------------------------------
    def _check_peer_access(self) -> None:
        if self.g.nprocs == 1:
            return
        # ranks on one node map 1:1 to devices (select_device); make sure NVLink/P2P is there
        me_dev = self.g.device.index
        devs = self.g.comm.allgather_object(me_dev)
        for r in self.neighbor_ranks:
            if r == self.g.me:
                continue
            can = C.c_int(0)
            native._check(self.lib.igg_can_access_peer(me_dev, devs[r], C.byref(can)))
            if not can.value:
                raise RuntimeError(f"GPU {me_dev} cannot access peer GPU {devs[r]} (rank {r}); "
                                   "use transport='nccl'")

    def zeros(self, shape: Sequence[int], dtype: torch.dtype = torch.float64) -> torch.Tensor:
        """Collective: allocate a zero-filled symmetric tensor, addressable by all neighbours."""
        shape = tuple(int(s) for s in shape)
        numel = 1
        for s in shape:
            numel *= s
        nbytes = max(numel * torch.empty((), dtype=dtype).element_size(), 16)
        ptr = C.c_void_p()
        native._check(self.lib.igg_fabric_alloc(nbytes, C.byref(ptr)))
        self._owned[ptr.value] = nbytes
        blk = _DeviceBlock(ptr.value, shape, _TYPESTR[dtype])
        t = torch.as_tensor(blk, device=self.g.device)
        if t.data_ptr() != ptr.value:
            raise RuntimeError("torch did not alias the fabric allocation")
        self._keepalive.append(blk)
        self.register(t)
        return t

    def register(self, t: torch.Tensor) -> None:
        """Collective: make ``t`` addressable by the neighbours (no-op if already registered).

        The fabric keeps ``t`` alive until :meth:`unregister` / :meth:`free` / :meth:`close`: a registration
        is a promise to the neighbours that the address stays valid.  Register long-lived fields (the
        solvers allocate theirs with :func:`fabric_zeros`), not per-step temporaries; every rank must
        register the same fields in the same order."""
        key = t.data_ptr()
        if key in self._peer_ptrs:
            return
        g = self.g
        self._registered[key] = t
        if g.nprocs == 1:
            self._peer_ptrs[key] = {0: key}
            return
        handle = (C.c_ubyte * 64)()
        off = C.c_uint64(0)
        try:
            native._check(self.lib.igg_ipc_export(key, handle, C.byref(off)))
        except native.NativeError as e:
            raise RuntimeError("cannot export this tensor over CUDA IPC (allocate it with "
                               "rocm_mpi_b200.zeros(..., symmetric=True), or do not use expandable "
                               f"segments): {e}") from e
        infos = g.comm.allgather_object((bytes(handle), int(off.value)))
        ptrs = {g.me: key}
        self._handles[key] = {}
        for r in self.neighbor_ranks:
            if r == g.me:
                continue
            h, o = infos[r]
            self._handles[key][r] = h
            base = self._opened.get(h)
            if base is None:
                out = C.c_void_p()
                native._check(self.lib.igg_ipc_open(h, C.byref(out)))
                base = out.value
                self._opened[h] = base
            ptrs[r] = base + o
        self._peer_ptrs[key] = ptrs

    def unregister(self, t: torch.Tensor) -> None:
        """Collective: drop the registration of ``t`` (all ranks, same order); the neighbours' mappings of
        its allocation are closed when no other registered field shares it."""
        key = t.data_ptr()
        if key not in self._peer_ptrs:
            return
        torch.cuda.synchronize()
        self.g.comm.barrier()                       # no kernel anywhere still stores into it
        mine = self._handles.pop(key, {})
        self._peer_ptrs.pop(key, None)
        self._registered.pop(key, None)
        still = {h for hs in self._handles.values() for h in hs.values()}
        for h in set(mine.values()) - still:
            base = self._opened.pop(h, None)
            if base is not None:
                self.lib.igg_ipc_close(base)
        self._forget_tables(key)
        self.g.comm.barrier()

    def root_buffer(self, shape: Sequence[int], dtype: torch.dtype = torch.float64, root: int = 0):
        """Collective: a zero-filled buffer that lives on ``root`` only and that EVERY rank can store into.

        Returns ``(tensor, ptr)``: the tensor on ``root`` (``None`` elsewhere) and this rank's pointer to the
        buffer (peer-mapped off the root).  Used by the device-side ``gather``: the gathered field exists once,
        in the root's memory, and the other ranks write their tiles straight into it.  Cached per
        ``(shape, dtype, root)``; released by :meth:`close`."""
        shape = tuple(int(x) for x in shape)
        key = (shape, dtype, int(root))
        hit = self._root_bufs.get(key)
        if hit is not None:
            return hit
        g = self.g
        t, info = None, None
        if g.me == root:
            numel = 1
            for x in shape:
                numel *= x
            nbytes = max(numel * torch.empty((), dtype=dtype).element_size(), 16)
            ptr = C.c_void_p()
            native._check(self.lib.igg_fabric_alloc(nbytes, C.byref(ptr)))
            self._owned[ptr.value] = nbytes
            blk = _DeviceBlock(ptr.value, shape, _TYPESTR[dtype])
            t = torch.as_tensor(blk, device=g.device)
            self._keepalive.append(blk)
            if g.nprocs > 1:
                handle = (C.c_ubyte * 64)()
                off = C.c_uint64(0)
                native._check(self.lib.igg_ipc_export(ptr.value, handle, C.byref(off)))
                info = (bytes(handle), int(off.value))
        my_ptr = t.data_ptr() if t is not None else 0
        if g.nprocs > 1:
            infos = g.comm.allgather_object(info)
            if g.me != root:
                h, o = infos[root]
                base = self._opened.get(h)
                if base is None:
                    out = C.c_void_p()
                    native._check(self.lib.igg_ipc_open(h, C.byref(out)))
                    base = out.value
                    self._opened[h] = base
                my_ptr = base + o
        self._root_bufs[key] = (t, my_ptr)
        return t, my_ptr

    def _forget_tables(self, data_ptr: int) -> None:
        eng = self.g.cache.get(("halo", "peer"))
        if eng is not None:
            eng.forget(data_ptr)

    def is_registered(self, t: torch.Tensor) -> bool:
        return t.data_ptr() in self._peer_ptrs

    def peer_ptr(self, t: torch.Tensor, rank: int) -> int:
        return self._peer_ptrs[t.data_ptr()][rank]

    # -- flag banks -----------------------------------------------------------------------------
    def new_bank(self) -> int:
        """Collective by convention: every rank takes (and releases) banks in the same order."""
        if self._free_banks:
            return self._free_banks.pop()
        if self._next_bank >= native.FLAG_BANKS:
            raise RuntimeError("out of flag banks (32 per grid): close() solvers / plans that are no longer used")
        b = self._next_bank
        self._next_bank += 1
        return b

    def release_bank(self, bank: int) -> None:
        """Collective: return ``bank`` to the free list with all its flags and counters cleared."""
        torch.cuda.synchronize()
        self.g.comm.barrier()                       # no neighbour is still signalling into it
        b = bank * native.BANK_STRIDE
        self.flag_pad[b:b + native.BANK_STRIDE].zero_()
        self.count_pad[b:b + native.BANK_STRIDE].zero_()
        torch.cuda.synchronize()
        self.g.comm.barrier()
        self._free_banks.append(bank)

    def check_all_errors(self) -> None:
        """Synchronise and raise if a flag wait on ANY bank timed out (called from ``toc``, ``gather`` and
        ``finalize_global_grid`` so that no application path can miss a stale halo)."""
        torch.cuda.current_stream().synchronize()
        words = self.flag_pad.view(native.FLAG_BANKS, native.BANK_STRIDE)[:, native.SLOT_ERR]
        bad = torch.nonzero(words).flatten().tolist()
        if bad:
            v = int(words[bad[0]].item())
            words[bad] = 0
            raise RuntimeError(f"peer flag wait timed out on bank(s) {bad} (word {v & ((1 << 63) - 1)}); "
                               "a neighbour did not reach the matching call -- halo contents are stale")

    def bank_ptr(self, bank: int, rank: Optional[int] = None) -> int:
        base = self.flag_pad.data_ptr() if rank is None else self.peer_ptr(self.flag_pad, rank)
        return base + bank * native.BANK_STRIDE * 8

    def count_bank_ptr(self, bank: int) -> int:
        return self.count_pad.data_ptr() + bank * native.BANK_STRIDE * 8

    def sync_struct(self, bank: int, epoch: int, offsets: Optional[Sequence] = None) -> native.HaloSync:
        s = native.HaloSync()
        for (o, r) in (self.neighbors if offsets is None else offsets):
            s.peer_bank[offset_slot(o)] = self.bank_ptr(bank, r)
        s.my_bank = self.bank_ptr(bank)
        s.my_count_bank = self.count_bank_ptr(bank)
        s.epoch = epoch
        return s

    def check_errors(self, bank: int) -> None:
        """Synchronise and raise if a flag wait on ``bank`` timed out."""
        torch.cuda.current_stream().synchronize()
        v = int(self.flag_pad[bank * native.BANK_STRIDE + native.SLOT_ERR].item())
        if v != 0:
            self.flag_pad[bank * native.BANK_STRIDE + native.SLOT_ERR] = 0
            raise RuntimeError(f"peer flag wait timed out on bank {bank} (word {v & ((1 << 63) - 1)}); "
                               "a neighbour did not reach the matching call")

    def free(self, t: torch.Tensor) -> None:
        """Collective: unmap ``t`` on every neighbour and release its memory.

        ``t`` (and every view of it) must not be used afterwards: the tensor only aliases the block."""
        key = t.data_ptr()
        if key not in self._owned:
            raise ValueError("tensor was not allocated by this fabric")
        torch.cuda.synchronize()
        self.g.comm.barrier()
        for h in set(self._handles.pop(key, {}).values()):
            base = self._opened.pop(h, None)
            if base is not None:
                self.lib.igg_ipc_close(base)
        self._peer_ptrs.pop(key, None)
        self._registered.pop(key, None)
        self._forget_tables(key)
        self.g.comm.barrier()                       # nobody maps the block any more
        native._check(self.lib.igg_fabric_free(key))
        del self._owned[key]

    def close(self) -> None:
        """Collective: unmap all peer memory, then free every block this fabric allocated.

        Tensors obtained from :meth:`zeros` alias those blocks and are invalid from here on."""
        torch.cuda.synchronize()
        self.g.comm.barrier()
        for base in self._opened.values():
            self.lib.igg_ipc_close(base)
        self._opened.clear()
        self._peer_ptrs.clear()
        self._handles.clear()
        self._registered.clear()
        self._root_bufs.clear()
        self.g.comm.barrier()                       # every rank has dropped its mappings
        for ptr in list(self._owned):
            self.lib.igg_fabric_free(ptr)
        self._owned.clear()
        self._keepalive.clear()


def get_fabric(g: Optional[GlobalGrid] = None) -> Fabric:
    g = g or global_grid()
    fab = g.cache.get("fabric")
    if fab is None:
        fab = Fabric(g)
        g.cache["fabric"] = fab
    return fab


def fabric_zeros(shape: Sequence[int], dtype: torch.dtype = torch.float64) -> torch.Tensor:
    """Collective allocation of a field in peer-addressable memory."""
    return get_fabric().zeros(shape, dtype)


# =====================================================================================================
class PeerHalo:
    """``update_halo`` over peer memory: one kernel launch per call (transport ``peer``)."""

    CHECK_EVERY = 256        # launches between two host-side looks at the error word

    def __init__(self, g: GlobalGrid) -> None:
        self.g = g
        self.fab = get_fabric(g)
        self.bank = self.fab.new_bank()
        self.epoch = 0
        self._tables: Dict[tuple, Tuple[torch.Tensor, int, int]] = {}
        self._sync = None
        self.launches = 0

    def _boxes(self, A: torch.Tensor):
        """``(offset, rank, src_start, dst_start, extent)`` of every neighbour box (see ``halo_boxes``)."""
        g = self.g
        ranks = dict(self.fab.neighbors)
        has_nb = lambda side, d: g.neighbors[side][d] != PROC_NULL
        for (o, src, dst, ext) in halo_boxes(g.nxyz, g.overlaps, A.shape, has_nb, list(ranks)):
            yield o, ranks[o], src, dst, ext

    def _table(self, fields: Sequence[torch.Tensor]):
        key = tuple((A.data_ptr(), tuple(A.shape), A.element_size()) for A in fields)
        hit = self._tables.get(key)
        if hit is not None:
            return hit
        for A in fields:
            if A.element_size() not in (1, 2, 4, 8):
                raise TypeError(f"the peer transport moves 1/2/4/8-byte elements, not {A.dtype}")
            self.fab.register(A)          # collective on first sight of a field
        xf: List[native.HaloXfer] = []
        for A in fields:
            shape = list(A.shape) + [1] * (NDIMS - A.dim())
            for (o, r, src, dst, ext) in self._boxes(A):
                x = native.HaloXfer()
                x.src = A.data_ptr()
                x.dst = self.fab.peer_ptr(A, r)
                for d in range(3):
                    x.n[d] = shape[d]
                    x.src_start[d] = src[d]
                    x.dst_start[d] = dst[d]
                    x.extent[d] = ext[d]
                x.elem_bytes = A.element_size()
                xf.append(x)
        n = len(xf)
        if n > 27 * native.MAX_HALO_FIELDS:
            raise ValueError(f"update_halo(peer): too many boxes in one call ({n}); pass at most "
                             f"{native.MAX_HALO_FIELDS} fields")
        # CTAs of the exchange kernel: one per 1024 copy units (16-byte vectors for contiguous runs, single
        # elements for strided planes) -- the grid follows the bytes moved, not the machine size
        units = 0
        for x in xf:
            rows, run, eb = x.extent[0] * x.extent[1], x.extent[2], x.elem_bytes
            vec = run * eb >= 64 and x.src_start[2] == x.dst_start[2] and (x.n[2] * eb) % 16 == 0
            units += rows * (1 + run * eb // 16 if vec else run)
        ctas = max(1, -(-units // 1024))
        if n:
            arr = (native.HaloXfer * n)(*xf)
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            dev = host.to(self.g.device)
        else:
            dev = torch.zeros(16, dtype=torch.uint8, device=self.g.device)
        self._tables[key] = (dev, n, ctas)
        if len(self._tables) > 64:
            self._tables.pop(next(iter(self._tables)))
        return dev, n, ctas

    def update(self, fields: Sequence[torch.Tensor]) -> None:
        dev, n, ctas = self._table(fields)
        if not self.fab.neighbors:
            return
        self.epoch += 1
        if self._sync is None:              # the neighbour banks never change: build the struct once
            self._sync = self.fab.sync_struct(self.bank, self.epoch)
            self._sync_ref = C.byref(self._sync)
        self._sync.epoch = self.epoch
        native._check(self.fab.lib.igg_halo_exchange_peer_ex(dev.data_ptr(), n, self._sync_ref, ctas,
                                                             native._stream()))
        self.launches += 1
        if self.launches % self.CHECK_EVERY == 0:
            # a flag wait that timed out leaves the halo stale (the kernel then stores nothing): do not let a
            # long loop run on silently -- the host looks at the error word now and then, and at every
            # toc() / gather() / finalize_global_grid()
            self.fab.check_errors(self.bank)

    def forget(self, data_ptr: int) -> None:
        """Drop cached transfer tables that mention the field at ``data_ptr`` (its registration ended)."""
        for key in [k for k in self._tables if any(f[0] == data_ptr for f in k)]:
            del self._tables[key]

    def close(self) -> None:
        self._tables.clear()


# =====================================================================================================
class PushPlan:
    """Descriptor factory for stencil kernels that push their own edge cells (2-D and 3-D).

    Usage per time loop::

        plan = PushPlan(T, T2)       # collective (registers the fields, takes a flag bank)
        plan.begin()                 # halos made consistent + neighbour barrier
        for it in range(nt):
            native.diffusion2d_step(T2, T, Cp, ..., push=plan.next(T2))   # pushes INTO T2's peers
            T, T2 = T2, T
        plan.finish()                # all incoming pushes have landed; error word checked
    """

    def __init__(self, *fields: torch.Tensor) -> None:
        g = global_grid()
        self.g = g
        self.fab = get_fabric(g)
        self.bank = self.fab.new_bank()
        self.epoch = 0
        self.fields = fields
        self.ndim = fields[0].dim()
        if self.ndim not in (2, 3):
            raise ValueError("PushPlan supports 2-D and 3-D fields")
        if any(g.overlaps[d] != 2 for d in range(self.ndim)):
            # the in-kernel push sends index 1 / n-2 into index n-1 / 0: overlap 2, halo width 1
            raise ValueError("the fused push needs overlaps == 2; use update_halo() for wider overlaps")
        for A in fields:
            if tuple(A.shape) != tuple(g.nxyz[:self.ndim]):
                raise ValueError("pushed fields must have the local grid size (no staggering)")
            self.fab.register(A)
        self._desc: Dict[int, object] = {}
        for A in fields:
            self._desc[A.data_ptr()] = self._build(A)
        self._active = False
        self._mid: Optional[native.Mid2D] = None
        self._desc2: Optional[Dict[int, object]] = None      # built by mid_lines()

    def mid_lines(self) -> "native.Mid2D":
        """Collective on first use: the mid-line buffers and flags of the two-steps-per-pass kernel
        (``igg_diffusion2d_midlines`` / ``igg_diffusion2d_step2``; sides x-low, x-high, y-low, y-high)."""
        if self._mid is None:
            if self.ndim != 2:
                raise ValueError("mid lines exist for 2-D fields only")
            fab, bank = self.fab, self.bank
            nx, ny = self.fields[0].shape
            buf = fab.zeros((2 * nx + 2 * ny,))          # [xlo | xhi | ylo | yhi]
            self._mid_buf = buf
            m = native.Mid2D()
            m.my_mid = buf.data_ptr()
            ranks = dict(fab.neighbors)
            for s, o in enumerate([(-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0)]):
                r = ranks.get(o)
                if r is None:
                    continue
                m.peer_mid[s] = fab.peer_ptr(buf, r)
                # the neighbour on side s sees this rank on its opposite side (s ^ 1)
                m.peer_flag[s] = fab.bank_ptr(bank, r) + (native.SLOT_MID + (s ^ 1)) * 8
                m.my_flag[s] = fab.bank_ptr(bank) + (native.SLOT_MID + s) * 8
                m.my_count[s] = fab.count_bank_ptr(bank) + (native.SLOT_MID + s) * 8
            self._mid = m
            # Arrival counters are compared modulo the number of edge tiles of the kernel that bumps them, and
            # the two-step kernel tiles the field differently from the one-step kernel: it gets its own
            # counters (counter-pad slots SLOT_READY.., unused otherwise) so the two can alternate.
            self._desc2 = {A.data_ptr(): self._build(A, count_base=native.SLOT_READY) for A in self.fields}
        return self._mid

    def _build(self, A: torch.Tensor, count_base: int = 0):
        fab, bank = self.fab, self.bank
        my_bank, cnt_bank = fab.bank_ptr(bank), fab.count_bank_ptr(bank)
        d = native.Push2D() if self.ndim == 2 else native.Push3D()
        for (o, r) in fab.neighbors:
            k3 = offset_slot(o)
            k = (o[0] + 1) * 3 + (o[1] + 1) if self.ndim == 2 else k3
            d.peer_dst[k] = fab.peer_ptr(A, r)
            d.peer_flag[k] = fab.bank_ptr(bank, r) + (native.SLOT_DATA + (26 - k3)) * 8
            d.my_flag[k] = my_bank + (native.SLOT_DATA + k3) * 8
            d.my_count[k] = cnt_bank + (count_base + k3) * 8
        d.err_word = my_bank + native.SLOT_ERR * 8
        d.epoch = 0
        return d

    def begin(self, sync_halos: bool = True) -> None:
        """Make the halos of all plan fields consistent, then barrier with the neighbours."""
        from .halo import update_halo
        if sync_halos:
            update_halo(*self.fields, transport="peer")
        b = self.bank * native.BANK_STRIDE
        self.fab.count_pad[b:b + native.BANK_STRIDE].zero_()
        if self.fab.neighbors:
            self.g.comm.barrier()      # hosts first, so the device barrier below never waits long
            self.epoch += 1
            sync = self.fab.sync_struct(self.bank, self.epoch)
            native._check(self.fab.lib.igg_neighbor_barrier(C.byref(sync), native._stream()))
        self._active = True

    def next(self, dst: torch.Tensor, two_step: bool = False):
        """Descriptor for the next step (or two-step pass), whose results are written to ``dst`` (and its peers)."""
        if not self._active:
            raise RuntimeError("PushPlan.begin() must be called before the time loop")
        if not self.fab.neighbors:
            return None
        if two_step and self._desc2 is None:
            raise RuntimeError("call PushPlan.mid_lines() (collective) before asking for two-step descriptors")
        self.epoch += 1
        d = (self._desc2 if two_step else self._desc)[dst.data_ptr()]
        d.epoch = self.epoch
        return d

    def close(self) -> None:
        """Collective: give the flag bank (and the mid-line buffer) back; the plan is unusable afterwards."""
        if self.bank is None:
            return
        if self._active:
            self.finish()
        if getattr(self, "_mid_buf", None) is not None:
            self.fab.free(self._mid_buf)
            self._mid_buf = None
        self.fab.release_bank(self.bank)
        self.bank = None

    def finish(self, check: bool = True) -> None:
        """Wait until every neighbour's pushes of the last epoch have landed here."""
        if self.fab.neighbors and self._active:
            sync = self.fab.sync_struct(self.bank, self.epoch)
            native._check(self.fab.lib.igg_wait_epoch(C.byref(sync), native._stream()))
        self._active = False
        if check:
            self.fab.check_errors(self.bank)
