"""``Diffusion2D``: the user-facing time stepper behind the perf / perf_hide applications.

One object owns the fields (``T``, ``T2``, ``Cp``), the communication plan and the streams, and
offers three ways to advance, all producing the same field:

``fused``  ONE kernel per step: boundary-frame tiles first, their edge cells stored straight into
           the neighbours' halo cells (NVLink peer stores) and the epoch flag released, while the
           same launch goes on to compute the interior.  No pack, no library call, no unpack.
``hide``   the reference's intended ``hide_communication`` structure
           (``scripts/diffusion_2D_perf_hide.jl:94-101``): boundary frame of width ``b_width`` on a
           high-priority stream (pushing edges as it goes), interior on a low-priority stream,
           join, swap.
``hide_exchange``  the same structure spelled the way the reference comments it (variant (3),
           ``scripts/diffusion_2D_perf_hide.jl:94-101``): boundary frame kernel, then ``update_halo(T2)`` over the
           chosen transport on the high-priority stream WHILE the interior kernel runs on the low-priority one,
           join, swap.  This is the overlap form for transports without peer stores (``nccl`` across nodes, ``host``)
           (``tests/test_hide_drivers_gpu.py``).
``persistent``  for small local grids: ``step(n)`` is ONE cooperative launch that runs all ``n`` steps with
           in-kernel grid barriers and the same in-kernel halo pushes (no per-step launch latency).
``resident``  for grids that fit in the shared memory of one thread-block cluster (up to ~256 x 256): the
           field stays in shared memory for all ``n`` steps of a ``step(n)`` call; CTAs exchange rows through
           distributed shared memory and cluster barriers, GPUs through peer stores and per-CTA flags.
``temporal``  two steps per pass over memory -- step k+1 lives in shared memory only -- with the neighbours'
           step-(k+1) edge lines sent half way by a small kernel; an odd last step falls back to the fused
           kernel.  The fastest mode on large grids (0.52 vs 1.01 ms per step at 16384^2; needs an even ny) and
           what ``bench.py`` times.
``plain``  the reference's serial structure (``scripts/diffusion_2D_perf.jl:49-51``): full kernel,
           swap, ``update_halo(T)`` over the chosen transport (``nccl`` = the same-box baseline).

``step_controlled`` is the end-to-end form used by ``bench.py``: per step the host hands the step
parameters over in pinned memory (asynchronous H2D copy, read by the kernel from device memory) and
reads back the step's diagnostic ``[max(T), sum(T)]`` (D2H into pinned memory).
"""
from __future__ import annotations

import os
from typing import Optional, Sequence, Tuple

import torch

from ..ops import native
from ..ops import reference as ref
from ..parallel import coords_g, global_grid, nx_g, ny_g, update_halo
from ..parallel.fabric import PushPlan, fabric_zeros
from ..parallel.halo import default_transport

__all__ = ["Diffusion2D"]


class Diffusion2D:
    def __init__(self, nx: int, ny: int, *, lx: float = 10.0, ly: float = 10.0, lam: float = 1.0, Cp0: float = 1.0,
                 mode: str = "fused", variant: str = "tma", b_width: Sequence[int] = (16, 4),
                 transport: Optional[str] = None, host_ic: bool = False, form: str = "perf") -> None:
        g = global_grid()
        if g.device_type != "cuda":
            raise RuntimeError("Diffusion2D runs on CUDA devices; use models.diffusion2d_ap on CPU")
        if (nx, ny) != tuple(g.nxyz[:2]):
            raise ValueError("nx, ny must be the local sizes given to init_global_grid")
        if mode not in ("fused", "hide", "hide_exchange", "plain", "persistent", "resident", "temporal"):
            raise ValueError("mode must be fused | hide | hide_exchange | plain | persistent | resident | temporal")
        if form not in ("perf", "kp"):
            raise ValueError("form must be 'perf' or 'kp'")
        if form == "kp" and mode not in ("persistent", "resident"):
            raise ValueError("form='kp' (the kp application's arithmetic) exists for mode='persistent' and 'resident'")
        self.form = form
        if mode == "temporal" and variant != "tma":
            raise ValueError("mode='temporal' is a TMA kernel (variant='tma', even ny)")
        native.require()
        self.g = g
        self.nx, self.ny = nx, ny
        self.lx, self.ly, self.lam, self.Cp0 = lx, ly, lam, Cp0
        self.dx, self.dy = lx / nx_g(), ly / ny_g()
        self._dx, self._dy = 1.0 / self.dx, 1.0 / self.dy
        self.dt = ref.time_step(self.dx, self.dy, Cp0, lam)
        self.mode, self.variant = mode, variant
        self.b_width = (int(b_width[0]), int(b_width[1]))
        dev = g.device
        self.transport = transport or default_transport(torch.empty(0, device=dev))
        self.uses_peer = mode in ("fused", "hide", "persistent", "resident", "temporal") or self.transport == "peer"
        if mode in ("persistent", "resident"):
            variant = self.variant = "direct"      # the multi-step kernel reads the field through L2 directly
        if variant == "tma" and ny % 2:
            raise ValueError("variant 'tma' needs an even ny (16-byte row pitch); use variant='direct'")
        # fields: peer-addressable when any peer path is in use
        if self.uses_peer:
            self.T = fabric_zeros((nx, ny))
            self.T2 = fabric_zeros((nx, ny))
        else:
            self.T = torch.zeros(nx, ny, dtype=torch.float64, device=dev)
            self.T2 = torch.zeros(nx, ny, dtype=torch.float64, device=dev)
        self.Cp = torch.full((nx, ny), float(Cp0), dtype=torch.float64, device=dev)
        self._init_condition(host_ic)
        self.T2.copy_(self.T)                      # reference scripts/diffusion_2D_perf.jl:36
        self.plan: Optional[PushPlan] = None
        if mode in ("fused", "hide", "persistent", "resident", "temporal"):
            self.plan = PushPlan(self.T, self.T2)
        self._mid = self.plan.mid_lines() if (mode == "temporal" and self.plan.fab.neighbors) else None
        self._rflags = None
        self._repoch = 0
        if mode == "resident":
            self._rflags = fabric_zeros((9 * 16 + 16,), torch.int64)      # per-(offset, CTA) epoch flags
            # landing buffers of the low-latency exchange: [parity][offset][index along the edge][2 words]
            self._rll = os.environ.get("IGG_RESIDENT_LL", "1") != "0"
            self._rland_len = max(nx, ny)
            self._rland = fabric_zeros((2 * 9 * self._rland_len * 2,), torch.int64) if self._rll else None
        self._closed = False
        self._gbar = torch.zeros(1, dtype=torch.int64, device=dev)
        self._gbar_base = 0
        self._running = False
        self.steps_done = 0
        # streams for the hide structure (reference perf_hide.jl:63-68: HIGH / LOW priority queues)
        self._s_hi = self._s_lo = None
        if mode in ("hide", "hide_exchange"):
            self._s_hi = torch.cuda.Stream(device=dev, priority=-1)      # -1 = high, 0 = default
            self._s_lo = torch.cuda.Stream(device=dev, priority=0)
            self._ev_hi = torch.cuda.Event()
            self._ev_lo = torch.cuda.Event()
            self._ev_main = torch.cuda.Event()
        # end-to-end control path
        self._params_dev = torch.zeros(4, dtype=torch.float64, device=dev)
        self._params_pin = torch.zeros(4, dtype=torch.float64).pin_memory()
        self._diag_dev = torch.zeros(2, dtype=torch.float64, device=dev)
        self._diag_pin = torch.zeros(2, dtype=torch.float64).pin_memory()
        # pipelined control loop (step_controlled(pipelined=True)): two slots of pinned parameters / diagnostics
        # and a completion event each, so that step k+1 is queued before the host waits for step k
        self._ctl_slots = None
        self._ctl_pending = []       # [(slot index, is_pair)] oldest first
        self._ctl_next = 0
        if mode == "temporal":       # a pass hands over two steps' parameters and returns two diagnostics
            self._params2_dev = torch.zeros(8, dtype=torch.float64, device=dev)
            self._params2_pin = torch.zeros(8, dtype=torch.float64).pin_memory()
            self._diag2_dev = torch.zeros(4, dtype=torch.float64, device=dev)
            self._diag2_pin = torch.zeros(4, dtype=torch.float64).pin_memory()

    # -- set-up -----------------------------------------------------------------------------------
    def _init_condition(self, host_ic: bool) -> None:
        g = self.g
        probe = torch.empty((self.nx, self.ny), device="meta")
        if host_ic or g.periods[0] or g.periods[1]:
            # the reference's way: host comprehension, then upload (scripts/diffusion_2D_ap.jl:28)
            xs = coords_g(0, self.dx, probe)
            ys = coords_g(1, self.dy, probe)
            self.T.copy_(ref.gaussian_ic(xs, ys, self.dx, self.dy, self.lx, self.ly))
        else:
            x0 = float(coords_g(0, self.dx, probe)[0])
            y0 = float(coords_g(1, self.dy, probe)[0])
            native.gaussian_ic_2d(self.T, x0, y0, self.dx, self.dy, self.lx, self.ly)

    def begin(self) -> None:
        """Collective: make halos consistent and open the push epoch sequence."""
        if self.plan is not None:
            self.plan.begin()
        elif self.g.nprocs > 1 or any(self.g.periods):
            update_halo(self.T, self.T2, transport=self.transport)
        self._running = True

    def finish(self) -> None:
        """Collective: all pushes have landed; after this the halos of ``T`` are valid."""
        self.drain_controlled()
        if self._rflags is not None and self._running:
            torch.cuda.current_stream().synchronize()
            if int(self._rflags[9 * 16].item()) != 0:
                self._rflags[9 * 16] = 0
                raise RuntimeError("a flag wait of the cluster-resident kernel timed out")
        if self._s_hi is not None:
            torch.cuda.current_stream().wait_stream(self._s_hi)
            torch.cuda.current_stream().wait_stream(self._s_lo)
        if self.plan is not None and self._running:
            self.plan.finish()
        self._running = False

    def close(self) -> None:
        """Collective: release what this solver took from the fabric (flag bank, mid lines, the peer-mapped
        fields ``T`` / ``T2`` and the resident-mode buffers).  The solver is unusable afterwards; a sweep that
        builds many solvers inside one grid calls this between them."""
        if self._closed:
            return
        if self._running:
            self.finish()
        self._closed = True
        if self.plan is not None:
            fab = self.plan.fab
            self.plan.close()
            self.plan = None
        if self.uses_peer:
            from ..parallel.fabric import get_fabric
            fab = get_fabric(self.g)
            for name in ("T", "T2", "_rflags", "_rland"):
                t = getattr(self, name, None)
                if t is not None and t.data_ptr() in fab._owned:
                    fab.free(t)
                setattr(self, name, None)
        self.T = self.T2 = self.Cp = None

    # -- stepping ---------------------------------------------------------------------------------
    def _kernel(self, mode: int, push, dev_params=None, diag_out=None) -> None:
        native.diffusion2d_step(self.T2, self.T, self.Cp, self.lam, self.dt, self._dx, self._dy, mode=mode,
                                b_width=self.b_width, variant=self.variant, push=push, dev_params=dev_params,
                                diag_out=diag_out)

    def step(self, n: int = 1, dev_params: Optional[torch.Tensor] = None,
             diag_out: Optional[torch.Tensor] = None) -> None:
        """Advance ``n`` steps.  ``diag_out`` (fused mode, TMA kernel): device ``[max, sum]`` of the cells
        updated by the step, computed inside the stencil kernel."""
        if not self._running:
            self.begin()
        if diag_out is not None and not self.fused_diag:
            raise ValueError("in-kernel diagnostics need mode='fused' or 'plain' with variant='tma'")
        if self.mode == "resident":
            if dev_params is not None or diag_out is not None:
                raise ValueError("mode='resident' takes its parameters by value and has no in-kernel diagnostic")
            if n <= 0:
                return
            fab = self.plan.fab
            pT = pT2 = pF = pL = None
            if fab.neighbors:
                pT, pT2, pF, pL = [0] * 9, [0] * 9, [0] * 9, [0] * 9
                for (o, r) in fab.neighbors:
                    k = (o[0] + 1) * 3 + (o[1] + 1)
                    pT[k], pT2[k] = fab.peer_ptr(self.T, r), fab.peer_ptr(self.T2, r)
                    pF[k] = fab.peer_ptr(self._rflags, r)
                    if self._rll:
                        pL[k] = fab.peer_ptr(self._rland, r)
                if not self._rll:
                    pL = None
            native.diffusion2d_resident(self.T, self.T2, self.Cp, self.lam, self.dt, self._dx, self._dy, n, pT, pT2,
                                        pF, self._rflags.data_ptr(), self._repoch + 1, pL,
                                        self._rland.data_ptr() if (self._rll and pL is not None) else 0,
                                        self._rland_len, form=self.form)
            self._repoch += n
            if n % 2:
                self.T, self.T2 = self.T2, self.T
            self.steps_done += n
            return
        if self.mode == "temporal":
            if diag_out is not None:
                raise ValueError("mode='temporal' has no in-kernel diagnostic")
            ph = (self.lam, self.dt, self._dx, self._dy)
            while n >= 2:
                push = self.plan.next(self.T2, two_step=True)     # ONE epoch per pass of two steps
                if push is not None:
                    native.diffusion2d_midlines(self.T, self.Cp, *ph, push=push, mid=self._mid, dev_params=dev_params)
                native.diffusion2d_step2(self.T2, self.T, self.Cp, *ph, push=push,
                                         mid=self._mid if push is not None else None, dev_params=dev_params)
                self.T, self.T2 = self.T2, self.T
                self.steps_done += 2
                n -= 2
            if n == 1:
                self._kernel(0, self.plan.next(self.T2), dev_params)
                self.T, self.T2 = self.T2, self.T
                self.steps_done += 1
            return
        if self.mode == "persistent":
            if dev_params is not None:
                raise ValueError("mode='persistent' takes its parameters by value")
            if n <= 0:
                return
            pe = po = None
            if self.plan.fab.neighbors:
                pe, po = self.plan._desc[self.T2.data_ptr()], self.plan._desc[self.T.data_ptr()]
                pe.epoch = self.plan.epoch + 1
            grid = native.diffusion2d_multistep(self.T, self.T2, self.Cp, self.lam, self.dt, self._dx, self._dy, n,
                                                self._gbar, self._gbar_base, pe, po, form=self.form)
            self._gbar_base += n * grid
            self.plan.epoch += n if self.plan.fab.neighbors else 0
            if n % 2:
                self.T, self.T2 = self.T2, self.T
            self.steps_done += n
            return
        if self.mode == "fused" and n > 1 and dev_params is None and diag_out is None:
            # native time loop: same launches, no per-step interpreter overhead
            pe = po = None
            if self.plan.fab.neighbors:
                pe, po = self.plan._desc[self.T2.data_ptr()], self.plan._desc[self.T.data_ptr()]
                pe.epoch = self.plan.epoch + 1
                self.plan.epoch += n
            native.diffusion2d_run(self.T, self.T2, self.Cp, self.lam, self.dt, self._dx, self._dy, n,
                                   b_width=self.b_width, variant=self.variant, push_even=pe, push_odd=po)
            if n % 2:
                self.T, self.T2 = self.T2, self.T
            self.steps_done += n
            return
        for _ in range(n):
            if self.mode == "fused":
                self._kernel(0, self.plan.next(self.T2), dev_params, diag_out)
                self.T, self.T2 = self.T2, self.T
            elif self.mode == "hide":
                main = torch.cuda.current_stream()
                self._ev_main.record(main)
                self._s_hi.wait_event(self._ev_main)
                self._s_lo.wait_event(self._ev_main)
                with torch.cuda.stream(self._s_hi):
                    self._kernel(1, self.plan.next(self.T2), dev_params)     # frame + push
                    self._ev_hi.record(self._s_hi)
                with torch.cuda.stream(self._s_lo):
                    self._kernel(2, None, dev_params)                       # interior
                    self._ev_lo.record(self._s_lo)
                main.wait_event(self._ev_hi)
                main.wait_event(self._ev_lo)
                self.T, self.T2 = self.T2, self.T
            elif self.mode == "hide_exchange":
                main = torch.cuda.current_stream()
                self._ev_main.record(main)
                self._s_hi.wait_event(self._ev_main)
                self._s_lo.wait_event(self._ev_main)
                with torch.cuda.stream(self._s_hi):
                    self._kernel(1, None, dev_params)                       # boundary frame
                    if self.g.nprocs > 1 or any(self.g.periods):
                        update_halo(self.T2, transport=self.transport)      # exchange of the NEW field ...
                    self._ev_hi.record(self._s_hi)
                with torch.cuda.stream(self._s_lo):
                    self._kernel(2, None, dev_params)                       # ... while the interior is computed
                    self._ev_lo.record(self._s_lo)
                main.wait_event(self._ev_hi)
                main.wait_event(self._ev_lo)
                self.T, self.T2 = self.T2, self.T
            else:
                self._kernel(0, None, dev_params, diag_out)
                self.T, self.T2 = self.T2, self.T
                if self.g.nprocs > 1 or any(self.g.periods):
                    update_halo(self.T, transport=self.transport)
            self.steps_done += 1

    def step_controlled(self, lam: Optional[float] = None, dt: Optional[float] = None, *,
                        pipelined: bool = False) -> Optional[Tuple[float, float]]:
        """One end-to-end step: parameters in through pinned memory, ``(max, sum)`` of the new field out.

        The diagnostic covers the cells this rank updated (its halo-free interior), i.e. the local part of
        ``maximum(T_v)`` / ``sum(T_v)`` of the gathered field.

        ``pipelined=True`` (fused TMA kernel): the step is queued and the call returns the diagnostic of the
        PREVIOUS pipelined step (``None`` the first time); ``drain_controlled()`` returns what is still in
        flight.  Every step still moves 32 B in and 16 B out, but the device never idles waiting for the host --
        the reference's loop does not block the device per step either (``scripts/diffusion_2D_perf.jl:47-52``)."""
        if pipelined:
            if not (self.mode in ("fused", "temporal") and self.variant == "tma"):
                raise ValueError("pipelined control needs mode='fused' or 'temporal' with variant='tma'")
            return self._step_controlled_pipelined(lam, dt, pair=False)
        if self._ctl_pending:
            raise RuntimeError("pipelined steps are in flight: call drain_controlled() first")
        p = self._params_pin
        p[0] = self.lam if lam is None else lam
        p[1] = self.dt if dt is None else dt
        p[2] = self._dx
        p[3] = self._dy
        if self.mode in ("fused", "temporal") and self.variant == "tma":
            # the whole step -- H2D 32 B, fused kernel with in-kernel diagnostic, D2H 16 B, sync -- is one
            # native call (in mode 'temporal' this is the single step that completes an odd count)
            if not self._running:
                self.begin()
            native.diffusion2d_step_controlled(self.T2, self.T, self.Cp, p, self._params_dev, self._diag_dev,
                                               self._diag_pin, b_width=self.b_width, push=self.plan.next(self.T2))
            self.T, self.T2 = self.T2, self.T
            self.steps_done += 1
            return float(self._diag_pin[0]), float(self._diag_pin[1])
        self._params_dev.copy_(p, non_blocking=True)                   # H2D, 32 bytes
        if self.fused_diag:
            self.step(1, dev_params=self._params_dev, diag_out=self._diag_dev)   # diagnostic inside the kernel
        else:
            self.step(1, dev_params=self._params_dev)
            native.reduce_max_sum(native.extract_inner(self.T).view(-1), self._diag_dev)
        self._diag_pin.copy_(self._diag_dev, non_blocking=True)        # D2H, 16 bytes
        torch.cuda.current_stream().synchronize()
        return float(self._diag_pin[0]), float(self._diag_pin[1])

    # -- pipelined control loop -------------------------------------------------------------------------------
    def _ctl_slot(self):
        if self._ctl_slots is None:
            self._ctl_slots = [(torch.zeros(8, dtype=torch.float64).pin_memory(),
                                torch.zeros(4, dtype=torch.float64).pin_memory(), native.DoneEvent())
                               for _ in range(2)]
        k = self._ctl_next
        self._ctl_next = (k + 1) % 2
        return k, self._ctl_slots[k]

    def _ctl_collect(self):
        k, pair = self._ctl_pending.pop(0)
        _, d, ev = self._ctl_slots[k]
        ev.wait()
        if pair:
            return (float(d[0]), float(d[1])), (float(d[2]), float(d[3]))
        return float(d[0]), float(d[1])

    def _step_controlled_pipelined(self, lam, dt, pair: bool):
        """Queue one controlled step (or pair), THEN wait for the previous one and return its diagnostic
        (``None`` on the first call).  Same H2D / D2H traffic per step as the blocking form; the host sees every
        step's ``(max, sum)`` one submission late, and the GPU never waits for the host."""
        if not self._running:
            self.begin()
        k, (p, d, ev) = self._ctl_slot()
        if pair:
            for q in range(2):
                p[4 * q + 0] = self.lam if lam is None else lam[q]
                p[4 * q + 1] = self.dt if dt is None else dt[q]
                p[4 * q + 2] = self._dx
                p[4 * q + 3] = self._dy
            push = self.plan.next(self.T2, two_step=True)
            native.diffusion2d_step2_controlled_submit(self.T2, self.T, self.Cp, p, self._params2_dev,
                                                       self._diag2_dev, d, ev, push=push,
                                                       mid=self._mid if push is not None else None)
            self.steps_done += 2
        else:
            p[0] = self.lam if lam is None else lam
            p[1] = self.dt if dt is None else dt
            p[2] = self._dx
            p[3] = self._dy
            native.diffusion2d_step_controlled_submit(self.T2, self.T, self.Cp, p, self._params_dev, self._diag_dev,
                                                      d, ev, b_width=self.b_width, push=self.plan.next(self.T2))
            self.steps_done += 1
        self.T, self.T2 = self.T2, self.T
        self._ctl_pending.append((k, pair))
        if len(self._ctl_pending) > 1:
            return self._ctl_collect()
        return None

    def drain_controlled(self):
        """Wait for every submitted controlled step; returns their diagnostics, oldest first."""
        out = []
        while self._ctl_pending:
            out.append(self._ctl_collect())
        return out

    def step_controlled_pair(self, lam: Optional[Sequence[float]] = None, dt: Optional[Sequence[float]] = None, *,
                             pipelined: bool = False):
        """``mode='temporal'``: one end-to-end PAIR of steps.  The host hands over both steps' parameters in pinned
        memory (64 B) and reads back ``((max, sum) after step k+1, (max, sum) after step k+2)`` (32 B); it decides
        every two steps instead of every step.  ``pipelined=True``: as in ``step_controlled``."""
        if self.mode != "temporal":
            raise ValueError("step_controlled_pair needs mode='temporal'")
        if pipelined:
            return self._step_controlled_pipelined(lam, dt, pair=True)
        if self._ctl_pending:
            raise RuntimeError("pipelined steps are in flight: call drain_controlled() first")
        if not self._running:
            self.begin()
        p = self._params2_pin
        for k in range(2):
            p[4 * k + 0] = self.lam if lam is None else lam[k]
            p[4 * k + 1] = self.dt if dt is None else dt[k]
            p[4 * k + 2] = self._dx
            p[4 * k + 3] = self._dy
        push = self.plan.next(self.T2, two_step=True)
        native.diffusion2d_step2_controlled(self.T2, self.T, self.Cp, p, self._params2_dev, self._diag2_dev,
                                            self._diag2_pin, push=push, mid=self._mid if push is not None else None)
        self.T, self.T2 = self.T2, self.T
        self.steps_done += 2
        d = self._diag2_pin
        return (float(d[0]), float(d[1])), (float(d[2]), float(d[3]))

    @property
    def fused_diag(self) -> bool:
        return self.variant == "tma" and self.mode in ("fused", "plain")

    h2d_bytes_per_step = 32
    d2h_bytes_per_step = 16

    def kernels_per_step(self) -> int:
        """Native kernel launches per ``step()`` (for launch accounting)."""
        if self.mode == "fused":
            return 1
        if self.mode in ("hide", "hide_exchange"):
            return 2
        return 1      # "persistent": one launch per step() call, however many steps

    # -- checkpoint / resume (the reference persists only the final PNG) -------------------
    def save(self, directory: str) -> str:
        """Collective: write ``T`` and ``Cp`` (halo included) of every rank; returns the checkpoint path.
        Call between ``finish()`` and the next ``begin()`` (pushes of the last step must have landed)."""
        from ..utils.checkpoint import save_checkpoint
        if self._running:
            self.finish()
        return save_checkpoint(directory, self.steps_done, T=self.T, Cp=self.Cp)

    def load(self, path: str) -> int:
        """Collective: restore the fields and the step counter from ``path``; returns the step."""
        from ..utils.checkpoint import load_checkpoint
        if self._running:
            self.finish()
        step, f = load_checkpoint(path, device=self.T.device)
        self.T.copy_(f["T"])
        self.T2.copy_(f["T"])
        self.Cp.copy_(f["Cp"])
        self.steps_done = step
        return step

    def max_sum(self) -> Tuple[float, float]:
        out = native.reduce_max_sum(self.T.view(-1))
        v = out.cpu()
        return float(v[0]), float(v[1])
