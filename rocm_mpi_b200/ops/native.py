"""ctypes bindings to the in-tree native library ``rocm_mpi_b200/_lib/libigg_b200.so``.

The library is plain C ABI (``include/igg_b200.h``); tensors cross the boundary as raw device
pointers and the current ``torch`` CUDA stream.  ``available()`` is the single switch the Python
layer consults: it is true when the shared object loads **and** a CUDA device is present.  On a
GPU machine a missing library is an error, not a silent fallback (``require()``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

# IGG_B200_LIB: load another build of the library (A/B runs of compile-time variants); default = the in-tree build
_LIB_PATH = os.environ.get("IGG_B200_LIB") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                            "_lib", "libigg_b200.so")
_lib = None
_load_error: Optional[str] = None

ERR_INVALID, ERR_UNSUPPORTED, ERR_TIMEOUT = -1, -2, -3
FLAG_BANKS, BANK_STRIDE = 32, 64
SLOT_DATA, SLOT_READY, SLOT_ERR, SLOT_CTA_COUNT = 0, 27, 54, 55
SLOT_MID = 56                  # mid flags (flag pad) / mid CTA counters (counter pad) of sides 0..3
MAX_HALO_FIELDS = 8


class Push2D(C.Structure):
    _fields_ = [("peer_dst", C.c_void_p * 9), ("peer_flag", C.c_void_p * 9), ("my_flag", C.c_void_p * 9),
                ("my_count", C.c_void_p * 9), ("err_word", C.c_void_p), ("epoch", C.c_uint64)]


class Push3D(C.Structure):
    _fields_ = [("peer_dst", C.c_void_p * 27), ("peer_flag", C.c_void_p * 27), ("my_flag", C.c_void_p * 27),
                ("my_count", C.c_void_p * 27), ("err_word", C.c_void_p), ("epoch", C.c_uint64)]


class Mid2D(C.Structure):
    """``igg_mid2d``: mid-line buffers and flags of the two-steps-per-pass kernel (sides: x-low, x-high, y-low, y-high)."""
    _fields_ = [("my_mid", C.c_void_p), ("peer_mid", C.c_void_p * 4), ("peer_flag", C.c_void_p * 4),
                ("my_flag", C.c_void_p * 4), ("my_count", C.c_void_p * 4)]


class StepOpts(C.Structure):
    _fields_ = [("dev_params", C.c_void_p), ("diag_out", C.c_void_p), ("diag_scratch", C.c_void_p)]


class HaloXfer(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("n", C.c_int * 3), ("src_start", C.c_int * 3),
                ("dst_start", C.c_int * 3), ("extent", C.c_int * 3), ("elem_bytes", C.c_int), ("reserved", C.c_int)]


class HaloSync(C.Structure):
    _fields_ = [("peer_bank", C.c_void_p * 27), ("my_bank", C.c_void_p), ("my_count_bank", C.c_void_p),
                ("epoch", C.c_uint64)]


class NativeError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def _declare(lib) -> None:
    i, d, vp, sz, u64 = C.c_int, C.c_double, C.c_void_p, C.c_size_t, C.c_uint64
    sig = {
        "igg_version": ([], i),
        "igg_last_error": ([], C.c_char_p),
        "igg_device_info": ([C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(sz)], i),
        "igg_launch_count": ([], C.c_ulonglong),
        "igg_diffusion2d_step": ([vp, vp, vp, i, i, d, d, d, d, C.POINTER(StepOpts), i, i, i, i,
                                 C.POINTER(Push2D), vp], i),
        "igg_diffusion2d_step_controlled": ([vp, vp, vp, i, i, vp, vp, vp, vp, vp, i, i, C.POINTER(Push2D), vp], i),
        "igg_diffusion2d_step_controlled_submit": ([vp, vp, vp, i, i, vp, vp, vp, vp, vp, i, i, C.POINTER(Push2D), vp,
                                                   vp], i),
        "igg_diffusion2d_step2_controlled_submit": ([vp, vp, vp, i, i, vp, vp, vp, vp, vp, C.POINTER(Push2D),
                                                    C.POINTER(Mid2D), vp, vp], i),
        "igg_event_create": ([C.POINTER(vp)], i),
        "igg_event_destroy": ([vp], i),
        "igg_event_wait": ([vp, i], i),
        "igg_diffusion2d_resident": ([vp, vp, vp, i, i, d, d, d, d, i, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), vp,
                                     u64, C.POINTER(vp), vp, i, C.POINTER(i), vp], i),
        "igg_diffusion2d_resident_ex": ([vp, vp, vp, i, i, d, d, d, d, i, i, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                        vp, u64, C.POINTER(vp), vp, i, C.POINTER(i), vp], i),
        "igg_diffusion2d_multistep_ex": ([vp, vp, vp, i, i, d, d, d, d, i, i, C.POINTER(Push2D), C.POINTER(vp), vp,
                                         C.c_ulonglong, C.POINTER(i), vp], i),
        "igg_diffusion2d_run": ([vp, vp, vp, i, i, d, d, d, d, i, i, i, i, C.POINTER(Push2D), C.POINTER(Push2D), vp], i),
        "igg_diffusion2d_multistep": ([vp, vp, vp, i, i, d, d, d, d, i, C.POINTER(Push2D), C.POINTER(vp), vp,
                                      C.c_ulonglong, C.POINTER(i), vp], i),
        "igg_flux2d": ([vp, vp, vp, i, i, d, d, d, vp], i),
        "igg_residual2d": ([vp, vp, vp, vp, i, i, d, d, vp], i),
        "igg_update2d": ([vp, vp, i, i, d, C.POINTER(Push2D), vp], i),
        "igg_diffusion3d_step": ([vp, vp, vp, i, i, i, d, d, d, d, d, i, i, i, i, i, C.POINTER(Push3D), vp], i),
        "igg_gaussian_ic2d": ([vp, i, i, d, d, d, d, d, d, vp], i),
        "igg_gaussian_ic3d": ([vp, i, i, i, d, d, d, d, d, d, d, d, d, vp], i),
        "igg_pack_plane": ([vp, i, i, i, i, i, vp, vp], i),
        "igg_unpack_plane": ([vp, i, i, i, i, i, vp, vp], i),
        "igg_extract_inner": ([vp, i, i, i, vp, vp], i),
        "igg_fabric_alloc": ([sz, C.POINTER(vp)], i),
        "igg_fabric_free": ([vp], i),
        "igg_ipc_export": ([vp, vp, C.POINTER(u64)], i),
        "igg_ipc_open": ([vp, C.POINTER(vp)], i),
        "igg_ipc_close": ([vp], i),
        "igg_can_access_peer": ([i, i, C.POINTER(i)], i),
        "igg_gather_inner": ([vp, i, i, i, vp, i, i, i, i, i, i, vp], i),
        "igg_peer_sendrecv": ([vp, vp, sz, vp, vp, u64, vp, vp, vp], i),
        "igg_halo_exchange_peer": ([vp, i, C.POINTER(HaloSync), vp], i),
        "igg_halo_exchange_peer_ex": ([vp, i, C.POINTER(HaloSync), i, vp], i),
        "igg_neighbor_barrier": ([C.POINTER(HaloSync), vp], i),
        "igg_wait_epoch": ([C.POINTER(HaloSync), vp], i),
        "igg_peer_copy": ([vp, vp, sz, vp, u64, vp, vp], i),
        "igg_wait_value": ([vp, u64, vp, vp], i),
        "igg_reduce_scratch_bytes": ([sz], sz),
        "igg_reduce_max_sum": ([vp, sz, vp, vp, vp], i),
        "igg_colormap": ([vp, sz, d, d, vp, vp, vp], i),
        "igg_cpu_diffusion2d_step": ([vp, vp, vp, i, i, d, d, d, d], i),
        "igg_cpu_diffusion2d_run": ([vp, vp, vp, i, i, d, d, d, d, i], i),
        "igg_cpu_flux2d": ([vp, vp, vp, i, i, d, d, d], i),
        "igg_cpu_residual2d": ([vp, vp, vp, vp, i, i, d, d], i),
        "igg_cpu_update2d": ([vp, vp, i, i, d], i),
        "igg_cpu_diffusion3d_step": ([vp, vp, vp, i, i, i, d, d, d, d, d], i),
        "igg_diffusion2d_step2": ([vp, vp, vp, i, i, d, d, d, d, C.POINTER(StepOpts), C.POINTER(Push2D),
                                  C.POINTER(Mid2D), vp], i),
        "igg_temporal2d_diag_scratch_bytes": ([], sz),
        "igg_diffusion2d_step2_controlled": ([vp, vp, vp, i, i, vp, vp, vp, vp, vp, C.POINTER(Push2D),
                                             C.POINTER(Mid2D), vp], i),
        "igg_diffusion2d_midlines": ([vp, vp, i, i, d, d, d, d, vp, C.POINTER(Push2D), C.POINTER(Mid2D), vp], i),
        "igg_temporal2d_default_rows": ([], i),
        "igg_cpu_temporal2d_emulate": ([vp, vp, vp, i, i, d, d, d, d, i, i, i, i, vp, i, vp, vp], i),
        "igg_cpu_temporal2d_midlines": ([vp, vp, i, i, d, d, d, d, C.POINTER(vp)], i),
    }
    for name, (argtypes, restype) in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype


def load(build_if_missing: bool = False):
    """Load (once) and return the ctypes library handle, or ``None`` if it cannot be loaded."""
    global _lib, _load_error
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH) and build_if_missing:
        from .. import build as _build
        _build.build()
    try:
        lib = C.CDLL(_LIB_PATH)
        _declare(lib)
        _lib = lib
    except (OSError, AttributeError) as e:  # missing file or stale library without a symbol
        _load_error = str(e)
        _lib = None
    return _lib


def loaded() -> bool:
    return load() is not None


def available() -> bool:
    """True when the CUDA kernels can run: library loaded and a CUDA device present."""
    return torch.cuda.is_available() and load() is not None


def require():
    """The native library, or a loud error (used on every GPU path: no silent PyTorch fallback)."""
    lib = load()
    if lib is None:
        raise NativeError(f"native library {_LIB_PATH} is not loadable ({_load_error}); "
                          "build it with `python -m rocm_mpi_b200.build`")
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        msg = require().igg_last_error().decode(errors="replace")
        raise NativeError(f"native call failed (rc={rc}): {msg}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _f64(t: torch.Tensor, name: str = "tensor") -> int:
    if t.dtype != torch.float64 or not t.is_cuda or not t.is_contiguous():
        raise TypeError(f"{name} must be a contiguous CUDA float64 tensor "
                        f"(got {t.dtype}, {t.device}, contiguous={t.is_contiguous()})")
    return t.data_ptr()


def launch_count() -> int:
    lib = load()
    return int(lib.igg_launch_count()) if lib is not None else 0


def device_info() -> dict:
    lib = require()
    sm, ma, mi, sh = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
    _check(lib.igg_device_info(C.byref(sm), C.byref(ma), C.byref(mi), C.byref(sh)))
    return {"sm_count": sm.value, "cc": (ma.value, mi.value), "smem_optin": sh.value}


# ---- stencils ---------------------------------------------------------------------------------------
VARIANTS = {"direct": 0, "tma": 1}
FORMS = {"perf": 0, "kp": 1}       # step arithmetic of the single-launch multi-step kernels


def diffusion2d_step(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, lam: float, dt: float, _dx: float,
                     _dy: float, mode: int = 0, b_width: Sequence[int] = (1, 1), variant: str = "tma",
                     push: Optional[Push2D] = None, dev_params: Optional[torch.Tensor] = None,
                     diag_out: Optional[torch.Tensor] = None) -> None:
    """``T2 <- step(T)``; see ``igg_diffusion2d_step`` in ``include/igg_b200.h``.

    ``dev_params``: optional device tensor ``[lam, dt, _dx, _dy]`` overriding the scalar arguments.
    ``diag_out``: optional device tensor of 2 doubles receiving ``[max, sum]`` over the cells written by
    this launch, computed inside the stencil kernel (TMA variant only)."""
    lib = require()
    nx, ny = T.shape
    if T2.shape != T.shape or Cp.shape != T.shape:
        raise ValueError("T, T2 and Cp must have the same shape")
    opts = None
    if dev_params is not None or diag_out is not None:
        opts = StepOpts()
        opts.dev_params = _f64(dev_params, "dev_params") if dev_params is not None else None
        if diag_out is not None:
            opts.diag_out = _f64(diag_out, "diag_out")
            opts.diag_scratch = _reduce_scratch(T.device).data_ptr()
    _check(lib.igg_diffusion2d_step(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny, lam, dt, _dx, _dy,
                                    C.byref(opts) if opts is not None else None,
                                    mode, int(b_width[0]), int(b_width[1]), VARIANTS[variant],
                                    C.byref(push) if push is not None else None, _stream()))


_t2_scratch = {}


def _temporal_scratch(device: torch.device) -> torch.Tensor:
    """Zero-filled once; the two-step kernel leaves it ready for the next launch."""
    key = device.index
    if key not in _t2_scratch:
        n = int(require().igg_temporal2d_diag_scratch_bytes())
        _t2_scratch[key] = torch.zeros((n + 7) // 8, dtype=torch.int64, device=device)
    return _t2_scratch[key]


def diffusion2d_step2(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, lam: float, dt: float, _dx: float,
                      _dy: float, push: Optional[Push2D] = None, mid: Optional["Mid2D"] = None,
                      dev_params: Optional[torch.Tensor] = None, diag_out: Optional[torch.Tensor] = None) -> None:
    """``T2 <- step(step(T))`` in one pass over memory (``igg_diffusion2d_step2``).  With neighbours, call
    :func:`diffusion2d_midlines` with the same descriptors first (``PushPlan.mid_lines()``).
    ``dev_params``: device tensor of 8 doubles (``[lam, dt, _dx, _dy]`` of step k+1, then of step k+2);
    ``diag_out``: device tensor of 4 doubles receiving ``[max, sum]`` of step k+1 and of step k+2.
    GPU-tested (``tests/test_temporal_gpu.py``); :func:`cpu_temporal2d_emulate` executes the same index code on the host."""
    lib = require()
    nx, ny = T.shape
    if T2.shape != T.shape or Cp.shape != T.shape:
        raise ValueError("T, T2 and Cp must have the same shape")
    opts = None
    if dev_params is not None or diag_out is not None:
        opts = StepOpts()
        if dev_params is not None:
            if dev_params.numel() < 8:
                raise ValueError("dev_params of the two-step kernel holds 8 doubles (two steps)")
            opts.dev_params = _f64(dev_params, "dev_params")
        if diag_out is not None:
            if diag_out.numel() < 4:
                raise ValueError("diag_out of the two-step kernel holds 4 doubles")
            opts.diag_out = _f64(diag_out, "diag_out")
            opts.diag_scratch = _temporal_scratch(T.device).data_ptr()
    _check(lib.igg_diffusion2d_step2(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny, lam, dt, _dx, _dy,
                                     C.byref(opts) if opts is not None else None,
                                     C.byref(push) if push is not None else None,
                                     C.byref(mid) if mid is not None else None, _stream()))


def diffusion2d_step2_controlled(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, params_pin: torch.Tensor,
                                 params_dev: torch.Tensor, diag_dev: torch.Tensor, diag_pin: torch.Tensor,
                                 push: Optional[Push2D] = None, mid: Optional["Mid2D"] = None) -> None:
    """One end-to-end pair of steps from native code (``igg_diffusion2d_step2_controlled``): H2D of 8 parameters,
    mid lines, two-step pass, D2H of the 4 diagnostic values; returns after they have arrived in ``diag_pin``."""
    lib = require()
    nx, ny = T.shape
    if not (params_pin.is_pinned() and diag_pin.is_pinned()):
        raise TypeError("params_pin and diag_pin must be pinned host tensors")
    if params_pin.numel() < 8 or params_dev.numel() < 8 or diag_dev.numel() < 4 or diag_pin.numel() < 4:
        raise ValueError("need 8 parameter doubles and 4 diagnostic doubles")
    _check(lib.igg_diffusion2d_step2_controlled(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny,
                                                params_pin.data_ptr(), _f64(params_dev), _f64(diag_dev),
                                                _temporal_scratch(T.device).data_ptr(), diag_pin.data_ptr(),
                                                C.byref(push) if push is not None else None,
                                                C.byref(mid) if mid is not None else None, _stream()))


class DoneEvent:
    """Completion event of a submitted controlled step (``igg_event_create``): ``wait()`` polls until the step's
    diagnostic has arrived in its pinned slot."""

    def __init__(self) -> None:
        h = C.c_void_p()
        _check(require().igg_event_create(C.byref(h)))
        self.handle = h.value

    def wait(self, spin: bool = True) -> None:
        _check(require().igg_event_wait(self.handle, 1 if spin else 0))

    def __del__(self) -> None:
        try:
            if self.handle and _lib is not None:
                _lib.igg_event_destroy(self.handle)
        except Exception:
            pass
        self.handle = None


def diffusion2d_step_controlled_submit(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, params_pin: torch.Tensor,
                                       params_dev: torch.Tensor, diag_dev: torch.Tensor, diag_pin: torch.Tensor,
                                       done: DoneEvent, b_width: Sequence[int] = (1, 1),
                                       push: Optional[Push2D] = None) -> None:
    """``igg_diffusion2d_step_controlled_submit``: queue H2D of the 4 parameters, fused step, D2H of ``[max, sum]``
    and record ``done``; does not wait.  ``params_pin`` / ``diag_pin`` are the slots of THIS step."""
    lib = require()
    nx, ny = T.shape
    if not (params_pin.is_pinned() and diag_pin.is_pinned()):
        raise TypeError("params_pin and diag_pin must be pinned host tensors")
    _check(lib.igg_diffusion2d_step_controlled_submit(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny,
                                                      params_pin.data_ptr(), _f64(params_dev), _f64(diag_dev),
                                                      _reduce_scratch(T.device).data_ptr(), diag_pin.data_ptr(),
                                                      int(b_width[0]), int(b_width[1]),
                                                      C.byref(push) if push is not None else None, done.handle,
                                                      _stream()))


def diffusion2d_step2_controlled_submit(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, params_pin: torch.Tensor,
                                        params_dev: torch.Tensor, diag_dev: torch.Tensor, diag_pin: torch.Tensor,
                                        done: DoneEvent, push: Optional[Push2D] = None,
                                        mid: Optional["Mid2D"] = None) -> None:
    """``igg_diffusion2d_step2_controlled_submit``: the pair form of ``diffusion2d_step_controlled_submit``
    (8 parameters in, 4 diagnostic values out)."""
    lib = require()
    nx, ny = T.shape
    if not (params_pin.is_pinned() and diag_pin.is_pinned()):
        raise TypeError("params_pin and diag_pin must be pinned host tensors")
    if params_pin.numel() < 8 or params_dev.numel() < 8 or diag_dev.numel() < 4 or diag_pin.numel() < 4:
        raise ValueError("need 8 parameter doubles and 4 diagnostic doubles")
    _check(lib.igg_diffusion2d_step2_controlled_submit(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny,
                                                       params_pin.data_ptr(), _f64(params_dev), _f64(diag_dev),
                                                       _temporal_scratch(T.device).data_ptr(), diag_pin.data_ptr(),
                                                       C.byref(push) if push is not None else None,
                                                       C.byref(mid) if mid is not None else None, done.handle,
                                                       _stream()))


def diffusion2d_midlines(T: torch.Tensor, Cp: torch.Tensor, lam: float, dt: float, _dx: float, _dy: float,
                         push: Push2D, mid: "Mid2D", dev_params: Optional[torch.Tensor] = None) -> None:
    """Send this rank's step-(k+1) edge lines to the neighbours (``igg_diffusion2d_midlines``)."""
    lib = require()
    nx, ny = T.shape
    _check(lib.igg_diffusion2d_midlines(_f64(T, "T"), _f64(Cp, "Cp"), nx, ny, lam, dt, _dx, _dy,
                                        _f64(dev_params, "dev_params") if dev_params is not None else None,
                                        C.byref(push), C.byref(mid), _stream()))


def diffusion2d_step_controlled(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, params_pin: torch.Tensor,
                                params_dev: torch.Tensor, diag_dev: torch.Tensor, diag_pin: torch.Tensor,
                                b_width: Sequence[int] = (1, 1), push: Optional[Push2D] = None) -> None:
    """One end-to-end controlled step from native code (``igg_diffusion2d_step_controlled``): H2D of the 4
    parameters, fused step, D2H of ``[max, sum]``; returns after the diagnostic has arrived in ``diag_pin``."""
    lib = require()
    nx, ny = T.shape
    if not (params_pin.is_pinned() and diag_pin.is_pinned()):
        raise TypeError("params_pin and diag_pin must be pinned host tensors")
    _check(lib.igg_diffusion2d_step_controlled(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny,
                                               params_pin.data_ptr(), _f64(params_dev), _f64(diag_dev),
                                               _reduce_scratch(T.device).data_ptr(), diag_pin.data_ptr(),
                                               int(b_width[0]), int(b_width[1]),
                                               C.byref(push) if push is not None else None, _stream()))


def diffusion2d_resident(T: torch.Tensor, T2: torch.Tensor, Cp: torch.Tensor, lam: float, dt: float, _dx: float,
                         _dy: float, nsteps: int, peer_T=None, peer_T2=None, peer_flags=None, my_flags: int = 0,
                         epoch0: int = 0, peer_land=None, my_land: int = 0, land_len: int = 0, form: str = "perf") -> int:
    """``nsteps`` steps with the field resident in one thread-block cluster (``igg_diffusion2d_resident``).

    ``peer_T`` / ``peer_T2`` / ``peer_flags``: sequences of 9 raw pointers (0 = no neighbour) indexed
    ``(ox+1)*3 + (oy+1)``, or ``None`` for a run without communication.  ``peer_land`` (9 raw pointers) +
    ``my_land`` + ``land_len`` select the low-latency exchange (values travel with their epoch tag into
    landing buffers).  Returns the cluster size used."""
    lib = require()
    nx, ny = T.shape
    csz = C.c_int(0)
    a = b = f = ld = None
    if peer_flags is not None:
        a = (C.c_void_p * 9)(*[x or None for x in peer_T])
        b = (C.c_void_p * 9)(*[x or None for x in peer_T2])
        f = (C.c_void_p * 9)(*[x or None for x in peer_flags])
    if peer_land is not None:
        ld = (C.c_void_p * 9)(*[x or None for x in peer_land])
    _check(lib.igg_diffusion2d_resident_ex(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny, lam, dt, _dx, _dy,
                                           int(nsteps), FORMS[form], a, b, f, my_flags or None, int(epoch0), ld,
                                           my_land or None, int(land_len), C.byref(csz), _stream()))
    return csz.value


def diffusion2d_run(T: torch.Tensor, T2: torch.Tensor, Cp: torch.Tensor, lam: float, dt: float, _dx: float, _dy: float,
                    nsteps: int, b_width: Sequence[int] = (1, 1), variant: str = "tma",
                    push_even: Optional[Push2D] = None, push_odd: Optional[Push2D] = None) -> None:
    """``nsteps`` fused steps launched from the native loop (``igg_diffusion2d_run``); the result is in ``T2``
    if ``nsteps`` is odd, else in ``T``."""
    lib = require()
    nx, ny = T.shape
    _check(lib.igg_diffusion2d_run(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny, lam, dt, _dx, _dy,
                                   int(b_width[0]), int(b_width[1]), VARIANTS[variant], int(nsteps),
                                   C.byref(push_even) if push_even is not None else None,
                                   C.byref(push_odd) if push_odd is not None else None, _stream()))


def diffusion2d_multistep(T: torch.Tensor, T2: torch.Tensor, Cp: torch.Tensor, lam: float, dt: float, _dx: float,
                          _dy: float, nsteps: int, grid_barrier: torch.Tensor, barrier_base: int,
                          push_even: Optional[Push2D] = None, push_odd: Optional[Push2D] = None,
                          form: str = "perf") -> int:
    """``nsteps`` fused steps in one cooperative launch (``igg_diffusion2d_multistep_ex``); ``form='kp'`` selects the
    arithmetic of the kp application (``Flux!`` / ``Residual!`` / ``Update!`` with their roundings, division by Cp).

    Returns the grid size used (the barrier counter advances by ``nsteps * grid``).  ``push_even`` /
    ``push_odd`` are the plan descriptors whose destinations are the neighbours' ``T2`` / ``T``."""
    lib = require()
    nx, ny = T.shape
    grid = C.c_int(0)
    odd = None
    if push_even is not None:
        odd = (C.c_void_p * 9)(*[push_odd.peer_dst[k] for k in range(9)])
    _check(lib.igg_diffusion2d_multistep_ex(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny, lam, dt, _dx, _dy,
                                            int(nsteps), FORMS[form],
                                            C.byref(push_even) if push_even is not None else None, odd,
                                            grid_barrier.data_ptr(), int(barrier_base), C.byref(grid), _stream()))
    return grid.value


def diffusion3d_step(T2, T, Cp, lam, dt, _dx, _dy, _dz, mode: int = 0, b_width=(1, 1, 1), variant: str = "tma",
                     push: Optional[Push3D] = None) -> None:
    lib = require()
    nx, ny, nz = T.shape
    if T2.shape != T.shape or Cp.shape != T.shape:
        raise ValueError("T, T2 and Cp must have the same shape")
    _check(lib.igg_diffusion3d_step(_f64(T, "T"), _f64(T2, "T2"), _f64(Cp, "Cp"), nx, ny, nz, lam, dt, _dx, _dy,
                                    _dz, mode, int(b_width[0]), int(b_width[1]), int(b_width[2]),
                                    VARIANTS[variant], C.byref(push) if push is not None else None, _stream()))


def flux2d(qx, qy, T, lam, _dx, _dy) -> None:
    nx, ny = T.shape
    if tuple(qx.shape) != (nx - 1, ny - 2) or tuple(qy.shape) != (nx - 2, ny - 1):
        raise ValueError("qx must be (nx-1, ny-2) and qy (nx-2, ny-1)")
    _check(require().igg_flux2d(_f64(qx), _f64(qy), _f64(T), nx, ny, lam, _dx, _dy, _stream()))


def residual2d(dTdt, qx, qy, Cp, _dx, _dy) -> None:
    nx, ny = Cp.shape
    if tuple(dTdt.shape) != (nx - 2, ny - 2):
        raise ValueError("dTdt must be (nx-2, ny-2)")
    _check(require().igg_residual2d(_f64(dTdt), _f64(qx), _f64(qy), _f64(Cp), nx, ny, _dx, _dy, _stream()))


def update2d(T, dTdt, dt, push: Optional[Push2D] = None) -> None:
    nx, ny = T.shape
    _check(require().igg_update2d(_f64(T), _f64(dTdt), nx, ny, dt, C.byref(push) if push is not None else None,
                                  _stream()))


def gaussian_ic_2d(T, x0, y0, dx, dy, lx, ly) -> None:
    nx, ny = T.shape
    _check(require().igg_gaussian_ic2d(_f64(T), nx, ny, x0, y0, dx, dy, lx, ly, _stream()))


def gaussian_ic_3d(T, x0, y0, z0, dx, dy, dz, lx, ly, lz) -> None:
    nx, ny, nz = T.shape
    _check(require().igg_gaussian_ic3d(_f64(T), nx, ny, nz, x0, y0, z0, dx, dy, dz, lx, ly, lz, _stream()))


# ---- halo helpers -----------------------------------------------------------------------------------
def _shape3(A: torch.Tensor):
    s = list(A.shape) + [1] * (3 - A.dim())
    return s[0], s[1], s[2]


def pack_plane(A, dim: int, idx: int, out) -> None:
    n0, n1, n2 = _shape3(A)
    _check(require().igg_pack_plane(_f64(A), n0, n1, n2, dim, idx, _f64(out), _stream()))


def unpack_plane(A, dim: int, idx: int, src) -> None:
    n0, n1, n2 = _shape3(A)
    _check(require().igg_unpack_plane(_f64(A), n0, n1, n2, dim, idx, _f64(src), _stream()))


def extract_inner(A, out=None) -> torch.Tensor:
    n0, n1, n2 = _shape3(A)
    shape = tuple(s - 2 if s > 2 else s for s in A.shape)
    if out is None:
        out = torch.empty(shape, dtype=A.dtype, device=A.device)
    _check(require().igg_extract_inner(_f64(A), n0, n1, n2, _f64(out), _stream()))
    return out


def gather_inner(A: torch.Tensor, G_ptr: int, dims: Sequence[int], coords: Sequence[int]) -> None:
    """Store the halo-free interior of ``A`` as block ``coords`` of the global array at ``G_ptr`` (usually a
    peer-mapped pointer into the root's memory); ``igg_gather_inner``."""
    n0, n1, n2 = _shape3(A)
    d = list(dims) + [1] * (3 - len(dims))
    c = list(coords) + [0] * (3 - len(coords))
    _check(require().igg_gather_inner(_f64(A), n0, n1, n2, G_ptr, int(d[0]), int(d[1]), int(d[2]), int(c[0]),
                                      int(c[1]), int(c[2]), _stream()))


# ---- diagnostics ------------------------------------------------------------------------------------
_reduce_ws = {}
_diag_ws = {}


def _reduce_scratch(device: torch.device) -> torch.Tensor:
    """Zero-initialised scratch for the in-kernel diagnostic (one per device, separate from the one used
    by :func:`reduce_max_sum` so the two kinds of launches may overlap)."""
    ws = _diag_ws.get(device.index)
    if ws is None:
        ws = torch.zeros(int(require().igg_reduce_scratch_bytes(0)), dtype=torch.uint8, device=device)
        _diag_ws[device.index] = ws
    return ws


def reduce_max_sum(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Device tensor ``[max, sum]`` of ``x`` (asynchronous; scratch is cached per device)."""
    lib = require()
    dev = x.device.index
    ws = _reduce_ws.get(dev)
    if ws is None:
        ws = torch.zeros(int(lib.igg_reduce_scratch_bytes(x.numel())), dtype=torch.uint8, device=x.device)
        _reduce_ws[dev] = ws
    if out is None:
        out = torch.empty(2, dtype=torch.float64, device=x.device)
    _check(lib.igg_reduce_max_sum(_f64(x), x.numel(), ws.data_ptr(), _f64(out), _stream()))
    return out


def colormap(field: torch.Tensor, vmin: float, vmax: float, lut: torch.Tensor) -> torch.Tensor:
    lib = require()
    lut_d = lut.to(device=field.device, dtype=torch.uint8).contiguous()
    rgb = torch.empty(tuple(field.shape) + (3,), dtype=torch.uint8, device=field.device)
    _check(lib.igg_colormap(_f64(field.contiguous()), field.numel(), vmin, vmax, lut_d.data_ptr(), rgb.data_ptr(),
                            _stream()))
    return rgb


# ---- CPU restatement --------------------------------------------------------------------------------
def cpu_diffusion2d_step(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, lam, dt, _dx, _dy) -> None:
    lib = require()
    for t in (T2, T, Cp):
        if t.is_cuda or t.dtype != torch.float64 or not t.is_contiguous():
            raise TypeError("cpu_diffusion2d_step needs contiguous CPU float64 tensors")
    nx, ny = T.shape
    _check(lib.igg_cpu_diffusion2d_step(T.data_ptr(), T2.data_ptr(), Cp.data_ptr(), nx, ny, lam, dt, _dx, _dy))


def _cpu_f64(*ts: torch.Tensor) -> None:
    for t in ts:
        if t.is_cuda or t.dtype != torch.float64 or not t.is_contiguous():
            raise TypeError("the CPU restatements need contiguous CPU float64 tensors")


def cpu_kp_step(T: torch.Tensor, Cp: torch.Tensor, qx: torch.Tensor, qy: torch.Tensor, dTdt: torch.Tensor, lam, dt,
                _dx, _dy) -> None:
    """``Flux!``, ``Residual!``, ``Update!`` on the CPU (OpenMP), in place on ``T`` like ``ops.reference.kp_step``."""
    lib = require()
    _cpu_f64(T, Cp, qx, qy, dTdt)
    nx, ny = T.shape
    if tuple(qx.shape) != (nx - 1, ny - 2) or tuple(qy.shape) != (nx - 2, ny - 1) or tuple(dTdt.shape) != (nx - 2, ny - 2):
        raise ValueError("qx, qy, dTdt must be (nx-1, ny-2), (nx-2, ny-1), (nx-2, ny-2)")
    _check(lib.igg_cpu_flux2d(qx.data_ptr(), qy.data_ptr(), T.data_ptr(), nx, ny, lam, _dx, _dy))
    _check(lib.igg_cpu_residual2d(dTdt.data_ptr(), qx.data_ptr(), qy.data_ptr(), Cp.data_ptr(), nx, ny, _dx, _dy))
    _check(lib.igg_cpu_update2d(T.data_ptr(), dTdt.data_ptr(), nx, ny, dt))


def cpu_diffusion3d_step(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, lam, dt, _dx, _dy, _dz) -> None:
    lib = require()
    _cpu_f64(T2, T, Cp)
    nx, ny, nz = T.shape
    _check(lib.igg_cpu_diffusion3d_step(T.data_ptr(), T2.data_ptr(), Cp.data_ptr(), nx, ny, nz, lam, dt, _dx, _dy, _dz))


def temporal2d_default_rows() -> int:
    return int(require().igg_temporal2d_default_rows())


def cpu_temporal2d_emulate(T2: torch.Tensor, T: torch.Tensor, Cp: torch.Tensor, lam, dt, _dx, _dy, scx: int = 252,
                           rpt: int = 2, workers: int = 148, mid_lines: Optional[torch.Tensor] = None,
                           side_mask: int = 0, phys2: Optional[Sequence[float]] = None,
                           diag_out: Optional[torch.Tensor] = None, srb: int = 16) -> None:
    """Run the index code of the two-steps-per-pass CUDA kernel on the host, thread by thread
    (``igg_cpu_temporal2d_emulate``): ``T2 <- step(step(T))`` on the cells that kernel writes."""
    lib = require()
    for t in (T2, T, Cp) + ((mid_lines,) if mid_lines is not None else ()):
        if t.is_cuda or t.dtype != torch.float64 or not t.is_contiguous():
            raise TypeError("cpu_temporal2d_emulate needs contiguous CPU float64 tensors")
    nx, ny = T.shape
    if mid_lines is not None and mid_lines.numel() < 2 * nx + 2 * ny:
        raise ValueError("mid_lines needs 2*nx + 2*ny elements")
    p2 = (C.c_double * 4)(*[float(v) for v in phys2]) if phys2 is not None else None
    if diag_out is not None and (diag_out.is_cuda or diag_out.dtype != torch.float64 or diag_out.numel() < 4):
        raise TypeError("diag_out must be a CPU float64 tensor of 4 elements")
    _check(lib.igg_cpu_temporal2d_emulate(T.data_ptr(), T2.data_ptr(), Cp.data_ptr(), nx, ny, lam, dt, _dx, _dy,
                                          srb, scx, rpt, workers,
                                          mid_lines.data_ptr() if mid_lines is not None else None, side_mask,
                                          C.cast(p2, C.c_void_p) if p2 is not None else None,
                                          diag_out.data_ptr() if diag_out is not None else None))


def cpu_temporal2d_midlines(T: torch.Tensor, Cp: torch.Tensor, lam, dt, _dx, _dy,
                            peer_mid: Sequence[Optional[torch.Tensor]]) -> None:
    """Host version of ``igg_diffusion2d_midlines``: ``peer_mid[s]`` is the mid-line buffer of the neighbour on
    side ``s`` (x-low, x-high, y-low, y-high) or ``None``."""
    lib = require()
    nx, ny = T.shape
    arr = (C.c_void_p * 4)(*[(m.data_ptr() if m is not None else None) for m in peer_mid])
    _check(lib.igg_cpu_temporal2d_midlines(T.data_ptr(), Cp.data_ptr(), nx, ny, lam, dt, _dx, _dy, arr))


def cpu_diffusion2d_run(T: torch.Tensor, T2: torch.Tensor, Cp: torch.Tensor, lam, dt, _dx, _dy, nt: int) -> None:
    lib = require()
    nx, ny = T.shape
    _check(lib.igg_cpu_diffusion2d_run(T.data_ptr(), T2.data_ptr(), Cp.data_ptr(), nx, ny, lam, dt, _dx, _dy, nt))
