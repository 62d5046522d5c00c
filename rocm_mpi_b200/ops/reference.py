"""Plain-PyTorch statements of the diffusion maths: oracle and array-programming path.

Three formulations exist in the reference and are kept apart here because they round
differently:

* ``ap``   -- broadcast form dividing by ``dx``/``Cp`` (reference ``scripts/diffusion_2D_ap.jl:38-41``
  with the difference helpers of ``:3-7``);
* ``kp``   -- three-kernel form multiplying by ``_dx = 1/dx`` and by ``1/Cp``
  (``scripts/diffusion_2D_kp.jl:16-54``);
* ``perf`` -- fused double-buffered form multiplying by ``Cp`` (``scripts/diffusion_2D_perf.jl:3-13``).

All functions work on any device / dtype; the oracles are run in fp64 on the CPU.  Index
order is ``T[ix, iy(, iz)]`` with the last dimension contiguous.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

__all__ = [
    "d_xa", "d_xi", "d_ya", "d_yi", "inn", "gaussian_ic", "gaussian_ic_3d",
    "ap_step", "kp_flux", "kp_residual", "kp_update", "kp_step", "perf_step", "perf_step_region",
    "perf_step_3d", "ap_step_3d", "time_step", "time_step_3d", "oracle_2d", "oracle_3d",
    "frame_mask",
]


# -- finite-difference helpers (reference scripts/diffusion_2D_ap.jl:3-7) ------------------
def d_xa(A): return A[1:, :] - A[:-1, :]
def d_xi(A): return A[1:, 1:-1] - A[:-1, 1:-1]
def d_ya(A): return A[:, 1:] - A[:, :-1]
def d_yi(A): return A[1:-1, 1:] - A[1:-1, :-1]
def inn(A): return A[1:-1, 1:-1]


def time_step(dx: float, dy: float, Cp0: float = 1.0, lam: float = 1.0) -> float:
    """``dt = min(dx^2, dy^2)*Cp0/lam/4.1`` (reference ``scripts/diffusion_2D_ap.jl:20``)."""
    return min(dx * dx, dy * dy) * Cp0 / lam / 4.1


def time_step_3d(dx: float, dy: float, dz: float, Cp0: float = 1.0, lam: float = 1.0) -> float:
    return min(dx * dx, dy * dy, dz * dz) * Cp0 / lam / 6.1


def gaussian_ic(xs: torch.Tensor, ys: torch.Tensor, dx: float, dy: float, lx: float, ly: float) -> torch.Tensor:
    """``exp(-(x+dx/2-lx/2)^2 - (y+dy/2-ly/2)^2)`` (reference ``scripts/diffusion_2D_ap.jl:28``)."""
    X = (xs + dx / 2 - lx / 2).reshape(-1, 1)
    Y = (ys + dy / 2 - ly / 2).reshape(1, -1)
    return torch.exp(-X ** 2 - Y ** 2)


def gaussian_ic_3d(xs, ys, zs, dx, dy, dz, lx, ly, lz) -> torch.Tensor:
    X = (xs + dx / 2 - lx / 2).reshape(-1, 1, 1)
    Y = (ys + dy / 2 - ly / 2).reshape(1, -1, 1)
    Z = (zs + dz / 2 - lz / 2).reshape(1, 1, -1)
    return torch.exp(-X ** 2 - Y ** 2 - Z ** 2)


# -- ap form --------------------------------------------------------------------------------
def ap_step(T, Cp, lam: float, dt: float, dx: float, dy: float, qx=None, qy=None, dTdt=None):
    """One in-place array-programming step on ``T`` (reference ``scripts/diffusion_2D_ap.jl:38-41``)."""
    qx_ = -lam * d_xi(T) / dx
    qy_ = -lam * d_yi(T) / dy
    dTdt_ = 1.0 / inn(Cp) * (-d_xa(qx_) / dx - d_ya(qy_) / dy)
    if qx is not None:
        qx.copy_(qx_)
    if qy is not None:
        qy.copy_(qy_)
    if dTdt is not None:
        dTdt.copy_(dTdt_)
    T[1:-1, 1:-1] = inn(T) + dt * dTdt_
    return T


# -- kp form --------------------------------------------------------------------------------
def kp_flux(qx, qy, T, lam: float, _dx: float, _dy: float):
    """``Flux!`` (reference ``scripts/diffusion_2D_kp.jl:16-26``)."""
    qx.copy_(-lam * d_xi(T) * _dx)
    qy.copy_(-lam * d_yi(T) * _dy)


def kp_residual(dTdt, qx, qy, Cp, _dx: float, _dy: float):
    """``Residual!`` (reference ``scripts/diffusion_2D_kp.jl:33-40``)."""
    dTdt.copy_(1.0 / inn(Cp) * -(d_xa(qx) * _dx + d_ya(qy) * _dy))


def kp_update(T, dt: float, dTdt):
    """``Update!`` (reference ``scripts/diffusion_2D_kp.jl:47-54``)."""
    T[1:-1, 1:-1] += dt * dTdt


def kp_step(T, Cp, qx, qy, dTdt, lam, dt, _dx, _dy):
    kp_flux(qx, qy, T, lam, _dx, _dy)
    kp_residual(dTdt, qx, qy, Cp, _dx, _dy)
    kp_update(T, dt, dTdt)
    return T


# -- perf form ------------------------------------------------------------------------------
def _perf_update(T, Cp, lam, dt, _dx, _dy, sx: slice, sy: slice):
    """Value of the fused update on the index box ``sx`` x ``sy`` (must be interior)."""
    def sh(s, k):
        return slice(s.start + k, s.stop + k)
    c = T[sx, sy]
    qxr = -lam * (T[sh(sx, 1), sy] - c) * _dx
    qxl = -lam * (c - T[sh(sx, -1), sy]) * _dx
    qyt = -lam * (T[sx, sh(sy, 1)] - c) * _dy
    qyb = -lam * (c - T[sx, sh(sy, -1)]) * _dy
    return c + dt * (Cp[sx, sy] * (-(qxr - qxl) * _dx - (qyt - qyb) * _dy))


def perf_step(T2, T, Cp, lam: float, dt: float, _dx: float, _dy: float):
    """Fused step ``T2 <- step(T)`` on the interior (reference ``scripts/diffusion_2D_perf.jl:3-13``).

    Boundary cells of ``T2`` are not written, as in the reference.
    """
    nx, ny = T.shape
    T2[1:-1, 1:-1] = _perf_update(T, Cp, lam, dt, _dx, _dy, slice(1, nx - 1), slice(1, ny - 1))
    return T2


def frame_mask(shape, b_width) -> torch.Tensor:
    """Boolean mask of the boundary frame of width ``b_width`` inside the interior.

    This is the *intended* ``hide_communication`` split: ``True`` cells are
    computed first (and contain every cell that is sent), ``False`` interior cells second.
    Cells on the array boundary itself are never computed and are ``False``.
    """
    m = torch.zeros(tuple(shape), dtype=torch.bool)
    nd = len(shape)
    inner_sl = tuple(slice(1, s - 1) for s in shape)
    m[inner_sl] = True
    core = tuple(slice(1 + b_width[d], shape[d] - 1 - b_width[d]) for d in range(nd))
    if all(c.stop > c.start for c in core):
        m[core] = False
    return m


def perf_step_region(T2, T, Cp, lam, dt, _dx, _dy, b_width: Tuple[int, int], istep: int):
    """Split step (reference ``scripts/diffusion_2D_perf_hide.jl:15-29``, intended semantics).

    ``istep == 1`` updates the boundary frame, ``istep == 2`` the remaining interior.
    """
    full = torch.empty_like(T2)
    full.copy_(T2)
    perf_step(full, T, Cp, lam, dt, _dx, _dy)
    m = frame_mask(T.shape, b_width).to(T.device)
    if istep == 2:
        inner_m = torch.zeros_like(m)
        inner_m[1:-1, 1:-1] = True
        m = inner_m & ~m
    T2[m] = full[m]
    return T2


# -- 3-D (IGG capability; the reference itself is 2-D only) ------------------
def perf_step_3d(T2, T, Cp, lam, dt, _dx, _dy, _dz):
    c = T[1:-1, 1:-1, 1:-1]
    qxr = -lam * (T[2:, 1:-1, 1:-1] - c) * _dx
    qxl = -lam * (c - T[:-2, 1:-1, 1:-1]) * _dx
    qyr = -lam * (T[1:-1, 2:, 1:-1] - c) * _dy
    qyl = -lam * (c - T[1:-1, :-2, 1:-1]) * _dy
    qzr = -lam * (T[1:-1, 1:-1, 2:] - c) * _dz
    qzl = -lam * (c - T[1:-1, 1:-1, :-2]) * _dz
    T2[1:-1, 1:-1, 1:-1] = c + dt * (Cp[1:-1, 1:-1, 1:-1] * (
        -(qxr - qxl) * _dx - (qyr - qyl) * _dy - (qzr - qzl) * _dz))
    return T2


def ap_step_3d(T, Cp, lam, dt, dx, dy, dz):
    qx = -lam * (T[1:, 1:-1, 1:-1] - T[:-1, 1:-1, 1:-1]) / dx
    qy = -lam * (T[1:-1, 1:, 1:-1] - T[1:-1, :-1, 1:-1]) / dy
    qz = -lam * (T[1:-1, 1:-1, 1:] - T[1:-1, 1:-1, :-1]) / dz
    dTdt = 1.0 / Cp[1:-1, 1:-1, 1:-1] * (-(qx[1:] - qx[:-1]) / dx - (qy[:, 1:] - qy[:, :-1]) / dy
                                           - (qz[:, :, 1:] - qz[:, :, :-1]) / dz)
    T[1:-1, 1:-1, 1:-1] = T[1:-1, 1:-1, 1:-1] + dt * dTdt
    return T


# -- single-domain oracles -------------------------------------------------------------------
def oracle_2d(nx_g: int, ny_g: int, nt: int, form: str = "ap", lx: float = 10.0, ly: float = 10.0,
              lam: float = 1.0, Cp0: float = 1.0, dtype=torch.float64):
    """Run ``nt`` steps on the whole ``nx_g x ny_g`` domain on the CPU; returns ``(T, dt)``.

    Because the decomposed run is bit-identical to the single-domain run, this
    is the oracle for every rank layout; the gathered field is ``T[1:-1, 1:-1]``.
    """
    dx, dy = lx / nx_g, ly / ny_g
    dt = time_step(dx, dy, Cp0, lam)
    xs = torch.arange(nx_g, dtype=dtype) * dx
    ys = torch.arange(ny_g, dtype=dtype) * dy
    T = gaussian_ic(xs, ys, dx, dy, lx, ly)
    Cp = Cp0 * torch.ones(nx_g, ny_g, dtype=dtype)
    if form == "ap":
        for _ in range(nt):
            ap_step(T, Cp, lam, dt, dx, dy)
    elif form == "kp":
        qx = torch.zeros(nx_g - 1, ny_g - 2, dtype=dtype)
        qy = torch.zeros(nx_g - 2, ny_g - 1, dtype=dtype)
        dTdt = torch.zeros(nx_g - 2, ny_g - 2, dtype=dtype)
        for _ in range(nt):
            kp_step(T, Cp, qx, qy, dTdt, lam, dt, 1.0 / dx, 1.0 / dy)
    elif form == "perf":
        T2 = T.clone()
        for _ in range(nt):
            perf_step(T2, T, Cp, lam, dt, 1.0 / dx, 1.0 / dy)
            T, T2 = T2, T
    else:
        raise ValueError(f"unknown form {form!r}")
    return T, dt


def oracle_3d(nx_g: int, ny_g: int, nz_g: int, nt: int, lx=10.0, ly=10.0, lz=10.0, lam=1.0, Cp0=1.0,
              dtype=torch.float64, form: str = "perf"):
    dx, dy, dz = lx / nx_g, ly / ny_g, lz / nz_g
    dt = time_step_3d(dx, dy, dz, Cp0, lam)
    xs = torch.arange(nx_g, dtype=dtype) * dx
    ys = torch.arange(ny_g, dtype=dtype) * dy
    zs = torch.arange(nz_g, dtype=dtype) * dz
    T = gaussian_ic_3d(xs, ys, zs, dx, dy, dz, lx, ly, lz)
    Cp = Cp0 * torch.ones_like(T)
    if form == "ap":
        for _ in range(nt):
            ap_step_3d(T, Cp, lam, dt, dx, dy, dz)
        return T, dt
    T2 = T.clone()
    for _ in range(nt):
        perf_step_3d(T2, T, Cp, lam, dt, 1.0 / dx, 1.0 / dy, 1.0 / dz)
        T, T2 = T2, T
    return T, dt
