"""Run configuration.

The reference hard-codes everything as script constants (``lx, ly, lam, Cp0`` at
``scripts/diffusion_2D_ap.jl:11-13``, ``nx, ny, nt`` at ``:15-16``, ``threads``/``grid`` at
``scripts/diffusion_2D_kp.jl:63-64``, ``fact`` at ``scripts/diffusion_2D_perf.jl:21``, ``b_width`` at
``scripts/diffusion_2D_perf_hide.jl:42``, ``do_vis`` kwarg at ``scripts/diffusion_2D_perf.jl:15``) and
selects variants by (un)commenting lines (``scripts/runme.sh:5-9``).  Here they are fields of a
dataclass that can be filled from the command line or from ``IGG_B200_*`` environment variables.
"""
from __future__ import annotations

import argparse
import os
from dataclasses import dataclass, field, fields
from typing import Optional, Tuple

__all__ = ["DiffusionConfig", "add_config_args", "config_from_args", "env_flag"]


def env_flag(name: str, default: bool = False) -> bool:
    v = os.environ.get(name)
    if v is None:
        return default
    return v.strip().lower() not in ("0", "false", "no", "off", "")


@dataclass
class DiffusionConfig:
    # physics (reference scripts/diffusion_2D_ap.jl:11-13)
    lx: float = 10.0
    ly: float = 10.0
    lz: float = 10.0
    lam: float = 1.0
    Cp0: float = 1.0
    # numerics: LOCAL sizes including the overlap (reference scripts/diffusion_2D_ap.jl:15-16)
    nx: int = 128
    ny: int = 128
    nz: int = 1
    nt: int = 1000
    warmup: int = 10              # iterations excluded from timing (reference perf.jl:48)
    # decomposition
    dimx: int = 0
    dimy: int = 0
    dimz: int = 0
    periodx: int = 0
    periody: int = 0
    periodz: int = 0
    # variant-specific
    b_width: Tuple[int, int, int] = (16, 4, 4)   # hide_communication boundary width
    transport: Optional[str] = None              # peer | nccl | host | gloo (None = default)
    kernel: str = "tma"                          # tma | direct  (stencil kernel flavour)
    fused: bool = True                           # fuse the halo push into the stencil kernel
    steps_per_pass: int = 2                      # perf app, fused mode: 2 = two time steps per pass over memory
                                                 # (Diffusion2D mode "temporal"; needs the TMA kernel and an even
                                                 # ny, else 1), 1 = one fused kernel per step
    kp_path: str = "three"                       # kp app on CUDA: three (Flux!/Residual!/Update! launches per step,
                                                 # the literal form) | persistent | resident | auto (ONE launch for
                                                 # all nt steps with the kp arithmetic; auto = resident if it fits)
    device_type: str = "auto"
    # checkpoint / resume (none in the reference)
    ckpt_every: int = 0                          # write a checkpoint every N steps (0 = never)
    ckpt_dir: str = "output/ckpt"
    resume: bool = False                         # start from the latest checkpoint in ckpt_dir, if there is one
    # timeline (Chrome / Perfetto trace, one file per rank: <trace>.rank<r>.json); trace_steps adds one span per step
    trace: Optional[str] = None
    trace_steps: bool = False
    # output
    do_vis: bool = True
    outdir: str = "output"
    quiet: bool = False

    def __post_init__(self):
        if self.transport is None:
            self.transport = os.environ.get("IGG_B200_TRANSPORT") or None
        bw = tuple(int(b) for b in self.b_width)
        self.b_width = bw + (bw[-1],) * (3 - len(bw))


def _to_bool(s: str) -> bool:
    return s.strip().lower() not in ("0", "false", "no", "off", "")


def _env_default(name: str, v):
    """``IGG_B200_<FIELD>`` overrides a field's default (the command line still wins)."""
    env = os.environ.get("IGG_B200_" + name.upper())
    if env is None or env == "":
        return v
    if name == "b_width":
        return [int(x) for x in env.replace(",", " ").split()]
    if isinstance(v, bool):
        return _to_bool(env)
    if v is None:
        return env
    return type(v)(env)


def add_config_args(p: argparse.ArgumentParser, defaults: Optional[DiffusionConfig] = None) -> None:
    d = defaults or DiffusionConfig()
    for f in fields(DiffusionConfig):
        v = _env_default(f.name, getattr(d, f.name))
        flag = "--" + f.name.replace("_", "-")
        if f.name == "b_width":
            p.add_argument(flag, type=int, nargs="+", default=list(v))
        elif isinstance(v, bool):
            p.add_argument(flag, type=_to_bool, default=v)
        elif v is None:
            p.add_argument(flag, type=str, default=None)
        else:
            p.add_argument(flag, type=type(v), default=v)


def config_from_args(ns: argparse.Namespace) -> DiffusionConfig:
    kw = {f.name: getattr(ns, f.name) for f in fields(DiffusionConfig) if hasattr(ns, f.name)}
    return DiffusionConfig(**kw)
