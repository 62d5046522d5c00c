"""Checkpoint / resume of distributed fields.

The reference persists nothing but the final PNG (``scripts/diffusion_2D_ap.jl:47``; SURVEY.md section 5:
"Checkpoint / resume: none").  Long runs at 16384 x 16384 per GPU want more: every rank writes its local
arrays (halo included, so a restart needs no exchange) plus the grid layout it was written with; loading
checks that the layout matches the current global grid.

Files: ``<dir>/ckpt_<step>/rank_<me>.pt`` and ``meta.json`` (rank 0).
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional, Tuple

import torch

from ..parallel.grid import global_grid

__all__ = ["save_checkpoint", "load_checkpoint", "latest_checkpoint"]


def _layout(g) -> dict:
    return {"nxyz": list(g.nxyz), "dims": list(g.dims), "periods": list(g.periods), "overlaps": list(g.overlaps),
            "nprocs": g.nprocs}


def save_checkpoint(directory: str, step: int, **fields: torch.Tensor) -> str:
    """Collective: every rank saves its local ``fields``; returns the checkpoint path."""
    g = global_grid()
    path = os.path.join(directory, f"ckpt_{int(step):08d}")
    os.makedirs(path, exist_ok=True)
    payload = {"step": int(step), "coords": list(g.coords), "layout": _layout(g),
               "fields": {k: v.detach().cpu() for k, v in fields.items()}}
    tmp = os.path.join(path, f"rank_{g.me}.pt.tmp")
    torch.save(payload, tmp)
    os.replace(tmp, os.path.join(path, f"rank_{g.me}.pt"))
    g.comm.barrier()
    if g.me == 0:
        with open(os.path.join(path, "meta.json"), "w") as f:
            json.dump({"step": int(step), "layout": _layout(g), "fields": sorted(fields)}, f, indent=1)
    g.comm.barrier()
    return path


def latest_checkpoint(directory: str) -> Optional[str]:
    if not os.path.isdir(directory):
        return None
    done = sorted(d for d in os.listdir(directory)
                  if d.startswith("ckpt_") and os.path.exists(os.path.join(directory, d, "meta.json")))
    return os.path.join(directory, done[-1]) if done else None


def load_checkpoint(path: str, device: Optional[torch.device] = None) -> Tuple[int, Dict[str, torch.Tensor]]:
    """Collective: load this rank's part of checkpoint ``path``; returns ``(step, fields)``."""
    g = global_grid()
    payload = torch.load(os.path.join(path, f"rank_{g.me}.pt"), map_location="cpu", weights_only=True)
    if payload["layout"] != _layout(g):
        raise ValueError(f"checkpoint layout {payload['layout']} does not match the current grid {_layout(g)}")
    if list(payload["coords"]) != list(g.coords):
        raise ValueError("checkpoint was written by a rank with different Cartesian coordinates")
    dev = device if device is not None else g.device
    return int(payload["step"]), {k: v.to(dev) for k, v in payload["fields"].items()}
