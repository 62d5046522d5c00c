"""Headless output: heat-map PNG and raw dumps of the gathered field.

The reference writes ``heatmap(transpose(T_v)); png("../output/Temp_ap_<nprocs>_<nx_g>_<ny_g>.png")``
through Plots.jl/GR with ``GKSwstype=nul`` (``scripts/diffusion_2D_ap.jl:30,47``).  There is no
graphics stack here (interactive GR/Qt windows: unavailable on this box), so the image is
rasterised from a host buffer: field -> 8-bit RGB through an inferno-like look-up table
(on the GPU with the ``colormap`` kernel when the field is a device tensor, otherwise in
numpy), composed with a colour bar and tick labels, and encoded as PNG with ``zlib`` only.
A raw ``.bin`` (+ ``.json`` header) / ``.npy`` dump of the same field is written next to it.
"""
from __future__ import annotations

import json
import os
import struct
import zlib
from typing import Optional, Tuple

import numpy as np
import torch

__all__ = ["inferno_lut", "colorize", "write_png", "save_heatmap", "save_raw", "output_name", "read_png_size"]

_STOPS = np.array([
    [0x00, 0x00, 0x04], [0x1B, 0x0C, 0x42], [0x4B, 0x0C, 0x6B], [0x78, 0x1C, 0x6D], [0xA5, 0x2C, 0x60],
    [0xCF, 0x44, 0x46], [0xED, 0x69, 0x25], [0xFB, 0x9A, 0x06], [0xF7, 0xD0, 0x3C], [0xFC, 0xFF, 0xA4],
], dtype=np.float64)


def inferno_lut(n: int = 256) -> np.ndarray:
    """``n x 3`` uint8 look-up table interpolated from ten inferno control colours."""
    pos = np.linspace(0.0, 1.0, len(_STOPS))
    x = np.linspace(0.0, 1.0, n)
    lut = np.stack([np.interp(x, pos, _STOPS[:, c]) for c in range(3)], axis=1)
    return np.clip(np.rint(lut), 0, 255).astype(np.uint8)


def colorize(field, vmin: Optional[float] = None, vmax: Optional[float] = None) -> np.ndarray:
    """Map a 2-D field to ``H x W x 3`` uint8 (rows = first index)."""
    if torch.is_tensor(field) and field.is_cuda:
        from ..ops import native
        if native.available():
            lo = float(field.min()) if vmin is None else vmin
            hi = float(field.max()) if vmax is None else vmax
            return native.colormap(field, lo, hi, torch.from_numpy(inferno_lut())).cpu().numpy()
        field = field.cpu()
    a = field.detach().cpu().numpy() if torch.is_tensor(field) else np.asarray(field)
    a = a.astype(np.float64)
    lo = float(a.min()) if vmin is None else vmin
    hi = float(a.max()) if vmax is None else vmax
    scale = 255.0 / (hi - lo) if hi > lo else 0.0
    idx = np.clip(np.floor((a - lo) * scale + 0.5), 0, 255).astype(np.int64)
    return inferno_lut()[idx]


def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png(path: str, rgb: np.ndarray) -> None:
    """Encode an ``H x W x 3`` uint8 array as an 8-bit RGB PNG (no external encoder)."""
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    h, w, c = rgb.shape
    assert c == 3
    raw = np.concatenate([np.zeros((h, 1), dtype=np.uint8), rgb.reshape(h, w * 3)], axis=1).tobytes()
    png = b"\x89PNG\r\n\x1a\n"
    png += _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
    png += _chunk(b"IDAT", zlib.compress(raw, 6))
    png += _chunk(b"IEND", b"")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(png)


def read_png_size(path: str) -> Tuple[int, int]:
    with open(path, "rb") as f:
        head = f.read(24)
    w, h = struct.unpack(">II", head[16:24])
    return w, h


def _resize_nearest(img: np.ndarray, h: int, w: int) -> np.ndarray:
    ys = (np.arange(h) * img.shape[0] // h).clip(0, img.shape[0] - 1)
    xs = (np.arange(w) * img.shape[1] // w).clip(0, img.shape[1] - 1)
    return img[ys][:, xs]


def _draw_text(canvas: np.ndarray, items) -> np.ndarray:
    """Draw ``(x, y, text)`` labels with Pillow's built-in bitmap font if Pillow is there."""
    try:
        from PIL import Image, ImageDraw
    except Exception:  # Pillow is optional; the image is complete without labels
        return canvas
    im = Image.fromarray(canvas)
    dr = ImageDraw.Draw(im)
    for (x, y, text) in items:
        dr.text((x, y), text, fill=(0, 0, 0))
    return np.asarray(im).copy()


def save_heatmap(path: str, field, size: Tuple[int, int] = (600, 400), vmin=None, vmax=None,
                 title: Optional[str] = None) -> str:
    """Write the heat map of ``field[ix, iy]`` (x to the right, y up, like ``heatmap(transpose(T_v))``)."""
    W, H = size
    left, right, top, bottom = 50, 110, 25, 35
    pw, ph = W - left - right, H - top - bottom
    nxv, nyv = int(field.shape[0]), int(field.shape[1])
    if torch.is_tensor(field) and field.is_cuda:
        # a gathered field can be many GB (12286^2 per rank in the reference's perf run): take min / max on the
        # device and colour-map only the pw x ph cells the picture shows (same nearest-neighbour picks as
        # _resize_nearest makes on the full image), instead of moving the whole field to the host
        lo_t, hi_t = torch.aminmax(field.detach())
        lo = float(lo_t) if vmin is None else vmin
        hi = float(hi_t) if vmax is None else vmax
        if nxv > pw or nyv > ph:
            xs = (torch.arange(pw) * nxv // pw).clamp(0, nxv - 1)
            ys = (torch.arange(ph) * nyv // ph).clamp(0, nyv - 1)
            yidx = (nyv - 1) - ys.flip(0)
            field = field.detach().index_select(0, xs.to(field.device)).index_select(1, yidx.to(field.device))
    else:
        a = field.detach() if torch.is_tensor(field) else torch.as_tensor(np.asarray(field))
        lo = float(a.min()) if vmin is None else vmin
        hi = float(a.max()) if vmax is None else vmax
    rgb = colorize(field, lo, hi)                 # [ix, iy, 3]
    img = np.flipud(np.transpose(rgb, (1, 0, 2)))  # rows = y (up), cols = x
    canvas = np.full((H, W, 3), 255, dtype=np.uint8)
    canvas[top:top + ph, left:left + pw] = _resize_nearest(img, ph, pw)
    # frame
    canvas[top - 1, left - 1:left + pw + 1] = 0
    canvas[top + ph, left - 1:left + pw + 1] = 0
    canvas[top - 1:top + ph + 1, left - 1] = 0
    canvas[top - 1:top + ph + 1, left + pw] = 0
    # colour bar
    bx0, bw = left + pw + 20, 18
    bar = inferno_lut()[np.linspace(255, 0, ph).astype(np.int64)]
    canvas[top:top + ph, bx0:bx0 + bw] = bar[:, None, :]
    canvas[top - 1, bx0 - 1:bx0 + bw + 1] = 0
    canvas[top + ph, bx0 - 1:bx0 + bw + 1] = 0
    canvas[top - 1:top + ph + 1, bx0 - 1] = 0
    canvas[top - 1:top + ph + 1, bx0 + bw] = 0
    labels = []
    for k in range(5):
        fx = k / 4.0
        px = left + int(fx * (pw - 1))
        canvas[top + ph:top + ph + 4, px] = 0
        labels.append((px - 8, top + ph + 6, str(int(round(fx * nxv)))))
        py = top + ph - 1 - int(fx * (ph - 1))
        canvas[py, left - 4:left] = 0
        labels.append((8, py - 5, str(int(round(fx * nyv)))))
        canvas[py, bx0 + bw:bx0 + bw + 4] = 0
        labels.append((bx0 + bw + 7, py - 5, "%.3g" % (lo + fx * (hi - lo))))
    if title:
        labels.append((left, 6, title))
    canvas = _draw_text(canvas, labels)
    write_png(path, canvas)
    return path


def save_raw(path_base: str, field, meta: Optional[dict] = None) -> Tuple[str, str]:
    """Raw little-endian dump ``<base>.bin`` with a ``<base>.json`` header, plus ``<base>.npy``."""
    a = field.detach().cpu().numpy() if torch.is_tensor(field) else np.asarray(field)
    a = np.ascontiguousarray(a)
    os.makedirs(os.path.dirname(os.path.abspath(path_base)), exist_ok=True)
    a.astype("<f8" if a.dtype == np.float64 else a.dtype).tofile(path_base + ".bin")
    hdr = {"shape": list(a.shape), "dtype": str(a.dtype), "order": "C", "index_order": "ix,iy,iz",
           "max": float(a.max()), "min": float(a.min()), "sum": float(a.sum())}
    if meta:
        hdr.update(meta)
    with open(path_base + ".json", "w") as f:
        json.dump(hdr, f, indent=1)
    np.save(path_base + ".npy", a)
    return path_base + ".bin", path_base + ".json"


def output_name(variant: str, nprocs: int, nx: int, ny: int, nz: Optional[int] = None) -> str:
    """``Temp_<variant>_<nprocs>_<nx>_<ny>`` (reference ``scripts/diffusion_2D_ap.jl:47``)."""
    if nz is None:
        return "Temp_%s_%d_%d_%d" % (variant, nprocs, nx, ny)
    return "Temp_%s_%d_%d_%d_%d" % (variant, nprocs, nx, ny, nz)
