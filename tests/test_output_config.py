"""Output (PNG / raw dump / colour map) and configuration plumbing -- no GPU needed."""
import argparse
import json
import os
import struct
import zlib

import numpy as np
import torch

from rocm_mpi_b200.utils import output
from rocm_mpi_b200.utils.config import DiffusionConfig, add_config_args, config_from_args, env_flag


def test_lut_and_colorize_endpoints():
    lut = output.inferno_lut()
    assert lut.shape == (256, 3) and lut.dtype == np.uint8
    assert tuple(lut[0]) == (0, 0, 4) and tuple(lut[-1]) == (0xFC, 0xFF, 0xA4)
    f = torch.tensor([[0.0, 0.5], [1.0, 2.0]], dtype=torch.float64)
    rgb = output.colorize(f, 0.0, 1.0)
    assert rgb.shape == (2, 2, 3)
    assert tuple(rgb[0, 0]) == tuple(lut[0]) and tuple(rgb[1, 0]) == tuple(lut[255]) and tuple(rgb[1, 1]) == tuple(lut[255])
    assert tuple(rgb[0, 1]) == tuple(lut[128])


def test_png_is_valid_and_sized(tmp_path):
    field = torch.rand(40, 30, dtype=torch.float64)
    p = output.save_heatmap(str(tmp_path / "a" / "Temp_test_1_40_30.png"), field, title="t")
    assert output.read_png_size(p) == (600, 400)
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    # walk the chunks and check CRCs; inflate IDAT to the expected scanline size
    pos, idat = 8, b""
    while pos < len(raw):
        n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
        data = raw[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])
        assert crc == (zlib.crc32(tag + data) & 0xFFFFFFFF)
        if tag == b"IDAT":
            idat += data
        pos += 12 + n
    assert len(zlib.decompress(idat)) == 400 * (1 + 600 * 3)


def test_raw_dump_roundtrip(tmp_path):
    f = torch.arange(12, dtype=torch.float64).reshape(3, 4)
    b, j = output.save_raw(str(tmp_path / "Temp_x"), f, {"variant": "x"})
    back = np.fromfile(b, dtype="<f8").reshape(3, 4)
    assert np.array_equal(back, f.numpy())
    meta = json.load(open(j))
    assert meta["shape"] == [3, 4] and meta["variant"] == "x" and meta["max"] == 11.0
    assert np.array_equal(np.load(str(tmp_path / "Temp_x.npy")), f.numpy())
    assert output.output_name("ap", 4, 254, 254) == "Temp_ap_4_254_254"


def test_config_cli_roundtrip(monkeypatch):
    p = argparse.ArgumentParser()
    add_config_args(p)
    cfg = config_from_args(p.parse_args(["--nx", "65", "--ny", "128", "--dimx", "4", "--dimy", "2", "--b-width", "16", "4",
                                         "--do-vis", "false", "--transport", "nccl", "--fused", "false"]))
    assert (cfg.nx, cfg.ny, cfg.dimx, cfg.dimy) == (65, 128, 4, 2)
    assert cfg.b_width == (16, 4, 4) and cfg.do_vis is False and cfg.transport == "nccl" and cfg.fused is False
    monkeypatch.setenv("IGG_B200_TRANSPORT", "host")
    assert DiffusionConfig().transport == "host"
    monkeypatch.setenv("X_FLAG", "0")
    assert env_flag("X_FLAG", True) is False and env_flag("X_MISSING", True) is True


def test_config_env_overrides_defaults_and_cli_wins(monkeypatch):
    """IGG_B200_<FIELD> environment variables replace the script constants of the reference (SURVEY.md section 5,
    "Config / flags"); explicit command-line flags take precedence."""
    import argparse
    from rocm_mpi_b200.utils.config import DiffusionConfig, add_config_args, config_from_args
    monkeypatch.setenv("IGG_B200_NX", "300")
    monkeypatch.setenv("IGG_B200_DO_VIS", "0")
    monkeypatch.setenv("IGG_B200_B_WIDTH", "8,2")
    monkeypatch.setenv("IGG_B200_LAM", "2.5")
    monkeypatch.setenv("IGG_B200_TRANSPORT", "nccl")
    p = argparse.ArgumentParser()
    add_config_args(p, DiffusionConfig(nx=128, ny=64))
    c = config_from_args(p.parse_args(["--ny", "70", "--lam", "3.0"]))
    assert (c.nx, c.ny, c.do_vis, c.b_width, c.lam, c.transport) == (300, 70, False, (8, 2, 2), 3.0, "nccl")
