"""The benchmark / example programs parse their command lines on a machine without GPUs (``--help``), and their flags
do not collide with torchrun's own abbreviations (``--n`` once did)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROGRAMS = ["bench.py", "benchmarks/halo_bw_sweep.py", "benchmarks/update_halo_sweep.py", "benchmarks/small_grid_latency.py",
            "benchmarks/gather_bench.py", "examples/heat2d_two_step_control.py", "examples/heat2d_fused.py",
            "scripts/check_temporal.py"]
TORCHRUN_PREFIXES = ("--n", "--m", "--r", "--s", "--l", "--t", "--no", "--mo", "--ma", "--lo", "--ro", "--rd")


@pytest.mark.parametrize("prog", PROGRAMS)
def test_help_runs_without_a_gpu(prog):
    r = subprocess.run([sys.executable, os.path.join(ROOT, prog), "--help"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-400:]
    assert "usage" in r.stdout.lower()


@pytest.mark.parametrize("prog", [p for p in PROGRAMS if p.startswith("benchmarks/")])
def test_flags_are_not_ambiguous_prefixes_for_torchrun(prog):
    """``torchrun prog --n 5`` is parsed by torchrun first: a flag that is a strict prefix of several torchrun options
    (``--n`` -> ``--nnodes``, ``--nproc-per-node``, ...) is rejected as ambiguous before the program ever sees it."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, prog), "--help"], capture_output=True, text=True, timeout=120)
    flags = {w.rstrip(",") for w in r.stdout.split() if w.startswith("--")}
    short = [f for f in flags if f in ("--n", "--nn", "--np", "--no", "--r", "--m", "--ma", "--max", "--mo", "--l", "--lo",
                                       "--s", "--st", "--t", "--rd")]
    assert not short, short
