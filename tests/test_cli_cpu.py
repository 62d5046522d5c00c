"""Command-line entry points of every application, run as a user would (``python -m ...``) on the CPU.

The reference is driven by (un)commenting lines of ``scripts/runme.sh:5-9``; here each variant is a module with
flags.  One process each, ``--device-type cpu``: argument parsing, set-up, time loop, output files.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = {**os.environ, "PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), "OMP_NUM_THREADS": "2"}
for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
    ENV.pop(k, None)


def _run(module, *args, cwd):
    r = subprocess.run([sys.executable, "-m", module, *args], capture_output=True, text=True, timeout=300, cwd=cwd,
                       env=ENV)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


@pytest.mark.parametrize("module,extra,tag", [
    ("rocm_mpi_b200.models.diffusion2d_ap", [], "ap"),
    ("rocm_mpi_b200.models.diffusion2d_kp", [], "kp"),
    ("rocm_mpi_b200.models.diffusion2d_perf", ["--do-vis", "true"], "perf"),
    ("rocm_mpi_b200.models.diffusion2d_perf_hide", ["--do-vis", "true", "--b-width", "4", "3"], "hide"),
])
def test_2d_apps_write_their_outputs(module, extra, tag, tmp_path):
    out = _run(module, "--nx", "34", "--ny", "30", "--nt", "20", "--device-type", "cpu", "--outdir", str(tmp_path),
               *extra, cwd=str(tmp_path))
    assert "Global grid: 34x30x1 (nprocs: 1, dims: 1x1x1)" in out          # IGG's banner, SURVEY G1
    assert "Process 0 selecting device" in out                              # reference scripts/diffusion_2D_ap.jl:18
    names = os.listdir(tmp_path)
    assert f"Temp_{tag}_1_34_30.png" in names, names
    if tag in ("perf", "hide"):
        assert "T_eff" in out and "Executed 20 steps" in out                # reference scripts/diffusion_2D_perf.jl:58


def test_profiled_app_writes_prof_txt(tmp_path):
    prof = tmp_path / "prof.txt"
    out = _run("rocm_mpi_b200.models.diffusion2d_perf_hide_prof", "--nx", "34", "--ny", "30", "--nt", "20",
               "--device-type", "cpu", "--b-width", "4", "3", "--prof-path", str(prof), "--outdir", str(tmp_path),
               cwd=str(tmp_path))
    assert "Executed 20 steps" in out
    assert prof.exists() and "cumulative" in prof.read_text()


def test_3d_app_runs_on_cpu(tmp_path):
    out = _run("rocm_mpi_b200.models.diffusion3d", "--nx", "12", "--ny", "10", "--nz", "9", "--nt", "5",
               "--device-type", "cpu", "--outdir", str(tmp_path), cwd=str(tmp_path))
    assert "Global grid: 12x10x9" in out


def test_trace_flag_writes_a_chrome_trace(tmp_path):
    import json
    _run("rocm_mpi_b200.models.diffusion2d_perf", "--nx", "34", "--ny", "30", "--nt", "12", "--warmup", "3",
         "--device-type", "cpu", "--do-vis", "false", "--trace", str(tmp_path / "tl"), "--trace-steps", "true",
         cwd=str(tmp_path))
    t = json.load(open(tmp_path / "tl.rank0.json"))
    spans = [e for e in t["traceEvents"] if e["ph"] == "X"]
    names = [e["name"] for e in spans]
    assert "warmup" in names and "timed_loop" in names and sum(n.startswith("step ") for n in names) == 12
    assert all(e["dur"] >= 0 and e["ts"] >= 0 for e in spans)
