"""The C ABI used from plain C: the public header is valid C99 and a C program linked against the library
reproduces the golden value of the reference configuration (254 x 254 global, 1000
steps, ``max(T_v) = 0.3978647302669094``)."""
import os
import re
import shutil
import subprocess

import pytest

from rocm_mpi_b200.ops import native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
gcc = shutil.which("gcc")
pytestmark = pytest.mark.skipif(gcc is None, reason="no C compiler")


def test_header_is_valid_c99():
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                        os.path.join(ROOT, "include", "igg_b200.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_c_program_reproduces_the_golden_value(tmp_path):
    exe = str(tmp_path / "heat_cpu")
    libdir = os.path.dirname(native.lib_path())
    r = subprocess.run([gcc, "-std=c99", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "examples", "c", "heat_cpu.c"), "-o", exe, "-L", libdir, "-ligg_b200",
                        f"-Wl,-rpath,{libdir}", "-lm"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([exe, "254", "254", "1000"], capture_output=True, text=True, timeout=120,
                         env={**os.environ, "OMP_NUM_THREADS": "2"})
    assert out.returncode == 0, out.stderr
    m = re.search(r"max\(T_v\) = ([0-9.]+)", out.stdout)
    assert m and abs(float(m.group(1)) - 0.3978647302669094) < 1e-15, out.stdout


@pytest.mark.skipif(not os.path.exists("/usr/local/cuda/include/cuda_runtime_api.h"), reason="no CUDA toolkit headers")
def test_gpu_c_example_compiles_and_links(tmp_path):
    """examples/c/heat_gpu.c (CUDA runtime + the library, no Python) must at least build against the header."""
    libdir = os.path.dirname(native.lib_path())
    r = subprocess.run([gcc, "-std=c99", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                        "-I", "/usr/local/cuda/include", os.path.join(ROOT, "examples", "c", "heat_gpu.c"),
                        "-o", str(tmp_path / "heat_gpu"), "-L", libdir, "-ligg_b200", f"-Wl,-rpath,{libdir}",
                        "-L", "/usr/local/cuda/lib64", "-lcudart"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
