"""Build the native library in-tree: ``csrc/*.cu`` -> ``rocm_mpi_b200/_lib/libigg_b200.so``.

``python -m rocm_mpi_b200.build [--force] [--verbose]``

nvcc cross-compiles for sm_100a without a GPU.  Objects are cached under ``build/`` and rebuilt
when a source or header is newer.  The shared object is git-ignored but travels to the GPU box
with the ``gpurun`` snapshot.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from typing import List

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
INC = os.path.join(ROOT, "include")
OBJ = os.path.join(ROOT, "build", "obj")
LIBDIR = os.path.join(ROOT, "rocm_mpi_b200", "_lib")
LIB = os.path.join(LIBDIR, "libigg_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
              "-Xptxas", "-v", "-cudart", "static", "--expt-relaxed-constexpr", "-I", INC]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _headers_mtime() -> float:
    m = 0.0
    for d in (CSRC, INC):
        for f in os.listdir(d):
            if f.endswith((".h", ".cuh")):
                m = max(m, os.path.getmtime(os.path.join(d, f)))
    return m


def _compile(src: str, force: bool, verbose: bool, hdr_m: float) -> str:
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(src), hdr_m):
        return obj
    cmd = [nvcc(), *ARCH, *NVCC_FLAGS, "-c", src, "-o", obj]
    if src.endswith(".cpp"):
        cmd = [nvcc(), *ARCH, "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
               "-Xcompiler", "-fopenmp", "-I", INC, "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(OBJ, os.path.basename(src) + ".log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(r.stderr, file=sys.stderr)
    return obj


def build_variant(name: str, defines: List[str]) -> str:
    """Compile-time variant of the library for A/B runs: ``_lib/libigg_b200_<name>.so`` built with extra ``-D``
    flags into its own object directory (load it with ``IGG_B200_LIB=<path>``)."""
    obj_dir = os.path.join(ROOT, "build", "obj_" + name)
    os.makedirs(obj_dir, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    for src in sources():
        obj = os.path.join(obj_dir, os.path.basename(src) + ".o")
        if src.endswith(".cpp"):
            cmd = [nvcc(), *ARCH, "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
                   "-Xcompiler", "-fopenmp", "-I", INC, *defines, "-c", src, "-o", obj]
        else:
            cmd = [nvcc(), *ARCH, *NVCC_FLAGS, *defines, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        with open(obj + ".log", "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        objs.append(obj)
    lib = os.path.join(LIBDIR, f"libigg_b200_{name}.so")
    r = subprocess.run([nvcc(), *ARCH, "-shared", "-cudart", "static", "-Xcompiler", "-fopenmp", "-o", lib, *objs],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return lib


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    srcs = sources()
    hdr_m = _headers_mtime()
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, force, verbose, hdr_m), srcs))
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
        cmd = [nvcc(), *ARCH, "-shared", "-cudart", "static", "-Xcompiler", "-fopenmp", "-o", LIB, *objs]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


def ptxas_report() -> str:
    """Concatenated ``-Xptxas -v`` output of the last build (registers, spills, shared memory)."""
    out = []
    for f in sorted(os.listdir(OBJ)) if os.path.isdir(OBJ) else []:
        if f.endswith(".log"):
            out.append(open(os.path.join(OBJ, f)).read())
    return "\n".join(out)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--variant", default=None, help="name of a compile-time variant to build next to the default library")
    ap.add_argument("-D", dest="defs", action="append", default=[], help="extra define for --variant, e.g. -D IGG_T2_CTAS_SRB4=5")
    a = ap.parse_args()
    if a.variant:
        print(build_variant(a.variant, ["-D" + d for d in a.defs]))
    else:
        print(build(a.force, a.verbose))
