"""Spawn N local worker processes running a ``tests/workers/*.py`` script under a torchrun-like env."""
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_workers(script: str, nprocs: int, args=(), timeout: int = 300, extra_env=None):
    port = free_port()
    procs = []
    for r in range(nprocs):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nprocs),
                    "LOCAL_WORLD_SIZE": str(nprocs),
                    "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                    "PYTHONPATH": ROOT + os.pathsep + env.get("PYTHONPATH", ""),
                    "OMP_NUM_THREADS": "1"})
        if extra_env:
            env.update(extra_env)
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "workers", script), *map(str, args)],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    failed = False
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            out, _ = p.communicate()
            failed = True
        outs.append(out)
        if p.returncode != 0:
            failed = True
    if failed:
        raise AssertionError("worker failure:\n" + "\n-----\n".join(outs))
    return outs
