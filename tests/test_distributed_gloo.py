"""Multi-process CPU tests (gloo): halo exchange, gather and the ap application."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import pytest

from mp_util import run_workers


@pytest.mark.parametrize("args,nprocs", [
    ((8, 7, 1, 0, 0, 0, 0, 0, 0), 4),      # 2x2
    ((8, 7, 1, 1, 2, 0, 1, 1, 0), 2),      # 1x2 periodic both (self-neighbour in x, same rank both sides in y)
    ((6, 5, 4, 0, 0, 0, 0, 1, 0), 4),      # 3-D 2x2x1... dims_create -> (2,2,1), periodic y
    ((6, 5, 4, 1, 1, 2, 0, 0, 1), 2),      # 3-D split in z, periodic z
])
def test_update_halo_gather_gloo(args, nprocs):
    outs = run_workers("halo_gloo.py", nprocs, args)
    assert all("ok" in o for o in outs)


def test_ap_app_4_ranks_bit_exact(tmp_path):
    # BASELINE.json config 1 at reduced step count (the full 1000-step golden run is in test_oracle)
    outs = run_workers("ap_gloo.py", 4, (128, 128, 0, 0, 100, str(tmp_path)))
    assert all("ok" in o for o in outs)


def test_ap_app_8_ranks_4x2_layout(tmp_path):
    outs = run_workers("ap_gloo.py", 8, (65, 128, 4, 2, 40, str(tmp_path)), timeout=600)
    assert all("ok" in o for o in outs)


def test_grid_can_be_reinitialised_in_one_process():
    outs = run_workers("reinit_gloo.py", 2)
    assert all("ok" in o for o in outs)


@pytest.mark.parametrize("nprocs", [4, 8])
def test_kp_3d_and_hide_communication_over_gloo(nprocs, tmp_path):
    outs = run_workers("apps_gloo.py", nprocs, (str(tmp_path), nprocs), timeout=600)
    assert all("ok" in o for o in outs)


@pytest.mark.parametrize("nprocs", [4, 8])
def test_perf_apps_on_cpu_over_gloo(nprocs, tmp_path):
    """perf / perf_hide / perf_hide_prof through their CPU stand-in stepper (models/solver_host.py)."""
    outs = run_workers("perf_gloo.py", nprocs, (str(tmp_path), nprocs), timeout=600)
    assert all("ok" in o for o in outs)


def test_per_rank_before_after_pictures(tmp_path):
    """The reference's proof-of-concept figure (two ranks, before / after, heat crossing the seam) as files."""
    outs = run_workers("poc_gloo.py", 2, (str(tmp_path),), timeout=300)
    assert all("ok" in o for o in outs)
    assert any("max |halo - neighbour's edge| across x seams = 0.0e+00" in o for o in outs)


def test_checkpoint_and_resume_through_the_perf_app(tmp_path):
    outs = run_workers("ckpt_gloo.py", 4, (str(tmp_path),), timeout=300)
    assert all("ok" in o for o in outs)


def test_ring_smoke_test_over_gloo():
    outs = run_workers("ring_gloo.py", 4)
    assert all("ok" in o for o in outs)


def test_a_dead_rank_is_detected_not_waited_for():
    """Fault injection: rank 1 of 2 exits abruptly; rank 0's next ``update_halo`` raises within the configured
    time-out (``IGG_B200_TIMEOUT_S``) instead of hanging."""
    import subprocess
    import sys as _sys
    from mp_util import ROOT, free_port
    import os as _os
    port = free_port()
    procs = []
    for r in range(2):
        env = dict(_os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": "2", "LOCAL_WORLD_SIZE": "2",
                    "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "OMP_NUM_THREADS": "1",
                    "IGG_B200_TIMEOUT_S": "20", "PYTHONPATH": ROOT + _os.pathsep + env.get("PYTHONPATH", "")})
        procs.append(subprocess.Popen([_sys.executable, _os.path.join(ROOT, "tests", "workers", "fault_gloo.py")],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=120)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise AssertionError("survivor hung on a dead neighbour")
        outs.append(out)
    assert procs[1].returncode == 7
    assert procs[0].returncode == 0 and "failure detected" in outs[0], outs[0]
