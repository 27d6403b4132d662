"""`bench.py --impl reference` (the CPU arm the driver runs beside the device arm): emits ONE JSON line with the contract's keys, on a
small refinement (without a GPU the arm assembles at most 1M dofs and says so)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line_is_well_formed():
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")     # what torchrun exports; the arm must override the thread count
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--dofs", "6e4", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "bpcg_krylov_iterations_per_s" and d["unit"] == "iterations/s"
    assert d["higher_is_better"] is True and d["dtype"] == "f64" and d["steps"] == 2 and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == d["value"] and cb["cores"] == len(os.sched_getaffinity(0))
    assert d["sample_dofs"] == d["config"]["dofs"] and "workload" in d["config"]


_WD_SCRIPT = r"""
import sys, time
sys.path.insert(0, %r)
import bench
line = {"metric": "m", "value": 1.0}
wd = bench.LegWatchdog(1.5, int(sys.argv[1]), line)
with wd.leg("fast"):
    time.sleep(0.1)
line["fast"] = True
if sys.argv[2] == "hang":
    with wd.leg("slow"):
        time.sleep(60)
    line["slow"] = True
wd.emit()
wd.emit()
"""


def _run_wd(rank, mode):
    return subprocess.run([sys.executable, "-c", _WD_SCRIPT % ROOT, str(rank), mode], capture_output=True, text=True, timeout=120, cwd=ROOT)


def test_leg_watchdog_saves_the_line_when_an_optional_leg_hangs():
    """The 4-GPU run of round 2 lost nothing only by luck: one rank failed in an optional leg, the others waited in a collective until the
    job was killed.  Rank 0 must print the line with what finished; every rank must leave with status 0 well before the hang ends."""
    r0 = _run_wd(0, "hang")
    assert r0.returncode == 0
    lines = [ln for ln in r0.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["value"] == 1.0 and d["fast"] is True and "slow" not in d and d["watchdog"]["leg"] == "slow"
    r1 = _run_wd(1, "hang")
    assert r1.returncode == 0 and r1.stdout.strip() == "" and "slow" in r1.stderr


def test_leg_watchdog_is_silent_on_the_normal_path():
    r0 = _run_wd(0, "ok")
    assert r0.returncode == 0
    lines = [ln for ln in r0.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1 and "watchdog" not in json.loads(lines[0])      # emit() twice prints once


def test_preconditioner_leg_failure_is_a_key_not_a_crash():
    """bench.py's `preconditioners` leg runs in its own process; here (no GPU) that process refuses to start, and the leg must come back
    with an error object instead of raising -- the same path a crash or a timeout of that process takes on the GPU box."""
    sys.path.insert(0, ROOT)
    import benchlegs
    env_backup = os.environ.get("CUDA_VISIBLE_DEVICES")
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    try:
        out = benchlegs.precond_leg(6e4, 300)
    finally:
        if env_backup is None:
            os.environ.pop("CUDA_VISIBLE_DEVICES", None)
        else:
            os.environ["CUDA_VISIBLE_DEVICES"] = env_backup
    assert "error" in out and "exited" in out["error"] and "CUDA" in out.get("stderr_tail", "")
    out = benchlegs.precond_leg(6e4, 0.05)
    assert "did not finish" in out["error"]


def test_reference_arm_on_a_rank_other_than_zero_exits_quietly():
    """Under torchrun (N > 1) rank 0 alone runs and prints the reference arm; the other ranks exit 0 without work or output."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", CUDA_VISIBLE_DEVICES="")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert res.returncode == 0 and res.stdout.strip() == ""
