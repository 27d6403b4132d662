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
