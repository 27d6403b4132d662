"""Statically condensed BramblePasciakCG row-partitioned over 2 GPUs (SURVEY.md 8(f) item 1 at N > 1) against the single-GPU condensed
solve of the same system: operator apply, Lanczos estimate, iteration count +-1, history 1e-10 over 50 iterations, solution 1e-7 (the
window the single-GPU condensed test grants this mesh).

STATUS: written after the round's GPU budget was spent -- the device side (csrc/cond.cuh exchanges, nsb_cond_create accepting a shared
halo plan) has never run on a GPU.  The host side (one shared velocity plan, the exchanged chain == the global operator) is verified on
CPU under gloo (tests/test_partition_condensed_gloo.py).  Until it has passed once on a device this test runs only on request
(NSB_RUN_UNVERIFIED_GPU_TESTS=1), so that an unverified path cannot stop the verified suite under `pytest -x`."""
import json
import os
import re
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.gpu
@pytest.mark.parametrize("sell", ["0", "200000"])
def test_two_rank_condensed_bpcg_matches_single_gpu(sell):
    import torch
    if os.environ.get("NSB_RUN_UNVERIFIED_GPU_TESTS", "0") != "1":
        pytest.skip("never run on a device (written after the round-2 GPU budget was spent); set NSB_RUN_UNVERIFIED_GPU_TESTS=1")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, NSB_SELL_MIN_NNZ=sell)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_condensed_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    rows = [json.loads(m) for m in re.findall(r"CONDRESULT (\{.*?\})", res.stdout)]
    assert len(rows) == 2
    for r in rows:
        assert r["halo_u"] > 0 and r["halo_p"] > 0, r
        assert r["apply_rel"] < 1e-12, r
        assert r["lam_rel"] < 1e-6, r
        assert abs(r["it"] - r["it1"]) <= 1, r
        assert r["hist_rel"] < 1e-10, r
        assert r["sol_rel"] < 1e-7, r


@pytest.mark.gpu
@pytest.mark.parametrize("sell", ["0", "200000"])
def test_two_rank_block_smoother_and_auxspace(sell):
    """Block smoother (Jacobi / Smooth / SmoothBack / symmetric GS) and the auxiliary-space composite MypreA row-partitioned over 2 GPUs
    against the same objects on one GPU (1e-12), then CG and BramblePasciakCG preconditioned by the composite against their single-GPU
    runs (counts +-1, history over 30 entries, the windows tests/test_gpu_parity.py grants these paths on one GPU).  Same status as the
    test above: host side verified under gloo (tests/test_partition_smoother_gloo.py), device side never run."""
    import torch
    if os.environ.get("NSB_RUN_UNVERIFIED_GPU_TESTS", "0") != "1":
        pytest.skip("never run on a device (written after the round-2 GPU budget was spent); set NSB_RUN_UNVERIFIED_GPU_TESTS=1")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, NSB_SELL_MIN_NNZ=sell)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_smoother_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    rows = [json.loads(m) for m in re.findall(r"SMOOTHRESULT (\{.*?\})", res.stdout)]
    assert len(rows) == 2
    for r in rows:
        assert r["local_blocks"] > 0 and r["colours"] >= 2, r
        for k in ("jacobi_rel", "symgs_rel", "smooth_rel", "smooth_back_rel", "aux_gs_rel", "aux_add_rel"):
            assert r[k] < 1e-12, (k, r)
        assert abs(r["cg_it"] - r["cg_it1"]) <= 1 and r["cg_hist_rel"] < 1e-9 and r["cg_sol_rel"] < 1e-8, r
        assert abs(r["it"] - r["it1"]) <= 2 and r["hist_rel"] < 1e-9 and r["sol_rel"] < 1e-7, r
        assert abs(r["minres_it"] - r["minres_it1"]) <= 1 and r["minres_hist_rel"] < 1e-9 and r["minres_sol_rel"] < 1e-7, r
