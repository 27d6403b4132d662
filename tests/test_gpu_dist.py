"""Row-partitioned solvers on 2 GPUs (NCCL halo exchange + scalar all-reduce) against the single-GPU
solve of the same global system: iteration counts +-1, histories 1e-10 over 50 iterations, solutions
1e-8 (north-star tolerances).  Skipped on a box with fewer than two GPUs."""
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
@pytest.mark.parametrize("sell,fused_halo", [("0", "1"), ("0", "0"), ("200000", "1")])
def test_two_rank_bpcg_and_minres_match_single_gpu(sell, fused_halo):
    """sell = NSB_SELL_MIN_NNZ ("0": every operator in SELL-LX storage, "200000": these small ones stay CSR);
    fused_halo = NSB_HALO_FUSED ("1": interface entries are pushed and the SELL SpMV waits for them inside the kernel,
    interior chunks first; "0": push + wait/copy kernels before the SpMV)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, NSB_SELL_MIN_NNZ=sell, NSB_HALO_FUSED=fused_halo)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    rows = [json.loads(m) for m in re.findall(r"DISTRESULT (\{.*?\})", res.stdout)]   # ranks may share a line
    assert len(rows) == 2
    for r in rows:
        assert abs(r["it"] - r["it1"]) <= 1, r
        assert r["hist_rel"] < 1e-10, r
        assert r["sol_rel"] < 1e-8, r
        assert r["lam_rel"] < 1e-6, r
        assert abs(r["minres_it"] - r["minres_it1"]) <= 1, r
        assert r["minres_hist_rel"] < 1e-10, r
        assert r["minres_sol_rel"] < 1e-8, r
        assert r["simple_gfu_rel"] < 1e-8 and r["simple_gfup_rel"] < 1e-6, r
        assert all(abs(a - b) <= 1 for a, b in zip(r["simple_its"], r["simple_its1"])), r


@pytest.mark.gpu
def test_dead_exchange_returns_an_error():
    """Failure detection of the peer-memory protocol: with one rank never publishing its interface entries the waits are abandoned
    after NSB_SPIN_TIMEOUT_MS and every rank's solver call returns NSB_ERR_NCCL (raised as NsbError) instead of hanging in a kernel."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, NSB_SELL_MIN_NNZ="0", NSB_FAULT_SKIP_PUBLISH="1", NSB_SPIN_TIMEOUT_MS="400")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_fault_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    rows = [json.loads(m) for m in re.findall(r"FAULTRESULT (\{.*?\})", res.stdout)]
    assert len(rows) == 2, res.stdout[-2000:] + res.stderr[-2000:]
    for r in rows:
        assert r["error"] is not None and "timed out" in r["error"], r
        assert r["seconds"] < 60, r
