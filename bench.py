#!/usr/bin/env python
"""Headline benchmark: Bramble-Pasciak CG (solvers/bramblepasciak_new.py semantics) on the 3D Stokes
channel of templates/NavierStokesSIMPLE_test_3D.py refined to ~20M Taylor-Hood unknowns
(BASELINE.json configs[4]), Krylov iterations/s, plus the SpMV roofline of the dominant kernel.

    python bench.py --gpus N --steps K --warmup W            (N>1 under torchrun, one rank per GPU)
    python bench.py --impl reference ...                      (CPU oracle port on the host cores)

A step = one BramblePasciakCG call of exactly --iters iterations (tol = 0 never triggers) from the same
start vector on the same assembled system, through the Python mirror of the reference signature
(navier_stokes_solver_b200.solvers.BramblePasciakCG, printrates=False).  `value` times K steps with every
input resident in HBM; `e2e` times the same call with host (pinned) f, g, sol buffers, copies inside the
timed region.  Beside the headline the line carries: the MINRES iteration on the merged saddle operator,
time to tolerance, a SIMPLE time step (DoTimeStep/Project) on the same mesh, Gram-Schmidt, a solution
fingerprint that is comparable across N, and parity objects (device vs the C oracle at N = 1, row-partitioned
vs single-GPU at N > 1).
"""
import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Algorithmic bytes per BPCG v2 iteration beyond the SpMVs' own x/y traffic (frozen):
#   B1 epilogue r(As0,z0,dA,s0) w(As0,tmp1,tmp4) = 7 n_u ; B2 epilogue r(s0,tmp0) = 2 n_u ; B3 epilogue r(s1) = 1 n_p
#   B4 velocity r8 w4 = 12 n_u, pressure r6 w3 = 9 n_p ; B5 r2 w1 = 3 n_u + 3 n_p
BPCG_VEC_PASSES_U = 24
BPCG_VEC_PASSES_P = 13
# MINRES iteration beyond the SpMV's own x/y: re-read z (1) + K2 r4 w2 (6) + K3 r4 w2 (6) = 13 passes of N
MINRES_VEC_PASSES = 13
# Jacobi-PCG iteration (NGSolve CGSolver) beyond the SpMV's own x/y: C1 epilogue re-reads s (1); C2 r(u,s,d,q,dinv) w(u,d,w) = 8;
# C3 r(s,w) w(s) = 3
CG_VEC_PASSES = 12
# modified Gram-Schmidt, m = 5, 3 sweeps: 195 passes of N
GS_BYTES_PER_ROW = 1560
WORKLOAD = ("3D Stokes channel [0,2.5]x[0,0.41]^2 (NavierStokesSIMPLE_test_3D.py), Taylor-Hood P2/P1, "
            "BramblePasciakCG (bramblepasciak_new.py) with Jacobi preA/preM, %d iterations per step")
LENGTHS = (2.5, 0.41, 0.41)
REF_ITERS_PER_STEP = 5   # reference arm: iterations of the CPU port per timed step (bounded sample of the 100-iteration step)


def spmv_bytes(rows, cols, nnz):
    return 12 * nnz + 4 * (rows + 1) + 8 * cols + 8 * rows


def bpcg_iter_bytes(n_u, n_p, nnzA, nnzB):
    return (spmv_bytes(n_u, n_u, nnzA) + spmv_bytes(n_p, n_u, nnzB) + spmv_bytes(n_u, n_p, nnzB)
            + 8 * (BPCG_VEC_PASSES_U * n_u + BPCG_VEC_PASSES_P * n_p))


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def config_dict(iters, cells, n_u, n_p, nnzA, nnzB, world):
    """The `config` object; both arms print the same one for the same command line."""
    return {"workload": WORKLOAD % iters, "cells": list(cells), "dofs": int(n_u + n_p), "n_u": int(n_u), "n_p": int(n_p),
            "nnz_A": int(nnzA), "nnz_B": int(nnzB), "iters_per_step": iters, "row_partition": "x-slabs" if world > 1 else "none",
            "l2_policy": "inputs larger than L2 (%.1f GB of CSR per iteration)" % (12e-9 * (nnzA + 2 * nnzB))}


@contextlib.contextmanager
def quiet():
    """The Python mirrors print the reference's own lines (condition, err0, warnings); this script prints ONE line."""
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        yield buf


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class LegWatchdog:
    """Bounds each OPTIONAL leg that follows the headline measurement.

    The legs after the timed region contain collectives; when one rank of N fails in such a leg (the 4-GPU weak-scaling run: device
    memory on rank 0 only) the others wait for it forever and the job dies without its JSON line.  The line is therefore assembled
    BEFORE these legs, each leg adds its key when it returns, and a leg that does not return within `limit` seconds ends the run:
    rank 0 prints the line with what finished plus a `watchdog` note, every rank leaves with os._exit(0) (a rank blocked inside a
    collective cannot be unwound any other way).  The headline leg itself is not covered: a hang there has no number to save."""

    def __init__(self, limit, rank, line):
        self.limit, self.rank, self.line = float(limit), rank, line
        self.lock = threading.Lock()
        self.name, self.deadline, self.printed = None, None, False
        if self.limit > 0:
            threading.Thread(target=self._watch, daemon=True).start()

    def arm(self, name):
        with self.lock:
            self.name, self.deadline = name, time.monotonic() + self.limit

    def disarm(self):
        with self.lock:
            self.name, self.deadline = None, None

    @contextlib.contextmanager
    def leg(self, name):
        self.arm(name)
        try:
            yield
        finally:
            self.disarm()

    def emit(self):
        """Print the line once (rank 0); used by the normal end of main() and by the expiry path."""
        with self.lock:
            if self.printed or self.rank != 0:
                return
            self.printed = True
            text = json.dumps(self.line)
        sys.stdout.write(text + "\n")
        sys.stdout.flush()

    def _watch(self):
        while True:
            time.sleep(1.0)
            with self.lock:
                expired = self.deadline is not None and time.monotonic() > self.deadline
                name = self.name
                if expired and not self.printed:
                    self.line["watchdog"] = {"leg": name, "limit_s": self.limit,
                                             "note": "this optional leg did not return; the keys before it are complete, it and later legs are absent"}
            if expired:
                print("bench.py: leg '%s' exceeded %.0f s on rank %d; leaving" % (name, self.limit, self.rank), file=sys.stderr, flush=True)
                self.emit()
                os._exit(0)


# ------------------------------------------------------------------------------------------------
# CPU side: the C + OpenMP oracle port (oracle/c/krylov_omp.c) on the SAME assembled blocks
# ------------------------------------------------------------------------------------------------
class HostSystem:
    """Host (numpy) copy of an assembled Stokes system in the layout the C oracle takes."""

    def __init__(self, S, syn):
        from oracle import cbuild

        def trip(csr):
            return (np.ascontiguousarray(csr.rowptr.cpu().numpy(), dtype=np.int64),
                    np.ascontiguousarray(csr.colidx.cpu().numpy(), dtype=np.int32),
                    np.ascontiguousarray(csr.values.cpu().numpy(), dtype=np.float64))
        self.n_u, self.n_p = S.n_u, S.n_p
        self.cells = tuple(S.cells)
        self.nnzA, self.nnzB = S.A.nnz, S.B.nnz
        self.blocks = cbuild.Blocks.__new__(cbuild.Blocks)
        self.blocks.n_u, self.blocks.n_p = S.n_u, S.n_p
        self.blocks.A, self.blocks.B, self.blocks.BT = trip(S.A), trip(S.B), trip(S.BT)
        self.dA = syn.jacobi_inverse(S.A, free=S.free_u).cpu().numpy()
        self.dM = syn.jacobi_inverse(S.Mp).cpu().numpy()
        self.f, self.g = S.f.cpu().numpy(), S.g.cpu().numpy()
        self.u_start = np.concatenate([S.u0.cpu().numpy(), np.zeros(S.n_p)])


def assemble_host(dofs, cells=None):
    import torch
    import navier_stokes_solver_b200.synthetic as syn
    cells = cells or syn.stokes_cells_for_dofs(dofs)
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    S = syn.assemble_stokes(cells, LENGTHS, device=dev)
    H = HostSystem(S, syn)
    del S
    if dev == "cuda":
        torch.cuda.empty_cache()
    return H


def cpu_bpcg(H, k_scale, iters, tol=0.0, threads=None):
    """One call of the C port of BramblePasciakCG on the host cores -> (it, history, loop_seconds, sol, threads)."""
    from oracle import cbuild
    nt = cbuild.set_threads(threads)
    sol = H.u_start.copy()
    it, hist, err0, loop_s = cbuild.bpcg_v2(H.blocks, H.f, H.g, H.dA, H.dM, k_scale, sol, tol=tol, maxsteps=iters, initialize=False)
    return it, hist, loop_s, sol, nt


def cpu_lambda(H, steps=12):
    """A safe under-estimate of lambda_min(dA A) for the CPU arm (the loop's speed does not depend on it): a few steps of the
    oracle's Lanczos with the threaded C SpMV, halved."""
    from oracle import cbuild
    from oracle import krylov_oracle as ko
    import ctypes as C
    lib = cbuild.load()
    rp, ci, va = H.blocks.A
    n = H.n_u

    def A_mul(x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(n)
        lib.ko_spmv(C.c_int64(n), C.c_int64(n), rp.ctypes.data_as(C.c_void_p), ci.ctypes.data_as(C.c_void_p),
                    va.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p))
        return y
    ritz = ko.lanczos_eigs(A_mul, H.dA, n, maxsteps=steps, tol=1e-2)
    return 0.5 * float(ritz[0])


def run_reference(args):
    """`--impl reference`: the reference's CPU implementation of the path (C + OpenMP port, one threaded primitive per
    reference statement; NGSolve itself is absent) on the SAME 20.4M-dof blocks, every host core, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cbuild
    from oracle import krylov_oracle as ko
    import torch
    full = torch.cuda.is_available()          # no GPU: assembling 20M dofs on the host is out of reach; say so
    dofs = args.dofs if full else min(args.dofs, 1.0e6)
    H = assemble_host(dofs)
    threads = cbuild.set_threads()
    k = ko.bpcg_scale(cpu_lambda(H))
    its = REF_ITERS_PER_STEP
    times = []
    t_all = time.perf_counter()
    for i in range(args.warmup + args.steps):
        _, _, loop_s, _, _ = cpu_bpcg(H, k, its, threads=threads)
        if i >= args.warmup:
            times.append(loop_s)
        if time.perf_counter() - t_all > 240 and len(times) >= 1:   # keep the whole arm within minutes
            break
    steps_done = len(times)
    v = its * steps_done / float(np.sum(times))
    sample = ("C+OpenMP port (oracle/c/krylov_omp.c: ko_bpcg_v2) of solvers/bramblepasciak_new.py, unfused like the reference's "
              "NGSolve statements; each timed step = %d iterations of the loop on the SAME %d-dof system as the device arm "
              "(no dof scaling), %d threads set explicitly; NGSolve absent" % (its, H.n_u + H.n_p, threads))
    cfg = config_dict(args.iters, H.cells, H.n_u, H.n_p, H.nnzA, H.nnzB, args.gpus)
    line = {"impl": "reference", "metric": "bpcg_krylov_iterations_per_s", "value": v, "unit": "iterations/s",
            "n_gpus": args.gpus, "steps": steps_done, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(times)),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg, "sample_dofs": int(H.n_u + H.n_p), "same_system_as_device_arm": bool(full),
            "cpu_baseline": {"value": v, "unit": "iterations/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
def sell_stream_bytes(fmt, rows, cols):
    """Bytes one SpMV of a SELL-LX operator moves by construction (sell.cu): 10 B per stored nonzero, 12 B per column run (or 4 B
    per distinct column), 4 B per slice offset, 2 B per row (perm), x once, y once."""
    if fmt["rows_per_chunk"] <= 0:
        return None
    slices = fmt["chunks"] * fmt["rows_per_chunk"] * fmt["lanes_per_row"] // 32
    idx = 12 * fmt["column_runs"] if fmt["column_runs"] > 0 else 4 * fmt["total_distinct"]
    return 10 * fmt["stored_nnz"] + idx + 4 * slices + 2 * fmt["chunks"] * fmt["rows_per_chunk"] + 8 * cols + 8 * rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--dofs", type=float, default=20e6)
    ap.add_argument("--iters", type=int, default=100, help="Krylov iterations per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lanczos-steps", type=int, default=1500)
    ap.add_argument("--no-minres", action="store_true")
    ap.add_argument("--minres-iters", type=int, default=60)
    ap.add_argument("--solve-to-tol", type=float, default=1e-10,
                    help="after the timed legs, run one full BramblePasciakCG solve to this relative tolerance (the reference's call "
                         "uses 1e-10, NavierStokesSIMPLE_iterative.py:397) and report time to tolerance (0 = skip)")
    ap.add_argument("--solve-maxsteps", type=int, default=30000)
    ap.add_argument("--no-parity", action="store_true", help="skip parity_at_sample (N = 1) / parity_vs_single (N > 1)")
    ap.add_argument("--parity-dofs", type=float, default=2.0e6)
    ap.add_argument("--no-simple", action="store_true", help="skip the SIMPLE time-step leg")
    ap.add_argument("--simple-steps", type=int, default=2)
    ap.add_argument("--simple-proj-maxsteps", type=int, default=200)
    ap.add_argument("--no-gs", action="store_true", help="skip the Gram-Schmidt leg")
    ap.add_argument("--weak", action="store_true",
                    help="N > 1: add the weak-scaling leg (--weak-dofs unknowns per GPU).  Off by default: its first run (4 GPUs) ran out of "
                         "device memory on one rank only and the other ranks then waited in a collective forever (DESIGN.md 6)")
    ap.add_argument("--weak-dofs", type=float, default=20e6, help="unknowns per GPU of the weak-scaling leg")
    ap.add_argument("--no-precond", action="store_true", help="skip the preconditioner comparison leg (N = 1)")
    ap.add_argument("--precond-dofs", type=float, default=1.0e6)
    ap.add_argument("--precond-timeout", type=float, default=150.0,
                    help="the preconditioner leg (own process) is abandoned after this many seconds: the default run stays within minutes")
    ap.add_argument("--leg-timeout", type=float, default=420.0,
                    help="seconds each optional leg after the headline may take before the run prints what it has and leaves (0 = no limit)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import navier_stokes_solver_b200 as nsb
    import navier_stokes_solver_b200.synthetic as syn
    from navier_stokes_solver_b200 import _lib
    import benchlegs

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ctx = nsb.Context.from_torch_distributed(device=local)
    else:
        ctx = nsb.Context(device=local)
    nsb.set_default_context(ctx)

    cells = syn.stokes_cells_for_dofs(args.dofs)
    t_asm = time.perf_counter()
    host = None
    from navier_stokes_solver_b200 import partition
    if world == 1:
        S = syn.assemble_stokes(cells, LENGTHS, device="cuda")
        if not args.no_cpu_baseline:
            host = HostSystem(S, syn)           # the CPU baseline runs on these very blocks
        A = nsb.SparseMatrix.from_csr(S.A, ctx)
        B = nsb.SparseMatrix.from_csr(S.B, ctx)
        BT = nsb.SparseMatrix.from_csr(S.BT, ctx)
        B.set_transpose(BT)
        dA = syn.jacobi_inverse(S.A, free=S.free_u)
        dM = syn.jacobi_inverse(S.Mp)
        n_u, n_p = S.n_u, S.n_p
        nnzA, nnzB = S.A.nnz, S.B.nnz
        f_d, g_d = S.f, S.g
        u_start = torch.cat([S.u0, torch.zeros(n_p, dtype=torch.float64, device="cuda")])
        S.A = S.B = S.BT = None
        start_vec = None
        space_u, space_p = partition.stokes_slab_spaces(cells, 1)
        P = None
    else:
        P = partition.build_distributed_stokes(cells, LENGTHS, ctx, rank, world, keep_local=not args.no_minres)
        A, B, BT, dA, dM = P.A, P.B, P.BT, P.dA, P.dM
        n_u, n_p, nnzA, nnzB = P.n_u_global, P.n_p_global, P.nnzA_global, P.nnzB_global
        f_d, g_d, u_start, start_vec = P.f, P.g, P.u_start, P.lanczos_start
        space_u, space_p = P.local.space_u, P.local.space_p
    torch.cuda.synchronize()
    t_asm = time.perf_counter() - t_asm
    pA = nsb.DiagonalPreconditioner.from_diagonal(dA, ctx)
    pM = nsb.DiagonalPreconditioner.from_diagonal(dM, ctx)
    torch.cuda.empty_cache()
    n_u_loc = int(f_d.numel())

    # scale factor k = 1/lambda_min + 1e-3 by device Lanczos (bramblepasciak_new.py:115-118), once per assembly
    t_l = time.perf_counter()
    lams = nsb.EigenValues_Preconditioner(A, pA, tol=1e-5, maxsteps=args.lanczos_steps, start=start_vec)
    lanczos_steps = nsb.solvers.last_info["iterations"]
    t_l = time.perf_counter() - t_l
    lam_min = min(lams)
    lib = ctx._lib
    import ctypes as C
    BTm = B.CreateTranspose()
    k_scale = 1.0 / lam_min + 1e-3

    def step(f_buf, g_buf, sol_buf, iters=args.iters, tol=0.0):
        """One solve through the reference-facing wrapper (solvers.BramblePasciakCG, printrates=False)."""
        with quiet():
            it, secs = nsb.BramblePasciakCG(A, B, None, f_buf, g_buf, pA, pM, sol_buf, tol=tol, maxsteps=iters, printrates=False,
                                            initialize=False, rel_err=True, lam_min=lam_min)
        if tol == 0.0:
            assert it == iters - 1, (it, nsb.solvers.last_info)
        return secs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    # every kernel and copy of the library runs on the context's own stream: the timed region is bracketed by CUDA
    # events recorded on THAT stream (torch.cuda.Event on torch's current stream would see none of it), after a
    # barrier + synchronize on both sides; max over ranks.  The host clock around the same region is kept as a cross-check.
    lib_stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", local))
    wall_check = {}

    def timed(nsteps, fn, tag=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(lib_stream)
        loop = 0.0
        for _ in range(nsteps):
            loop += fn()
        e1.record(lib_stream)
        barrier()
        wall = time.perf_counter() - t0
        dt = e0.elapsed_time(e1) * 1e-3
        if world > 1:
            t = torch.tensor([dt, loop, wall], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt, loop, wall = float(t[0]), float(t[1]), float(t[2])
        if tag:
            wall_check[tag] = wall
        return dt, loop

    env = benchlegs.Env(args=args, nsb=nsb, syn=syn, partition=partition, ctx=ctx, world=world, rank=rank, local=local, dist=dist,
                        timed=timed, barrier=barrier, quiet=quiet, lib_stream=lib_stream, cells=cells, lengths=LENGTHS)

    # ---- device-resident leg
    sol_d = u_start.clone()

    def dev_step():
        sol_d.copy_(u_start)
        torch.cuda.current_stream().synchronize()
        return step(f_d, g_d, sol_d)

    timed(args.warmup, dev_step)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launch_count()
    ctx.sync_stats(reset=True)
    dt, loop = timed(args.steps, dev_step, 'device')
    launches = ctx.launch_count() - l0
    sync_stats = ctx.sync_stats()
    if world > 1 and any(v["count"] for v in sync_stats.values()):   # -DNSB_SYNC_STATS builds (tools/sync_stats.sh)
        print("SYNCSTATS " + json.dumps({"rank": rank, "loop_s": loop, "iterations": args.iters * args.steps,
                                         "us_per_iteration": {k: 1e-3 * v["ns"] / (args.iters * args.steps) for k, v in sync_stats.items()},
                                         "counts": {k: v["count"] for k, v in sync_stats.items()}}), file=sys.stderr, flush=True)
    clocks = sampler.stop() if rank == 0 else None
    its = args.iters * args.steps
    value = its / dt
    hist_dev = nsb.solvers.last_info["history"].copy()
    finger_100 = benchlegs.fingerprint(env, sol_d, n_u_loc, space_u, space_p, n_u, n_p)

    # ---- end-to-end leg: host pinned f, g, sol; H2D/D2H inside the timed region
    f_h = torch.empty(f_d.numel(), dtype=torch.float64).pin_memory()
    g_h = torch.empty(g_d.numel(), dtype=torch.float64).pin_memory()
    s_h = torch.empty(u_start.numel(), dtype=torch.float64).pin_memory()
    u_h = u_start.cpu()
    f_h.copy_(f_d)
    g_h.copy_(g_d)

    def e2e_step():
        return step(f_h, g_h, s_h)   # sol carries over between steps (tol = 0: always --iters iterations)

    s_h.copy_(u_h)
    timed(1, e2e_step)
    same = np.allclose(nsb.solvers.last_info["history"], hist_dev, rtol=1e-12)   # first host step == device leg
    dt_e, _ = timed(args.steps, e2e_step, 'e2e')
    e2e_value = its / dt_e
    h2d = 8 * (f_h.numel() + g_h.numel() + s_h.numel())
    d2h = 8 * s_h.numel() + 8 * (args.iters + 1)
    del f_h, g_h, s_h

    # ---- roofline of the dominant kernel (SpMV with A): CUDA events on the context's stream
    ms_A = A.spmv_bench(warmup=20, reps=200)   # 200 timed repetitions after 20 warm-up
    ms_B = B.spmv_bench(warmup=5, reps=20)
    ms_BT = BTm.spmv_bench(warmup=5, reps=20)
    peak, peak_src = measured_peaks()
    bytes_A = A.spmv_bytes()
    ach_A = bytes_A / (ms_A * 1e-3) / 1e9
    iter_bytes = bpcg_iter_bytes(n_u, n_p, nnzA, nnzB)
    iter_gbs = iter_bytes * (its / loop) / 1e9 / world if loop > 0 else None
    fmtA = A.format()
    stream_A = sell_stream_bytes(fmtA, A.shape[0], A.shape[1])

    # DRAM bytes per launch of the same kernel from a committed ncu --set full capture (never measured under this run); used only
    # when the capture was taken on the same storage format (fingerprint), else null -- the by-construction stream bytes stand in
    traffic, traffic_src = None, None
    for name in ("r2_spmv_A_traffic.json", "r1c_spmv_A_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if world == 1 and traffic is None and os.path.exists(tpath):
            tj = json.load(open(tpath))
            fp_ok = tj.get("format") is None or all(tj["format"].get(k) == fmtA.get(k) for k in ("stored_nnz", "column_runs", "chunks"))
            if tj.get("dofs") == int(n_u + n_p) and fp_ok:
                traffic, traffic_src = tj["dram_bytes_read"] + tj["dram_bytes_write"], "profiles/" + name
    line = {
        "metric": "bpcg_krylov_iterations_per_s", "value": value, "unit": "iterations/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args.iters, cells, n_u, n_p, nnzA, nnzB, world),
        "setup": {"lambda_min": lam_min, "lanczos_steps": lanczos_steps, "assembly_s": t_asm, "lanczos_s": t_l},
        "iteration_loop": {"iterations_per_s": its / loop if loop else None, "algorithmic_bytes_per_iteration": iter_bytes,
                           "achieved_gbs_per_gpu": iter_gbs, "frac_of_measured_peak": iter_gbs / peak if iter_gbs else None,
                           "frac_of_nominal_8TBs": iter_gbs / 8000.0 if iter_gbs else None},
        "roofline": {"bound": "hbm", "kernel": "spmv_sell_kernel (A, %d nnz%s)" % (nnzA, "" if world == 1 else ", this rank's rows"),
                     "achieved": ach_A, "peak": peak,
                     "peak_source": peak_src, "unit": "GB/s", "frac": ach_A / peak, "frac_of_nominal_8TBs": ach_A / 8000.0,
                     "algorithmic_bytes": bytes_A, "ms": ms_A, "traffic": traffic, "traffic_source": traffic_src,
                     # algorithmic bytes follow the 12 B/nnz CSR model; the kernel's SELL-LX stream is ~10 B/nnz, so `frac` can
                     # exceed 1 -- the same launch time against the bytes the format moves by construction / ncu counted:
                     "format": fmtA, "stream_bytes_by_construction": stream_A,
                     "frac_by_stream_bytes": (stream_A / (ms_A * 1e-3) / 1e9 / peak) if stream_A else None,
                     "frac_by_traffic": (traffic / (ms_A * 1e-3) / 1e9 / peak) if traffic else None,
                     "other": {"spmv_B_gbs": spmv_bytes(*_shape_nnz(n_p, n_u, nnzB, world)) / ms_B / 1e6 if world == 1 else None,
                               "spmv_BT_gbs": spmv_bytes(*_shape_nnz(n_u, n_p, nnzB, world)) / ms_BT / 1e6 if world == 1 else None,
                               "spmv_B_ms": ms_B, "spmv_BT_ms": ms_BT}},
        "e2e": {"value": e2e_value, "unit": "iterations/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "history_matches_device_leg": bool(same), "through": "solvers.BramblePasciakCG(..., printrates=False) with pinned host f, g, sol"},
        # residual-history fingerprint of the device leg: equal (to reduction-order rounding) at every N for the same --dofs
        "history_checkpoints": {"err0": float(hist_dev[0]), "err_%d" % (args.iters // 2): float(hist_dev[args.iters // 2]),
                                "err_%d" % args.iters: float(hist_dev[args.iters])},
        "solution_fingerprint_after_%d_iterations" % args.iters: finger_100,
        "gpu_launches": int(launches), "clocks": clocks,
        "timing": {"how": "CUDA events on the library's stream around the K steps, barrier + synchronize on both sides, max over ranks",
                   "host_clock_ms_per_step": 1e3 * wall_check.get("device", 0.0) / args.steps,
                   "host_clock_ms_per_step_e2e": 1e3 * wall_check.get("e2e", 0.0) / args.steps},
    }
    # The headline line is complete here.  Every leg below is optional, adds its key when it returns, and is bounded by the
    # watchdog: an asymmetric failure in one of them (one rank out of memory, the others waiting in a collective) costs that leg,
    # not the line.
    wd = LegWatchdog(args.leg_timeout, rank, line)

    # ---- MINRES on the merged saddle-point operator K = [[A, B^T], [B, 0]] (minres.py; BASELINE metric names both), every N
    if not args.no_minres:
        with wd.leg("minres"):
            minres_info = benchlegs.minres_leg(env, A, B, BTm, pA, pM, dA, dM, f_d, g_d, u_start, P, n_u, n_p, nnzA, nnzB, peak,
                                               spmv_bytes, MINRES_VEC_PASSES)
        if minres_info:
            line["minres"] = minres_info
    if P is not None:
        P.local.A = P.local.B = P.local.BT = None
        torch.cuda.empty_cache()

    # ---- time to tolerance (the reference's call: tol 1e-10 relative), every N; the solution's fingerprint
    if args.solve_to_tol > 0:
        tsol = u_start.clone()
        with wd.leg("time_to_tolerance"):
            barrier()
            t0 = time.perf_counter()
            try:   # an extra leg after the timed ones: a failure here is reported, it does not take the headline line down
                with quiet():
                    it_t, secs_t = nsb.BramblePasciakCG(A, B, None, f_d, g_d, pA, pM, tsol, tol=args.solve_to_tol,
                                                        maxsteps=args.solve_maxsteps, printrates=False, initialize=False, rel_err=True,
                                                        lam_min=lam_min)
                barrier()
                info = dict(nsb.solvers.last_info)
                th = info["history"]
                ttt = {"tol": args.solve_to_tol, "rel_err": True, "preA": "Jacobi", "iterations": int(info["updates"]),
                       "loop_seconds": secs_t, "call_seconds": time.perf_counter() - t0, "status": info["status"],
                       "final_relative_error": float(th[-1] / th[0]) if th[0] > 0 else None,
                       "solution_fingerprint": benchlegs.fingerprint(env, tsol, n_u_loc, space_u, space_p, n_u, n_p)}
            except Exception as ex:   # noqa: BLE001  (NsbError or a CUDA error surfaced by torch)
                ttt = {"tol": args.solve_to_tol, "error": str(ex)[:200]}
        line["time_to_tolerance"] = ttt
        del tsol

    # ---- Gram-Schmidt (orthonormalization.py) at the bench's vector length
    if not args.no_gs:
        with wd.leg("gram_schmidt"):
            gs_info = benchlegs.guarded(benchlegs.gram_schmidt_leg, env, n_u_loc + int(g_d.numel()), n_u + n_p, peak, GS_BYTES_PER_ROW)
        if gs_info:
            line["gram_schmidt"] = gs_info

    # ---- the operators of the headline are no longer needed: free them before the legs that assemble their own
    del A, B, BT, BTm, pA, pM, sol_d
    if P is not None:
        P.A = P.B = P.BT = None
    P = None
    torch.cuda.empty_cache()

    # ---- SIMPLE time step (DoTimeStep / Project, NavierStokesSIMPLE_iterative.py:427-443) on the same mesh, every N
    if not args.no_simple:
        with wd.leg("simple_step"):
            simple_info = benchlegs.guarded(benchlegs.simple_step_leg, env, peak, spmv_bytes, CG_VEC_PASSES)
        if simple_info:
            line["simple_step"] = simple_info

    # ---- weak scaling beside the strong-scaling headline: --weak-dofs unknowns PER GPU (N > 1)
    if world > 1 and args.weak:
        with wd.leg("weak_scaling"):
            weak_info = benchlegs.guarded(benchlegs.weak_scaling_leg, env, args.weak_dofs, args.iters, peak, bpcg_iter_bytes)
        if weak_info:
            line["weak_scaling"] = weak_info

    # ---- parity where the driver can see it
    if not args.no_parity:
        with wd.leg("parity"):
            if world == 1:
                parity = benchlegs.guarded(benchlegs.parity_at_sample, env, assemble_host, cpu_bpcg)
            else:
                parity = benchlegs.guarded(benchlegs.parity_vs_single, env)
        if parity:
            line["parity_at_sample" if world == 1 else "parity_vs_single"] = parity

    if host is not None and rank == 0:
        from oracle import krylov_oracle as ko   # noqa: F401  (builds / loads the C oracle before the timed CPU runs)
        best, nt = None, None
        for _ in range(3):
            _, _, loop_s, _, nt = cpu_bpcg(host, k_scale, 20)
            best = loop_s if best is None or loop_s < best else best
        line["cpu_baseline"] = {"value": 20 / best, "unit": "iterations/s", "cores": nt, "kind": "port",
                                "sample": "C+OpenMP oracle port of bramblepasciak_new.py (oracle/c/krylov_omp.c), 20 iterations on the SAME "
                                          "%d-dof blocks the device solves (no dof scaling), best of 3, %d threads set explicitly"
                                          % (n_u + n_p, nt)}
    # ---- the preconditioners that cut the iteration count (block smoother, auxiliary-space MypreA) beside Jacobi: time to tolerance
    # at --precond-dofs, N = 1, in a process of its own, last -- whatever happens there costs only this key
    if world == 1 and not args.no_precond:
        torch.cuda.empty_cache()
        with wd.leg("preconditioners"):
            line["preconditioners"] = benchlegs.precond_leg(args.precond_dofs, args.precond_timeout)
    wd.emit()
    if world > 1:
        with wd.leg("teardown"):
            dist.destroy_process_group()


def _shape_nnz(r, c, z, world):
    return r, c, z


if __name__ == "__main__":
    main()
