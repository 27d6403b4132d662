#!/usr/bin/env python
"""Preconditioner leg of bench.py, run in its OWN process (bench.py: `preconditioners`): time to tolerance of BramblePasciakCG
(solvers/bramblepasciak_new.py, tol 1e-10 relative as at NavierStokesSIMPLE_iterative.py:397) on the 3D channel at --dofs unknowns
with each device preconditioner for A:

    jacobi        Preconditioner(a, 'local')                                       (run.py:62)
    block_jacobi  a.mat.CreateBlockSmoother(blocks) applied additively             (NavierStokesSIMPLE_iterative.py:373,:383)
    block_gs      y = 0; Smooth(y, x); SmoothBack(y, x)                            (:375-381 without the auxiliary term)
    aux_gs        MypreA.Mult with GS = True                                       (:374-381)

The reference's facet blocks, H(div) <- H1 transform and h1amg coarse solver live in NGSolve; the stand-ins here are blocks of four
consecutive free dofs, a piecewise-constant aggregation of six, and Jacobi on the Galerkin coarse matrix -- so the iteration counts say
what THESE preconditioners buy, the times what their kernels cost.  Separate process: this leg has run on small meshes only
(tests/test_gpu_parity.py); whatever happens to it cannot take bench.py's line down.  Prints one JSON line."""
import argparse
import contextlib
import io
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
LENGTHS = (2.5, 0.41, 0.41)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dofs", type=float, default=2.0e6)
    ap.add_argument("--tol", type=float, default=1e-10)
    ap.add_argument("--maxsteps", type=int, default=20000)
    ap.add_argument("--lanczos-steps", type=int, default=1500)
    ap.add_argument("--safety", type=float, default=0.9, help="lambda_min estimate is multiplied by this before k = 1/lambda + 1e-3")
    ap.add_argument("--which", default="jacobi,block_jacobi,block_gs,aux_gs")
    args = ap.parse_args()

    import scipy.sparse as sp
    import torch
    import navier_stokes_solver_b200 as nsb
    import navier_stokes_solver_b200.synthetic as syn
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device")
    ctx = nsb.Context(device=0)
    nsb.set_default_context(ctx)
    cells = syn.stokes_cells_for_dofs(args.dofs)
    t0 = time.perf_counter()
    S = syn.assemble_stokes(cells, LENGTHS, device="cuda")
    A_host = S.A.to_scipy()                                   # colouring and the Galerkin diagonal are host set-up work
    free = S.free_u.cpu().numpy().astype(bool)
    A, B, BT = (nsb.SparseMatrix.from_csr(m, ctx) for m in (S.A, S.B, S.BT))
    B.set_transpose(BT)
    pM = nsb.DiagonalPreconditioner.from_diagonal(syn.jacobi_inverse(S.Mp), ctx)
    t_asm = time.perf_counter() - t0
    start = torch.cat([S.u0, torch.zeros(S.n_p, dtype=torch.float64, device="cuda")])
    lz = (1.0 + 0.5 * torch.sin(torch.arange(1, S.n_u + 1, dtype=torch.float64, device="cuda"))) * S.free_u.to(torch.float64)
    out = {"dofs": int(S.n_u + S.n_p), "cells": list(cells), "tol": args.tol, "lambda_safety": args.safety, "assembly_s": t_asm,
           "stand_ins": "blocks of 4 consecutive free dofs; aggregation of 6 with Jacobi on the Galerkin matrix for transform / h1amg"}

    fi = np.nonzero(free)[0]
    blocks = [fi[i:i + 4] for i in range(0, len(fi), 4)]
    built = {}

    def smoother(gs):
        key = "gs" if gs else "add"
        if key not in built:
            t = time.perf_counter()
            built[key] = nsb.BlockSmoother(A, blocks, pattern=A_host if gs else None, gauss_seidel=gs)   # (colouring: sweeps only)
            out.setdefault("setup_s", {})["block_smoother_" + key] = time.perf_counter() - t
            out["blocks"] = len(built[key].blocks)
            if gs:
                out["colours"] = int(len(built[key].color_ptr) - 1)
        return built[key]

    def make(name):
        if name == "jacobi":
            return nsb.DiagonalPreconditioner.from_diagonal(syn.jacobi_inverse(S.A, free=S.free_u), ctx)
        if name == "block_jacobi":
            return smoother(False)
        if name == "block_gs":
            return smoother(True)
        if name == "aux_gs":
            t = time.perf_counter()
            agg = np.arange(len(fi)) // 6
            T = sp.csr_matrix((np.ones(len(fi)), (fi, agg)), shape=(S.n_u, int(agg.max()) + 1))
            dc = 1.0 / (T.T @ A_host @ T).diagonal()
            pre = nsb.AuxiliarySpacePreconditioner(smoother(True), nsb.SparseMatrix.from_scipy(T, ctx),
                                                   nsb.DiagonalPreconditioner.from_diagonal(dc, ctx), GS=True)
            out.setdefault("setup_s", {})["aux_space"] = time.perf_counter() - t
            return pre
        raise SystemExit("unknown preconditioner " + name)

    for name in args.which.split(","):
        row = {}
        try:
            pre = make(name)
            t = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                lams = nsb.EigenValues_Preconditioner(A, pre, tol=1e-5, maxsteps=args.lanczos_steps, start=lz)
            row["lanczos_s"], row["lanczos_steps"] = time.perf_counter() - t, int(nsb.solvers.last_info["iterations"])
            lam = args.safety * min(lams)
            sol = start.clone()
            l0 = ctx.launch_count()
            t = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                it, secs = nsb.BramblePasciakCG(A, B, None, S.f, S.g, pre, pM, sol, tol=args.tol, maxsteps=args.maxsteps, printrates=False,
                                                initialize=False, rel_err=True, lam_min=lam)
            info = dict(nsb.solvers.last_info)
            h = info["history"]
            row.update(lambda_min=float(min(lams)), iterations=int(info["updates"]), loop_seconds=float(secs),
                       call_seconds=time.perf_counter() - t, status=info["status"], ms_per_iteration=1e3 * float(secs) / max(int(info["updates"]), 1),
                       launches_per_iteration=(ctx.launch_count() - l0) / max(int(info["updates"]), 1),
                       final_relative_error=float(h[-1] / h[0]) if h[0] > 0 else None,
                       norm_u=float(sol[:S.n_u].norm()), norm_p=float(sol[S.n_u:].norm()))
        except Exception as ex:   # noqa: BLE001
            row["error"] = str(ex)[:300]
        out[name] = row
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
