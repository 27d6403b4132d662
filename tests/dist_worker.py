"""Worker of tests/test_gpu_dist.py (launched under torchrun, one rank per GPU): row-partitioned BPCG and
MinRes against the single-GPU solve of the same global system; prints one JSON line per rank."""
import contextlib
import io
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import navier_stokes_solver_b200 as nsb  # noqa: E402
import navier_stokes_solver_b200.partition as P  # noqa: E402
import navier_stokes_solver_b200.synthetic as syn  # noqa: E402

CELLS, LENGTHS = (12, 3, 3), (2.5, 0.41, 0.41)


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    os.environ.setdefault("NSB_SELL_MIN_NNZ", os.environ.get("NSB_SELL_MIN_NNZ", "200000"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = "cuda:%d" % local
    # ---- single-GPU reference of the global system (every rank, on its own GPU)
    ctx1 = nsb.Context(device=local)
    G = syn.assemble_stokes(CELLS, LENGTHS, device=dev)
    A1, B1 = nsb.SparseMatrix.from_csr(G.A, ctx1), nsb.SparseMatrix.from_csr(G.B, ctx1)
    dA = syn.jacobi_inverse(G.A, free=G.free_u)
    dM = syn.jacobi_inverse(G.Mp)
    pA1, pM1 = nsb.DiagonalPreconditioner.from_diagonal(dA, ctx1), nsb.DiagonalPreconditioner.from_diagonal(dM, ctx1)
    lam = min(quiet(nsb.EigenValues_Preconditioner, A1, pA1, tol=1e-8, maxsteps=400))
    sol1 = torch.cat([G.u0, torch.zeros(G.n_p, dtype=torch.float64, device=dev)])
    it1, _ = quiet(nsb.BramblePasciakCG, A1, B1, None, G.f, G.g, pA1, pM1, sol1, tol=1e-10, maxsteps=5000, initialize=False,
                   lam_min=lam)
    h1 = nsb.solvers.last_info["history"].copy()
    K1 = nsb.BlockMatrix([[A1, B1.T], [B1, None]])
    C1 = nsb.BlockMatrix([[pA1, None], [None, pM1]])
    msol1 = torch.cat([G.u0, torch.zeros(G.n_p, dtype=torch.float64, device=dev)])
    _, merr1 = quiet(nsb.MinRes, K1, torch.cat([G.f, G.g]), C1, msol1, 3000, False, False, 1e-8)

    # ---- row-partitioned solve
    ctx = nsb.Context.from_torch_distributed(device=local)
    D = P.build_distributed_stokes(CELLS, LENGTHS, ctx, rank, world, device=dev)
    pA = nsb.DiagonalPreconditioner.from_diagonal(D.dA, ctx)
    pM = nsb.DiagonalPreconditioner.from_diagonal(D.dM, ctx)
    sol = D.u_start.clone()
    it, _ = quiet(nsb.BramblePasciakCG, D.A, D.B, None, D.f, D.g, pA, pM, sol, tol=1e-10, maxsteps=5000, initialize=False,
                  lam_min=lam)
    h = nsb.solvers.last_info["history"].copy()
    ou, op_ = D.local.owned_u, D.local.owned_p
    ref_own = torch.cat([sol1[:G.n_u][ou], sol1[G.n_u:][op_]])
    m = min(50, len(h), len(h1))
    res = dict(rank=rank, it=int(it), it1=int(it1),
               hist_rel=float(np.max(np.abs(h[:m] - h1[:m]) / np.abs(h1[:m]))),
               sol_rel=float((sol - ref_own).norm() / ref_own.norm()))
    # device Lanczos on the partitioned operator reproduces the single-GPU estimate (same global start vector)
    lam_d = min(quiet(nsb.EigenValues_Preconditioner, D.A, pA, tol=1e-8, maxsteps=400, start=D.lanczos_start))
    res["lam_rel"] = abs(lam_d - lam) / lam

    # MinRes on the stacked local saddle operator
    Kl, plan = P.local_saddle(P.localize_stokes(syn.assemble_stokes(CELLS, LENGTHS, device=dev, slab=rank, nslabs=world),
                                                rank, world))
    hk = P.upload_halo(ctx, plan)
    K = nsb.SparseMatrix.from_csr(Kl, ctx)
    K.set_row_panels([0, D.local.n_u, D.local.n_u + D.local.n_p])
    hk.attach(K)
    C = nsb.DiagonalPreconditioner.from_diagonal(torch.cat([D.dA, D.dM]), ctx)
    msol = D.u_start.clone()
    _, merr = quiet(nsb.MinRes, K, torch.cat([D.f, D.g]), C, msol, 3000, False, False, 1e-8)
    mref = torch.cat([msol1[:G.n_u][ou], msol1[G.n_u:][op_]])
    mm = min(50, len(merr), len(merr1))
    res.update(minres_it=len(merr) - 1, minres_it1=len(merr1) - 1,
               minres_hist_rel=float(np.max(np.abs(np.array(merr[:mm]) - np.array(merr1[:mm])) / np.abs(np.array(merr1[:mm])))),
               minres_sol_rel=float((msol - mref).norm() / mref.norm()))
    # SIMPLE time step (nsb_simple_step) row-partitioned vs the same steps on one GPU
    import navier_stokes_solver_b200.simple_synth as ss
    kw = dict(proj_precision=1e-10, proj_maxsteps=3000, mstar_precision=1e-10, mstar_maxsteps=500)
    D1 = ss.build_device_simple(CELLS, LENGTHS, 1e-3, ctx1, device=dev, **kw)
    DN = ss.build_device_simple(CELLS, LENGTHS, 1e-3, ctx, rank, world, device=dev, **kw)
    for _ in range(2):
        D1.stepper.DoTimeStep()
        DN.stepper.DoTimeStep()
    own = DN.sys.owned_u
    g1 = D1.stepper.gfu[own]
    p1 = D1.stepper.gfup[DN.sys.owned_p]
    res.update(simple_gfu_rel=float((DN.stepper.gfu - g1).norm() / g1.norm()),
               simple_gfup_rel=float((DN.stepper.gfup - p1).norm() / max(float(p1.norm()), 1e-300)),
               simple_its=list(DN.stepper.cg_iterations), simple_its1=list(D1.stepper.cg_iterations))
    print("DISTRESULT " + json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
