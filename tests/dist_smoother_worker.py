"""Worker of tests/test_gpu_dist_condensed.py::test_two_rank_block_smoother_and_auxspace (torchrun, one rank per GPU): the block smoother
(Jacobi, Smooth, SmoothBack, symmetric Gauss-Seidel) and the auxiliary-space composite MypreA (templates/NavierStokesSIMPLE_iterative.py
:360-383) row-partitioned over the ranks against the same objects on one GPU, then CG and BramblePasciakCG preconditioned by the
composite (:397) against their single-GPU runs.  One JSON line per rank."""
import contextlib
import io
import json
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import navier_stokes_solver_b200 as nsb  # noqa: E402
import navier_stokes_solver_b200.partition as P  # noqa: E402
import navier_stokes_solver_b200.synthetic as syn  # noqa: E402
from test_partition_smoother_gloo import CELLS, LENGTHS, slab_blocks  # noqa: E402


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def rel(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    G = syn.assemble_stokes(CELLS, LENGTHS, device="cpu")
    A, B, BT = G.A.to_scipy(), G.B.to_scipy(), G.BT.to_scipy()
    free = G.free_u.numpy().astype(bool)
    dM = syn.jacobi_inverse(G.Mp).numpy()
    f, g = G.f.numpy(), G.g.numpy()
    start = np.concatenate([G.u0.numpy(), np.zeros(G.n_p)])
    su, sp_ = P.stokes_slab_spaces(CELLS, world)
    blocks = slab_blocks(su, free)
    free_idx = np.nonzero(free)[0]
    agg = np.arange(len(free_idx)) // 6
    nc = int(agg.max()) + 1
    T = sp.csr_matrix((np.ones(len(free_idx)), (free_idx, agg)), shape=(G.n_u, nc))
    dc = 1.0 / (T.T @ A @ T).diagonal()
    rng = np.random.default_rng(31)
    x = rng.uniform(-1, 1, G.n_u) * free
    y0 = rng.uniform(-1, 1, G.n_u) * free

    # ---- one GPU
    ctx1 = nsb.Context(device=local)
    A1, B1 = nsb.SparseMatrix.from_scipy(A, ctx1), nsb.SparseMatrix.from_scipy(B, ctx1)
    sm1 = {gs: nsb.BlockSmoother(A1, blocks, pattern=A, gauss_seidel=gs) for gs in (False, True)}
    cm_blocks, cptr = sm1[True].blocks, sm1[True].color_ptr
    T1 = nsb.SparseMatrix.from_scipy(T, ctx1)
    co1 = nsb.DiagonalPreconditioner.from_diagonal(dc, ctx1)
    pre1 = {gs: nsb.AuxiliarySpacePreconditioner(sm1[gs], T1, co1, GS=gs) for gs in (False, True)}
    pM1 = nsb.DiagonalPreconditioner.from_diagonal(dM, ctx1)

    # ---- row-partitioned
    ctx = nsb.Context.from_torch_distributed(device=local)
    sa = P.IndexSpace([[(nc * r // world, nc * (r + 1) // world)] for r in range(world)])
    Lu = P.distribute_operators([A, B], [su, sp_], su, rank)
    Lp = P.distribute_operators([BT], [su], sp_, rank)
    Lt = P.distribute_operators([T], [su], sa, rank)
    Ltt = P.distribute_operators([T.T.tocsr()], [sa], su, rank)
    hu, hp, ha, hv = (P.upload_halo(ctx, L.plan) for L in (Lu, Lp, Lt, Ltt))
    AN, BN = nsb.SparseMatrix.from_csr(Lu.ops[0], ctx), nsb.SparseMatrix.from_csr(Lu.ops[1], ctx)
    BTN = nsb.SparseMatrix.from_csr(Lp.ops[0], ctx)
    TN, TTN = nsb.SparseMatrix.from_csr(Lt.ops[0], ctx), nsb.SparseMatrix.from_csr(Ltt.ops[0], ctx)
    hu.attach(AN)
    hu.attach(BN)
    hp.attach(BTN)
    ha.attach(TN)
    hv.attach(TTN)
    BN.set_transpose(BTN)
    lb, lptr = P.localize_blocks(cm_blocks, cptr, su, rank)
    ou, op_, oa = su.owned_ids(rank).numpy(), sp_.owned_ids(rank).numpy(), sa.owned_ids(rank).numpy()
    smN = {gs: nsb.BlockSmoother(AN, lb, gauss_seidel=gs, color_ptr=lptr) for gs in (False, True)}
    coN = nsb.DiagonalPreconditioner.from_diagonal(np.ascontiguousarray(dc[oa]), ctx)
    preN = {gs: nsb.AuxiliarySpacePreconditioner(smN[gs], TN, coN, GS=gs, transform_t=TTN) for gs in (False, True)}
    pMN = nsb.DiagonalPreconditioner.from_diagonal(np.ascontiguousarray(dM[op_]), ctx)
    xo = np.ascontiguousarray(x[ou])

    res = dict(rank=rank, local_blocks=len(lb), colours=int(len(lptr) - 1))
    res["jacobi_rel"] = rel((smN[False] * xo).NumPy(), (sm1[False] * x).NumPy()[ou])
    res["symgs_rel"] = rel((smN[True] * xo).NumPy(), (sm1[True] * x).NumPy()[ou])
    for back in (False, True):
        y1, yN = y0.copy(), np.ascontiguousarray(y0[ou])
        (sm1[True].SmoothBack if back else sm1[True].Smooth)(y1, x)
        (smN[True].SmoothBack if back else smN[True].Smooth)(yN, xo)
        res["smooth_back_rel" if back else "smooth_rel"] = rel(yN, y1[ou])
    for gs in (False, True):
        res["aux_gs_rel" if gs else "aux_add_rel"] = rel((preN[gs] * xo).NumPy(), (pre1[gs] * x).NumPy()[ou])

    # CG on A with the multiplicative composite
    rhs = (A @ (free * rng.uniform(-1, 1, G.n_u))) * free
    u1, e1 = nsb.CGSolve(A1, rhs, pre1[True], None, 1e-10, 2000)
    uN, eN = nsb.CGSolve(AN, np.ascontiguousarray(rhs[ou]), preN[True], None, 1e-10, 2000)
    m = min(30, len(e1), len(eN))
    res.update(cg_it=len(eN), cg_it1=len(e1), cg_hist_rel=float(np.max(np.abs(np.array(eN[:m]) - np.array(e1[:m])) / np.abs(np.array(e1[:m])))),
               cg_sol_rel=float(np.linalg.norm(np.asarray(uN) - np.asarray(u1)[ou]) / np.linalg.norm(np.asarray(u1)[ou])))

    # BramblePasciakCG with the composite as preA (:397)
    lam = 0.5 * min(quiet(nsb.EigenValues_Preconditioner, A1, pre1[True], tol=1e-9, maxsteps=2000))
    sol1 = start.copy()
    it1, _ = quiet(nsb.BramblePasciakCG, A1, B1, None, f, g, pre1[True], pM1, sol1, tol=1e-10, maxsteps=20000, initialize=False, lam_min=lam)
    h1 = nsb.solvers.last_info["history"].copy()
    sol = np.concatenate([start[:G.n_u][ou], start[G.n_u:][op_]])
    it, _ = quiet(nsb.BramblePasciakCG, AN, BN, None, np.ascontiguousarray(f[ou]), np.ascontiguousarray(g[op_]), preN[True], pMN, sol,
                  tol=1e-10, maxsteps=20000, initialize=False, lam_min=lam)
    h = nsb.solvers.last_info["history"].copy()
    ref_own = np.concatenate([sol1[:G.n_u][ou], sol1[G.n_u:][op_]])
    m = min(30, len(h), len(h1))
    res.update(it=int(it), it1=int(it1), hist_rel=float(np.max(np.abs(h[:m] - h1[:m]) / np.abs(h1[:m]))),
               sol_rel=float(np.linalg.norm(sol - ref_own) / np.linalg.norm(ref_own)))
    # MinRes on the stacked saddle operator with the block-diagonal preconditioner [[MypreA, 0], [0, Jacobi(Mp)]] (run.py:45-52)
    import types
    K1 = nsb.BlockMatrix([[A1, B1.T], [B1, None]])
    C1 = nsb.BlockMatrix([[pre1[True], None], [None, pM1]])
    msol1 = start.copy()
    _, merr1 = quiet(nsb.MinRes, K1, np.concatenate([f, g]), C1, msol1, 3000, False, False, 1e-8)
    Lloc = types.SimpleNamespace(plan_u=Lu.plan, plan_p=Lp.plan, A=Lu.ops[0], B=Lu.ops[1], BT=Lp.ops[0], n_u=Lu.plan.n_owned,
                                 n_p=Lp.plan.n_owned)
    Kl, kplan = P.local_saddle(Lloc)
    hk = P.upload_halo(ctx, kplan)
    KN = nsb.SparseMatrix.from_csr(Kl, ctx)
    KN.set_row_panels([0, Lloc.n_u, Lloc.n_u + Lloc.n_p])
    hk.attach(KN)
    CN = nsb.BlockMatrix([[preN[True], None], [None, pMN]])
    msol = np.concatenate([start[:G.n_u][ou], start[G.n_u:][op_]])
    _, merr = quiet(nsb.MinRes, KN, np.concatenate([f[ou], g[op_]]), CN, msol, 3000, False, False, 1e-8)
    mref = np.concatenate([msol1[:G.n_u][ou], msol1[G.n_u:][op_]])
    mm = min(30, len(merr), len(merr1))
    res.update(minres_it=len(merr) - 1, minres_it1=len(merr1) - 1,
               minres_hist_rel=float(np.max(np.abs(np.array(merr[:mm]) - np.array(merr1[:mm])) / np.abs(np.array(merr1[:mm])))),
               minres_sol_rel=float(np.linalg.norm(msol - mref) / np.linalg.norm(mref)))
    print("SMOOTHRESULT " + json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
