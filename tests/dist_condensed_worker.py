"""Worker of tests/test_gpu_dist_condensed.py (torchrun, one rank per GPU): the statically condensed BramblePasciakCG
(solvers/bramblepasciak_new.py:8-21,84-106,:188 call site) row-partitioned over the ranks against the single-GPU condensed solve of
the same global system, the condensed operator apply and the condensed Lanczos estimate beside it.  One JSON line per rank."""
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

# the 2D channel of oracle.cases "chan2d" (run.py:22-25 without the cylinder): on the 3D box the condensed Bramble-Pasciak form is only
# semi-definite and two roundings of the recurrence decorrelate after ~60 iterations (tests/test_gpu_parity.py), so 1 GPU vs N GPUs
# could only be compared loosely there
CELLS, LENGTHS = (24, 6), (2.0, 0.41)


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


class Blf:
    condense = True


def main():
    from oracle.cases import synthetic_condensation
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    G = syn.assemble_stokes(CELLS, LENGTHS, device="cpu")              # every rank holds the global system (a few thousand dofs)
    A, B, BT = G.A.to_scipy(), G.B.to_scipy(), G.BT.to_scipy()
    free = G.free_u.numpy().astype(bool)
    cd = synthetic_condensation(A, free)
    dS = cd["mat"].diagonal()
    dAc = np.where(free & ~cd["interior"] & (dS != 0), 1.0 / np.where(dS != 0, dS, 1.0), 0.0)     # Jacobi of the condensed block
    dM = syn.jacobi_inverse(G.Mp).numpy()
    f, g = G.f.numpy(), G.g.numpy()
    start = np.concatenate([G.u0.numpy(), np.zeros(G.n_p)])
    lz = 1.0 + 0.5 * np.sin(np.arange(G.n_u, dtype=np.float64) + 1.0)

    # ---- single-GPU condensed reference (every rank, on its own GPU)
    ctx1 = nsb.Context(device=local)
    blf1 = Blf()
    for k, name in (("mat", "mat"), ("H", "harmonic_extension"), ("HT", "harmonic_extension_trans"), ("inner_solve", "inner_solve"),
                    ("inner_matrix", "inner_matrix")):
        setattr(blf1, name, nsb.SparseMatrix.from_scipy(cd[k], ctx1))
    B1 = nsb.SparseMatrix.from_scipy(B, ctx1)
    pA1, pM1 = nsb.DiagonalPreconditioner.from_diagonal(dAc, ctx1), nsb.DiagonalPreconditioner.from_diagonal(dM, ctx1)
    K1 = nsb.CondensedOperator.from_bilinear_form(blf1)
    lam_1 = min(quiet(nsb.EigenValues_Preconditioner, K1, pA1, tol=1e-8, maxsteps=400, start=lz))
    # the solves use a safe scale factor from the Schur complement, the same on both sides (as tests/test_gpu_parity.py does)
    g_ = dAc > 0
    Sg = cd["mat"][g_][:, g_].toarray()
    lam = 0.9 * float(np.linalg.eigvalsh(np.sqrt(dAc[g_])[:, None] * Sg * np.sqrt(dAc[g_])[None, :])[0])
    sol1 = start.copy()
    it1, _ = quiet(nsb.BramblePasciakCG, blf1, B1, None, f, g, pA1, pM1, sol1, tol=1e-10, maxsteps=5000, initialize=False, lam_min=lam)
    h1 = nsb.solvers.last_info["history"].copy()
    x = np.random.default_rng(5).uniform(-1, 1, G.n_u)
    y1 = (K1 * x).NumPy().copy()

    # ---- row-partitioned
    ctx = nsb.Context.from_torch_distributed(device=local)
    su, sp_ = P.stokes_slab_spaces(CELLS, world)
    D = P.build_distributed_condensed(ctx, rank, su, sp_, cd["mat"], cd["H"], cd["HT"], cd["inner_solve"], cd["inner_matrix"], B, BT)
    ou, op_ = su.owned_ids(rank).numpy(), sp_.owned_ids(rank).numpy()
    pA, pM = nsb.DiagonalPreconditioner.from_diagonal(dAc[ou], ctx), nsb.DiagonalPreconditioner.from_diagonal(dM[op_], ctx)
    KN = nsb.CondensedOperator.from_bilinear_form(D)
    yN = (KN * np.ascontiguousarray(x[ou])).NumPy()
    lam_d = min(quiet(nsb.EigenValues_Preconditioner, KN, pA, tol=1e-8, maxsteps=400, start=np.ascontiguousarray(lz[ou])))
    sol = np.concatenate([start[:G.n_u][ou], start[G.n_u:][op_]])
    it, _ = quiet(nsb.BramblePasciakCG, D, D.B, None, np.ascontiguousarray(f[ou]), np.ascontiguousarray(g[op_]), pA, pM, sol, tol=1e-10,
                  maxsteps=5000, initialize=False, lam_min=lam)
    h = nsb.solvers.last_info["history"].copy()
    ref_own = np.concatenate([sol1[:G.n_u][ou], sol1[G.n_u:][op_]])
    m = min(50, len(h), len(h1))
    res = dict(rank=rank, it=int(it), it1=int(it1), apply_rel=float(np.abs(yN - y1[ou]).max() / np.abs(y1).max()),
               lam_rel=float(abs(lam_d - lam_1) / lam_1), hist_rel=float(np.max(np.abs(h[:m] - h1[:m]) / np.abs(h1[:m]))),
               sol_rel=float(np.linalg.norm(sol - ref_own) / np.linalg.norm(ref_own)),
               halo_u=int(D.plan_u.n_halo), halo_p=int(D.plan_p.n_halo))
    print("CONDRESULT " + json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
