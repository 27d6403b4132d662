import os, sys, io, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import navier_stokes_solver_b200 as nsb
from oracle import krylov_oracle as ko
from oracle.cases import Case, synthetic_condensation
for mesh in ("chan3d",):
    c = Case(mesh)
    cd = synthetic_condensation(c.A, c.free_u)
    dS = cd["mat"].diagonal()
    dAc = np.where(c.free_u & ~cd["interior"] & (dS != 0), 1.0 / np.where(dS != 0, dS, 1.0), 0.0)
    g_ = dAc > 0
    Sg = cd["mat"][g_][:, g_].toarray()
    ev = np.linalg.eigvalsh(np.sqrt(dAc[g_])[:, None] * Sg * np.sqrt(dAc[g_])[None, :])
    print("eig range", ev[0], ev[-1], "interior", cd["interior"].sum(), "of", c.n_u)
    lam = 0.9 * float(ev[0])
    cond = dict(H=cd["H"], HT=cd["HT"], inner_solve=cd["inner_solve"], inner_matrix=cd["inner_matrix"])
    sol_ref = c.start()
    _, h_ref, info = ko.bramble_pasciak_cg_v2(cd["mat"], c.B, c.f, c.g, dAc, c.dM, sol=sol_ref, tol=1e-10, maxsteps=20000,
                                              initialize=False, lam_min=lam, BT=c.BT, cond=cond)
    parts = [nsb.SparseMatrix.from_scipy(cd[k]) for k in ("mat", "H", "HT", "inner_solve", "inner_matrix")]
    matA = nsb.CondensedOperator(*parts)
    pA, pM = nsb.DiagonalPreconditioner.from_diagonal(dAc), nsb.DiagonalPreconditioner.from_diagonal(c.dM)
    B = nsb.SparseMatrix.from_scipy(c.B)
    sol = c.start()
    with contextlib.redirect_stdout(io.StringIO()):
        ret = nsb.BramblePasciakCG(matA, B, None, c.f, c.g, pA, pM, sol, tol=1e-10, maxsteps=20000, initialize=False, lam_min=lam)
    h = nsb.solvers.last_info["history"]; h_ref = np.array(h_ref)
    m = min(len(h), len(h_ref))
    print(mesh, ret, info["it"], len(h), len(h_ref))
    print("rel diff at", [(i, float(abs(h[i] - h_ref[i]) / h_ref[i])) for i in range(0, m, 15)])
    print("sol rel", np.linalg.norm(sol - sol_ref) / np.linalg.norm(sol_ref), "err0", h[0], h_ref[0], "tail", h[-3:], h_ref[-3:])
