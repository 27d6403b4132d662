import os, sys, io, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import navier_stokes_solver_b200 as nsb
from oracle import krylov_oracle as ko
from oracle.cases import Case
c = Case("chan2d")
fi = np.nonzero(c.free_u)[0]; blocks = [fi[i:i + 4] for i in range(0, len(fi), 4)]
A, B = nsb.SparseMatrix.from_scipy(c.A), nsb.SparseMatrix.from_scipy(c.B)
for gs in (False, True):
    pre = A.CreateBlockSmoother(blocks, pattern=c.A, gauss_seidel=gs)
    ref = ko.BlockSmoother(c.A, pre.blocks)
    pre_ref = ref.symmetric_gs if gs else ref.jacobi
    pM = nsb.DiagonalPreconditioner.from_diagonal(c.dM)
    lam = 0.5 * min(nsb.EigenValues_Preconditioner(A, pre, tol=1e-9, maxsteps=2000))
    sol_ref = c.start()
    _, h_ref, info = ko.bramble_pasciak_cg_v2(c.A, c.B, c.f, c.g, pre_ref, c.dM, sol=sol_ref, tol=1e-10, maxsteps=20000, initialize=False, lam_min=lam, BT=c.BT)
    sol = c.start()
    with contextlib.redirect_stdout(io.StringIO()):
        ret = nsb.BramblePasciakCG(A, B, None, c.f, c.g, pre, pM, sol, tol=1e-10, maxsteps=20000, initialize=False, lam_min=lam)
    h = nsb.solvers.last_info["history"]; h_ref = np.array(h_ref); m = min(len(h), len(h_ref))
    rel = np.abs(h[:m] - h_ref[:m]) / h_ref[:m]
    print("gs", gs, ret, info["it"], [(i, float("%.2e" % rel[i])) for i in range(0, min(m, 80), 5)], "sol", np.linalg.norm(sol - sol_ref) / np.linalg.norm(sol_ref))
    x = np.random.default_rng(1).uniform(-1, 1, c.n_u)
    print("apply diff", np.abs((pre * x).NumPy() - pre_ref(x)).max())
