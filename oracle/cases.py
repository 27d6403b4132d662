"""Seeded parity cases shared by oracle/make_golden.py and tests/ (test infrastructure).

Each case is a small assembled saddle-point system from the synthetic Taylor-Hood assembler plus the
solver arguments the reference call sites use (run.py:52, run.py:36, stokes_hcurldiv.py:76,
templates/NavierStokesSIMPLE_iterative.py:397) with Jacobi preconditioners on both sides
(BASELINE.md 4)."""
from __future__ import annotations

import os
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

import navier_stokes_solver_b200.synthetic as syn  # noqa: E402

GOLDEN_DIR = os.path.join(_ROOT, "tests", "golden")

# name -> (cells, lengths)
MESHES = {
    "sq8": ((8, 8), (1.0, 1.0)),                 # config 1 shape (unit square Taylor-Hood), coarse
    "sq20": ((20, 20), (1.0, 1.0)),              # config 1: h = 0.05
    "chan2d": ((24, 6), (2.0, 0.41)),            # run.py:22-25 channel without the cylinder
    "chan3d": ((6, 2, 2), (2.5, 0.41, 0.41)),    # NavierStokesSIMPLE_test_3D.py:9-13 box, coarse
}


class Case:
    def __init__(self, mesh):
        cells, lengths = MESHES[mesh]
        S = syn.assemble_stokes(cells, lengths)
        self.mesh = mesh
        self.S = S
        self.A = S.A.to_scipy()
        self.B = S.B.to_scipy()
        self.BT = S.BT.to_scipy()
        self.Mp = S.Mp.to_scipy()
        self.f = S.f.numpy().copy()
        self.g = S.g.numpy().copy()
        self.u0 = S.u0.numpy().copy()
        self.free_u = S.free_u.numpy().copy()
        self.n_u, self.n_p = S.n_u, S.n_p
        self.dA = syn.jacobi_inverse(S.A, free=S.free_u).numpy()     # 'local' on A, free dofs only
        self.dM = syn.jacobi_inverse(S.Mp).numpy()                   # 'local' on the pressure mass
        self._lam = None

    @property
    def N(self):
        return self.n_u + self.n_p

    def start(self):
        """sol carrying the inlet Dirichlet data (run.py:102-103), pressure zero."""
        return np.concatenate([self.u0, np.zeros(self.n_p)])

    def rhs(self):
        return np.concatenate([self.f, self.g])

    def lam_min(self):
        """Smallest eigenvalue of dA*A on the free subspace (dense; cases are small)."""
        if self._lam is None:
            fr = self.free_u
            Aff = self.A[fr][:, fr].toarray()
            d = self.dA[fr]
            sym = np.sqrt(d)[:, None] * Aff * np.sqrt(d)[None, :]
            self._lam = float(np.linalg.eigvalsh(sym)[0])
        return self._lam

    def fingerprint(self):
        return np.array([self.n_u, self.n_p, self.A.nnz, self.B.nnz, np.abs(self.A.data).sum(),
                         np.abs(self.B.data).sum(), np.abs(self.f).sum(), np.abs(self.u0).sum()])


def load_golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"))


def synthetic_condensation(A, free):
    """Static condensation of an independent set I of free dofs (A_II diagonal), laid out like a condensed NGSolve
    BilinearForm (solvers/bramblepasciak_new.py:8-21,84-103): returns dict(mat=S embedded, H, HT, inner_solve,
    inner_matrix, interior mask) with (I - HT)(S + inner_matrix)(I - H) == A."""
    import scipy.sparse as sp
    A = A.tocsr()
    n = A.shape[0]
    interior = np.zeros(n, dtype=bool)
    blocked = ~np.asarray(free, dtype=bool)
    indptr, indices = A.indptr, A.indices
    for i in range(n):
        if blocked[i]:
            continue
        nb = indices[indptr[i]:indptr[i + 1]]
        if interior[nb].any():
            continue
        interior[i] = True
        blocked[nb] = True          # neighbours stay on the interface
        blocked[i] = True
    d = A.diagonal()
    PI = sp.diags(interior.astype(float))
    PG = sp.diags((~interior).astype(float))
    Dinv = sp.diags(np.where(interior, 1.0 / np.where(d != 0, d, 1.0), 0.0))
    A_IG = PI @ A @ PG
    A_GI = PG @ A @ PI
    H = (-(Dinv @ A_IG)).tocsr()
    S = (PG @ A @ PG - A_GI @ Dinv @ A_IG).tocsr()
    inner_matrix = sp.diags(np.where(interior, d, 0.0)).tocsr()
    out = dict(mat=S, H=H, HT=H.T.tocsr(), inner_solve=Dinv.tocsr(), inner_matrix=inner_matrix, interior=interior)
    for k in ("mat", "H", "HT", "inner_solve", "inner_matrix"):
        out[k].eliminate_zeros()
        out[k].sort_indices()
    return out
