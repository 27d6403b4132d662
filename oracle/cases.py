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


def heat_case(cells=(10, 10)):
    """2D heat problem of heat.py on the unit square: P2 mass / stiffness on a structured mesh, zero Dirichlet boundary,
    initial temperature = nodal values of the sine series of heat.py:13-18,150-151.  Returns K, M, free, (x, y), u0."""
    import navier_stokes_solver_b200.synthetic as syn
    S = syn.assemble_stokes(cells, (1.0, 1.0), with_mass2=True)
    K, M = S.extras["K2"].to_scipy(), S.extras["M2"].to_scipy()
    nx, ny = 2 * cells[0] + 1, 2 * cells[1] + 1
    ix, iy = np.divmod(np.arange(nx * ny), ny)
    free = (ix > 0) & (ix < nx - 1) & (iy > 0) & (iy < ny - 1)
    x, y = ix / (nx - 1), iy / (ny - 1)
    kl = [(1, 1), (2, 1), (1, 3), (3, 3), (2, 3), (4, 5), (5, 2)]
    u0 = sum(2 * np.sin(k * np.pi * x) * np.sin(l * np.pi * y) for k, l in kl) * free
    return K, M, free, (x, y), u0


def simple_case(mesh_cells=(12, 4), lengths=(2.0, 0.41), dt=1e-3, seed=5, perm="random"):
    """Synthetic operators with the shapes of NavierStokesSIMPLE_iterative.py:79-144: a, mstar = M + dt a, the mixed
    projection system [[-Ml, B^T], [B, 0]] condensed onto the pressure (diagonal mass Ml), bproj = [0; B], mapV a
    permutation between the continuous (V) and the 'broken' (V2) velocity numbering."""
    import scipy.sparse as sp
    import navier_stokes_solver_b200.synthetic as syn
    S = syn.assemble_stokes(mesh_cells, lengths, with_mass2=True)
    d = S.dim
    A = S.A.to_scipy()
    B = S.B.to_scipy()
    free = S.free_u.numpy()
    M = sp.kron(sp.identity(d), S.extras["M2"].to_scipy()).tocsr()
    mstar = (M + dt * A).tocsr()
    n_u, n_p = S.n_u, S.n_p
    if isinstance(perm, str) and perm == "rotate":      # V dof (c, node) lives at V2 dof ((c + 1) % d, node): node-local
        n2 = n_u // d
        perm = (np.arange(n_u) + n2) % n_u
    else:
        rng = np.random.default_rng(seed)
        perm = rng.permutation(n_u)                     # V dof i lives at V2 dof perm[i]
    Ml = M.diagonal()                                   # (row sums vanish at P2 vertices; the diagonal is positive)
    Bf = (B @ sp.diags(free.astype(float))).tocsr()
    P = sp.csr_matrix((np.ones(n_u), (perm, np.arange(n_u))), shape=(n_u, n_u))   # V -> V2
    B2 = (Bf @ P.T).tocsr()
    Ml2 = P @ Ml
    N = n_u + n_p
    Z = sp.csr_matrix((n_u, n_u))
    Smat = sp.bmat([[Z, None], [None, B2 @ sp.diags(1.0 / Ml2) @ B2.T]]).tocsr()
    H = sp.bmat([[Z, sp.diags(1.0 / Ml2) @ B2.T], [sp.csr_matrix((n_p, n_u)), sp.csr_matrix((n_p, n_p))]]).tocsr()
    inner = sp.bmat([[sp.diags(-1.0 / Ml2), None], [None, sp.csr_matrix((n_p, n_p))]]).tocsr()
    bproj = sp.bmat([[sp.csr_matrix((n_u, n_u))], [Bf]]).tocsr()
    dS = Smat.diagonal()
    return dict(A=A, mstar=mstar, free=free, Bf=Bf, S=Smat, H=H, HT=H.T.tocsr(), inner=inner, bproj=bproj, perm=perm,
                dS=np.where(dS > 0, 1.0 / np.where(dS > 0, dS, 1.0), 0.0), n_u=n_u, n_p=n_p, N=N, f=S.f.numpy(), u0=S.u0.numpy(), dt=dt)


def dof_blocks(c, size=5):
    """Disjoint blocks of consecutive free velocity dofs: stand where the reference has its facet blocks
    (templates/NavierStokesSIMPLE_iterative.py:360-362)."""
    free_idx = np.nonzero(c.free_u)[0]
    return [free_idx[i:i + size] for i in range(0, len(free_idx), size)]


def aggregation_transform(c, size=6):
    """Piecewise-constant transform from an auxiliary space of aggregates (consecutive free dofs) to the velocity space and the
    Galerkin coarse matrix; stands where the reference has the H(div) <- H1 `transform` (:282-290)."""
    import scipy.sparse as sp
    free_idx = np.nonzero(c.free_u)[0]
    agg = np.arange(len(free_idx)) // size
    T = sp.csr_matrix((np.ones(len(free_idx)), (free_idx, agg)), shape=(c.n_u, int(agg.max()) + 1))
    Ac = (T.T @ c.A @ T).tocsr()
    return T, Ac


def direct_stokes_solution(c):
    """Sparse direct solve of the saddle-point system of a Case on its free dofs, Dirichlet values from ``c.u0`` (the
    reference's own cross-check idea: umfpack at run.py:205, sparsecholesky at NavierStokesSIMPLE_iterative.py:403).
    Returns the (u | p) vector in the layout of ``c.start()``."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    fr = c.free_u
    nf = int(fr.sum())
    A, B = c.A.tocsr(), c.B.tocsr()
    Aff, Afc, Bf, Bc = A[fr][:, fr], A[fr][:, ~fr], B[:, fr], B[:, ~fr]
    uc = c.u0[~fr]
    K = sp.bmat([[Aff, Bf.T], [Bf, None]]).tocsc()
    x = spl.spsolve(K, np.concatenate([c.f[fr] - Afc @ uc, c.g - Bc @ uc]))
    u = c.u0.copy()
    u[fr] = x[:nf]
    return np.concatenate([u, x[nf:]])
