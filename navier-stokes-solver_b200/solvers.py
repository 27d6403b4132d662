"""Device solvers behind the reference's own call signatures (names, argument order, defaults, return
shapes, in-place ``sol``).  Switching a reference script to the B200 path is a change
of import line plus building the operators with ``la.SparseMatrix`` instead of ``bf.mat``:

    from navier_stokes_solver_b200.solvers import MinRes                 # minres.py:12
    from navier_stokes_solver_b200.solvers import bramble_pasciak_cg     # bramble_pasciak_cg.py:65
    from navier_stokes_solver_b200.solvers import BramblePasciakCG       # solvers/bramblepasciak_new.py:24
    from navier_stokes_solver_b200.solvers import orthonormalize         # orthonormalization.py:5

Every function calls the C ABI (include/nsb200.h) and nothing else; unsupported operator types raise.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib
from .la import (AuxiliarySpacePreconditioner, BaseVector, BlockMatrix, BlockSmoother, BlockVector, CondensedOperator, DiagonalPreconditioner, SparseMatrix,
                 _is_torch,
                 buffer_of, default_context)

last_info = {}   # diagnostics of the most recent solve (status, seconds, iterations)


def _flat(v, n, what):
    """Caller-owned flat storage of a vector argument (no copy)."""
    p, ln, keep = buffer_of(v)
    if ln != n:
        raise ValueError("%s has length %d, operator needs %d" % (what, ln, n))
    return p, keep


def _pre_handle(pre, n):
    if pre is None:
        return None, None
    if isinstance(pre, BlockMatrix):
        pre = pre.blockdiag_preconditioner()
    if not isinstance(pre, (DiagonalPreconditioner, BlockSmoother, AuxiliarySpacePreconditioner)):
        raise TypeError("preconditioner %r cannot run on the device: only diagonal ('local'/Jacobi) preconditioners, "
                        "2x2 block-diagonals of them, block smoothers and auxiliary-space composites are available; there is no CPU fallback" % (type(pre),))
    if pre.n != n:
        raise ValueError("preconditioner size %d != %d" % (pre.n, n))
    return pre.h, pre


def _new_like(rhs, n):
    if isinstance(rhs, BaseVector):
        return rhs.CreateVector()
    if _is_torch(rhs):
        return rhs.new_zeros(n)
    return np.zeros(n)


def _has_entries(v):
    """``sol if sol else ...`` of the reference (minres.py:50) is a length test on NGSolve vectors."""
    if v is None:
        return False
    _, ln, _ = buffer_of(v)
    return ln > 0


# ------------------------------------------------------------------------------------------------
def MinRes(mat, rhs, pre=None, sol=None, maxsteps=100, printrates=True, initialize=True, tol=1e-7):
    """Preconditioned MINRES on the device; signature and semantics of minres.py:12-149.

    mat : device SparseMatrix (square) or la.BlockMatrix [[A, B.T], [B, C|None]] (run.py:45)
    pre : None, la.DiagonalPreconditioner, or la.BlockMatrix [[preA, None], [None, preS]] (run.py:46)
    Returns (u, errors) with errors[0] == 1.0 and one relative entry per iteration (minres.py:95,125,149);
    ``sol`` is overwritten in place and returned as ``u`` when given (minres.py:50).
    """
    K = mat.saddle() if isinstance(mat, BlockMatrix) else mat
    if not isinstance(K, SparseMatrix):
        raise TypeError("MinRes needs a device operator (la.SparseMatrix / la.BlockMatrix), got %r" % (type(mat),))
    ctx = K.ctx
    n = K._halo.n_owned if K._halo is not None else K.height
    pre_h, keep_pre = _pre_handle(pre, n)
    u = sol if _has_entries(sol) else _new_like(rhs, n)                  # minres.py:50
    p_rhs, k1 = _flat(rhs, n, "rhs")
    p_sol, k2 = _flat(u, n, "sol")
    hist = np.zeros(int(maxsteps) + 2)
    its, secs, status = C.c_int(), C.c_double(), C.c_int()
    _lib.check(ctx._lib.nsb_minres(ctx.h, K.h, pre_h, p_rhs, p_sol, int(maxsteps), float(tol), int(bool(initialize)),
                                   hist.ctypes.data, C.byref(its), C.byref(secs), C.byref(status)), "nsb_minres")
    errors = hist[:its.value + 1].tolist()
    last_info.clear()
    last_info.update(solver="MinRes", iterations=its.value, seconds=secs.value, status=_lib.STATUS_NAMES[status.value])
    if status.value == _lib.BREAKDOWN:
        raise ZeroDivisionError("MinRes breakdown: gamma = sqrt(<C r, r>) is zero or not finite (minres.py:73)")
    if printrates:
        # same lines the reference prints (minres.py:81,124), emitted after the device loop has run
        err0 = 1.0
        print("\rit = ", 0, " err = ", "(relative)", errors[0], " " * 20, end="")
        for k in range(1, len(errors)):
            print("\rit =", k, "rel err =", errors[k] / err0, " " * 20, end="")
    if status.value in (_lib.MAXSTEPS, _lib.ABSTOL):
        print("\nWarning: MinRes did not converge to TOL")               # minres.py:145-146
    print("")                                                            # minres.py:147
    return (u, errors)


# ------------------------------------------------------------------------------------------------
def EigenValues_Preconditioner(mat, pre, tol=1e-10, maxsteps=2000, start=None):
    """[lambda_min, lambda_max] of pre*mat by device CG-Lanczos; stands where NGSolve's
    EigenValues_Preconditioner stands (bramble_pasciak_cg.py:70, solvers/bramblepasciak_new.py:115).
    The reference only uses min() and max() of the returned list."""
    if not isinstance(mat, (SparseMatrix, CondensedOperator)):
        raise TypeError("EigenValues_Preconditioner needs a device SparseMatrix or CondensedOperator")
    ctx = mat.ctx
    n = mat._halo.n_owned if getattr(mat, "_halo", None) is not None else mat.height
    pre_h, keep = _pre_handle(pre, n)
    lmin, lmax, steps = C.c_double(), C.c_double(), C.c_int()
    p_start = None
    if start is not None:
        p_start, keep2 = _flat(start, n, "start")
    fn = ctx._lib.nsb_lanczos_eigs_condensed if isinstance(mat, CondensedOperator) else ctx._lib.nsb_lanczos_eigs
    _lib.check(fn(ctx.h, mat.h, pre_h, p_start, int(maxsteps), float(tol), C.byref(lmin), C.byref(lmax), C.byref(steps)),
               "nsb_lanczos_eigs")
    last_info.clear()
    last_info.update(solver="lanczos", iterations=steps.value)
    return [lmin.value, lmax.value]


def _bpcg_engine(A, B, preA, preM, f, g, sol_flat, n_u, n_p, k, tol, maxsteps, initialize, rel_err, mode):
    if not (isinstance(A, (SparseMatrix, CondensedOperator)) and isinstance(B, SparseMatrix)):
        raise TypeError("Bramble-Pasciak CG needs device SparseMatrix blocks (got %r, %r)" % (type(A), type(B)))
    ctx = A.ctx
    BT = B.CreateTranspose()                                             # bramblepasciak_new.py:198
    preA_h, k1 = _pre_handle(preA, n_u)
    preM_h, k2 = _pre_handle(preM, n_p)
    if preA_h is None:
        raise TypeError("Bramble-Pasciak CG needs a preconditioner for A")
    p_f, k3 = _flat(f, n_u, "f")
    p_g, k4 = _flat(g, n_p, "g")
    p_sol, k5 = _flat(sol_flat, n_u + n_p, "sol")
    hist = np.zeros(int(maxsteps) + 2)
    upd, secs, status = C.c_int(), C.c_double(), C.c_int()
    if isinstance(A, CondensedOperator):
        if mode != 2:
            raise NotImplementedError("static condensation exists only in BramblePasciakCG (bramblepasciak_new.py)")
        _lib.check(ctx._lib.nsb_bpcg_condensed(ctx.h, A.h, B.h, BT.h, preA_h, preM_h, float(k), p_f, p_g, p_sol, int(maxsteps),
                                               float(tol), int(bool(initialize)), int(bool(rel_err)), hist.ctypes.data,
                                               C.byref(upd), C.byref(secs), C.byref(status)), "nsb_bpcg_condensed")
    else:
        _lib.check(ctx._lib.nsb_bpcg(ctx.h, A.h, B.h, BT.h, preA_h, preM_h, float(k), p_f, p_g, p_sol, int(maxsteps), float(tol),
                                     int(bool(initialize)), int(bool(rel_err)), int(mode), hist.ctypes.data, C.byref(upd),
                                     C.byref(secs), C.byref(status)), "nsb_bpcg")
    last_info.clear()
    last_info.update(solver="bpcg", updates=upd.value, seconds=secs.value, status=_lib.STATUS_NAMES[status.value], k=k,
                     history=hist[:upd.value + 1].copy(), err0=float(hist[0]))
    return hist[:upd.value + 1], upd.value, secs.value, status.value


def _require_device_blocks(A, B):
    """No CPU fallback: anything that is not a device operator is refused before any of its attributes is used."""
    if not (isinstance(A, (SparseMatrix, CondensedOperator)) and isinstance(B, SparseMatrix)):
        raise TypeError("Bramble-Pasciak CG needs device SparseMatrix blocks (got %r, %r)" % (type(A), type(B)))


def _owned(m):
    return m._halo.n_owned if getattr(m, "_halo", None) is not None else None


def bramble_pasciak_cg(a_matrix, b_matrix, c_matrix, pre_a, pre_schur_complement,
                       upper_rhs, lower_rhs, solution=None,
                       tolerance=1e-12, max_steps=1000, print_rates=True, lam_min=None):
    """Bramble-Pasciak CG, signature of bramble_pasciak_cg.py:65-67; returns (solution, errors) (:148).

    ``c_matrix`` must be None (it is at every reference call site: run.py:36, stokes_hcurldiv.py:76).
    ``lam_min`` (extra keyword) injects min eig(pre_a * A); by default it is computed by device Lanczos
    in place of NGSolve's EigenValues_Preconditioner (:70)."""
    if c_matrix is not None:
        raise NotImplementedError("c_matrix is None at every reference call site; a (2,2) block is not supported")
    _require_device_blocks(a_matrix, b_matrix)
    n_u, n_p = a_matrix.height, b_matrix.height
    if lam_min is None:
        lams = EigenValues_Preconditioner(a_matrix, pre_a)                # :70
        lam_min, lam_max = min(lams), max(lams)
    else:
        lam_max = float("nan")
    k = 1 / lam_min + 1e-3                                               # :71
    print("scale factor: ", k)                                           # :73-74
    print("condition number: ", lam_max / lam_min)
    if not _has_entries(solution):                                       # :88-90
        solution = BlockVector([BaseVector(n_u), BaseVector(n_p)])
    hist, upd, secs, status = _bpcg_engine(a_matrix, b_matrix, pre_a, pre_schur_complement, upper_rhs, lower_rhs,
                                           solution, n_u, n_p, k, tolerance, max_steps, False, True, 1)
    err0 = hist[0] if hist[0] != 0 else 1.0
    nerr = upd + 1 if status == _lib.CONVERGED else upd                  # :116-121: the breaking pass is recorded
    errors = (hist[:nerr] / err0).tolist()
    if print_rates:
        for it, e in enumerate(errors):
            print("\rit =", it, "rel err =", e, "abs err =", e * err0, " " * 20, end="")   # :117
    if status != _lib.CONVERGED:
        print("\nWarning: CG did not converge to TOL")                   # :144-145
    print("")
    return (solution, errors)


def BramblePasciakCG(blfA, blfB, matC, f, g, preA_unscaled, preM, sol=None, tol=1e-6, maxsteps=100, printrates=True,
                     initialize=True, rel_err=True, lam_min=None):
    """Signature of solvers/bramblepasciak_new.py:24-25; returns (it, seconds) (:253).  ``blfA`` / ``blfB`` are
    objects with ``.mat`` (device SparseMatrix) and ``.condense`` like NGSolve BilinearForms, or the
    matrices themselves.  ``matC`` is accepted and unused, as in the reference.  The residual history
    the reference prints (:245) is also kept in ``solvers.last_info['history']``."""
    matB = getattr(blfB, "mat", blfB)
    if getattr(blfA, "condense", False):
        # matA = (I - HT)(A + inner_matrix)(I - H), preconditioner through harmonic_extension (:8-21, :84-106)
        matA = blfA if isinstance(blfA, CondensedOperator) else CondensedOperator.from_bilinear_form(blfA)
    else:
        matA = getattr(blfA, "mat", blfA)
    _require_device_blocks(matA, matB)
    n_u, n_p = matA.height, matB.height
    if lam_min is None:
        lams = EigenValues_Preconditioner(matA, preA_unscaled, tol=1e-3)  # :115
        lam_min, lam_max = min(lams), max(lams)
    else:
        lam_max = float("nan")
    k = 1. / lam_min + 1e-3                                              # :118
    print("condition", lam_max / lam_min)                                # :120
    u = sol if _has_entries(sol) else BlockVector([BaseVector(n_u), BaseVector(n_p)])   # :137
    hist, upd, secs, status = _bpcg_engine(matA, matB, preA_unscaled, preM, f, g, u, n_u, n_p, k, tol, maxsteps,
                                           initialize, rel_err, 2)
    print("err0", hist[0])                                               # :188
    if status == _lib.ZERO_RESIDUAL:
        return u                                                         # :191-192
    if printrates:
        for it in range(upd):
            print("it = ", it, " err = ", hist[it], " " * 20)            # :245
    if status != _lib.CONVERGED:
        print("Warning: BPCG did not converge to TOL")                   # :248-249
    print("\n")
    return (upd - 1, secs)                                               # :253


# ------------------------------------------------------------------------------------------------
def CGSolve(mat, rhs, pre=None, sol=None, tol=1e-12, maxsteps=1000, initialize=True, printrates=False):
    """Preconditioned CG with NGSolve CGSolver's stopping rule (templates/NavierStokesSIMPLE_iterative.py:93,128),
    also the Jacobi-CG replacement of the direct inverse of heat.py:72.  Returns (u, errors_abs)."""
    if not isinstance(mat, SparseMatrix):
        raise TypeError("CGSolve needs a device SparseMatrix")
    ctx = mat.ctx
    n = mat._halo.n_owned if mat._halo is not None else mat.height
    pre_h, keep = _pre_handle(pre, n)
    u = sol if _has_entries(sol) else _new_like(rhs, n)
    p_rhs, k1 = _flat(rhs, n, "rhs")
    p_sol, k2 = _flat(u, n, "sol")
    hist = np.zeros(int(maxsteps) + 2)
    its, secs, status = C.c_int(), C.c_double(), C.c_int()
    _lib.check(ctx._lib.nsb_cg(ctx.h, mat.h, pre_h, p_rhs, p_sol, int(maxsteps), float(tol), int(bool(initialize)),
                               hist.ctypes.data, C.byref(its), C.byref(secs), C.byref(status)), "nsb_cg")
    last_info.clear()
    last_info.update(solver="cg", iterations=its.value, seconds=secs.value, status=_lib.STATUS_NAMES[status.value])
    errors = hist[:its.value + 1].tolist()
    if printrates:
        for k, e in enumerate(errors):
            print("it = ", k, " err = ", e)
    return (u, errors)


class CGSolver:
    """Operator form ``inv = CGSolver(mat, pre, precision=...)``; ``inv * rhs`` solves
    (templates/NavierStokesSIMPLE_iterative.py:93,128)."""

    def __init__(self, mat, pre=None, precision=1e-8, maxsteps=200, printrates=False):
        self.mat, self.pre, self.precision, self.maxsteps, self.printrates = mat, pre, precision, maxsteps, printrates
        self.iterations = 0

    @property
    def height(self):
        return self.mat.height

    @property
    def width(self):
        return self.mat.width

    def __mul__(self, rhs):
        u, errs = CGSolve(self.mat, rhs, self.pre, None, self.precision, self.maxsteps, True, self.printrates)
        self.iterations = len(errs) - 1
        return u if isinstance(u, BaseVector) else BaseVector(u)


# ------------------------------------------------------------------------------------------------
def orthonormalize(basis, tries=3):
    """Modified Gram-Schmidt with ``tries`` sweeps on the device, in place; orthonormalization.py:5-16.
    ``basis`` is a list of equally long vectors (returned, mutated) or one (m, n) array / CUDA tensor."""
    ctx = default_context()
    if isinstance(basis, (list, tuple)):
        m = len(basis)
        if m == 0:
            return basis
        bufs = [buffer_of(b) for b in basis]
        n = bufs[0][1]
        if any(b[1] != n for b in bufs):
            raise ValueError("basis vectors differ in length")
        if _is_torch(bufs[0][2]):
            import torch
            Q = torch.stack([b[2] for b in bufs]).contiguous()
        else:
            Q = np.ascontiguousarray(np.stack([b[2] for b in bufs]))
        p, _, keep = buffer_of(Q.reshape(-1))
        _lib.check(ctx._lib.nsb_orthonormalize(ctx.h, p, n, m, int(tries)), "nsb_orthonormalize")
        for i, b in enumerate(bufs):
            b[2][...] = Q[i]
        return basis
    m, n = basis.shape
    p, _, keep = buffer_of(basis.reshape(-1))
    _lib.check(ctx._lib.nsb_orthonormalize(ctx.h, p, n, m, int(tries)), "nsb_orthonormalize")
    return basis


# ------------------------------------------------------------------------------------------------
def heat_large_step(diffusion, mass, heat, pre_heat, temperature, time_step, weights, subspace_dimension=5,
                    cg_tol=1e-12, cg_maxsteps=2000):
    """One pass of the time-loop body of heat.py:81-146 on the device: Krylov-like basis from m-1 implicit Euler
    sub-steps (:95-98), orthonormalize (:100), projected m x m matrices (:109-118), Gauss collocation step of the
    projected system (:135) and the lift (:140-142).  ``diffusion`` / ``mass`` / ``heat`` are device SparseMatrix
    (diffusion.mat, mass.mat, heat.mat = M + time_step K, heat.py:51-52); the direct ``heat.mat.Inverse(freedofs)`` of heat.py:72 is
    replaced by Jacobi-CG with ``pre_heat`` (zero on constrained dofs) to ``cg_tol``.  ``temperature`` is updated
    in place and returned; diagnostics in ``last_info``."""
    for m_ in (diffusion, mass, heat):
        if not isinstance(m_, SparseMatrix):
            raise TypeError("heat_large_step needs device SparseMatrix operators")
    ctx = diffusion.ctx
    n = diffusion._halo.n_owned if diffusion._halo is not None else diffusion.height
    pre_h, keep = _pre_handle(pre_heat, n)
    p_u, keep_u = _flat(temperature, n, "temperature")
    m = int(subspace_dimension)
    a = np.ascontiguousarray(weights.a, dtype=np.float64)
    b = np.ascontiguousarray(weights.b, dtype=np.float64)
    Kh, Mh = np.zeros((m, m)), np.zeros((m, m))
    its = C.c_int()
    _lib.check(ctx._lib.nsb_heat_step(ctx.h, diffusion.h, mass.h, heat.h, pre_h, p_u, float(time_step), m, a.shape[0],
                                      a.ctypes.data, b.ctypes.data, float(cg_tol), int(cg_maxsteps), C.byref(its),
                                      Kh.ctypes.data, Mh.ctypes.data), "nsb_heat_step")
    last_info.clear()
    last_info.update(solver="heat_large_step", cg_iterations=its.value, subspace_diffusion=Kh, subspace_mass=Mh)
    return temperature


# ------------------------------------------------------------------------------------------------
# Gauss collocation on the projected (m x m, m = 5) system: host-side dense algebra ("negligible; stays on host").
class ImplicitRungeKuttaMethodWeights:
    """Butcher tableau of the ``deg``-stage Gauss collocation method on [0, 1] (the object runge_kutta_method.py:10-23 builds):
    nodes ``c`` = Gauss-Legendre points, ``a[i, j]`` = integral of the j-th Lagrange basis polynomial over [0, c_i], ``b[j]`` the
    same over [0, 1].  The integrals are exact: a ``deg``-point Gauss rule mapped onto [0, c_i] integrates the degree ``deg - 1``
    basis polynomials exactly (no adaptive quadrature)."""

    def __init__(self, deg=3):
        from numpy.polynomial.legendre import leggauss
        nodes, gauss_w = leggauss(deg)
        self.c = 0.5 * (nodes + 1.0)
        self.b_numpy = gauss_w / deg
        # basis[j](x) in product form (stable for deg = 10, where the power basis loses five digits), for an array of points x
        denom = np.array([np.prod(self.c[j] - np.delete(self.c, j)) for j in range(deg)])

        def basis(x):                                    # -> (deg, len(x))
            diff = x[None, :] - self.c[:, None]          # (node, point)
            return np.stack([np.prod(np.delete(diff, j, axis=0), axis=0) / denom[j] for j in range(deg)])

        def integrate_to(upper):                         # every basis polynomial over [0, upper]: deg Gauss points are exact (degree deg-1)
            x = 0.5 * upper * (nodes + 1.0)
            return 0.5 * upper * (basis(x) @ gauss_w)
        self.a = np.stack([integrate_to(ci) for ci in self.c])      # a[i, j] = int_0^{c_i} basis_j
        self.b = integrate_to(1.0)


def linear_implicit_runge_kutta_step(weights, matrix, current_value, step_width):
    """runge_kutta_method.py:26-59 with the quadruple Python loop written as a Kronecker product."""
    matrix = np.asarray(matrix, dtype=float)
    y = np.asarray(current_value, dtype=float)
    m, s = matrix.shape[0], weights.a.shape[0]
    coef = np.eye(m * s) - step_width * np.kron(weights.a, matrix)
    k = np.linalg.solve(coef, np.tile(matrix @ y, s))
    return y + step_width * (weights.b[None, :] @ k.reshape(s, m)).ravel()
