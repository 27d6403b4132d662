"""ngsolve.la stand-in (test infrastructure).  ``from ngsolve.la import *`` in
solvers/bramblepasciak_new.py:1 must yield the vector/matrix names."""
import numpy as _np

from . import (BaseVector, BlockVector, BaseMatrix, BlockMatrix, IdentityMatrix,  # noqa: F401
               SparseMatrix, DiagonalMatrix, InnerProduct, Norm, Projector, CreateVVector)

__all__ = ["BaseVector", "BlockVector", "BaseMatrix", "BlockMatrix", "IdentityMatrix", "SparseMatrix",
           "DiagonalMatrix", "InnerProduct", "Norm", "Projector", "CreateVVector",
           "EigenValues_Preconditioner"]

# Parity rule: lambda_min of pre*mat is the one number NGSolve computes with an
# unpinned Lanczos; the harness injects the value so that oracle and device use the same k.
_injected = []


def inject_eigenvalues(lams):
    """Queue the list returned by the next EigenValues_Preconditioner call."""
    _injected.append([float(x) for x in lams])


def EigenValues_Preconditioner(mat, pre, tol=1e-10):
    if _injected:
        return _injected.pop(0)
    # dense fallback for small systems: eigenvalues of pre*mat restricted to range(pre)
    n = mat.height
    if n > 4000:
        raise RuntimeError("inject eigenvalues for large systems")
    eye = _np.eye(n)
    A = _np.zeros((n, n))
    P = _np.zeros((n, n))
    for j in range(n):
        x = BaseVector(eye[:, j].copy())
        y = BaseVector(n)
        mat.MultAdd(1.0, x, y)
        A[:, j] = y._a
        y = BaseVector(n)
        pre.MultAdd(1.0, x, y)
        P[:, j] = y._a
    lam = _np.linalg.eigvals(P @ A).real
    lam = lam[_np.abs(lam) > 1e-12 * _np.abs(lam).max()]
    return sorted(float(x) for x in lam)
