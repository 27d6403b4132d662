"""Device-resident SIMPLE / projection time-step chain: the host-side mirror of ``NavierStokes.DoTimeStep`` and
``NavierStokes.Project`` (templates/NavierStokesSIMPLE_iterative.py:427-443) and of the operator compositions
they use (``invmstar``, ``invproj``: :93-98, :128-131).  Every vector is a CUDA tensor that stays on the device
across time steps; every operation is a call into libnsb200 (SpMV, Jacobi-CG, axpby, gather) -- PyTorch only owns
the storage.  The matrix-free convection operator (:106-113) stays in NGSolve on the host: it enters through the
optional ``convection`` callback of DoTimeStep.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .la import BaseVector, SparseMatrix, buffer_of, default_context
from .solvers import CGSolver


def _t(v):
    return v._a if isinstance(v, BaseVector) else v


def axpby(y, a, x, b=1.0, ctx=None):
    """y = a*x + b*y on the device (in place)."""
    ctx = ctx or default_context()
    py, n, _ = buffer_of(y)
    px, nx, _ = buffer_of(x)
    if n != nx:
        raise ValueError("axpby: lengths differ")
    _lib.check(ctx._lib.nsb_vec_axpby(ctx.h, py, float(a), px, float(b), n), "nsb_vec_axpby")
    return y


class PermutationMatrix:
    """``PermutationMatrix(width, ind)``: (P x)[i] = x[ind[i]]  (:138-144)."""

    def __init__(self, width, ind, device="cuda"):
        import torch
        self.width = int(width)
        self.ind = torch.as_tensor(np.asarray(ind, dtype=np.int32), device=device).contiguous()
        self.height = int(self.ind.numel())

    def sub_from(self, y, x, ctx=None):
        """y -= P x   (vel.data -= mapV * tmp_projection, :442)."""
        ctx = ctx or default_context()
        py, n, _ = buffer_of(y)
        px, nx, _ = buffer_of(x)
        if n != self.height or nx != self.width:
            raise ValueError("PermutationMatrix: size mismatch")
        _lib.check(ctx._lib.nsb_vec_gather_axpy(ctx.h, py, -1.0, px, self.ind.data_ptr(), n), "nsb_vec_gather_axpy")
        return y


class ExtendedInverse:
    """``ext @ inv1 @ extT + inner_solve`` with ext = I + harmonic_extension, extT = I + harmonic_extension_trans
    (:96-98, :129-131): the inverse of a statically condensed form.  ``inv1`` is anything with ``inv1 * vec`` (a device
    CGSolver); with ``H is None`` (uncondensed form) it is just ``inv1``."""

    def __init__(self, inv1, harmonic_extension=None, harmonic_extension_trans=None, inner_solve=None):
        self.inv1, self.H, self.HT, self.inner_solve = inv1, harmonic_extension, harmonic_extension_trans, inner_solve
        for m in (self.H, self.HT, self.inner_solve):
            if m is not None and not isinstance(m, SparseMatrix):
                raise TypeError("ExtendedInverse needs device SparseMatrix parts")

    def __mul__(self, rhs):
        rhs = _t(rhs)
        if self.H is None:
            return self.inv1 * rhs
        r = _t(self.HT * rhs)
        axpby(r, 1.0, rhs)                       # r = (I + HT) rhs
        y = _t(self.inv1 * r)
        hy = _t(self.H * y)
        axpby(y, 1.0, hy)                        # y = (I + H) y
        axpby(y, 1.0, _t(self.inner_solve * rhs))
        return BaseVector(y)


class SimpleStepper:
    """State and operators of the explicit-convection / implicit-Stokes splitting step.

    a         device SparseMatrix, ``self.a.mat`` (:79-80)
    invmstar  operator with ``* vec``: ``self.invmstar`` (:93-98) -- ExtendedInverse(CGSolver(mstar, pre, precision=1e-4), ...)
    bproj     device SparseMatrix (Xproj.ndof x V.ndof), ``self.bproj.mat`` (:133-135)
    invproj   operator with ``* vec``: ``self.invproj`` (:128-131)
    mapV      PermutationMatrix V.ndof <- Xproj.ndof (:138-144)
    gfu       CUDA tensor of length X.ndof, the first ``n_vel`` entries are the velocity ``gfu.vec[0:V.ndof]``
    p_range   (start, stop) of the pressure inside the projection vector: ``[V2.ndof, V2.ndof + Q.ndof)`` (:441)
    """

    def __init__(self, a, invmstar, bproj, invproj, mapV, gfu, n_vel, p_range, timestep, f=None):
        self.a, self.invmstar, self.bproj, self.invproj, self.mapV = a, invmstar, bproj, invproj, mapV
        self.gfu, self.n_vel, self.p_range, self.timestep, self.f = gfu, int(n_vel), p_range, float(timestep), f
        self.gfup = None

    def Project(self, vel):
        """:440-443.  ``vel`` is a view of length V.ndof, updated in place; the pressure lands in ``self.gfup``."""
        tmp_projection = _t(self.invproj * _t(self.bproj * vel))
        self.gfup = tmp_projection[self.p_range[0]:self.p_range[1]].clone()
        self.mapV.sub_from(vel, tmp_projection)
        return vel

    def DoTimeStep(self, convection=None):
        """:427-438.  ``convection(gfu) -> tensor`` supplies ``conv_operator * gfu.vec`` (host NGSolve, matrix-free)."""
        temp = _t(self.a * self.gfu)
        axpby(temp, 0.0, temp, -1.0)                     # temp = -a.mat * gfu.vec
        if convection is not None:
            axpby(temp, 1.0, _t(convection(self.gfu)))       # += conv_operator * gfu.vec
        if self.f is not None:
            axpby(temp, 1.0, self.f)                     # += f.vec
        temp2 = _t(self.invmstar * temp)                 # :436
        self.Project(temp2[0:self.n_vel])                # :437
        axpby(self.gfu, self.timestep, temp2)            # :438
        return self.gfu
