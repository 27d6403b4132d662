"""Device-resident SIMPLE / projection time-step chain: the host-side mirror of ``NavierStokes.DoTimeStep`` and
``NavierStokes.Project`` (templates/NavierStokesSIMPLE_iterative.py:427-443) and of the operator compositions
they use (``invmstar``, ``invproj``: :93-98, :128-131).  A time step is ONE call into libnsb200 (nsb_simple_step:
SpMVs with the vector statements fused into their row epilogues, two device CG solves whose captured batches are kept
across steps, one gather/update kernel); the state vector stays on the device across time steps -- PyTorch only owns
the storage.  The matrix-free convection operator (:106-113) stays in NGSolve on the host: it enters through the
optional ``convection`` callback of DoTimeStep.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .la import BaseVector, SparseMatrix, _is_torch, buffer_of, default_context
from .solvers import CGSolver, _pre_handle


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
        if _is_torch(ind):
            self.ind = ind.to(device=device, dtype=torch.int32).contiguous()
        else:
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


def _inverse_parts(inv, what):
    """(CGSolver, H, HT, inner_solve) of an ``invmstar`` / ``invproj`` argument."""
    if isinstance(inv, ExtendedInverse):
        cg, parts = inv.inv1, (inv.H, inv.HT, inv.inner_solve)
    else:
        cg, parts = inv, (None, None, None)
    if not isinstance(cg, CGSolver) or not isinstance(cg.mat, SparseMatrix):
        raise TypeError("%s must be a device CGSolver, or an ExtendedInverse around one (got %r); there is no CPU fallback"
                        % (what, type(inv)))
    if any(p is None for p in parts) and not all(p is None for p in parts):
        raise TypeError("%s: give harmonic_extension, harmonic_extension_trans and inner_solve together" % what)
    return cg, parts


class SimpleStepper:
    """State and operators of the explicit-convection / implicit-Stokes splitting step; ``DoTimeStep`` is one device call.

    a         device SparseMatrix, ``self.a.mat`` (:79-80)
    invmstar  ``self.invmstar`` (:93-98): ExtendedInverse(CGSolver(mstar, pre, precision=1e-4), H, HT, inner_solve) or the CGSolver alone
    bproj     device SparseMatrix (Xproj.ndof x V.ndof), ``self.bproj.mat`` (:133-135)
    invproj   ``self.invproj`` (:128-131), same shapes as invmstar
    mapV      PermutationMatrix V.ndof <- Xproj.ndof (:138-144)
    gfu       vector of length X.ndof (CUDA tensor: stays on the device; numpy: copied per step); the first ``n_vel`` entries are
              the velocity ``gfu.vec[0:V.ndof]``
    p_range   (start, stop) of the pressure inside the projection vector: ``[V2.ndof, V2.ndof + Q.ndof)`` (:441)
    Row-partitioned contexts: operators carry their halo plans, vectors hold the owned rows, mapV indexes [owned | halo of aproj].
    """

    def __init__(self, a, invmstar, bproj, invproj, mapV, gfu, n_vel, p_range, timestep, f=None, ctx=None):
        if not isinstance(a, SparseMatrix) or not isinstance(bproj, SparseMatrix):
            raise TypeError("SimpleStepper needs device SparseMatrix operators")
        self.ctx = ctx or a.ctx
        self.a, self.invmstar, self.bproj, self.invproj, self.mapV = a, invmstar, bproj, invproj, mapV
        self.gfu, self.n_vel, self.p_range, self.timestep, self.f = gfu, int(n_vel), (int(p_range[0]), int(p_range[1])), float(timestep), f
        cg_m, (mH, mHT, mI) = _inverse_parts(invmstar, "invmstar")
        cg_p, (pH, pHT, pI) = _inverse_parts(invproj, "invproj")
        n_x = a._halo.n_owned if a._halo is not None else a.height
        n_proj = cg_p.mat._halo.n_owned if cg_p.mat._halo is not None else cg_p.mat.height
        pm_h, self._keep_pm = _pre_handle(cg_m.pre, n_x)
        pp_h, self._keep_pp = _pre_handle(cg_p.pre, n_proj)
        self._cg = (cg_m, cg_p)
        h = C.c_void_p()

        def hm(m):
            return m.h if m is not None else None
        _lib.check(self.ctx._lib.nsb_simple_create(self.ctx.h, a.h, cg_m.mat.h, pm_h, hm(mH), hm(mHT), hm(mI), bproj.h, cg_p.mat.h, pp_h,
                                                   hm(pH), hm(pHT), hm(pI), mapV.ind.data_ptr(), self.n_vel, self.p_range[0], self.p_range[1],
                                                   self.timestep, float(cg_m.precision), int(cg_m.maxsteps), float(cg_p.precision),
                                                   int(cg_p.maxsteps), C.byref(h)), "nsb_simple_create")
        self.h = h
        npq = self.p_range[1] - self.p_range[0]
        self.gfup = gfu.new_zeros(npq) if _is_torch(gfu) else np.zeros(npq)
        self.cg_iterations, self.seconds = (0, 0), 0.0

    def Project(self, vel):
        """:440-443.  ``vel`` (length V.ndof) is updated in place; the pressure lands in ``self.gfup``."""
        pv, n, _ = buffer_of(vel)
        if n != self.n_vel:
            raise ValueError("Project: velocity has length %d, expected %d" % (n, self.n_vel))
        pp, _, _ = buffer_of(self.gfup)
        its = C.c_int()
        _lib.check(self.ctx._lib.nsb_simple_project(self.ctx.h, self.h, pv, pp, C.byref(its)), "nsb_simple_project")
        self.cg_iterations = (0, its.value)
        return vel

    def DoTimeStep(self, convection=None):
        """:427-438.  ``convection(gfu) -> vector`` supplies ``conv_operator * gfu.vec`` (host NGSolve, matrix-free)."""
        pc = None
        if convection is not None:
            conv = _t(convection(self.gfu))
            pc, _, keep_c = buffer_of(conv)
        pf = None
        if self.f is not None:
            pf, _, _ = buffer_of(self.f)
        pg, _, _ = buffer_of(self.gfu)
        pp, _, _ = buffer_of(self.gfup)
        its = (C.c_int * 2)()
        secs = C.c_double()
        _lib.check(self.ctx._lib.nsb_simple_step(self.ctx.h, self.h, pg, pf, pc, pp, its, C.byref(secs)), "nsb_simple_step")
        self.cg_iterations, self.seconds = (its[0], its[1]), secs.value
        self._cg[0].iterations, self._cg[1].iterations = its[0], its[1]
        return self.gfu

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_simple_destroy(self.h)
        except Exception:
            pass
