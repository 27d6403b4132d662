"""Stand-in for the finite-element names heat.py pulls out of ``from ngsolve import *``.

TEST INFRASTRUCTURE ONLY (used by oracle/make_golden.py::gold_heat to run /root/reference/heat.py unmodified).
It does not discretise anything: the mass and stiffness matrices, the free-dof mask and the node coordinates of ONE
fixed nodal space are installed beforehand with ``install(K, M, free, nodes)``; the symbolic layer only has to recognise
which of the two bilinear terms a form is made of (``grad(u)*grad(v)`` -> K, ``u*v`` -> M) and with which factor.
``H1(mesh, order=..)`` ignores mesh and order -- the time loop of heat.py:80-146, which is what the golden vectors pin,
does not depend on them.  CoefficientFunctions are evaluated at the nodes only (nodal interpolation in ``Set``, a
trapezoid sum in ``Integrate``).
"""
import math as _math

import numpy as _np
import scipy.sparse as _sp
import scipy.sparse.linalg as _spla

from . import BaseMatrix, BaseVector, SparseMatrix

_problem = {}
redraw_hooks = []        # callables run by Redraw(): heat.py calls it once per time step (:146)
gridfunctions = []       # every GridFunction created, in order


def install(K, M, free, nodes):
    _problem.update(K=K.tocsr(), M=M.tocsr(), free=_np.asarray(free, dtype=bool), nodes=nodes)
    gridfunctions.clear()


# ---------------------------------------------------------------------------------------- coefficient functions
class CoefficientFunction:
    def __init__(self, f=0):
        if callable(f):
            self._f = f
        else:
            c = float(f)
            self._f = lambda X, Y: _np.full_like(X, c, dtype=float)

    def at_nodes(self):
        X, Y = _problem["nodes"]
        return _np.asarray(self._f(X, Y), dtype=float)

    @staticmethod
    def _lift(o):
        return o if isinstance(o, CoefficientFunction) else CoefficientFunction(o)

    def __add__(self, o):
        o = self._lift(o)
        return CoefficientFunction(lambda X, Y: self._f(X, Y) + o._f(X, Y))

    __radd__ = __add__

    def __sub__(self, o):
        o = self._lift(o)
        return CoefficientFunction(lambda X, Y: self._f(X, Y) - o._f(X, Y))

    def __rsub__(self, o):
        return self._lift(o) - self

    def __mul__(self, o):
        if isinstance(o, (_Term, _Symbol)):
            return NotImplemented
        o = self._lift(o)
        return CoefficientFunction(lambda X, Y: self._f(X, Y) * o._f(X, Y))

    __rmul__ = __mul__

    def __neg__(self):
        return CoefficientFunction(lambda X, Y: -self._f(X, Y))


x = CoefficientFunction(lambda X, Y: X)
y = CoefficientFunction(lambda X, Y: Y)


def _unary(np_fn, math_fn):
    def fn(a):
        if isinstance(a, CoefficientFunction):
            return CoefficientFunction(lambda X, Y: np_fn(a._f(X, Y)))
        return math_fn(a)
    return fn


sin = _unary(_np.sin, _math.sin)
cos = _unary(_np.cos, _math.cos)
exp = _unary(_np.exp, _math.exp)
sqrt = _unary(_np.sqrt, _math.sqrt)


# ---------------------------------------------------------------------------------------- symbolic forms
class _Symbol:
    """Trial / test function or its gradient."""

    def __init__(self, kind, is_grad=False):
        self.kind, self.is_grad = kind, is_grad

    def __mul__(self, o):
        if isinstance(o, _Symbol):
            if self.is_grad != o.is_grad or self.kind == o.kind:
                raise NotImplementedError("stand-in knows grad(u)*grad(v) and u*v only")
            return _Term({"K" if self.is_grad else "M": 1.0})
        if isinstance(o, (int, float)):
            return _Term({"src": float(o)})          # 0 * test_function (heat.py:54)
        return NotImplemented

    __rmul__ = __mul__


class _Term:
    def __init__(self, coef):
        self.coef = dict(coef)

    def __mul__(self, o):
        if o is dx:
            return self
        if isinstance(o, (int, float)):
            return _Term({k: v * float(o) for k, v in self.coef.items()})
        return NotImplemented

    __rmul__ = __mul__

    def __add__(self, o):
        out = dict(self.coef)
        for k, v in o.coef.items():
            out[k] = out.get(k, 0.0) + v
        return _Term(out)


class _Dx:
    pass


dx = _Dx()


def grad(s):
    return _Symbol(s.kind, is_grad=True)


# ---------------------------------------------------------------------------------------- mesh, space, forms
class Mesh:
    def __init__(self, *a, **k):
        pass


class H1:
    def __init__(self, mesh, order=1, dirichlet=None, **k):
        self.mesh, self.order, self.dirichlet = mesh, order, dirichlet
        self.ndof = _problem["K"].shape[0]

    def FreeDofs(self, *a):
        return _problem["free"]

    def TrialFunction(self):
        return _Symbol("trial")

    def TestFunction(self):
        return _Symbol("test")

    def __str__(self):
        return "stand-in nodal space, ndof = %d (order argument %s ignored)" % (self.ndof, self.order)


class _DirectInverse(BaseMatrix):
    """``mat.Inverse(freedofs)``: sparse LU of the free-free block, zero elsewhere."""

    def __init__(self, csr, free):
        super().__init__()
        self._n = csr.shape[0]
        self._idx = _np.flatnonzero(free)
        self._lu = _spla.splu(csr[self._idx][:, self._idx].tocsc())

    def MultAdd(self, s, xv, yv):
        yv._a[self._idx] += s * self._lu.solve(xv._a[self._idx])

    def Height(self):
        return self._n

    def Width(self):
        return self._n


class _FormMatrix(SparseMatrix):
    def Inverse(self, freedofs=None, inverse=None):
        free = _np.ones(self.Height(), dtype=bool) if freedofs is None else _np.asarray(freedofs, dtype=bool)
        return _DirectInverse(self.csr, free)


class BilinearForm:
    def __init__(self, space, **k):
        self.space, self.coef, self.mat, self.condense = space, {}, None, False

    def __iadd__(self, term):
        for k, v in term.coef.items():
            self.coef[k] = self.coef.get(k, 0.0) + v
        return self

    def Assemble(self):
        n = self.space.ndof
        m = _sp.csr_matrix((n, n))
        for k, v in self.coef.items():
            m = m + v * _problem[k]
        self.mat = _FormMatrix(m.tocsr())
        return self


class LinearForm:
    def __init__(self, space):
        self.space, self.vec = space, BaseVector(space.ndof)

    def __iadd__(self, term):
        return self

    def Assemble(self):
        return self


class GridFunction(CoefficientFunction):
    def __init__(self, space, name=""):
        self.space, self.name = space, name
        self.vec = BaseVector(space.ndof)
        CoefficientFunction.__init__(self, lambda X, Y: self.vec._a)      # nodal values (evaluation at the nodes only)
        gridfunctions.append(self)

    def Set(self, cf):
        self.vec._a[:] = CoefficientFunction._lift(cf).at_nodes()


def Integrate(cf, mesh, **k):
    """Trapezoid sum over the (uniform, tensor-product) node grid."""
    X, Y = _problem["nodes"]
    v = CoefficientFunction._lift(cf).at_nodes()
    nx, ny = len(_np.unique(X)), len(_np.unique(Y))
    wx = _np.full(nx, 1.0 / (nx - 1)); wx[[0, -1]] *= 0.5
    wy = _np.full(ny, 1.0 / (ny - 1)); wy[[0, -1]] *= 0.5
    ix = _np.rint(X * (nx - 1)).astype(int)
    iy = _np.rint(Y * (ny - 1)).astype(int)
    return float(_np.sum(v * wx[ix] * wy[iy]))


class TaskManager:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def Draw(*a, **k):
    pass


def Redraw(*a, **k):
    for h in redraw_hooks:
        h()


FEM_NAMES = ["CoefficientFunction", "x", "y", "sin", "cos", "exp", "sqrt", "dx", "grad", "Mesh", "H1", "BilinearForm",
             "LinearForm", "GridFunction", "Integrate", "TaskManager", "Draw", "Redraw"]
