"""numpy/scipy stand-in for the slice of the NGSolve linear-algebra API that the
reference's Krylov files touch.

TEST INFRASTRUCTURE ONLY.  Purpose: run /root/reference/{minres,bramble_pasciak_cg,
orthonormalization,runge_kutta_method}.py and solvers/bramblepasciak_new.py
*unmodified* in the build container (NGSolve itself is not installable here) to
produce the golden vectors under tests/golden/ (see oracle/make_golden.py).
Nothing in the product package imports this.

Semantics mirrored from NGSolve (third-party, version unpinned by the reference):
  * ``mat * vec`` is a *lazy* expression that is evaluated into the left-hand
    side of ``x.data = ...`` / ``x.data += ...``  (needed by
    bramble_pasciak_cg.py:45-47,58-59 -- MatrixAB indexes blocks of the target).
  * BlockVector components share storage with the flat block vector.
  * ``Preconditioner``-like operators are plain BaseMatrix objects.
"""
import math as _math
import time as _time

import numpy as _np
import scipy.sparse as _sp

__all__ = [
    "BaseVector", "BlockVector", "BaseMatrix", "BlockMatrix", "IdentityMatrix",
    "SparseMatrix", "DiagonalMatrix", "InnerProduct", "Norm", "Projector", "Vector", "Matrix",
    "CreateVVector", "PermutationMatrix",
]


# --------------------------------------------------------------------------- vectors
class _Expr:
    """Lazy linear combination  sum_k  c_k * (Op_k * v_k | v_k)."""

    def __init__(self, terms):
        self.terms = terms  # list of (coef, op_or_None, vec)

    def __add__(self, o):
        return _Expr(self.terms + _as_expr(o).terms)

    __radd__ = __add__

    def __sub__(self, o):
        return _Expr(self.terms + [(-c, m, v) for c, m, v in _as_expr(o).terms])

    def __rsub__(self, o):
        return _as_expr(o) - self

    def __neg__(self):
        return _Expr([(-c, m, v) for c, m, v in self.terms])

    def __mul__(self, s):
        return _Expr([(c * s, m, v) for c, m, v in self.terms])

    __rmul__ = __mul__

    def Evaluate(self):
        """Materialise as a fresh flat vector (used when the left side is not a .data target)."""
        n = None
        for _, m, v in self.terms:
            n = m.height if m is not None else len(v)
            break
        out = BaseVector(n)
        _assign(out, self, add=False)
        return out


def _as_expr(o):
    if isinstance(o, _Expr):
        return o
    if isinstance(o, BaseVector):
        return _Expr([(1.0, None, o)])
    raise TypeError("cannot use %r in a vector expression" % (type(o),))


def _assign(target, expr, add):
    expr = _as_expr(expr)
    acc = None
    ops = []
    for c, m, v in expr.terms:
        if m is None:
            acc = c * v._a if acc is None else acc + c * v._a      # reads old values
        else:
            if _np.shares_memory(v._a, target._a):
                v = v.Copy()                                          # x aliases y: snapshot first
            ops.append((c, m, v))
    if add:
        if acc is not None:
            target._a += acc
    else:
        target._a[:] = 0.0 if acc is None else acc
    for c, m, v in ops:
        m.MultAdd(c, v, target)


class BaseVector:
    def __init__(self, n_or_array):
        if isinstance(n_or_array, _np.ndarray):
            self._a = n_or_array
        else:
            self._a = _np.zeros(int(n_or_array))

    # -- NGSolve API
    def CreateVector(self):
        return BaseVector(len(self._a))

    def Copy(self):
        return BaseVector(self._a.copy())

    def FV(self):
        return self

    def NumPy(self):
        return self._a

    def __len__(self):
        return len(self._a)

    @property
    def size(self):
        return len(self._a)

    @property
    def data(self):
        return self

    @data.setter
    def data(self, expr):
        if expr is self:
            return  # result of  x.data += ...
        _assign(self, expr, add=False)

    def __iadd__(self, expr):
        _assign(self, expr, add=True)
        return self

    def __isub__(self, expr):
        _assign(self, -_as_expr(expr), add=True)
        return self

    def __imul__(self, s):
        self._a *= s
        return self

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return BaseVector(self._a[idx])
        return float(self._a[idx])

    def __setitem__(self, idx, val):
        if isinstance(val, BaseVector):
            self._a[idx] = val._a
        else:
            self._a[idx] = val

    # -- expression building
    def __add__(self, o):
        return _as_expr(self) + o

    def __radd__(self, o):
        return _as_expr(self) + o

    def __sub__(self, o):
        return _as_expr(self) - o

    def __rsub__(self, o):
        return _as_expr(o) - self

    def __neg__(self):
        return -_as_expr(self)

    def __mul__(self, s):
        return _as_expr(self) * s

    __rmul__ = __mul__

    def InnerProduct(self, o):
        return InnerProduct(self, o)

    def Norm(self):
        return Norm(self)


def CreateVVector(n):
    return BaseVector(n)


class BlockVector(BaseVector):
    """Block vector; components are views into one flat array."""

    def __init__(self, comps):
        sizes = [len(c) for c in comps]
        offs = _np.concatenate([[0], _np.cumsum(sizes)]).astype(int)
        # NGSolve's BlockVector references the component vectors (GridFunction storage).
        # To share storage both ways we re-home each component into the flat array.
        flat = _np.concatenate([c._a for c in comps]) if comps else _np.zeros(0)
        super().__init__(flat)
        self._offs = offs
        self._comps = []
        for i, c in enumerate(comps):
            view = self._a[offs[i]:offs[i + 1]]
            c._a = view  # caller's vector now aliases the block storage (in-place sol semantics)
            self._comps.append(c)

    def CreateVector(self):
        return BlockVector([BaseVector(int(self._offs[i + 1] - self._offs[i])) for i in range(len(self._comps))])

    def Copy(self):
        out = self.CreateVector()
        out._a[:] = self._a
        return out

    @property
    def nblocks(self):
        return len(self._comps)

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return BaseVector(self._a[idx])
        return self._comps[idx]

    def __setitem__(self, idx, val):
        if isinstance(idx, slice):
            self._a[idx] = val._a if isinstance(val, BaseVector) else val
        else:
            self._comps[idx]._a[:] = val._a if isinstance(val, BaseVector) else val


def InnerProduct(a, b):
    return float(_np.dot(a._a, b._a))


def Norm(a):
    return float(_math.sqrt(_np.dot(a._a, a._a)))


# --------------------------------------------------------------------------- matrices
class BaseMatrix:
    """Operator protocol.  Subclasses override Mult and/or MultAdd (python subclasses in the
    reference: bramble_pasciak_cg.py:9-62, solvers/bramblepasciak_new.py:84-103)."""

    def __init__(self):
        pass

    # defaults in terms of each other, like NGSolve
    def Mult(self, x, y):
        y._a[:] = 0.0
        self.MultAdd(1.0, x, y)

    def MultAdd(self, s, x, y):
        if type(self).Mult is BaseMatrix.Mult:
            raise NotImplementedError
        t = y.CreateVector()
        self.Mult(x, t)
        y._a += s * t._a

    def MultTrans(self, x, y):
        y._a[:] = 0.0
        self.MultTransAdd(1.0, x, y)

    def MultTransAdd(self, s, x, y):
        raise NotImplementedError

    def Height(self):
        raise NotImplementedError

    def Width(self):
        raise NotImplementedError

    @property
    def height(self):
        return self.Height()

    @property
    def width(self):
        return self.Width()

    def CreateColVector(self):
        return BaseVector(self.height)

    def CreateRowVector(self):
        return BaseVector(self.width)

    def CreateVector(self):
        return self.CreateColVector()

    @property
    def T(self):
        return _Transposed(self)

    def CreateTranspose(self):
        return _Transposed(self)

    def __mul__(self, v):
        if isinstance(v, BaseVector):
            return _Expr([(1.0, self, v)])
        if isinstance(v, _Expr):
            return _Expr([(1.0, self, v.Evaluate())])
        return NotImplemented

    def __rmul__(self, s):
        return _Scaled(float(s), self)

    def __neg__(self):
        return _Scaled(-1.0, self)

    def __matmul__(self, o):
        return _Product(self, o)

    def __add__(self, o):
        return _Sum(self, o, 1.0)

    def __sub__(self, o):
        return _Sum(self, o, -1.0)


class _Scaled(BaseMatrix):
    def __init__(self, s, m):
        self.s, self.m = s, m

    def MultAdd(self, s, x, y):
        self.m.MultAdd(s * self.s, x, y)

    def MultTransAdd(self, s, x, y):
        self.m.MultTransAdd(s * self.s, x, y)

    def Height(self):
        return self.m.height

    def Width(self):
        return self.m.width

    def CreateColVector(self):
        return self.m.CreateColVector()

    def CreateRowVector(self):
        return self.m.CreateRowVector()


class _Transposed(BaseMatrix):
    def __init__(self, m):
        self.m = m

    def MultAdd(self, s, x, y):
        self.m.MultTransAdd(s, x, y)

    def MultTransAdd(self, s, x, y):
        self.m.MultAdd(s, x, y)

    def Height(self):
        return self.m.width

    def Width(self):
        return self.m.height


class _Product(BaseMatrix):
    def __init__(self, a, b):
        self.a, self.b = a, b

    def MultAdd(self, s, x, y):
        t = self.b.CreateColVector() if self.b.height >= 0 else BaseVector(len(x))
        t._a[:] = 0.0
        self.b.MultAdd(1.0, x, t)
        self.a.MultAdd(s, t, y)

    def MultTransAdd(self, s, x, y):
        t = self.a.CreateRowVector() if self.a.width >= 0 else BaseVector(len(x))
        t._a[:] = 0.0
        self.a.MultTransAdd(1.0, x, t)
        self.b.MultTransAdd(s, t, y)

    def Height(self):
        h = self.a.height
        return h if h >= 0 else self.b.height

    def Width(self):
        w = self.b.width
        return w if w >= 0 else self.a.width


class _Sum(BaseMatrix):
    def __init__(self, a, b, sb):
        self.a, self.b, self.sb = a, b, sb

    def MultAdd(self, s, x, y):
        self.a.MultAdd(s, x, y)
        self.b.MultAdd(s * self.sb, x, y)

    def MultTransAdd(self, s, x, y):
        self.a.MultTransAdd(s, x, y)
        self.b.MultTransAdd(s * self.sb, x, y)

    def Height(self):
        return max(self.a.height, self.b.height)

    def Width(self):
        return max(self.a.width, self.b.width)


class IdentityMatrix(BaseMatrix):
    def __init__(self, n=None):
        self.n = -1 if n is None else int(n)

    def MultAdd(self, s, x, y):
        y._a += s * x._a

    MultTransAdd = MultAdd

    def Height(self):
        return self.n

    def Width(self):
        return self.n

    def CreateColVector(self):
        return BaseVector(max(self.n, 0))


class SparseMatrix(BaseMatrix):
    """Wraps a scipy CSR (what ``BilinearForm.mat`` is in NGSolve)."""

    def __init__(self, csr):
        self.csr = _sp.csr_matrix(csr)
        self._csrT = None

    def MultAdd(self, s, x, y):
        y._a += s * (self.csr @ x._a)

    def MultTransAdd(self, s, x, y):
        if self._csrT is None:
            self._csrT = self.csr.T.tocsr()
        y._a += s * (self._csrT @ x._a)

    def Height(self):
        return self.csr.shape[0]

    def Width(self):
        return self.csr.shape[1]

    def CreateTranspose(self):
        return SparseMatrix(self.csr.T.tocsr())

    @property
    def T(self):
        return _Transposed(self)

    def CreateBlockSmoother(self, blocks):
        return _BlockSmoother(self.csr, blocks)


class _BlockSmoother(BaseMatrix):
    """``mat.CreateBlockSmoother(blocks)`` (NGSolve, third-party: restated).  As an operator it is additive block Jacobi,
    y = sum_b P_b^T A_bb^-1 P_b x; ``Smooth(u, f)`` is one block Gauss-Seidel sweep over the blocks in list order,
    u_b += A_bb^-1 (f - A u)_b, ``SmoothBack`` the same sweep in reverse order."""

    def __init__(self, csr, blocks):
        self.csr = csr
        self.blocks = [_np.asarray(b, dtype=_np.int64) for b in blocks if len(b)]
        self.inv = [_np.linalg.inv(csr[b][:, b].toarray()) for b in self.blocks]

    def MultAdd(self, s, x, y):
        for b, inv in zip(self.blocks, self.inv):
            y._a[b] += s * (inv @ x._a[b])

    MultTransAdd = MultAdd

    def _sweep(self, u, f, order):
        for k in order:
            b = self.blocks[k]
            u._a[b] += self.inv[k] @ (f._a[b] - self.csr[b] @ u._a)

    def Smooth(self, u, f):
        self._sweep(u, f, range(len(self.blocks)))

    def SmoothBack(self, u, f):
        self._sweep(u, f, range(len(self.blocks) - 1, -1, -1))

    def Height(self):
        return self.csr.shape[0]

    def Width(self):
        return self.csr.shape[1]


class DiagonalMatrix(BaseMatrix):
    """Jacobi ('local') preconditioner: y = d .* x, with d = 0 on constrained dofs."""

    def __init__(self, d):
        self.d = _np.asarray(d, dtype=float)

    def MultAdd(self, s, x, y):
        y._a += s * (self.d * x._a)

    MultTransAdd = MultAdd

    def Height(self):
        return len(self.d)

    def Width(self):
        return len(self.d)


class Projector(BaseMatrix):
    def __init__(self, mask, range=True):
        m = _np.asarray(mask, dtype=bool)
        self.d = (m if range else ~m).astype(float)

    def MultAdd(self, s, x, y):
        y._a += s * self.d * x._a

    MultTransAdd = MultAdd

    def Height(self):
        return len(self.d)

    def Width(self):
        return len(self.d)


class BlockMatrix(BaseMatrix):
    def __init__(self, rows):
        self.rows = rows
        self.nr = len(rows)
        self.nc = len(rows[0])

    def _row_h(self, i):
        for m in self.rows[i]:
            if m is not None and m.height >= 0:
                return m.height
        raise ValueError("block row without sized block")

    def _col_w(self, j):
        for i in range(self.nr):
            m = self.rows[i][j]
            if m is not None and m.width >= 0:
                return m.width
        raise ValueError("block col without sized block")

    def MultAdd(self, s, x, y):
        for i in range(self.nr):
            for j in range(self.nc):
                m = self.rows[i][j]
                if m is not None:
                    m.MultAdd(s, x[j], y[i])

    def MultTransAdd(self, s, x, y):
        for i in range(self.nr):
            for j in range(self.nc):
                m = self.rows[i][j]
                if m is not None:
                    m.MultTransAdd(s, x[i], y[j])

    def Height(self):
        return sum(self._row_h(i) for i in range(self.nr))

    def Width(self):
        return sum(self._col_w(j) for j in range(self.nc))

    def CreateColVector(self):
        return BlockVector([BaseVector(self._row_h(i)) for i in range(self.nr)])

    def CreateRowVector(self):
        return BlockVector([BaseVector(self._col_w(j)) for j in range(self.nc)])

    def __getitem__(self, ij):
        return self.rows[ij[0]][ij[1]]


class PermutationMatrix(BaseMatrix):
    """``PermutationMatrix(width, ind)``: (P x)[i] = x[ind[i]]  (templates/NavierStokesSIMPLE_iterative.py:138-144)."""

    def __init__(self, width, ind):
        self._w, self._ind = int(width), _np.asarray(ind, dtype=_np.int64)

    def MultAdd(self, s, x, y):
        y._a += s * x._a[self._ind]

    def Height(self):
        return len(self._ind)

    def Width(self):
        return self._w


# --------------------------------------------------------------------------- dense (runge_kutta_method.py)
class Vector:
    """Dense vector (ngsolve.bla.Vector)."""

    def __new__(cls, n):
        return _DVec(_np.zeros(int(n)))


class _DVec(BaseVector):
    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return _DVec(self._a[idx])
        return float(self._a[idx])

    def __add__(self, o):
        return _DVec(self._a + (o._a if isinstance(o, BaseVector) else o))

    def __iadd__(self, o):
        self._a += o._a if isinstance(o, BaseVector) else o
        return self

    def __rmul__(self, s):
        return _DVec(self._a * s)

    __mul__ = __rmul__


class Matrix:
    """Dense matrix (ngsolve.bla.Matrix)."""

    def __init__(self, h, w=None):
        if isinstance(h, _np.ndarray):
            self._m = h
        else:
            self._m = _np.zeros((int(h), int(w)))

    def Height(self):
        return self._m.shape[0]

    def Width(self):
        return self._m.shape[1]

    @property
    def h(self):
        return self._m.shape[0]

    @property
    def w(self):
        return self._m.shape[1]

    def __getitem__(self, ij):
        return float(self._m[ij])

    def __setitem__(self, ij, v):
        self._m[ij] = v

    def Inverse(self, out):
        out._m[:] = _np.linalg.inv(self._m)

    def __mul__(self, o):
        if isinstance(o, Matrix):
            return Matrix(self._m @ o._m)
        if isinstance(o, BaseVector):
            return _DVec(self._m @ o._a)
        return Matrix(self._m * o)

    def __rmul__(self, s):
        return Matrix(self._m * s)

    def __neg__(self):
        return Matrix(-self._m)

    def NumPy(self):
        return self._m


from . import la, ngstd  # noqa: E402,F401
from . import fem  # noqa: E402
from .fem import (CoefficientFunction, x, y, sin, cos, exp, sqrt, dx, grad, Mesh, H1, BilinearForm, LinearForm,  # noqa: E402,F401
                  GridFunction, Integrate, TaskManager, Draw, Redraw)

__all__ += fem.FEM_NAMES
