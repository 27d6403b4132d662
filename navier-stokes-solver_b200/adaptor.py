"""Glue between live NGSolve objects and the device solvers, plus the result tables the reference scripts write
(SURVEY.md 8(f) item 4).  NGSolve itself is not imported here: everything goes through the public accessors of the
objects handed in (``bf.mat.CSR()``, ``vec.FV().NumPy()``, ``fes.FreeDofs()``), so the module also works on any
stand-in exposing the same accessors (tests use one).

    A, B = export_bilinear_form(a), export_bilinear_form(b)          # once per Assemble()  (run.py:105-106)
    preA = jacobi_from_space(A, X)                                    # Preconditioner(a, 'local')  (run.py:62)
    rhs, sol = block_vector(f.vec, g.vec), block_vector(gfu.vec, gfp.vec)   # zero-copy views (run.py:47-48)
    errors_frame(...).to_csv('errors.csv')                           # columns of run.py:244-259
"""
from __future__ import annotations

import numpy as np

from .la import BaseVector, BlockVector, CondensedOperator, DiagonalPreconditioner, SparseMatrix


def csr_of(ngs_matrix):
    """(rowptr int64, colidx int32, values fp64, shape) of an assembled NGSolve SparseMatrix.  NGSolve's ``CSR()``
    returns (values, column indices, row pointers)."""
    vals, cols, rowptr = ngs_matrix.CSR()
    return (np.asarray(rowptr, dtype=np.int64), np.asarray(cols, dtype=np.int32), np.asarray(vals, dtype=np.float64),
            (int(ngs_matrix.height), int(ngs_matrix.width)))


def export_matrix(ngs_matrix, ctx=None) -> SparseMatrix:
    rp, ci, va, shape = csr_of(ngs_matrix)
    return SparseMatrix(rp, ci, va, shape, ctx)


class ExportedForm:
    """What the device solvers need from a BilinearForm: ``.mat`` and, when statically condensed, the four extra
    operators of solvers/bramblepasciak_new.py:8-21,84-103 -- all uploaded once per assembly."""

    def __init__(self, blf, ctx=None):
        self.mat = export_matrix(blf.mat, ctx)
        self.condense = bool(getattr(blf, "condense", False))
        if self.condense:
            for name in ("harmonic_extension", "harmonic_extension_trans", "inner_solve", "inner_matrix"):
                op = getattr(blf, name)
                if not hasattr(op, "CSR"):
                    raise TypeError("blf.%s is not a sparse matrix with a CSR() accessor: assemble it explicitly (NGSolve keeps "
                                    "it element by element) before exporting; the device has no element-matrix path" % name)
                setattr(self, name, export_matrix(op, ctx))

    def condensed_operator(self) -> CondensedOperator:
        return CondensedOperator.from_bilinear_form(self)


def export_bilinear_form(blf, ctx=None):
    """``SparseMatrix`` of an uncondensed form, ``ExportedForm`` (with ``.condense = True``) of a condensed one."""
    return ExportedForm(blf, ctx) if getattr(blf, "condense", False) else export_matrix(blf.mat, ctx)


def freedofs_mask(space_or_bitarray, *args) -> np.ndarray:
    """uint8 mask from ``fes.FreeDofs()`` (a BitArray) or from the space itself."""
    bits = space_or_bitarray.FreeDofs(*args) if hasattr(space_or_bitarray, "FreeDofs") else space_or_bitarray
    return np.fromiter((1 if b else 0 for b in bits), dtype=np.uint8, count=len(bits))


def jacobi_from_space(mat: SparseMatrix, space_or_bitarray=None) -> DiagonalPreconditioner:
    """``Preconditioner(bf, 'local')`` restricted to the free dofs (run.py:61-62, stokes_hcurldiv.py:53)."""
    mask = freedofs_mask(space_or_bitarray) if space_or_bitarray is not None else None
    return DiagonalPreconditioner.jacobi(mat, mask)


def vector_view(ngs_vector) -> BaseVector:
    """Zero-copy view of an NGSolve BaseVector: the solver writes straight into the GridFunction storage."""
    return BaseVector(ngs_vector.FV().NumPy())


def block_vector(*ngs_vectors) -> BlockVector:
    """BlockVector([gfu.vec, gfp.vec]) (run.py:47-48).  The device call works on one flat (u | p) array; ``sync_back``
    copies the components into the NGSolve vectors afterwards."""
    bv = BlockVector([np.array(v.FV().NumPy(), dtype=np.float64, copy=True) for v in ngs_vectors])
    bv._ngs = list(ngs_vectors)
    return bv


def sync_back(bv: BlockVector):
    """Write the components of ``bv`` back into the NGSolve vectors it was built from (in-place ``sol`` semantics)."""
    for comp, v in zip(bv._comps, getattr(bv, "_ngs", [])):
        v.FV().NumPy()[:] = comp._a


# ---- result tables ---------------------------------------------------------------------------------------
def errors_frame(errors, solver_time, mesh_size, discretization, order, solver, method, ndofs, nvertices=0, nedges=0,
                 nfaces=0, nfacets=0, nelements=0):
    """One block of ``errors.csv`` with the columns of run.py:244-259 (so plot.py reads it unchanged)."""
    import pandas as pd
    return pd.DataFrame({
        'mesh_size': mesh_size, 'discretization': discretization, 'order': order, 'solver': solver,
        'iteration': range(len(errors)), 'error': list(errors), 'solver_time': solver_time, 'nvertices': nvertices,
        'nedges': nedges, 'nfaces': nfaces, 'nfacets': nfacets, 'nelements': nelements, 'ndofs': ndofs, 'method': method})


def sweep_frame(mesh_size, order, iterations, time, gauss_seidel_enabled):
    """One row of ``data.csv`` (templates/run_navier_stokes_parameter_sweep.py:58-64)."""
    import pandas as pd
    return pd.DataFrame({'mesh_size': mesh_size, 'order': order, 'iterations': iterations, 'time': time,
                         'gauss_seidel_enabled': gauss_seidel_enabled}, index=[0])


def write_csv(frames, path):
    """``pd.concat(frames, ignore_index=True).to_csv(path)`` (run.py:261-262)."""
    import pandas as pd
    data = pd.concat(list(frames), ignore_index=True)
    data.to_csv(path)
    return data
