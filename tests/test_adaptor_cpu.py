"""NGSolve export adaptor and CSV writers on an accessor-compatible stand-in; no GPU."""
import io
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class FakeNgsMatrix:
    def __init__(self, m):
        self.m = m.tocsr().copy()
        self.m.sort_indices()                # NGSolve rows are sorted by column
        self.height, self.width = self.m.shape

    def CSR(self):                       # NGSolve order: values, columns, row pointers
        return self.m.data, self.m.indices, self.m.indptr


class FakeFV:
    def __init__(self, a):
        self.a = a

    def NumPy(self):
        return self.a


class FakeVec:
    def __init__(self, a):
        self.a = np.asarray(a, dtype=np.float64)

    def FV(self):
        return FakeFV(self.a)


class FakeSpace:
    def __init__(self, bits):
        self.bits = bits

    def FreeDofs(self, *a):
        return self.bits


def test_csr_export_order_and_types():
    from navier_stokes_solver_b200 import adaptor
    m = sp.random(7, 5, density=0.4, random_state=1, format="csr")
    rp, ci, va, shape = adaptor.csr_of(FakeNgsMatrix(m))
    assert shape == (7, 5) and rp.dtype == np.int64 and ci.dtype == np.int32 and va.dtype == np.float64
    back = sp.csr_matrix((va, ci, rp), shape=shape)
    assert abs(back - m).max() == 0


def test_freedofs_and_vectors_are_views():
    from navier_stokes_solver_b200 import adaptor
    mask = adaptor.freedofs_mask(FakeSpace([True, False, True]))
    assert mask.tolist() == [1, 0, 1] and mask.dtype == np.uint8
    gfu, gfp = FakeVec([1.0, 2.0, 3.0]), FakeVec([4.0, 5.0])
    v = adaptor.vector_view(gfu)
    v.NumPy()[0] = 9.0
    assert gfu.a[0] == 9.0                               # zero-copy
    bv = adaptor.block_vector(gfu, gfp)
    assert len(bv) == 5 and bv[1].NumPy().tolist() == [4.0, 5.0]
    bv.NumPy()[:] = np.arange(5.0)
    adaptor.sync_back(bv)
    assert gfu.a.tolist() == [0.0, 1.0, 2.0] and gfp.a.tolist() == [3.0, 4.0]


def test_condensed_form_needs_sparse_parts():
    from navier_stokes_solver_b200 import adaptor

    class Blf:
        condense = True
        mat = None
        harmonic_extension = object()
    with pytest.raises(Exception):
        adaptor.ExportedForm(Blf())


def test_csv_tables_have_the_reference_columns(tmp_path):
    from navier_stokes_solver_b200 import adaptor
    f = adaptor.errors_frame([1.0, 0.1, 0.01], 0.5, 0.05, "taylor hood 2", 2, "minres", "mixed", 4321, 10, 20, 0, 20, 30)
    assert list(f.columns) == ['mesh_size', 'discretization', 'order', 'solver', 'iteration', 'error', 'solver_time',
                               'nvertices', 'nedges', 'nfaces', 'nfacets', 'nelements', 'ndofs', 'method']   # run.py:244-259
    assert f['iteration'].tolist() == [0, 1, 2]
    s = adaptor.sweep_frame(0.25, 3, 17, 0.2, True)
    assert list(s.columns) == ['mesh_size', 'order', 'iterations', 'time', 'gauss_seidel_enabled']   # parameter_sweep.py:58-64
    out = adaptor.write_csv([f, f], tmp_path / "errors.csv")
    assert len(out) == 6 and (tmp_path / "errors.csv").exists()
