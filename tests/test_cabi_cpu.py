"""CPU checks of the boundary: the C-ABI library loads, exports every symbol include/nsb200.h declares,
the ctypes table matches the header, and the product path refuses to run without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "nsb200.h")).read()
    return re.findall(r"^NSB_API\s+[\w\s\*]+?\b(nsb_\w+)\s*\(", src, flags=re.M)


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__
    __graft_entry__.build()
    from navier_stokes_solver_b200 import _lib
    return _lib.load()


def test_header_declares_expected_count():
    syms = header_symbols()
    assert len(syms) == len(set(syms)) and len(syms) >= 30


def test_library_exports_every_declared_symbol(lib):
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_ctypes_table_matches_header(lib):
    from navier_stokes_solver_b200 import _lib
    assert sorted(_lib.SIGNATURES) == sorted(header_symbols())


def test_version_and_error_string(lib):
    assert lib.nsb_version() >= 100
    assert isinstance(lib.nsb_last_error(), bytes)


def test_no_cpu_fallback(lib):
    """Without a CUDA device context creation fails loudly (there is no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from navier_stokes_solver_b200 import _lib
    h = ctypes.c_void_p()
    rc = lib.nsb_ctx_create(0, ctypes.byref(h))
    assert rc != 0
    assert b"no CUDA device" in lib.nsb_last_error() or b"CUDA" in lib.nsb_last_error()
    from navier_stokes_solver_b200 import la
    with pytest.raises(_lib.NsbError):
        la.Context()


def test_product_package_does_not_import_oracle():
    pkg = os.path.join(ROOT, "navier-stokes-solver_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f
                assert "krylov_oracle" not in txt or f == "cg.cu", f


def test_rk_host_algebra_matches_golden():
    from oracle.cases import load_golden
    from navier_stokes_solver_b200.solvers import ImplicitRungeKuttaMethodWeights, linear_implicit_runge_kutta_step
    gold = load_golden("runge_kutta")
    for deg in (3, 10):
        w = ImplicitRungeKuttaMethodWeights(deg)
        np.testing.assert_allclose(w.a, gold["a%d" % deg], rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(w.b, gold["b%d" % deg], rtol=1e-12, atol=1e-15)
        nxt = linear_implicit_runge_kutta_step(w, gold["evo%d" % deg], gold["y%d" % deg], 0.01)
        np.testing.assert_allclose(nxt, gold["next%d" % deg], rtol=1e-10)
