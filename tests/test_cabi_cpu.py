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


def _header_params():
    """name -> list of parameter type strings of every NSB_API declaration in include/nsb200.h."""
    src = open(os.path.join(ROOT, "include", "nsb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"^NSB_API\s+([\w\s\*]+?)\b(nsb_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.M | re.S):
        params = [p.strip() for p in " ".join(m.group(3).split()).split(",")]
        if params == ["void"] or params == [""]:
            params = []
        out[m.group(2)] = (m.group(1).strip(), params)
    return out


def _klass_of_c(decl):
    """pointer / float / int class of a C parameter or return declaration."""
    handles = ("nsb_ctx", "nsb_mat", "nsb_pre", "nsb_halo", "nsb_cond", "nsb_simple")
    if "*" in decl or "[" in decl or any(re.search(r"\b%s\b" % h, decl) for h in handles):
        return "ptr"
    if re.search(r"\bdouble\b", decl):
        return "float"
    return "int"


def _klass_of_ctypes(t):
    if t in (ctypes.c_double, ctypes.c_float):
        return "float"
    if t in (ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_int32):
        return "int"
    return "ptr"        # c_void_p, c_char_p, POINTER(...)


def test_ctypes_argument_lists_match_the_header_parameter_by_parameter():
    """Arity and pointer / integer / double class of every parameter: an ABI drift between include/nsb200.h and the ctypes table (a
    forgotten argument, a double passed as an int) shows up here instead of as a crash on the GPU box."""
    from navier_stokes_solver_b200 import _lib
    decls = _header_params()
    assert sorted(decls) == sorted(_lib.SIGNATURES)
    for name, (ret, params) in decls.items():
        res, args = _lib.SIGNATURES[name]
        assert len(args) == len(params), (name, params, args)
        for k, (p, a) in enumerate(zip(params, args)):
            assert _klass_of_c(p) == _klass_of_ctypes(a), (name, k, p, a)
        # 64-bit integers must not be declared as c_int (and vice versa)
        for k, (p, a) in enumerate(zip(params, args)):
            if _klass_of_c(p) == "int":
                assert (a is ctypes.c_int64) == bool(re.search(r"\bint64_t\b", p)), (name, k, p, a)
        want_ret = "ptr" if ("*" in ret or "char" in ret) else ("int" if re.search(r"\bint(64_t)?\b", ret) else "ptr")
        assert _klass_of_ctypes(res) == want_ret, (name, ret, res)


def test_integration_stub_matches_the_header():
    """The binding INTEGRATION.md shows a reference maintainer (section 2) declares the same arities as the header."""
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    decls = _header_params()
    found = re.findall(r"lib\.(nsb_\w+)\.argtypes\s*=\s*\[(.*?)\]", text, flags=re.S)
    assert len(found) >= 5
    for name, body in found:
        n = len([x for x in re.sub(r"C\.POINTER\(([^)]*)\)", r"PTR", body).split(",") if x.strip()])
        assert name in decls and n == len(decls[name][1]), (name, n, decls.get(name))


def test_every_python_call_into_the_library_passes_the_declared_number_of_arguments():
    """Static pass over the host mirror: each `<lib>.nsb_*(...)` call site passes as many positional arguments as the ctypes table
    declares (ctypes only complains at call time, i.e. on the GPU box)."""
    import ast
    from navier_stokes_solver_b200 import _lib
    pkg = os.path.join(ROOT, "navier-stokes-solver_b200")
    files = [os.path.join(pkg, f) for f in os.listdir(pkg) if f.endswith(".py")] + [os.path.join(ROOT, f) for f in ("bench.py", "benchlegs.py")]
    seen = 0
    for path in files:
        tree = ast.parse(open(path).read())
        for node in ast.walk(tree):
            if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr.startswith("nsb_"):
                name = node.func.attr
                assert name in _lib.SIGNATURES, (path, node.lineno, name)
                if any(isinstance(a, ast.Starred) for a in node.args):
                    fixed = sum(1 for a in node.args if not isinstance(a, ast.Starred))
                    assert fixed <= len(_lib.SIGNATURES[name][1])
                else:
                    assert len(node.args) == len(_lib.SIGNATURES[name][1]) and not node.keywords, (path, node.lineno, name, len(node.args))
                seen += 1
    assert seen >= 40
