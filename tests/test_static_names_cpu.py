"""Undefined-name pass over the Python that only ever runs on the GPU box (bench legs, torchrun workers, GPU tests, the device side of
the host mirror): a name that is read in a function must be bound in that function, an enclosing one, the module, or builtins.  No
linter is installed in the image; this is the part of one that matters for code the CPU suite cannot execute."""
import ast
import builtins
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bound_in(node):
    """Names bound anywhere inside `node` without descending into nested function / class bodies (their names are their own)."""
    out = set()

    def visit(n, top):
        if isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)) and not top:
            out.add(n.name)
            return
        if isinstance(n, ast.Lambda) and not top:
            return
        if isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
            out.add(n.id)
        elif isinstance(n, (ast.Import, ast.ImportFrom)):
            for a in n.names:
                out.add((a.asname or a.name).split(".")[0])
        elif isinstance(n, ast.ExceptHandler) and n.name:
            out.add(n.name)
        elif isinstance(n, (ast.Global, ast.Nonlocal)):
            out.update(n.names)
        elif isinstance(n, ast.arguments):
            for a in n.posonlyargs + n.args + n.kwonlyargs + ([n.vararg] if n.vararg else []) + ([n.kwarg] if n.kwarg else []):
                out.add(a.arg)
        for c in ast.iter_child_nodes(n):
            visit(c, False)
    visit(node, True)
    return out


def undefined_names(path):
    tree = ast.parse(open(path).read(), path)
    module = _bound_in(tree) | set(dir(builtins)) | {"__file__", "__name__", "__doc__"}
    bad = []

    def walk(node, scopes):
        for child in ast.iter_child_nodes(node):
            if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
                inner = scopes + [_bound_in(child)]
                for d in getattr(child, "decorator_list", []):
                    walk_expr(d, scopes)
                for dflt in child.args.defaults + [k for k in child.args.kw_defaults if k is not None]:
                    walk_expr(dflt, scopes)
                walk(child, inner)
            elif isinstance(child, ast.ClassDef):
                walk(child, scopes + [_bound_in(child)])
            elif isinstance(child, ast.Name) and isinstance(child.ctx, ast.Load):
                if not any(child.id in s for s in scopes):
                    bad.append((child.lineno, child.id))
            else:
                walk(child, scopes)

    def walk_expr(expr, scopes):
        for n in ast.walk(expr):
            if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Load) and not any(n.id in s for s in scopes):
                bad.append((n.lineno, n.id))
    walk(tree, [module])
    return bad


def _targets():
    pkg = os.path.join(ROOT, "navier-stokes-solver_b200")
    files = [os.path.join(pkg, f) for f in sorted(os.listdir(pkg)) if f.endswith(".py")]
    files += [os.path.join(ROOT, f) for f in ("bench.py", "benchlegs.py", "benchlegs_precond.py", "__graft_entry__.py")]
    files += [os.path.join(ROOT, "tests", f) for f in sorted(os.listdir(os.path.join(ROOT, "tests")))
              if f.endswith(".py") and (f.startswith(("dist_", "test_gpu_", "test_zz_")) or f == "conftest.py")]
    files += [os.path.join(ROOT, "tools", f) for f in sorted(os.listdir(os.path.join(ROOT, "tools"))) if f.endswith(".py")]
    files += [os.path.join(ROOT, "examples", f) for f in sorted(os.listdir(os.path.join(ROOT, "examples"))) if f.endswith(".py")]
    return files


@pytest.mark.parametrize("path", _targets(), ids=lambda p: os.path.relpath(p, ROOT))
def test_no_undefined_names(path):
    assert undefined_names(path) == []


def test_the_pass_sees_what_it_should():
    import tempfile
    src = "import os\ndef f(a):\n    b = a + missing\n    return [q for q in range(b)] + [os.sep, later]\ndef g():\n    return f(undefined_too)\nlater = 1\n"
    with tempfile.NamedTemporaryFile("w", suffix=".py", delete=False) as t:
        t.write(src)
    try:
        assert sorted(n for _, n in undefined_names(t.name)) == ["missing", "undefined_too"]
    finally:
        os.unlink(t.name)


def _own_module_attribute_misses(path):
    """`alias.attr` where alias is one of this repo's modules imported in the file: the attribute must exist on the module."""
    import importlib
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    tree = ast.parse(open(path).read(), path)
    own = ("navier_stokes_solver_b200", "benchlegs", "bench", "oracle", "test_partition_smoother_gloo", "test_partition_condensed_gloo")
    alias = {}
    for n in ast.walk(tree):
        if isinstance(n, ast.Import):
            for a in n.names:
                if a.name.split(".")[0] in own:
                    alias[a.asname or a.name.split(".")[0]] = a.name if a.asname else a.name.split(".")[0]
        elif isinstance(n, ast.ImportFrom) and n.module and n.module.split(".")[0] in own and n.level == 0:
            for a in n.names:
                try:
                    sub = importlib.import_module(n.module + "." + a.name)
                    alias[a.asname or a.name] = sub.__name__
                except ImportError:
                    mod = importlib.import_module(n.module)
                    if not hasattr(mod, a.name):
                        return [(n.lineno, n.module + "." + a.name)]
    misses = []
    for n in ast.walk(tree):
        if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name) and n.value.id in alias:
            mod = importlib.import_module(alias[n.value.id])
            if not hasattr(mod, n.attr):
                try:
                    importlib.import_module(alias[n.value.id] + "." + n.attr)       # a submodule not yet imported
                except ImportError:
                    misses.append((n.lineno, n.value.id + "." + n.attr))
    return misses


@pytest.mark.parametrize("path", _targets(), ids=lambda p: os.path.relpath(p, ROOT))
def test_attributes_of_own_modules_exist(path):
    if os.path.basename(path) in ("build.py",):
        pytest.skip("no own-module attribute access")
    assert _own_module_attribute_misses(path) == []


def _own_call_signature_misses(path):
    """Calls `alias.fn(...)` / `alias.Class(...)` into this repo's own modules must bind to the callee's signature (number of positional
    arguments, keyword names)."""
    import importlib
    import inspect
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    tree = ast.parse(open(path).read(), path)
    own = ("navier_stokes_solver_b200", "benchlegs", "bench", "oracle", "test_partition_smoother_gloo", "test_partition_condensed_gloo")
    alias, direct = {}, {}
    for n in ast.walk(tree):
        if isinstance(n, ast.Import):
            for a in n.names:
                if a.name.split(".")[0] in own:
                    alias[a.asname or a.name.split(".")[0]] = a.name if a.asname else a.name.split(".")[0]
        elif isinstance(n, ast.ImportFrom) and n.module and n.module.split(".")[0] in own and n.level == 0:
            for a in n.names:
                try:
                    alias[a.asname or a.name] = importlib.import_module(n.module + "." + a.name).__name__
                except ImportError:
                    direct[a.asname or a.name] = getattr(importlib.import_module(n.module), a.name, None)
    misses = []
    for n in ast.walk(tree):
        if not isinstance(n, ast.Call):
            continue
        callee = None
        if isinstance(n.func, ast.Attribute) and isinstance(n.func.value, ast.Name) and n.func.value.id in alias:
            callee = getattr(importlib.import_module(alias[n.func.value.id]), n.func.attr, None)
        elif isinstance(n.func, ast.Name) and n.func.id in direct:
            callee = direct[n.func.id]
        if callee is None or not (inspect.isfunction(callee) or inspect.isclass(callee)):
            continue
        if any(isinstance(a, ast.Starred) for a in n.args) or any(k.arg is None for k in n.keywords):
            continue
        try:
            sig = inspect.signature(callee)
            sig.bind(*[None] * len(n.args), **{k.arg: None for k in n.keywords})
        except (TypeError, ValueError) as ex:
            if isinstance(ex, TypeError):
                misses.append((n.lineno, getattr(callee, "__name__", "?"), str(ex)))
    return misses


@pytest.mark.parametrize("path", _targets(), ids=lambda p: os.path.relpath(p, ROOT))
def test_calls_into_own_modules_bind_to_their_signatures(path):
    assert _own_call_signature_misses(path) == []
