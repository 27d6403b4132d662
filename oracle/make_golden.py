"""Generate tests/golden/*.npz by running the reference's own Krylov files UNMODIFIED
(/root/reference, imported by path) over the numpy stand-in of the NGSolve la layer
(oracle/ngs_standin).  Run in the build container only; the GPU box has no /root/reference.

    python oracle/make_golden.py

Fixtures hold: solver outputs (histories, iteration counts, solutions), the injected lambda_min,
and a fingerprint of the assembled inputs so a drifting assembler is detected by the tests.
"""
import contextlib
import importlib.util
import io
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "ngs_standin"))   # the stand-in named `ngsolve`
sys.path.insert(0, os.path.dirname(HERE))

import ngsolve  # noqa: E402  (stand-in)
from ngsolve import BaseVector, BlockVector, BlockMatrix, SparseMatrix, DiagonalMatrix  # noqa: E402
from ngsolve import la as ngla  # noqa: E402
from oracle.cases import Case, GOLDEN_DIR  # noqa: E402

REF = os.environ.get("NSB_REFERENCE", "/root/reference")


def load_ref(relpath, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ref_minres = load_ref("minres.py", "ref_minres")
ref_bpcg1 = load_ref("bramble_pasciak_cg.py", "ref_bpcg1")
ref_bpcg2 = load_ref("solvers/bramblepasciak_new.py", "ref_bpcg2")
ref_orth = load_ref("orthonormalization.py", "ref_orth")
ref_rk = load_ref("runge_kutta_method.py", "ref_rk")


def quiet(fn, *a, **k):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        out = fn(*a, **k)
    return out, buf.getvalue()


def gold_minres(mesh, initialize, tol, maxsteps, use_pre=True, tag=""):
    c = Case(mesh)
    A, B = SparseMatrix(c.A), SparseMatrix(c.B)
    K = BlockMatrix([[A, B.T], [B, None]])                               # run.py:45
    C = BlockMatrix([[DiagonalMatrix(c.dA), None], [None, DiagonalMatrix(c.dM)]]) if use_pre else None
    rhs = BlockVector([BaseVector(c.f.copy()), BaseVector(c.g.copy())])
    gfu, gfp = BaseVector(c.u0.copy()), BaseVector(np.zeros(c.n_p))
    sol = BlockVector([gfu, gfp])
    (u, errors), _ = quiet(ref_minres.MinRes, mat=K, pre=C, rhs=rhs, sol=sol, initialize=initialize,
                           tol=tol, maxsteps=maxsteps)                   # run.py:52
    assert u is sol
    name = "minres_%s%s" % (mesh, tag)
    np.savez_compressed(os.path.join(GOLDEN_DIR, name), errors=np.array(errors), sol=sol.NumPy().copy(),
                        gfu=gfu.NumPy().copy(), fingerprint=c.fingerprint(),
                        args=np.array([float(initialize), tol, maxsteps, float(use_pre)]))
    print(name, "iterations", len(errors) - 1, "final", errors[-1])


def gold_bpcg(mesh, tol, maxsteps, tag=""):
    c = Case(mesh)
    lam = c.lam_min()
    A, B = SparseMatrix(c.A), SparseMatrix(c.B)
    preA, preM = DiagonalMatrix(c.dA), DiagonalMatrix(c.dM)
    # ---- v1  (stokes_hcurldiv.py:75-77, run.py:33-38)
    gfu, gfp = BaseVector(c.u0.copy()), BaseVector(np.zeros(c.n_p))
    solution = BlockVector([gfu, gfp])
    ngla.inject_eigenvalues([lam, 1.0])
    (sol1, errors1), _ = quiet(ref_bpcg1.bramble_pasciak_cg, A, B, None, preA, preM,
                               BaseVector(c.f.copy()), BaseVector(c.g.copy()), solution=solution,
                               tolerance=tol, max_steps=maxsteps)
    # ---- v2  (templates/NavierStokesSIMPLE_iterative.py:397)
    class Blf:
        def __init__(self, m):
            self.mat = m
            self.condense = False
    gfu2, gfp2 = BaseVector(c.u0.copy()), BaseVector(np.zeros(c.n_p))
    sol2 = BlockVector([gfu2, gfp2])
    ngla.inject_eigenvalues([lam, 1.0])
    (it2, secs), out = quiet(ref_bpcg2.BramblePasciakCG, Blf(A), Blf(B), None, BaseVector(c.f.copy()),
                             BaseVector(c.g.copy()), preA, preM, sol2, initialize=False, tol=tol,
                             maxsteps=maxsteps, rel_err=True)
    hist2 = [float(m.group(2)) for m in re.finditer(r"it =\s+(\d+)\s+err =\s+(\S+)", out)]
    err0 = float(re.search(r"err0 (\S+)", out).group(1))
    name = "bpcg_%s%s" % (mesh, tag)
    np.savez_compressed(os.path.join(GOLDEN_DIR, name), lam_min=lam, errors_v1=np.array(errors1),
                        sol_v1=sol1.NumPy().copy(), it_v2=it2, hist_v2=np.array(hist2), err0_v2=err0,
                        sol_v2=sol2.NumPy().copy(), fingerprint=c.fingerprint(),
                        args=np.array([tol, maxsteps]))
    print(name, "v1 passes", len(errors1), "v2 it", it2, "hist", len(hist2), "lam_min", lam)


def gold_bpcg_condensed(mesh, tol, maxsteps):
    """solvers/bramblepasciak_new.py with blfA.condense = True (:8-21, :84-103) on the synthetic condensation of
    oracle/cases.py: myAmatrix and harmonic_extension are the reference's own code, only the five operators are ours."""
    from oracle.cases import synthetic_condensation
    c = Case(mesh)
    cd = synthetic_condensation(c.A, c.free_u)
    dS = cd["mat"].diagonal()
    g_ = c.free_u & ~cd["interior"] & (dS != 0)
    dAc = np.where(g_, 1.0 / np.where(dS != 0, dS, 1.0), 0.0)
    Sg = cd["mat"][g_][:, g_].toarray()
    lam = 0.9 * float(np.linalg.eigvalsh(np.sqrt(dAc[g_])[:, None] * Sg * np.sqrt(dAc[g_])[None, :])[0])

    class Blf:
        pass
    blfA, blfB = Blf(), Blf()
    blfA.condense = True
    blfA.mat = SparseMatrix(cd["mat"])
    blfA.harmonic_extension = SparseMatrix(cd["H"])
    blfA.harmonic_extension_trans = SparseMatrix(cd["HT"])
    blfA.inner_solve = SparseMatrix(cd["inner_solve"])
    blfA.inner_matrix = SparseMatrix(cd["inner_matrix"])
    blfB.mat, blfB.condense = SparseMatrix(c.B), False
    sol = BlockVector([BaseVector(c.u0.copy()), BaseVector(np.zeros(c.n_p))])
    ngla.inject_eigenvalues([lam, 1.0])
    (it, secs), out = quiet(ref_bpcg2.BramblePasciakCG, blfA, blfB, None, BaseVector(c.f.copy()), BaseVector(c.g.copy()),
                            DiagonalMatrix(dAc), DiagonalMatrix(c.dM), sol, initialize=False, tol=tol,
                            maxsteps=maxsteps, rel_err=True)
    hist = [float(m.group(2)) for m in re.finditer(r"it =\s+(\d+)\s+err =\s+(\S+)", out)]
    err0 = float(re.search(r"err0 (\S+)", out).group(1))
    name = "bpcg_condensed_%s" % mesh
    np.savez_compressed(os.path.join(GOLDEN_DIR, name), lam_min=lam, it_v2=it, hist_v2=np.array(hist), err0_v2=err0,
                        sol_v2=sol.NumPy().copy(), dAc=dAc, fingerprint=c.fingerprint(), args=np.array([tol, maxsteps]))
    print(name, "it", it, "hist", len(hist), "lam_min", lam)


def gold_heat():
    """Runs /root/reference/heat.py as it is (module body included: the sweep over 7 time steps of :150-163) over the
    FEM stand-in of ngs_standin/ngsolve/fem.py with the matrices of oracle.cases.heat_case, and records the temperature
    vector after every pass of the time loop (heat.py calls Redraw() once per pass, :146)."""
    import builtins
    import runpy
    import tempfile
    import pandas as pd
    from ngsolve import fem
    from oracle.cases import heat_case
    K, M, free, nodes, u0 = heat_case()
    fem.install(K, M, free, nodes)
    runs = []                       # one list of states per solve() call

    def on_redraw():
        gf = fem.gridfunctions[-1]
        if not runs or runs[-1][0] is not gf:
            runs.append((gf, []))
        runs[-1][1].append(gf.vec.NumPy().copy())
    fem.redraw_hooks[:] = [on_redraw]
    cwd, old_input = os.getcwd(), builtins.input
    sys.path.insert(0, REF)         # heat.py:7-8 import its sibling modules by bare name
    try:
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            builtins.input = lambda *a: ""                     # heat.py:165
            _, out = quiet(runpy.run_path, os.path.join(REF, "heat.py"), run_name="ref_heat")
            errors = pd.read_csv("heat_errors.csv")
    finally:
        os.chdir(cwd)
        builtins.input = old_input
        sys.path.remove(REF)
        fem.redraw_hooks[:] = []
    out = dict(fingerprint=np.array([K.nnz, float(abs(K).sum()), float(abs(M).sum()), float(np.abs(u0).sum())]),
               time_steps=errors["time_step"].to_numpy(), errors=errors["error"].to_numpy(),
               n_steps=np.array([len(st) for _, st in runs]))
    assert len(runs) == len(out["time_steps"]) == 7
    for k, (_, st) in enumerate(runs):
        out["final_%d" % k] = st[-1]
        out["first_%d" % k] = np.stack(st[:3])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "heat_sweep"), **out)
    print("heat_sweep steps", out["n_steps"], "errors", out["errors"])


def gold_simple():
    """DoTimeStep / Project of templates/NavierStokesSIMPLE_iterative.py:427-443, the reference's own method bodies,
    called unbound on a namespace that carries the attributes they read (operators of oracle.cases.simple_case composed
    exactly as :93-98 and :128-131 compose them, with the commented-out sparse direct inverses of :90,127 as inv1)."""
    import types
    import scipy.sparse as sp
    from ngsolve import IdentityMatrix, PermutationMatrix
    from ngsolve.fem import _DirectInverse
    from oracle.cases import simple_case, synthetic_condensation
    sys.path.insert(0, REF)          # :7 imports solvers.bramblepasciak_new relative to the repository root
    try:
        ref_simple = load_ref("templates/NavierStokesSIMPLE_iterative.py", "ref_simple")
    finally:
        sys.path.remove(REF)
    NS = ref_simple.NavierStokes
    T = simple_case()
    n_u, n_p, N, dt = T["n_u"], T["n_p"], T["N"], T["dt"]
    conv = (0.3 * T["mstar"] - 0.3 * dt * T["A"]).tocsr()      # any linear operator in place of conv_operator: 0.3 M
    out = dict(conv_scale=0.3, fingerprint=np.array([T["A"].nnz, float(abs(T["A"]).sum()), float(abs(T["mstar"]).sum()),
                                                     float(abs(T["bproj"]).sum()), float(T["perm"][:7].sum())]))
    for tag, condense in (("plain", False), ("cond", True)):
        if condense:
            cd = synthetic_condensation(T["mstar"], T["free"])
            inv1 = _DirectInverse(cd["mat"], T["free"] & ~cd["interior"])
            ext = IdentityMatrix(n_u) + SparseMatrix(cd["H"])                    # :94-96
            extT = IdentityMatrix(n_u) + SparseMatrix(cd["HT"])
            invmstar = ext @ inv1 @ extT + SparseMatrix(cd["inner_solve"])
        else:
            invmstar = _DirectInverse(T["mstar"], T["free"])
        invproj1 = _DirectInverse(T["S"], T["dS"] > 0)
        ext = IdentityMatrix() + SparseMatrix(T["H"])                            # :129-131
        extT = IdentityMatrix() + SparseMatrix(T["HT"])
        invproj = ext @ invproj1 @ extT + SparseMatrix(T["inner"])
        me = types.SimpleNamespace()
        me.a = types.SimpleNamespace(mat=SparseMatrix(T["A"]))
        me.f = types.SimpleNamespace(vec=BaseVector(T["f"].copy()), Assemble=lambda: None)
        me.conv_operator = SparseMatrix(conv)
        me.gfu = types.SimpleNamespace(vec=BaseVector(T["u0"].copy()))
        me.gfup = types.SimpleNamespace(vec=BaseVector(n_p))
        me.invmstar, me.invproj = invmstar, invproj
        me.bproj = types.SimpleNamespace(mat=SparseMatrix(T["bproj"]))
        me.tmp_projection = BaseVector(N)
        me.mapV = PermutationMatrix(N, T["perm"])
        me.V, me.V2, me.Q = (types.SimpleNamespace(ndof=n) for n in (n_u, n_u, n_p))
        me.timestep = dt
        me.Project = lambda vel, me=me: NS.Project(me, vel)
        us, ps = [], []
        for _ in range(3):
            NS.DoTimeStep(me)
            us.append(me.gfu.vec.NumPy().copy())
            ps.append(me.gfup.vec.NumPy().copy())
        out["gfu_" + tag], out["gfup_" + tag] = np.stack(us), np.stack(ps)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "simple_steps"), **out)
    print("simple_steps |gfu|", [float(np.linalg.norm(u)) for u in out["gfu_plain"]],
          "plain vs cond", float(np.linalg.norm(out["gfu_plain"][-1] - out["gfu_cond"][-1])))


def gold_mypreA():
    """MypreA.Mult (templates/NavierStokesSIMPLE_iterative.py:364-391): the class is defined inside SolveInitial, so its source
    is cut out of the reference file with ast and executed, unmodified, over the stand-in with `transform` and `preAh1` in scope
    (as they are in the enclosing function).  Both variants (GS True / False), two inputs each."""
    import ast
    import textwrap
    from navier_stokes_solver_b200.la import color_blocks
    from oracle.cases import aggregation_transform, dof_blocks
    path = os.path.join(REF, "templates", "NavierStokesSIMPLE_iterative.py")
    src = open(path).read()
    node = next(n for n in ast.walk(ast.parse(src)) if isinstance(n, ast.ClassDef) and n.name == "MypreA")
    lines = src.splitlines()[node.lineno - 1:node.end_lineno]
    code = textwrap.dedent("\n".join(lines))
    c = Case("chan2d")
    T, Ac = aggregation_transform(c)
    dc = 1.0 / Ac.diagonal()
    blocks = dof_blocks(c, 4)
    order, color_ptr = color_blocks(c.A, blocks)          # the device smoother sweeps colour by colour: fix that block order
    blocks = [blocks[i] for i in order]
    ns = {"BaseMatrix": ngsolve.BaseMatrix, "transform": SparseMatrix(T), "preAh1": DiagonalMatrix(dc)}
    exec(compile(code, path + ":MypreA", "exec"), ns)
    space = type("Space", (), {"ndof": c.n_u})()
    blf = type("Blf", (), {"mat": SparseMatrix(c.A), "space": space})()
    rng = np.random.default_rng(11)
    xs = np.stack([rng.uniform(-1, 1, c.n_u) * c.free_u for _ in range(2)])
    out = dict(x=xs, block_ptr=np.concatenate([[0], np.cumsum([len(b) for b in blocks])]), block_dofs=np.concatenate(blocks),
               fingerprint=np.append(c.fingerprint(), [T.nnz, float(dc.sum())]), source_lines=np.array([node.lineno, node.end_lineno]))
    for GS in (True, False):
        pre = ns["MypreA"](space, blf, blocks, GS=GS)
        assert pre.Height() == c.n_u and pre.Width() == c.n_u
        ys = []
        for x in xs:
            y = BaseVector(np.full(c.n_u, 7.0))           # Mult must overwrite, not accumulate
            pre.Mult(BaseVector(x.copy()), y)
            ys.append(y.NumPy().copy())
        out["y_gs" if GS else "y_add"] = np.stack(ys)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "mypreA"), **out)
    print("mypreA lines", node.lineno, node.end_lineno, "|y_gs|", float(np.linalg.norm(out["y_gs"])), "|y_add|", float(np.linalg.norm(out["y_add"])))


def _reference_mypreA(c, GS):
    """An instance of the reference's MypreA class (source cut out of the reference file, see gold_mypreA) on Case `c` with the
    stand-in blocks / transform / coarse preconditioner; returns (instance, blocks in the colour-major order the device sweeps)."""
    import ast
    import textwrap
    from navier_stokes_solver_b200.la import color_blocks
    from oracle.cases import aggregation_transform, dof_blocks
    path = os.path.join(REF, "templates", "NavierStokesSIMPLE_iterative.py")
    src = open(path).read()
    node = next(n for n in ast.walk(ast.parse(src)) if isinstance(n, ast.ClassDef) and n.name == "MypreA")
    code = textwrap.dedent("\n".join(src.splitlines()[node.lineno - 1:node.end_lineno]))
    T, Ac = aggregation_transform(c)
    dc = 1.0 / Ac.diagonal()
    blocks = dof_blocks(c, 4)
    order, _ = color_blocks(c.A, blocks)
    blocks = [blocks[i] for i in order]
    ns = {"BaseMatrix": ngsolve.BaseMatrix, "transform": SparseMatrix(T), "preAh1": DiagonalMatrix(dc)}
    exec(compile(code, path + ":MypreA", "exec"), ns)
    space = type("Space", (), {"ndof": c.n_u})()
    blf = type("Blf", (), {"mat": SparseMatrix(c.A), "space": space})()
    return ns["MypreA"](space, blf, blocks, GS=GS), blocks, T, dc


def gold_solvers_with_mypreA():
    """The reference's solver loops with the reference's own MypreA as the velocity preconditioner, all unmodified:
    BramblePasciakCG(..., preA, preM, ...) exactly as called at templates/NavierStokesSIMPLE_iterative.py:391-397, and MinRes with
    BlockMatrix([[preA, None], [None, preM]]) (the run.py:46,52 pattern).  lambda_min of preA*A is computed densely (1274 free+fixed
    velocity dofs) and injected with the factor 0.5 the device test uses."""
    c = Case("chan2d")
    out = dict(fingerprint=c.fingerprint())
    for GS in (True, False):
        tag = "gs" if GS else "add"
        preA, blocks, T, dc = _reference_mypreA(c, GS)
        n = c.n_u
        P = np.zeros((n, n))
        for j in range(n):
            e = np.zeros(n)
            e[j] = 1.0
            y = BaseVector(np.zeros(n))
            preA.Mult(BaseVector(e), y)
            P[:, j] = y.NumPy()
        lam_all = np.linalg.eigvals(P @ c.A.toarray()).real
        lam = 0.5 * float(lam_all[np.abs(lam_all) > 1e-10 * np.abs(lam_all).max()].min())
        A, B = SparseMatrix(c.A), SparseMatrix(c.B)

        class Blf:
            def __init__(self, m):
                self.mat = m
                self.condense = False
        sol = BlockVector([BaseVector(c.u0.copy()), BaseVector(np.zeros(c.n_p))])
        ngla.inject_eigenvalues([lam, 1.0])
        (it, secs), txt = quiet(ref_bpcg2.BramblePasciakCG, Blf(A), Blf(B), None, BaseVector(c.f.copy()), BaseVector(c.g.copy()), preA,
                                DiagonalMatrix(c.dM), sol, initialize=False, tol=1e-10, maxsteps=20000, rel_err=True)
        hist = [float(m.group(2)) for m in re.finditer(r"it =\s+(\d+)\s+err =\s+(\S+)", txt)]
        out.update({"lam_" + tag: lam, "bpcg_it_" + tag: it, "bpcg_hist_" + tag: np.array(hist), "bpcg_sol_" + tag: sol.NumPy().copy(),
                    "bpcg_err0_" + tag: float(re.search(r"err0 (\S+)", txt).group(1))})
        # MinRes on K with the block-diagonal preconditioner
        K = BlockMatrix([[A, B.T], [B, None]])
        C = BlockMatrix([[preA, None], [None, DiagonalMatrix(c.dM)]])
        msol = BlockVector([BaseVector(c.u0.copy()), BaseVector(np.zeros(c.n_p))])
        (u, errors), _ = quiet(ref_minres.MinRes, mat=K, pre=C, rhs=BlockVector([BaseVector(c.f.copy()), BaseVector(c.g.copy())]),
                               sol=msol, initialize=False, tol=1e-8, maxsteps=5000)
        out.update({"minres_errors_" + tag: np.array(errors), "minres_sol_" + tag: msol.NumPy().copy()})
        print("solvers_mypreA", tag, "lam", lam, "bpcg it", it, "minres iterations", len(errors) - 1, "final", errors[-1])
    out["block_ptr"] = np.concatenate([[0], np.cumsum([len(b) for b in blocks])])
    out["block_dofs"] = np.concatenate(blocks)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "solvers_mypreA"), **out)


def gold_orth():
    rng = np.random.default_rng(7)
    n, m = 777, 5
    base = rng.standard_normal((m, n))
    base[3] = base[0] * 0.5 + 1e-3 * base[3]      # nearly dependent column: exercises the 3 sweeps
    basis = [BaseVector(base[i].copy()) for i in range(m)]
    out = ref_orth.orthonormalize(basis)
    assert out is basis
    Q = np.stack([b.NumPy() for b in basis])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "orthonormalize"), input=base, Q=Q)
    print("orthonormalize", np.abs(Q @ Q.T - np.eye(m)).max())


def gold_rk():
    out = {}
    for deg in (3, 10):
        w = ref_rk.ImplicitRungeKuttaMethodWeights(deg)                  # heat.py:76 uses 10
        out["a%d" % deg] = w.a.NumPy().copy()
        out["b%d" % deg] = w.b.NumPy().copy()
        out["c%d" % deg] = np.array(w.c)
        rng = np.random.default_rng(deg)
        Msym = rng.standard_normal((5, 5))
        evo = -(Msym @ Msym.T) - np.eye(5)
        y = rng.standard_normal(5)
        nxt = ref_rk.linear_implicit_runge_kutta_step(w, ngsolve.Matrix(evo.copy()), ngsolve._DVec(y.copy()), 0.01)
        out["evo%d" % deg] = evo
        out["y%d" % deg] = y
        out["next%d" % deg] = nxt.NumPy().copy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "runge_kutta"), **out)
    print("runge_kutta ok")


if __name__ == "__main__":
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    gold_minres("sq8", initialize=False, tol=1e-7, maxsteps=10000)
    gold_minres("sq8", initialize=True, tol=1e-7, maxsteps=10000, tag="_init")
    gold_minres("sq8", initialize=False, tol=1e-7, maxsteps=40, tag="_cap40")          # non-convergence exit
    gold_minres("sq8", initialize=True, tol=1e-3, maxsteps=10000, use_pre=False, tag="_nopre")
    gold_minres("sq20", initialize=False, tol=1e-7, maxsteps=10000)                    # config 1
    gold_minres("chan2d", initialize=False, tol=1e-7, maxsteps=10000)
    gold_minres("chan3d", initialize=False, tol=1e-7, maxsteps=10000)
    gold_bpcg("sq8", tol=1e-10, maxsteps=10000)
    gold_bpcg("sq8", tol=1e-12, maxsteps=10000, tag="_floor")   # tolerance at the rounding floor: count is chaotic
    gold_bpcg("chan2d", tol=1e-10, maxsteps=10000)
    gold_bpcg("chan3d", tol=1e-10, maxsteps=100000)
    gold_bpcg("sq8", tol=1e-12, maxsteps=25, tag="_cap25")
    gold_bpcg_condensed("chan2d", tol=1e-10, maxsteps=20000)
    gold_orth()
    gold_rk()
    gold_heat()
    gold_simple()
    gold_mypreA()
    gold_solvers_with_mypreA()
