"""Row f4 on the GPU: operators and vectors exported through the NGSolve adaptor (accessor-compatible stand-in: NGSolve is absent)
feed the device solvers, and the results are held to the reference-generated goldens -- run.py:44-56 (MinRes call site) and
templates/NavierStokesSIMPLE_iterative.py:188,397 (condensed BramblePasciakCG call site)."""
import contextlib
import io

import numpy as np
import pytest

from test_adaptor_cpu import FakeNgsMatrix, FakeSpace, FakeVec

pytestmark = pytest.mark.gpu


def _quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def test_minres_on_adaptor_exports_matches_reference_golden(cases):
    import navier_stokes_solver_b200 as nsb
    from navier_stokes_solver_b200 import adaptor
    from oracle.cases import load_golden
    gold = load_golden("minres_sq8")
    c = cases("sq8")
    # what run.py:77-90,105 holds after Assemble(): a.mat, b.mat, mp.mat, f.vec, g.vec, gfu.vec, gfp.vec, V.FreeDofs()
    blf_a = type("Blf", (), {"mat": FakeNgsMatrix(c.A), "condense": False})()
    blf_b = type("Blf", (), {"mat": FakeNgsMatrix(c.B), "condense": False})()
    blf_m = type("Blf", (), {"mat": FakeNgsMatrix(c.Mp), "condense": False})()
    A, B, Mp = (adaptor.export_bilinear_form(x) for x in (blf_a, blf_b, blf_m))
    preA = adaptor.jacobi_from_space(A, FakeSpace([bool(b) for b in c.free_u]))        # Preconditioner(a, 'local') on the free dofs
    preS = adaptor.jacobi_from_space(Mp)
    gfu, gfp = FakeVec(c.u0.copy()), FakeVec(np.zeros(c.n_p))
    rhs = adaptor.block_vector(FakeVec(c.f.copy()), FakeVec(c.g.copy()))
    sol = adaptor.block_vector(gfu, gfp)
    K = nsb.BlockMatrix([[A, B.T], [B, None]])                                          # run.py:45-46
    C = nsb.BlockMatrix([[preA, None], [None, preS]])
    u, errors = _quiet(nsb.MinRes, mat=K, pre=C, rhs=rhs, sol=sol, initialize=False, tol=1e-7, maxsteps=10000)   # run.py:52
    adaptor.sync_back(sol)
    assert abs(len(errors) - len(gold["errors"])) <= 1
    m = min(50, len(errors), len(gold["errors"]))
    np.testing.assert_allclose(errors[:m], gold["errors"][:m], rtol=1e-10)
    got = np.concatenate([gfu.FV().NumPy(), gfp.FV().NumPy()])                          # the caller reads the grid functions (run.py:107-109)
    # (the envelope of tests/test_gpu_parity.py: two roundings of the recurrence agree down to the level the solve reached)
    assert np.linalg.norm(got - gold["sol"]) / np.linalg.norm(gold["sol"]) < max(1e-8, 10 * float(gold["errors"][-1]))
    frame = adaptor.errors_frame(errors, nsb.solvers.last_info["seconds"], 0.125, "taylor_hood", 2, "minres", "device", c.N)
    assert list(frame["error"]) == list(errors) and frame["ndofs"].iloc[0] == c.N


def test_condensed_bpcg_on_adaptor_exports_matches_reference_golden(cases):
    import navier_stokes_solver_b200 as nsb
    from navier_stokes_solver_b200 import adaptor
    from oracle.cases import load_golden, synthetic_condensation
    gold = load_golden("bpcg_condensed_chan2d")
    c = cases("chan2d")
    cd = synthetic_condensation(c.A, c.free_u)
    blfA = type("Blf", (), dict(condense=True, mat=FakeNgsMatrix(cd["mat"]), harmonic_extension=FakeNgsMatrix(cd["H"]),
                                harmonic_extension_trans=FakeNgsMatrix(cd["HT"]), inner_solve=FakeNgsMatrix(cd["inner_solve"]),
                                inner_matrix=FakeNgsMatrix(cd["inner_matrix"])))()
    blfB = type("Blf", (), {"mat": FakeNgsMatrix(c.B), "condense": False})()
    ea = adaptor.export_bilinear_form(blfA)                                             # ExportedForm with the four extra operators
    eb = adaptor.export_bilinear_form(blfB)
    assert ea.condense and isinstance(eb, nsb.SparseMatrix)
    preA = nsb.DiagonalPreconditioner.from_diagonal(gold["dAc"])
    preM = nsb.DiagonalPreconditioner.from_diagonal(c.dM)
    gfu, gfp = FakeVec(c.u0.copy()), FakeVec(np.zeros(c.n_p))
    sol = adaptor.block_vector(gfu, gfp)
    it, secs = _quiet(nsb.BramblePasciakCG, ea, eb, None, c.f, c.g, preA, preM, sol, initialize=False, tol=1e-10, maxsteps=20000,
                      rel_err=True, lam_min=float(gold["lam_min"]))                     # :397
    adaptor.sync_back(sol)
    h = nsb.solvers.last_info["history"]
    m = min(50, len(h), len(gold["hist_v2"]))
    np.testing.assert_allclose(h[:m], gold["hist_v2"][:m], rtol=1e-10)
    got = np.concatenate([gfu.FV().NumPy(), gfp.FV().NumPy()])
    assert np.linalg.norm(got - gold["sol_v2"]) / np.linalg.norm(gold["sol_v2"]) < 1e-7
