"""The drop-in boundary of SURVEY.md 8(b) checked without a GPU: the Python mirrors keep the reference's signatures verbatim (names,
order, defaults -- a caller changes only the import line), and what cannot run on the device is refused with a clear exception
instead of being computed on the CPU (8(b) "Argument protocol": no CPU fallback)."""
import inspect
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# (name, parameters as the reference declares them) -- minres.py:12, bramble_pasciak_cg.py:65-67, solvers/bramblepasciak_new.py:24-25,
# orthonormalization.py:5, runge_kutta_method.py:10,26
REFERENCE_SIGNATURES = {
    "MinRes": [("mat", inspect.Parameter.empty), ("rhs", inspect.Parameter.empty), ("pre", None), ("sol", None), ("maxsteps", 100),
               ("printrates", True), ("initialize", True), ("tol", 1e-7)],
    "bramble_pasciak_cg": [("a_matrix", inspect.Parameter.empty), ("b_matrix", inspect.Parameter.empty), ("c_matrix", inspect.Parameter.empty),
                           ("pre_a", inspect.Parameter.empty), ("pre_schur_complement", inspect.Parameter.empty),
                           ("upper_rhs", inspect.Parameter.empty), ("lower_rhs", inspect.Parameter.empty), ("solution", None),
                           ("tolerance", 1e-12), ("max_steps", 1000), ("print_rates", True)],
    "BramblePasciakCG": [("blfA", inspect.Parameter.empty), ("blfB", inspect.Parameter.empty), ("matC", inspect.Parameter.empty),
                         ("f", inspect.Parameter.empty), ("g", inspect.Parameter.empty), ("preA_unscaled", inspect.Parameter.empty),
                         ("preM", inspect.Parameter.empty), ("sol", None), ("tol", 1e-6), ("maxsteps", 100), ("printrates", True),
                         ("initialize", True), ("rel_err", True)],
    "orthonormalize": [("basis", inspect.Parameter.empty), ("tries", 3)],
    "linear_implicit_runge_kutta_step": [("weights", inspect.Parameter.empty), ("matrix", inspect.Parameter.empty),
                                         ("current_value", inspect.Parameter.empty), ("step_width", inspect.Parameter.empty)],
}


@pytest.mark.parametrize("name", sorted(REFERENCE_SIGNATURES))
def test_mirror_keeps_the_reference_signature(name):
    import navier_stokes_solver_b200 as nsb
    got = [(p.name, p.default) for p in inspect.signature(getattr(nsb, name)).parameters.values()]
    want = REFERENCE_SIGNATURES[name]
    # the reference's parameters come first, in order, with its defaults; extra keywords (lam_min) may only follow, with defaults
    assert got[:len(want)] == want
    assert all(d is not inspect.Parameter.empty for _, d in got[len(want):])


def test_operators_that_cannot_run_on_the_device_are_refused():
    import navier_stokes_solver_b200 as nsb
    K = np.eye(4)
    rhs = np.ones(4)
    with pytest.raises(TypeError, match="device operator"):
        nsb.MinRes(K, rhs)                                          # a host matrix: no CPU fallback
    with pytest.raises(TypeError):
        nsb.CGSolve(K, rhs)
    with pytest.raises(TypeError):
        nsb.EigenValues_Preconditioner(K, None)
    with pytest.raises(TypeError):
        nsb.BramblePasciakCG(K, K, None, rhs, rhs, None, None, lam_min=1.0)
    with pytest.raises(NotImplementedError):                        # c_matrix is None at every reference call site
        nsb.bramble_pasciak_cg(K, K, K, None, None, rhs, rhs)
    with pytest.raises(NotImplementedError, match="NGSolve"):       # BDDC / h1amg live inside NGSolve
        nsb.Preconditioner(K, "bddc")
    with pytest.raises(TypeError):
        nsb.BlockSmoother(K, [[0, 1]])
    with pytest.raises(TypeError):
        nsb.CondensedOperator(K, K, K, K, K)


def test_vector_protocol_on_the_host():
    """BaseVector / BlockVector mirror the NGSolve protocol the reference uses at run.py:47-48,107-109: components are views of one
    flat (u | p) array, the caller's grid-function vectors see what the solver wrote."""
    import navier_stokes_solver_b200 as nsb
    gfu, gfp = nsb.BaseVector(np.arange(3.0)), nsb.BaseVector(np.array([7.0, 8.0]))
    sol = nsb.BlockVector([gfu, gfp])
    assert len(sol) == 5 and sol.nblocks == 2
    sol.NumPy()[:] = [1, 2, 3, 4, 5]
    assert gfu.NumPy().tolist() == [1, 2, 3] and gfp.NumPy().tolist() == [4, 5]
    w = sol.CreateVector()
    assert len(w) == 5 and w.nblocks == 2 and nsb.InnerProduct(sol, sol) == 55.0 and abs(nsb.Norm(sol) - 55.0 ** 0.5) < 1e-15
