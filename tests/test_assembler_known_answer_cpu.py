"""Known answers for the synthetic Taylor-Hood assembler (navier_stokes_solver_b200.synthetic.assemble_stokes): the inputs of every
parity case and of the bench come from it, so it is checked against exact solutions that P2/P1 represents exactly."""
import numpy as np
import pytest

import navier_stokes_solver_b200.synthetic as syn


def _coords(S):
    d = S.dim
    dims2 = [2 * c + 1 for c in S.cells]
    idx = np.stack(np.unravel_index(np.arange(S.n2), dims2), axis=1)
    return idx * (np.array(S.lengths) / (np.array(dims2) - 1))


@pytest.mark.parametrize("cells,lengths", [((12, 5), (2.0, 0.41)), ((7, 4), (1.0, 1.0))])
def test_poiseuille_flow_solves_the_discrete_stokes_system(cells, lengths):
    """Plane Poiseuille flow u = (U 4 y (H - y) / H^2, 0), p = 8 nu U / H^2 (x - L) (sign convention of run.py:77-81: a(u,v) + b(v,p) = f(v),
    natural outflow p = 0 at x = L) is in the discrete spaces, so it satisfies the assembled equations on the free dofs to rounding."""
    S = syn.assemble_stokes(cells, lengths)
    A, B, BT = S.A.to_scipy(), S.B.to_scipy(), S.BT.to_scipy()
    X2 = _coords(S)
    L, H, U = lengths[0], lengths[1], 1.5
    ux = U * 4.0 * X2[:, 1] * (H - X2[:, 1]) / H ** 2
    u = np.concatenate([ux, np.zeros(S.n2)])
    dims1 = [c + 1 for c in cells]
    idx1 = np.stack(np.unravel_index(np.arange(S.n1), dims1), axis=1) * (np.array(lengths) / np.array(cells))
    p = 8.0 * U / H ** 2 * (idx1[:, 0] - L)
    free = S.free_u.numpy()
    res = A @ u + BT @ p
    scale = np.abs(A @ u).max()
    assert np.abs(res[free]).max() < 1e-10 * scale
    assert np.abs(B @ u).max() < 1e-12 * np.abs(B).sum(axis=1).max() * U
    # the inlet data of the assembler is this profile, the walls are no-slip
    inlet = (~free) & (np.concatenate([X2[:, 0], X2[:, 0]]) == 0.0)
    np.testing.assert_allclose(S.u0.numpy()[inlet], u[inlet], rtol=1e-13, atol=1e-15)
    # B^T is the transpose of B; A is symmetric; the pressure mass matrix integrates 1 to the area
    assert abs(BT - B.T).max() < 1e-15
    assert abs(A - A.T).max() < 1e-13 * abs(A).max()
    Mp = S.Mp.to_scipy()
    assert abs(Mp.sum() - L * H) < 1e-12 * L * H


def test_3d_blocks_reproduce_polynomial_fields():
    """3D: grad-grad form of a quadratic field against a linear test function, divergence of a quadratic field, masses = volume."""
    cells, lengths = (4, 3, 2), (2.5, 0.41, 0.41)
    S = syn.assemble_stokes(cells, lengths, with_mass2=True)
    A, B = S.A.to_scipy(), S.B.to_scipy()
    X2 = _coords(S)
    x, y, z = X2.T
    vol = float(np.prod(lengths))
    # u = (x^2, -2 x y + z^2, 0): div u = 0 pointwise -> B u = 0
    u = np.concatenate([x * x, -2 * x * y + z * z, np.zeros(S.n2)])
    assert np.abs(B @ u).max() < 1e-12 * np.abs(B).sum(axis=1).max() * np.abs(u).max()
    # u = (x, 2 y, 3 z): div u = 6 -> 1^T B u = 6 |Omega| (partition of unity of the P1 basis)
    u = np.concatenate([x, 2 * y, 3 * z])
    assert abs((B @ u).sum() - 6.0 * vol) < 1e-11 * vol
    # a(u, v) with u_0 = x^2, v_0 = x: int grad(x^2).grad(x) = int 2x = L_x^2 L_y L_z
    u = np.concatenate([x * x, np.zeros(2 * S.n2)])
    v = np.concatenate([x, np.zeros(2 * S.n2)])
    assert abs(v @ (A @ u) - lengths[0] ** 2 * lengths[1] * lengths[2]) < 1e-11 * vol
    M2 = S.extras["M2"].to_scipy()
    assert abs(M2.sum() - vol) < 1e-12 * vol
    assert abs(x @ (M2 @ x) - lengths[0] ** 3 / 3 * lengths[1] * lengths[2]) < 1e-11 * vol
