#!/usr/bin/env python
"""Secondary measurements: BASELINE.json configs[0..3] (the small / latency-bound cases) on one GPU, each with
time-to-tolerance, iterations, iterations/s and the same run of the C+OpenMP oracle on the host cores.
These are parity-test sized; the headline bench line stays bench.py (configs[4]).

    python tools/bench_configs.py            -> one JSON object per config on stdout
"""
import contextlib
import io
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import navier_stokes_solver_b200 as nsb  # noqa: E402
import navier_stokes_solver_b200.synthetic as syn  # noqa: E402
from oracle import cbuild, krylov_oracle as ko  # noqa: E402


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def stokes(cells, lengths):
    S = syn.assemble_stokes(cells, lengths)
    A, B, BT, Mp = S.A.to_scipy(), S.B.to_scipy(), S.BT.to_scipy(), S.Mp.to_scipy()
    dA = syn.jacobi_inverse(S.A, free=S.free_u).numpy()
    dM = syn.jacobi_inverse(S.Mp).numpy()
    start = np.concatenate([S.u0.numpy(), np.zeros(S.n_p)])
    return S, A, B, BT, dA, dM, start


def best_of(fn, n=3):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        out = fn()
        ts.append(time.perf_counter() - t0)
    return min(ts), out


def config1():
    """2D Stokes, unit square h = 0.05, Taylor-Hood, MinRes + block-diagonal Jacobi, tol 1e-7 (run.py:292-295)."""
    S, A, B, BT, dA, dM, start = stokes((20, 20), (1.0, 1.0))
    dA_, dB_ = nsb.SparseMatrix.from_scipy(A), nsb.SparseMatrix.from_scipy(B)
    K = nsb.BlockMatrix([[dA_, dB_.T], [dB_, None]])
    Cp = nsb.BlockMatrix([[nsb.DiagonalPreconditioner.from_diagonal(dA), None], [None, nsb.DiagonalPreconditioner.from_diagonal(dM)]])
    rhs = np.concatenate([S.f.numpy(), S.g.numpy()])
    K.saddle(); Cp.blockdiag_preconditioner()

    def dev():
        sol = start.copy()
        return quiet(nsb.MinRes, K, rhs, Cp, sol, 10000, False, False, 1e-7)
    t_dev, (_, errs) = best_of(dev)
    loop_dev = nsb.solvers.last_info["seconds"]
    bl = cbuild.Blocks(A, B, BT)

    def cpu():
        sol = start.copy()
        return cbuild.minres(bl, np.concatenate([dA, dM]), rhs, sol, tol=1e-7, maxsteps=10000, initialize=False)
    t_cpu, (e_cpu, _, loop_cpu) = best_of(cpu)
    return dict(config="1: 2D Stokes h=0.05 P2/P1, MinRes + Jacobi, tol 1e-7", dofs=int(S.n_u + S.n_p), iterations=len(errs) - 1,
                device_call_s=t_dev, device_loop_s=loop_dev, device_its_per_s=(len(errs) - 1) / loop_dev,
                cpu_iterations=len(e_cpu) - 1, cpu_loop_s=loop_cpu, cpu_threads=cbuild.num_threads())


def config2():
    """2D heat, time loop of heat.py with 5-dimensional subspace, 10 Gauss stages, Jacobi-CG to 1e-12."""
    cells = (20, 20)
    S = syn.assemble_stokes(cells, (1.0, 1.0), with_mass2=True)
    K, M = S.extras["K2"].to_scipy(), S.extras["M2"].to_scipy()
    nx, ny = 2 * cells[0] + 1, 2 * cells[1] + 1
    ix, iy = np.divmod(np.arange(nx * ny), ny)
    free = (ix > 0) & (ix < nx - 1) & (iy > 0) & (iy < ny - 1)
    x, y = ix / (nx - 1), iy / (ny - 1)
    kl = [(1, 1), (2, 1), (1, 3), (3, 3), (2, 3), (4, 5), (5, 2)]
    u0 = sum(2 * np.sin(k * np.pi * x) * np.sin(l * np.pi * y) for k, l in kl) * free
    exact = lambda t: sum(2 * np.exp(-(k * k + l * l) * np.pi ** 2 * t) * np.sin(k * np.pi * x) * np.sin(l * np.pi * y) for k, l in kl)
    out = []
    w = nsb.ImplicitRungeKuttaMethodWeights(10)
    for time_step in (1e-2, 1e-3):
        m, end = 5, 0.05
        H = (M + time_step * K).tocsr()                       # heat.py:51-52
        dH = np.where(free, 1.0 / H.diagonal(), 0.0)
        dK, dM_, dHm = (nsb.SparseMatrix.from_scipy(a) for a in (K, M, H))
        pre = nsb.DiagonalPreconditioner.from_diagonal(dH)
        u = u0.copy()
        nsteps = int(round(end / time_step))
        t0 = time.perf_counter()
        cg = 0
        for _ in range(nsteps):
            nsb.heat_large_step(dK, dM_, dHm, pre, u, time_step, w, m, cg_tol=1e-12)
            cg += nsb.solvers.last_info["cg_iterations"]
        t = time.perf_counter() - t0
        ex = exact(nsteps * time_step) * free
        out.append(dict(time_step=time_step, steps=nsteps, seconds=t, cg_iterations=cg,
                        rel_l2_error_vs_analytic=float(np.sqrt((u - ex) @ (M @ (u - ex)) / (ex @ (M @ ex))))))
    return dict(config="2: 2D heat, subspace 5 x Gauss 10 (heat.py), Jacobi-CG 1e-12", dofs=int(K.shape[0]), runs=out)


def config3():
    """2D channel saddle point, Bramble-Pasciak CG v1 signature, tolerance 1e-12, max_steps 10000 (stokes_hcurldiv.py:76-77);
    Taylor-Hood blocks stand in for the MCS H(div) spaces."""
    S, A, B, BT, dA, dM, start = stokes((48, 12), (2.0, 0.41))
    dA_, dB_ = nsb.SparseMatrix.from_scipy(A), nsb.SparseMatrix.from_scipy(B)
    pA, pM = nsb.DiagonalPreconditioner.from_diagonal(dA), nsb.DiagonalPreconditioner.from_diagonal(dM)
    lam = min(quiet(nsb.EigenValues_Preconditioner, dA_, pA, tol=1e-8, maxsteps=3000))

    def dev():
        sol = start.copy()
        return quiet(nsb.bramble_pasciak_cg, dA_, dB_, None, pA, pM, S.f.numpy(), S.g.numpy(), sol, 1e-12, 10000, False, lam_min=0.98 * lam)
    t_dev, (_, errs) = best_of(dev)
    loop_dev = nsb.solvers.last_info["seconds"]
    bl = cbuild.Blocks(A, B, BT)

    def cpu():
        sol = start.copy()
        return cbuild.bpcg_v2(bl, S.f.numpy(), S.g.numpy(), dA, dM, ko.bpcg_scale(0.98 * lam), sol, tol=1e-12, maxsteps=10000, initialize=False)
    t_cpu, (it_cpu, h_cpu, _, loop_cpu) = best_of(cpu)
    return dict(config="3: 2D channel saddle point, Bramble-Pasciak CG, tol 1e-12", dofs=int(S.n_u + S.n_p), iterations=len(errs) - 1,
                device_call_s=t_dev, device_loop_s=loop_dev, device_its_per_s=(len(errs) - 1) / loop_dev,
                cpu_iterations=int(it_cpu) + 1, cpu_loop_s=loop_cpu, cpu_threads=cbuild.num_threads())


def config4():
    """200 explicit-convection / implicit-Stokes splitting steps (DoTimeStep + Project, NavierStokesSIMPLE_iterative.py:427-443)
    on a 2D channel, device-resident: per step one M* CG solve (precision 1e-4, :93), one projection CG solve (:128), 3 SpMVs."""
    import importlib.util
    import torch
    spec = importlib.util.spec_from_file_location("tgp", os.path.join(ROOT, "tests", "test_gpu_parity.py"))
    tgp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tgp)
    T = tgp._simple_case(mesh_cells=(80, 20), lengths=(2.0, 0.41), dt=1e-3)
    n_u, n_p, dt = T["n_u"], T["n_p"], T["dt"]
    up = nsb.SparseMatrix.from_scipy
    dm = np.where(T["free"], 1.0 / T["mstar"].diagonal(), 0.0)
    inv1 = nsb.CGSolver(up(T["mstar"]), nsb.DiagonalPreconditioner.from_diagonal(dm), precision=1e-4, maxsteps=2000)
    invp1 = nsb.CGSolver(up(T["S"]), nsb.DiagonalPreconditioner.from_diagonal(T["dS"]), precision=1e-8, maxsteps=4000)
    invproj = nsb.ExtendedInverse(invp1, up(T["H"]), up(T["HT"]), up(T["inner"]))
    gfu = torch.as_tensor(T["u0"].copy(), device="cuda")
    st = nsb.SimpleStepper(up(T["A"]), nsb.ExtendedInverse(inv1), up(T["bproj"]), invproj, nsb.PermutationMatrix(T["N"], T["perm"]),
                           gfu, n_u, (n_u, n_u + n_p), dt, f=torch.as_tensor(T["f"], device="cuda"))
    st.DoTimeStep()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    its = [0, 0]
    for _ in range(200):
        st.DoTimeStep()
        its[0] += inv1.iterations
        its[1] += invp1.iterations
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    div = T["Bf"] @ (gfu.cpu().numpy() - T["u0"])
    return dict(config="4: 200 splitting steps (DoTimeStep/Project), 2D channel, CG 1e-4 / 1e-8", dofs=int(n_u + n_p), seconds=t,
                ms_per_step=1e3 * t / 200, mstar_cg_iterations=its[0], projection_cg_iterations=its[1],
                max_abs_divergence_of_increment=float(np.abs(div).max()))


def main():
    for fn in (config1, config2, config3, config4):
        print(json.dumps(fn()))


if __name__ == "__main__":
    main()
