#!/usr/bin/env python
"""The solve part of the reference's run.py (solve_with_min_res, :41-58; solve_with_bramble_pasciak_cg, :28-38) with the device
solvers dropped in: same calls, same arguments, same return values -- only the import line and where the matrices come from differ.
NGSolve is not on this box, so the assembled blocks come from the synthetic Taylor-Hood assembler (the channel of run.py:22-25 without
the cylinder); with NGSolve they come from adaptor.export_bilinear_form(a), ....

    python examples/stokes_channel.py [--cells 48 12] [--csv errors.csv]         (needs a B200; there is no CPU fallback)
"""
import argparse
import os
import sys
from time import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import navier_stokes_solver_b200 as nsb                                    # noqa: E402
import navier_stokes_solver_b200.synthetic as syn                          # noqa: E402
from navier_stokes_solver_b200 import adaptor                              # noqa: E402
from navier_stokes_solver_b200 import MinRes, bramble_pasciak_cg           # noqa: E402   (reference: from minres import MinRes ...)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=int, nargs="+", default=[48, 12])
    ap.add_argument("--csv", default=None)
    args = ap.parse_args()
    lengths = (2.0, 0.41) if len(args.cells) == 2 else (2.5, 0.41, 0.41)
    S = syn.assemble_stokes(tuple(args.cells), lengths)                    # stands where a.Assemble(), b.Assemble(), ... stand
    a_mat, b_mat = nsb.SparseMatrix.from_csr(S.A), nsb.SparseMatrix.from_csr(S.B)
    pre_a = nsb.Preconditioner(a_mat, "local", freedofs=S.free_u)          # Preconditioner(a, 'local') on V.FreeDofs()   (run.py:61-62)
    pre_s = nsb.DiagonalPreconditioner.from_diagonal(syn.jacobi_inverse(S.Mp))   # 'local' on the pressure mass matrix
    f, g = S.f.numpy(), S.g.numpy()
    frames = []

    # ---- solve_with_min_res (run.py:41-58)
    gfu, gfp = nsb.BaseVector(S.u0.numpy().copy()), nsb.BaseVector(np.zeros(S.n_p))
    K = nsb.BlockMatrix([[a_mat, b_mat.T], [b_mat, None]])                 # :45
    C = nsb.BlockMatrix([[pre_a, None], [None, pre_s]])                    # :46
    rhs = nsb.BlockVector([nsb.BaseVector(f.copy()), nsb.BaseVector(g.copy())])
    sol = nsb.BlockVector([gfu, gfp])
    t0 = time()
    u, errors = MinRes(mat=K, pre=C, rhs=rhs, sol=sol, initialize=False, tol=1e-7, maxsteps=10000, printrates=False)   # :52
    t_minres = time() - t0
    print("MinRes: %d iterations, %.3f s (device loop %.3f s), |u| = %.6e" % (len(errors) - 1, t_minres, nsb.solvers.last_info["seconds"],
                                                                              float(np.linalg.norm(gfu.NumPy()))))
    frames.append(adaptor.errors_frame(errors, t_minres, 2.0 / args.cells[0], "taylor_hood", 2, "minres", "device", S.n_u + S.n_p))

    # ---- solve_with_bramble_pasciak_cg (run.py:28-38)
    gfu2, gfp2 = nsb.BaseVector(S.u0.numpy().copy()), nsb.BaseVector(np.zeros(S.n_p))
    sol2 = nsb.BlockVector([gfu2, gfp2])
    t0 = time()
    sol2, errors2 = bramble_pasciak_cg(a_mat, b_mat, None, pre_a, pre_s, nsb.BaseVector(f.copy()), nsb.BaseVector(g.copy()),
                                       solution=sol2, tolerance=1e-7, max_steps=10000, print_rates=False)              # :36
    t_bpcg = time() - t0
    print("bramble_pasciak_cg: %d iterations, %.3f s, |u_minres - u_bpcg| / |u| = %.2e"
          % (len(errors2) - 1, t_bpcg, float(np.linalg.norm(gfu.NumPy() - gfu2.NumPy()) / np.linalg.norm(gfu.NumPy()))))
    frames.append(adaptor.errors_frame(errors2, t_bpcg, 2.0 / args.cells[0], "taylor_hood", 2, "bramble pasciak cg", "device", S.n_u + S.n_p))
    if args.csv:
        adaptor.write_csv(frames, args.csv)                                # run.py:261-262
        print("wrote", args.csv)


if __name__ == "__main__":
    main()
