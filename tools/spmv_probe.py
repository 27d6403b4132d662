"""SpMV micro-benchmark on the 3D Stokes blocks: device-timed GB/s of A, B, B^T
and the merged K.  Also the command profiled under ncu (profiles/)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import navier_stokes_solver_b200 as nsb  # noqa: E402
import navier_stokes_solver_b200.synthetic as syn  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--dofs", type=float, default=2e6)
ap.add_argument("--reps", type=int, default=50)
ap.add_argument("--warmup", type=int, default=10)
ap.add_argument("--which", default="A,B,BT,K")
args = ap.parse_args()
cells = syn.stokes_cells_for_dofs(args.dofs)
S = syn.assemble_stokes(cells, (2.5, 0.41, 0.41), device="cuda")
ctx = nsb.default_context()
mats = {"A": nsb.SparseMatrix.from_csr(S.A), "B": nsb.SparseMatrix.from_csr(S.B), "BT": nsb.SparseMatrix.from_csr(S.BT)}
if "K" in args.which:
    mats["K"] = nsb.BlockMatrix([[mats["A"], mats["BT"]], [mats["B"], None]]).saddle()
peak = 6550.1
out = {"cells": cells, "n_u": S.n_u, "n_p": S.n_p}
for name in args.which.split(","):
    m = mats[name]
    ms = m.spmv_bench(args.warmup, args.reps)
    gbs = m.spmv_bytes() / ms / 1e6
    out[name + "_format"] = m.format()
    out[name] = {"rows": m.height, "nnz": m.nnz, "nnz_per_row": m.nnz / m.height, "ms": ms, "GBs": gbs, "frac_measured": gbs / peak}
print(json.dumps(out))
