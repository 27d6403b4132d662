"""Worker of tests/test_gpu_dist.py::test_dead_exchange_returns_an_error: rank 1 never publishes its halo entries
(NSB_FAULT_SKIP_PUBLISH=1); every rank's BramblePasciakCG call must come back with an error instead of spinning forever."""
import contextlib
import io
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import navier_stokes_solver_b200 as nsb  # noqa: E402
import navier_stokes_solver_b200.partition as P  # noqa: E402

CELLS, LENGTHS = (12, 3, 3), (2.5, 0.41, 0.41)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = nsb.Context.from_torch_distributed(device=local)
    D = P.build_distributed_stokes(CELLS, LENGTHS, ctx, rank, world, device="cuda:%d" % local)
    pA = nsb.DiagonalPreconditioner.from_diagonal(D.dA, ctx)
    pM = nsb.DiagonalPreconditioner.from_diagonal(D.dM, ctx)
    sol = D.u_start.clone()
    t0 = time.perf_counter()
    err = None
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            nsb.BramblePasciakCG(D.A, D.B, None, D.f, D.g, pA, pM, sol, tol=1e-10, maxsteps=400, initialize=False, lam_min=1e-3)
    except nsb._lib.NsbError as ex:
        err = str(ex)
    print("FAULTRESULT " + json.dumps(dict(rank=rank, error=err, seconds=time.perf_counter() - t0)), flush=True)
    os._exit(0)   # the exchange state is broken by design: no collective teardown


if __name__ == "__main__":
    main()
