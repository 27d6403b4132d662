"""Row-partition host logic under torch.distributed/gloo, world_size 2 and 3, on CPU:
halo plans built collectively, halo exchange restated with point-to-point calls, local SpMV + halo ==
global SpMV rows, all-reduced dot == global dot, stacked saddle plan."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CELLS, LENGTHS = (6, 2, 2), (2.5, 0.41, 0.41)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import navier_stokes_solver_b200.partition as P
        import navier_stokes_solver_b200.synthetic as syn
        G = syn.assemble_stokes(CELLS, LENGTHS)                       # every rank holds the global reference
        S = syn.assemble_stokes(CELLS, LENGTHS, slab=rank, nslabs=world)
        L = P.localize_stokes(S, rank, world)
        rng = np.random.default_rng(3)
        xu = torch.as_tensor(rng.uniform(-1, 1, G.n_u))
        xp = torch.as_tensor(rng.uniform(-1, 1, G.n_p))
        A, B, BT = G.A.to_scipy(), G.B.to_scipy(), G.BT.to_scipy()
        xu_loc = P.host_halo_exchange(L.plan_u, xu[L.owned_u])
        xp_loc = P.host_halo_exchange(L.plan_p, xp[L.owned_p])
        ou, op_ = L.owned_u.numpy(), L.owned_p.numpy()
        e = [np.abs(L.A.to_scipy() @ xu_loc.numpy() - (A @ xu.numpy())[ou]).max(),
             np.abs(L.B.to_scipy() @ xu_loc.numpy() - (B @ xu.numpy())[op_]).max(),
             np.abs(L.BT.to_scipy() @ xp_loc.numpy() - (BT @ xp.numpy())[ou]).max()]
        # stacked saddle operator: rows of [[A, B^T], [B, 0]] owned here
        K, plan = P.local_saddle(L)
        x_own = torch.cat([xu[L.owned_u], xp[L.owned_p]])
        x_loc = P.host_halo_exchange(plan, x_own)
        yk = K.to_scipy() @ x_loc.numpy()
        yref = np.concatenate([(A @ xu.numpy() + BT @ xp.numpy())[ou], (B @ xu.numpy())[op_]])
        e.append(np.abs(yk - yref).max())
        # dot product: local partial + all_reduce == global
        d = torch.tensor([float(xu[L.owned_u] @ xu[L.owned_u])], dtype=torch.float64)
        dist.all_reduce(d)
        e.append(abs(float(d) - float(xu @ xu)) / float(xu @ xu))
        # Jacobi diagonals are the owned slices of the global ones
        dA = syn.jacobi_inverse(G.A, free=G.free_u)
        e.append(float((L.dA - dA[L.owned_u]).abs().max()))
        counts = torch.tensor([L.n_u, L.n_p], dtype=torch.int64)
        dist.all_reduce(counts)
        ok = counts.tolist() == [G.n_u, G.n_p]
        out[rank] = (e, ok, [int(x) for x in L.plan_u.neighbors])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_row_partition_matches_global_operator(world):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        for r in range(world):
            e, ok, nb = out[r]
            assert ok
            assert max(e) < 1e-12, (r, e)
            assert nb == [x for x in (r - 1, r + 1) if 0 <= x < world]   # x-slabs: two slab neighbours


def test_index_space_and_plans_single_process():
    import navier_stokes_solver_b200.partition as P
    sp = P.IndexSpace([[(0, 4), (10, 12)], [(4, 10)], [(12, 15)]])
    owner, loc = sp.lookup(torch.tensor([0, 3, 4, 9, 10, 11, 12, 14]))
    assert owner.tolist() == [0, 0, 1, 1, 0, 0, 2, 2]
    assert loc.tolist() == [0, 3, 0, 5, 4, 5, 0, 2]
    with pytest.raises(ValueError):
        sp.lookup(torch.tensor([15]))
    with pytest.raises(ValueError):
        P.IndexSpace([[(0, 5)], [(4, 8)]])
    plans = P.make_plans_single_process(sp, [torch.tensor([0, 5, 13]), torch.tensor([4, 11]), torch.tensor([12])])
    assert plans[0].halo_gids.tolist() == [5, 13] and plans[0].recv_counts.tolist() == [1, 1]
    assert plans[1].send_idx.tolist() == [1] and plans[1].halo_gids.tolist() == [11]
    assert plans[0].send_idx.tolist() == [5]          # gid 11 is local entry 5 of rank 0
    assert plans[2].n_halo == 0 and plans[2].send_idx.tolist() == [1]
    assert plans[0].localize(sp, torch.tensor([0, 5, 13, 11])).tolist() == [0, 6, 7, 5]
    with pytest.raises(ValueError):
        plans[0].localize(sp, torch.tensor([6]))
