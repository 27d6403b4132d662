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


def _exchange_in_process(plans, owned_vectors):
    """The halo exchange of all ranks emulated in one process: rank r's tail is filled neighbour by neighbour with the
    entries the neighbour's send list names (the order nsb_halo_create / k_halo_push use)."""
    out = []
    for r, plan in enumerate(plans):
        tail = []
        for k, nb in enumerate(plan.neighbors):
            nb = int(nb)
            other = plans[nb]
            j = list(other.neighbors).index(r)
            so = int(other.send_counts[:j].sum())
            sc = int(other.send_counts[j])
            assert sc == int(plan.recv_counts[k])               # what nb ships to r is what r expects from nb
            idx = torch.as_tensor(other.send_idx[so:so + sc].astype(np.int64))
            tail.append(owned_vectors[nb][idx])
        out.append(torch.cat([owned_vectors[r]] + tail) if tail else owned_vectors[r].clone())
    return out


@pytest.mark.parametrize("world", [4, 8])
def test_row_partition_plans_at_the_bench_world_sizes(world):
    """World sizes 4 and 8 (what the scaling bench runs) planned in one process: every rank's local blocks applied to
    [owned | exchanged halo] reproduce its rows of the global operators, slabs have two neighbours, send and receive
    counts pair up, and the owned ranges tile the index spaces."""
    import navier_stokes_solver_b200.partition as P
    import navier_stokes_solver_b200.synthetic as syn
    cells = (2 * world + 1, 2, 2)        # at least two vertex planes per slab
    G = syn.assemble_stokes(cells, LENGTHS)
    slabs = [syn.assemble_stokes(cells, LENGTHS, slab=r, nslabs=world) for r in range(world)]
    su, sp = P.stokes_slab_spaces(cells, world)
    plans_u = P.make_plans_single_process(su, [torch.cat([S.A.colidx, S.B.colidx]) for S in slabs])
    plans_p = P.make_plans_single_process(sp, [S.BT.colidx for S in slabs])
    locs = [P.localize_stokes(slabs[r], r, world, plans=(plans_u[r], plans_p[r])) for r in range(world)]
    assert sum(L.n_u for L in locs) == G.n_u and sum(L.n_p for L in locs) == G.n_p
    assert torch.equal(torch.sort(torch.cat([L.owned_u for L in locs]))[0], torch.arange(G.n_u))
    assert torch.equal(torch.sort(torch.cat([L.owned_p for L in locs]))[0], torch.arange(G.n_p))
    rng = np.random.default_rng(11)
    xu = torch.as_tensor(rng.uniform(-1, 1, G.n_u))
    xp = torch.as_tensor(rng.uniform(-1, 1, G.n_p))
    xu_loc = _exchange_in_process(plans_u, [xu[L.owned_u] for L in locs])
    xp_loc = _exchange_in_process(plans_p, [xp[L.owned_p] for L in locs])
    A, B, BT = G.A.to_scipy(), G.B.to_scipy(), G.BT.to_scipy()
    yA, yB, yBT = A @ xu.numpy(), B @ xu.numpy(), BT @ xp.numpy()
    for r, L in enumerate(locs):
        assert [int(x) for x in L.plan_u.neighbors] == [x for x in (r - 1, r + 1) if 0 <= x < world]
        ou, op_ = L.owned_u.numpy(), L.owned_p.numpy()
        assert np.abs(L.A.to_scipy() @ xu_loc[r].numpy() - yA[ou]).max() < 1e-12
        assert np.abs(L.B.to_scipy() @ xu_loc[r].numpy() - yB[op_]).max() < 1e-12
        assert np.abs(L.BT.to_scipy() @ xp_loc[r].numpy() - yBT[ou]).max() < 1e-12
        # the stacked saddle rows with the stacked plan
        K, plan = P.local_saddle(L)
        assert K.shape[0] == L.n_u + L.n_p
    # a partial dot product per rank sums to the global one
    assert abs(sum(float(xu[L.owned_u] @ xu[L.owned_u]) for L in locs) - float(xu @ xu)) < 1e-10 * float(xu @ xu)
