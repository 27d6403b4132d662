"""Row partition of a statically condensed velocity form (SURVEY.md 8(f) item 1 at N > 1; solvers/bramblepasciak_new.py:8-21,84-103)
under torch.distributed/gloo on CPU: the five operators and B share ONE velocity halo plan (partition.distribute_operators), and the
chains of csrc/cond.cuh -- every SpMV preceded by the exchange of its input -- restated with host point-to-point calls reproduce
the global operator and the global harmonic-extension preconditioner on the owned rows."""
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


def _global_case(cells=CELLS):
    import navier_stokes_solver_b200.synthetic as syn
    from oracle.cases import synthetic_condensation
    G = syn.assemble_stokes(cells, LENGTHS)
    A, B, BT = G.A.to_scipy(), G.B.to_scipy(), G.BT.to_scipy()
    free = G.free_u.numpy().astype(bool)
    cd = synthetic_condensation(A, free)
    return G, A, B, BT, free, cd


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import navier_stokes_solver_b200.partition as P
        G, A, B, BT, free, cd = _global_case()
        su, sp_ = P.stokes_slab_spaces(CELLS, world)
        L = P.distribute_operators([cd["mat"], cd["H"], cd["HT"], cd["inner_solve"], cd["inner_matrix"], B], [su] * 5 + [sp_], su, rank,
                                   compact=[False, False, False, True, True, False])
        S, H, HT, isolve, imat, Bl = [m.to_scipy() for m in L.ops]
        n, nx = L.plan.n_owned, L.nx
        assert S.shape == (n, nx) and H.shape == (n, nx) and HT.shape == (n, nx) and isolve.shape == (n, n) and imat.shape == (n, n)
        assert Bl.shape == (sp_.n_owned(rank), nx) and L.touches_halo[3:5] == [False, False]
        ou = su.owned_ids(rank).numpy()
        opp = sp_.owned_ids(rank).numpy()
        rng = np.random.default_rng(11)
        x = rng.uniform(-1, 1, G.n_u)
        f = rng.uniform(-1, 1, G.n_u)
        d = rng.uniform(0.5, 2.0, G.n_u)
        s = 0.37

        def exch(v_owned):      # halo_exchange of dist.cu / p2p.cu: [owned | tail filled from the neighbours]
            return P.host_halo_exchange(L.plan, torch.as_tensor(np.ascontiguousarray(v_owned))).numpy()

        # cond_apply_A (cond.cuh): t = x - H x ; q = A t + inner_matrix t ; out = q - HT q, each SpMV input exchanged first
        xl = exch(x[ou])
        t = xl[:n] - H @ xl
        tl = exch(t)
        q = S @ tl + imat @ tl[:n]
        ql = exch(q)
        y = ql[:n] - HT @ ql
        y_ref = (A @ x)[ou]            # (I - HT)(S + inner_matrix)(I - H) == A  (oracle.cases.synthetic_condensation)
        e_A = np.abs(y - y_ref).max() / np.abs(y_ref).max()
        # cond_apply_pre: r = f + HT f ; y = s d r ; out = y + H y + inner_solve r
        fl = exch(f[ou])
        r = fl[:n] + HT @ fl
        yv = s * (d[ou] * r)
        yl = exch(yv)
        pre = yl[:n] + H @ yl + isolve @ r
        rg = f + cd["HT"] @ f
        yg = s * (d * rg)
        pre_ref = (yg + cd["H"] @ yg + cd["inner_solve"] @ rg)[ou]
        e_pre = np.abs(pre - pre_ref).max() / np.abs(pre_ref).max()
        # B reads the same exchanged velocity vector (one plan for everything that reads a velocity vector)
        e_B = np.abs(Bl @ xl - (B @ x)[opp]).max()
        # B^T over the pressure plan
        Lp = P.distribute_operators([BT], [su], sp_, rank)
        p = rng.uniform(-1, 1, G.n_p)
        pl = P.host_halo_exchange(Lp.plan, torch.as_tensor(p[opp])).numpy()
        e_BT = np.abs(Lp.ops[0].to_scipy() @ pl - (BT @ p)[ou]).max()
        out[rank] = (float(e_A), float(e_pre), float(e_B), float(e_BT), [int(v) for v in L.plan.neighbors], bool(any(L.touches_halo[:3])))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_condensed_chain_row_partitioned_matches_global(world):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(240)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        for r in range(world):
            e_A, e_pre, e_B, e_BT, nb, touches = out[r]
            assert e_A < 1e-12 and e_pre < 1e-12 and e_B < 1e-12 and e_BT < 1e-12, (r, e_A, e_pre, e_B, e_BT)
            assert nb == [q for q in (r - 1, r + 1) if 0 <= q < world]          # slabs: the plan is symmetric, nearest neighbours only
            assert touches                                                       # the harmonic extension does cross the slab interface


def test_take_rows_and_compact_contract_single_process():
    import navier_stokes_solver_b200.partition as P
    G, A, B, BT, free, cd = _global_case()
    su, _ = P.stokes_slab_spaces(CELLS, 2)
    ids = su.owned_ids(1)
    sub = P.take_rows(P._as_csr(cd["H"], torch.device("cpu")), ids).to_scipy()
    assert abs(sub - cd["H"][ids.numpy()]).max() == 0
    plans = P.make_plans_single_process(su, [torch.as_tensor(np.concatenate([(m[su.owned_ids(r).numpy()]).indices for m in
                                                                               (cd["mat"], cd["H"], cd["HT"])]).astype(np.int64))
                                             for r in range(2)])
    # the diagonal inner operators are element-local: compact, n_owned columns, no plan needed
    # (the greedy interior set of the synthetic condensation is one-sided at a slab interface: the interior dofs of slab 0 see
    #  interface dofs of slab 1, so H crosses on rank 0 and H^T on rank 1)
    L = P.distribute_operators([cd["inner_solve"], cd["H"], cd["HT"]], [su] * 3, su, 1, compact=[True, False, False], plan=plans[1])
    assert L.ops[0].shape == (su.n_owned(1), su.n_owned(1)) and L.ops[1].shape == (su.n_owned(1), L.nx)
    assert L.touches_halo == [False, False, True]
    # declaring an operator compact whose rows reference dofs of the other slab must be refused
    with pytest.raises(ValueError):
        P.distribute_operators([cd["HT"]], [su], su, 1, compact=[True], plan=plans[1])
    with pytest.raises(ValueError):
        P.distribute_operators([cd["H"]], [su], su, 0, compact=[True], plan=plans[0])


def _simulate_exchange(plans, owned):
    """The exchange of dist.cu / p2p.cu played in one process for every rank at once: rank q's tail is filled neighbour by neighbour
    (q.neighbors order) with what that neighbour's send list for q selects from the neighbour's owned entries."""
    out = []
    for q, pq in enumerate(plans):
        tail = []
        for k, r in enumerate(pq.neighbors):
            pr = plans[int(r)]
            j = int(np.nonzero(pr.neighbors == q)[0][0])            # q must be in r's list: the plans are symmetric
            s0 = int(pr.send_counts[:j].sum())
            seg = pr.send_idx[s0:s0 + int(pr.send_counts[j])]
            assert len(seg) == int(pq.recv_counts[k])
            tail.append(owned[int(r)][seg])
        out.append(np.concatenate([owned[q]] + tail) if tail else owned[q].copy())
    return out


@pytest.mark.parametrize("world", [4, 8])
def test_condensed_chain_at_the_bench_world_sizes_planned_in_one_process(world):
    """World sizes 4 and 8 (ranks with two neighbours) without spawning processes: the plans of every rank are built in one process
    (partition.make_plans_single_process) and the exchange is played for all ranks at once."""
    import navier_stokes_solver_b200.partition as P
    cells = (16, 2, 2)                                              # 17 vertex planes: every one of 8 slabs owns rows
    G, A, B, BT, free, cd = _global_case(cells)
    su, sp_ = P.stokes_slab_spaces(cells, world)
    names = ("mat", "H", "HT", "inner_solve", "inner_matrix")
    need = [torch.as_tensor(np.concatenate([cd[k][su.owned_ids(r).numpy()].indices for k in names[:3]] +
                                           [B[sp_.owned_ids(r).numpy()].indices]).astype(np.int64)) for r in range(world)]
    plans = P.make_plans_single_process(su, need)
    L = [P.distribute_operators([cd[k] for k in names] + [B], [su] * 5 + [sp_], su, r, compact=[False, False, False, True, True, False],
                                plan=plans[r]) for r in range(world)]
    ou = [su.owned_ids(r).numpy() for r in range(world)]
    rng = np.random.default_rng(17)
    x = rng.uniform(-1, 1, G.n_u)
    ops = [[m.to_scipy() for m in Lr.ops] for Lr in L]
    n = [p.n_owned for p in plans]
    xl = _simulate_exchange(plans, [x[o] for o in ou])
    t = [xl[r][:n[r]] - ops[r][1] @ xl[r] for r in range(world)]
    tl = _simulate_exchange(plans, t)
    q = [ops[r][0] @ tl[r] + ops[r][4] @ tl[r][:n[r]] for r in range(world)]
    ql = _simulate_exchange(plans, q)
    ref = A @ x
    for r in range(world):
        y = ql[r][:n[r]] - ops[r][2] @ ql[r]
        assert np.abs(y - ref[ou[r]]).max() < 1e-12 * np.abs(ref).max()
        nb = [int(v) for v in plans[r].neighbors]
        assert nb == [v for v in (r - 1, r + 1) if 0 <= v < world]
        opp = sp_.owned_ids(r).numpy()
        assert np.abs(ops[r][5] @ xl[r] - (B @ x)[opp]).max() < 1e-12


def test_more_slabs_than_vertex_planes_is_refused():
    import navier_stokes_solver_b200.partition as P
    with pytest.raises(ValueError, match="no rows"):
        P.stokes_slab_spaces((6, 2, 2), 8)
    P.stokes_slab_spaces((7, 2, 2), 8)
