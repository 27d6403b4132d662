"""Row-partitioned block smoother and auxiliary-space preconditioner (MypreA, templates/NavierStokesSIMPLE_iterative.py:360-383;
SURVEY.md 8(f) item 2 at N > 1) under torch.distributed/gloo on CPU.  The sequence of csrc/blockpre.cu -- blocks kept inside a slab,
colouring of the GLOBAL block graph, the iterate's halo exchanged before every colour, the input of every SpMV of the composite
exchanged first -- restated with host point-to-point calls reproduces the oracle's sequential sweeps and MypreA.Mult on the owned rows."""
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

CELLS, LENGTHS = (24, 6), (2.0, 0.41)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def slab_blocks(space, free, size=4):
    """Blocks of consecutive free dofs that never cross an owned range: stand where the reference has its facet blocks (:360-362),
    with the partition keeping a facet's dofs together."""
    blocks = []
    for rr in space.ranges:
        for a, b in rr:
            idx = np.arange(a, b)[free[a:b]]
            blocks += [idx[i:i + size] for i in range(0, len(idx), size)]
    return blocks


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import scipy.sparse as sp
        import navier_stokes_solver_b200.partition as P
        import navier_stokes_solver_b200.synthetic as syn
        from navier_stokes_solver_b200.la import color_blocks
        from oracle import krylov_oracle as ko
        G = syn.assemble_stokes(CELLS, LENGTHS)
        A = G.A.to_scipy()
        free = G.free_u.numpy().astype(bool)
        su, _ = P.stokes_slab_spaces(CELLS, world)
        blocks = slab_blocks(su, free)
        order, cptr = color_blocks(A, blocks)
        blocks = [blocks[i] for i in order]                       # colour-major, the order of the sequential reference sweep
        ref = ko.BlockSmoother(A, blocks)
        lb, lptr = P.localize_blocks(blocks, cptr, su, rank)
        assert len(lptr) == len(cptr)                              # every rank walks every colour
        L = P.distribute_operators([A], [su], su, rank)
        Al = L.ops[0].to_scipy()
        n = L.plan.n_owned
        inv = [np.linalg.inv(Al[b][:, b].toarray()) for b in lb]  # k_block_invert: owned columns carry the owned rows' numbering
        ou = su.owned_ids(rank).numpy()
        rng = np.random.default_rng(21)
        x = rng.uniform(-1, 1, G.n_u) * free
        y0 = rng.uniform(-1, 1, G.n_u) * free

        def exch(plan, v):
            return P.host_halo_exchange(plan, torch.as_tensor(np.ascontiguousarray(v))).numpy()

        def sweep(y, xo, backward):                                # block_gs_sweep (blockpre.cu), row-partitioned
            y = y.copy()
            ncol = len(lptr) - 1
            for k in range(ncol):
                c = ncol - 1 - k if backward else k
                yl = exch(L.plan, y)                               # also on a rank without blocks of this colour
                for b in range(int(lptr[c]), int(lptr[c + 1])):
                    r = xo[lb[b]] - Al[lb[b]] @ yl
                    y[lb[b]] += inv[b] @ r
            return y

        def jacobi(xo):
            y = np.zeros(n)
            for b, iv in zip(lb, inv):
                y[b] = iv @ xo[b]
            return y

        e = {}
        e["jacobi"] = np.abs(jacobi(x[ou]) - ref.jacobi(x)[ou]).max()
        yf = ref.smooth(y0.copy(), x)
        e["smooth"] = np.abs(sweep(y0[ou], x[ou], False) - yf[ou]).max()
        yb = ref.smooth(y0.copy(), x, backward=True)
        e["smooth_back"] = np.abs(sweep(y0[ou], x[ou], True) - yb[ou]).max()

        # ---- auxiliary-space composite: transform over an aggregate space split into contiguous chunks
        free_idx = np.nonzero(free)[0]
        agg = np.arange(len(free_idx)) // 6
        nc = int(agg.max()) + 1
        T = sp.csr_matrix((np.ones(len(free_idx)), (free_idx, agg)), shape=(G.n_u, nc))
        dc = 1.0 / (T.T @ A @ T).diagonal()
        sa = P.IndexSpace([[(nc * r // world, nc * (r + 1) // world)] for r in range(world)])
        Lt = P.distribute_operators([T], [su], sa, rank)                    # velocity rows, auxiliary columns
        Ltt = P.distribute_operators([T.T.tocsr()], [sa], su, rank)         # auxiliary rows, velocity columns
        Tl, TTl = Lt.ops[0].to_scipy(), Ltt.ops[0].to_scipy()
        oa = sa.owned_ids(rank).numpy()

        def aux(xo, GS):                                                     # auxspace_apply (blockpre.cu), row-partitioned
            if GS:
                y = sweep(np.zeros(n), xo, False)
                t = xo - Al @ exch(L.plan, y)
            else:
                y = jacobi(xo)
                t = xo
            cvec = TTl @ exch(Ltt.plan, t)
            z = dc[oa] * cvec
            y = y + Tl @ exch(Lt.plan, z)
            return sweep(y, xo, True) if GS else y

        for GS in (True, False):
            want = ko.aux_space_pre(ref, A, T, dc, x, GS=GS)
            e["aux_gs" if GS else "aux_add"] = np.abs(aux(x[ou], GS) - want[ou]).max() / np.abs(want).max()
        out[rank] = ({k: float(v) for k, v in e.items()}, int(lptr[-1]), len(blocks), int(len(cptr) - 1))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_row_partitioned_smoother_and_auxspace_match_the_sequential_reference(world):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        nblocks = 0
        for r in range(world):
            e, nb_local, nb_global, ncol = out[r]
            assert all(v < 1e-12 for v in e.values()), (r, e)
            assert ncol >= 2
            nblocks += nb_local
        assert nblocks == nb_global                                 # every block lives on exactly one rank


def test_a_block_cut_by_the_partition_is_refused():
    import navier_stokes_solver_b200.partition as P
    su, _ = P.stokes_slab_spaces(CELLS, 2)
    a, b = su.ranges[0][0]
    straddling = [np.array([b - 2, b - 1, b, b + 1])]               # two dofs of slab 0, two of slab 1 (same component)
    with pytest.raises(ValueError):
        P.localize_blocks(straddling, np.array([0, 1]), su, 0)
    inside = [np.array([a, a + 1]), np.array([b, b + 1])]
    lb, ptr = P.localize_blocks(inside, np.array([0, 1, 2]), su, 1)
    assert [x.tolist() for x in lb] == [[0, 1]] and ptr.tolist() == [0, 0, 1]      # colour 0 is empty on rank 1 and still listed
