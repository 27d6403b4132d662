"""Property tests of the row-partition planner on random ownership and random sparsity, one process: whatever the
contiguous ranges and the sparsity pattern, (a) neighbour lists are symmetric and what a rank receives from a neighbour is exactly what
that neighbour's send list selects, (b) local rows x exchanged vector == global rows, (c) the stacked plan of two spaces delivers the
same entries as the two plans apart."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp
import torch
from hypothesis import given, settings, strategies as st

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from test_partition_condensed_gloo import _simulate_exchange  # noqa: E402


def random_space(P, rng, n, world, pieces):
    """`world` ranks own `pieces` contiguous ranges each, in a random interleaving (component-major numbering owns several)."""
    cuts = np.sort(rng.choice(np.arange(1, n), size=world * pieces - 1, replace=False)) if n > world * pieces else None
    if cuts is None:
        pytest.skip("too few indices")
    bounds = np.concatenate([[0], cuts, [n]])
    owner = rng.permutation(np.repeat(np.arange(world), pieces))
    ranges = [[] for _ in range(world)]
    for k, r in enumerate(owner):
        ranges[int(r)].append((int(bounds[k]), int(bounds[k + 1])))
    return P.IndexSpace(ranges)


@settings(max_examples=25, deadline=None)
@given(seed=st.integers(0, 10_000), world=st.integers(2, 6), pieces=st.integers(1, 3), density=st.floats(0.01, 0.2))
def test_random_partition_reproduces_the_global_product(seed, world, pieces, density):
    import navier_stokes_solver_b200.partition as P
    rng = np.random.default_rng(seed)
    n, m = 97, 61                                   # columns space / rows space of a rectangular operator (B-like)
    cs, rs = random_space(P, rng, n, world, pieces), random_space(P, rng, m, world, pieces)
    M = sp.random(m, n, density=density, random_state=seed, format="csr")
    M.sort_indices()
    x = rng.uniform(-1, 1, n)
    need = [torch.as_tensor(M[rs.owned_ids(r).numpy()].indices.astype(np.int64)) for r in range(world)]
    plans = P.make_plans_single_process(cs, need)
    for r, p in enumerate(plans):
        for k, q in enumerate(p.neighbors):             # (a) symmetry
            assert r in [int(v) for v in plans[int(q)].neighbors]
        assert p.n_owned == cs.n_owned(r) and int(p.recv_counts.sum()) == p.n_halo
    xl = _simulate_exchange(plans, [x[cs.owned_ids(r).numpy()] for r in range(world)])
    for r in range(world):
        # (a) the tail holds the global entries the plan says it holds
        np.testing.assert_array_equal(xl[r][plans[r].n_owned:], x[plans[r].halo_gids.numpy()])
        L = P.distribute_operators([M], [rs], cs, r, plan=plans[r])
        y = L.ops[0].to_scipy() @ xl[r]
        np.testing.assert_allclose(y, (M @ x)[rs.owned_ids(r).numpy()], rtol=0, atol=1e-13)     # (b)
        assert L.touches_halo[0] == bool(plans[r].n_halo > 0 and len(need[r]) > 0 and
                                         (cs.lookup(need[r])[0] != r).any())


@settings(max_examples=15, deadline=None)
@given(seed=st.integers(0, 10_000), world=st.integers(2, 5))
def test_stacked_plan_delivers_what_the_two_plans_deliver(seed, world):
    import navier_stokes_solver_b200.partition as P
    rng = np.random.default_rng(seed)
    nu, npn = 83, 41
    su, sp_ = random_space(P, rng, nu, world, 2), random_space(P, rng, npn, world, 1)
    A = sp.random(nu, nu, density=0.08, random_state=seed, format="csr")
    BT = sp.random(nu, npn, density=0.1, random_state=seed + 1, format="csr")
    need_u = [torch.as_tensor(A[su.owned_ids(r).numpy()].indices.astype(np.int64)) for r in range(world)]
    need_p = [torch.as_tensor(BT[su.owned_ids(r).numpy()].indices.astype(np.int64)) for r in range(world)]
    pu, pp = P.make_plans_single_process(su, need_u), P.make_plans_single_process(sp_, need_p)
    xu, xp = rng.uniform(-1, 1, nu), rng.uniform(-1, 1, npn)
    stacked = [P.stack_plans([pu[r], pp[r]]) for r in range(world)]

    class Pl:                                           # the dict of stack_plans as an object _simulate_exchange reads
        def __init__(self, d):
            self.neighbors, self.send_counts, self.send_idx, self.recv_counts = d["neighbors"], d["send_counts"], d["send_idx"], d["recv_counts"]
            self.n_owned, self.n_halo = d["n_owned"], d["n_halo"]
    owned = [np.concatenate([xu[su.owned_ids(r).numpy()], xp[sp_.owned_ids(r).numpy()]]) for r in range(world)]
    xl = _simulate_exchange([Pl(d) for d, _ in stacked], owned)
    for r in range(world):
        d, tails = stacked[r]
        tail = xl[r][d["n_owned"]:]
        np.testing.assert_array_equal(tail[tails[0]], xu[pu[r].halo_gids.numpy()])
        np.testing.assert_array_equal(tail[tails[1]], xp[pp[r].halo_gids.numpy()])


def test_unstructured_numbering_nnz_balance_and_bandwidth_reduction():
    """Operators from a mesh generator come in an arbitrary numbering: partition.bandwidth_reducing_order (RCM) + balanced_row_space
    give contiguous ranges with balanced nonzeros and a small interface, and the exchanged product still equals the global one."""
    import navier_stokes_solver_b200.partition as P
    import navier_stokes_solver_b200.synthetic as syn
    G = syn.assemble_stokes((10, 4, 4), (2.5, 0.41, 0.41))
    A = G.A.to_scipy()
    n = A.shape[0]
    rng = np.random.default_rng(3)
    shuffle = rng.permutation(n)                                    # "mesh generator" numbering
    As = A[shuffle][:, shuffle].tocsr()
    world = 4
    halo = {}
    for name, M in (("shuffled", As), ("rcm", None)):
        if M is None:
            perm = P.bandwidth_reducing_order(As)
            M = As[perm][:, perm].tocsr()
        M.sort_indices()
        space = P.balanced_row_space(M.indptr, world)
        need = [torch.as_tensor(M[space.owned_ids(r).numpy()].indices.astype(np.int64)) for r in range(world)]
        plans = P.make_plans_single_process(space, need)
        nnz = [int(M[space.owned_ids(r).numpy()].nnz) for r in range(world)]
        assert max(nnz) < 1.1 * M.nnz / world                        # balanced by nonzeros
        halo[name] = sum(p.n_halo for p in plans)
        x = rng.uniform(-1, 1, n)
        xl = _simulate_exchange(plans, [x[space.owned_ids(r).numpy()] for r in range(world)])
        for r in range(world):
            L = P.distribute_operators([M], [space], space, r, plan=plans[r])
            np.testing.assert_allclose(L.ops[0].to_scipy() @ xl[r], (M @ x)[space.owned_ids(r).numpy()], rtol=0, atol=1e-10)
    assert halo["rcm"] < 0.25 * halo["shuffled"], halo                # the ordering is what makes contiguous ranges cheap
    with pytest.raises(ValueError):
        P.balanced_row_space(np.arange(4), 5)
