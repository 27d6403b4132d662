"""Row partition of the saddle-point blocks across the GPUs of one box.

The reference is single-process; this module is the host-side plan for the 1-D row partition the
north star asks for: every rank owns a set of contiguous global row ranges of each index space
(velocity dofs, pressure dofs), its rows of every operator with the column indices renumbered to
``[owned | halo]``, and a halo plan (who sends which owned entries to whom) executed on the device
by libnsb200 (dist.cu: pack -> ncclSend/ncclRecv -> SpMV).

Everything here is index arithmetic on torch tensors (CPU or CUDA) plus one small
``all_gather_object`` of the interface lists, so it runs unchanged under the ``gloo`` backend on CPU
(tests/test_partition_gloo.py) and under ``nccl`` on the GPUs.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch


# ------------------------------------------------------------------------------------------------
# ownership of one global index space
# ------------------------------------------------------------------------------------------------
class IndexSpace:
    """Ownership of a global index space ``0..n`` by contiguous ranges.  ``ranges[r]`` is the list of
    ``(start, stop)`` global ranges owned by rank r, in the order of r's local numbering (so local index
    = offset of the range + position inside it)."""

    def __init__(self, ranges):
        self.ranges = [[(int(a), int(b)) for a, b in rr] for rr in ranges]
        starts, stops, owner, loc = [], [], [], []
        for r, rr in enumerate(self.ranges):
            off = 0
            for a, b in rr:
                if b > a:
                    starts.append(a); stops.append(b); owner.append(r); loc.append(off)
                off += b - a
        order = np.argsort(np.asarray(starts, dtype=np.int64), kind="stable")
        self._starts = np.asarray(starts, dtype=np.int64)[order]
        self._stops = np.asarray(stops, dtype=np.int64)[order]
        self._owner = np.asarray(owner, dtype=np.int64)[order]
        self._loc = np.asarray(loc, dtype=np.int64)[order]
        if len(self._starts) > 1 and np.any(self._starts[1:] < self._stops[:-1]):
            raise ValueError("IndexSpace: owned ranges overlap")
        self._cache = {}

    @property
    def nranks(self):
        return len(self.ranges)

    def n_owned(self, rank):
        return sum(b - a for a, b in self.ranges[rank])

    def n_global(self):
        return int(sum(self.n_owned(r) for r in range(self.nranks)))

    def owned_ids(self, rank, device="cpu"):
        parts = [torch.arange(a, b, dtype=torch.int64, device=device) for a, b in self.ranges[rank]]
        return torch.cat(parts) if parts else torch.zeros(0, dtype=torch.int64, device=device)

    def _tables(self, device):
        key = str(device)
        if key not in self._cache:
            self._cache[key] = tuple(torch.as_tensor(t, device=device) for t in
                                     (self._starts, self._stops, self._owner, self._loc))
        return self._cache[key]

    def lookup(self, gids: torch.Tensor):
        """(owner rank, local index in the owner's numbering) of every global id."""
        starts, stops, owner, loc = self._tables(gids.device)
        gids = gids.to(torch.int64)
        k = torch.searchsorted(starts, gids, right=True) - 1
        if gids.numel() and (int(k.min()) < 0 or bool((gids >= stops[k]).any())):
            raise ValueError("IndexSpace.lookup: id outside every owned range")
        return owner[k], loc[k] + (gids - starts[k])


# ------------------------------------------------------------------------------------------------
# halo plan of one rank for one index space
# ------------------------------------------------------------------------------------------------
@dataclass
class Plan:
    """What nsb_halo_create takes (include/nsb200.h): the halo tail of a vector is filled neighbour by
    neighbour in ``neighbors`` order; ``send_idx`` lists, neighbour by neighbour, the OWNED local entries to
    ship.  ``halo_gids`` are the global ids of the tail entries, in tail order."""
    rank: int
    n_owned: int
    neighbors: np.ndarray      # int32, ascending
    send_counts: np.ndarray    # int64 per neighbour
    send_idx: np.ndarray       # int32, concatenated
    recv_counts: np.ndarray    # int64 per neighbour
    halo_gids: torch.Tensor    # int64 (device of the request)
    _sorted: tuple = field(default=None, repr=False)

    @property
    def n_halo(self):
        return int(self.halo_gids.numel())

    def localize(self, space: IndexSpace, gids: torch.Tensor) -> torch.Tensor:
        """Global column ids -> local ``[owned | halo]`` numbering (int32)."""
        owner, loc = space.lookup(gids)
        mine = owner == self.rank
        out = torch.empty_like(loc)
        out[mine] = loc[mine]
        if not bool(mine.all()):
            if self._sorted is None:
                hs, perm = torch.sort(self.halo_gids.to(gids.device))
                self._sorted = (hs, perm)
            hs, perm = self._sorted
            g = gids[~mine].to(torch.int64)
            pos = torch.searchsorted(hs, g)
            if bool((pos >= hs.numel()).any()) or bool((hs[pos.clamp(max=hs.numel() - 1)] != g).any()):
                raise ValueError("Plan.localize: column outside owned + halo")
            out[~mine] = self.n_owned + perm[pos]
        return out.to(torch.int32)


def plan_requests(space: IndexSpace, rank: int, needed_gids: torch.Tensor):
    """Phase 1 (local): the off-rank global ids ``rank`` needs, in tail order (by owner, then global id),
    and the request lists {owner: local indices in the owner's numbering}."""
    need = torch.unique(needed_gids.to(torch.int64))
    owner, loc = space.lookup(need)
    off = owner != rank
    need, owner, loc = need[off], owner[off], loc[off]
    key = torch.argsort(owner * (space.n_global() + 1) + need)
    need, owner, loc = need[key], owner[key], loc[key]
    owner_h = owner.cpu().numpy()
    loc_h = loc.cpu().numpy()
    requests = {int(r): loc_h[owner_h == r].astype(np.int64) for r in np.unique(owner_h)}
    return need, requests


def finish_plan(space: IndexSpace, rank: int, need: torch.Tensor, requests: dict, allreq: list) -> Plan:
    """Phase 2 (after the request lists of every rank are known): send lists = what the others asked
    of this rank."""
    sends = {r: np.asarray(req[rank], dtype=np.int64) for r, req in enumerate(allreq) if r != rank and rank in req}
    nb = sorted(set(requests) | set(sends))
    empty = np.zeros(0, dtype=np.int64)
    send_counts = np.array([len(sends.get(r, empty)) for r in nb], dtype=np.int64)
    recv_counts = np.array([len(requests.get(r, empty)) for r in nb], dtype=np.int64)
    send_idx = np.concatenate([sends.get(r, empty) for r in nb]).astype(np.int32) if nb else np.zeros(0, np.int32)
    if send_idx.size and (send_idx.min() < 0 or send_idx.max() >= space.n_owned(rank)):
        raise ValueError("finish_plan: a neighbour requested an entry this rank does not own")
    return Plan(rank, space.n_owned(rank), np.asarray(nb, dtype=np.int32), send_counts, send_idx, recv_counts, need)


def make_plan(space: IndexSpace, rank: int, needed_gids: torch.Tensor, group=None) -> Plan:
    """Halo plan of ``rank``: ``needed_gids`` are the global column ids its rows touch (any order,
    duplicates allowed).  Collective over ``group`` (every rank calls it): one all_gather_object of
    the per-neighbour request lists (interface entries only)."""
    import torch.distributed as dist
    need, requests = plan_requests(space, rank, needed_gids)
    allreq = [None] * space.nranks
    dist.all_gather_object(allreq, requests, group=group)
    return finish_plan(space, rank, need, requests, allreq)


def make_plans_single_process(space: IndexSpace, needed_per_rank):
    """All ranks' plans computed in one process (tests, and planning on the host before a launch)."""
    ph1 = [plan_requests(space, r, needed) for r, needed in enumerate(needed_per_rank)]
    allreq = [req for _, req in ph1]
    return [finish_plan(space, r, ph1[r][0], ph1[r][1], allreq) for r in range(space.nranks)]


def union_space_plan_offsets(n_owned_list, n_halo_list):
    """Offsets of a stacked vector [owned_0 | owned_1 | ... | halo_0 | halo_1 | ...] (the saddle-point
    vector (u, p) of a rank: run.py:47-48)."""
    own_off = np.concatenate([[0], np.cumsum(n_owned_list)]).astype(np.int64)
    halo_off = own_off[-1] + np.concatenate([[0], np.cumsum(n_halo_list)]).astype(np.int64)
    return own_off, halo_off


def stack_plans(plans):
    """Halo plan of the stacked vector of several index spaces (velocity, pressure) from their
    individual plans.  Tail order is neighbour-major (the C ABI fills the tail neighbour by
    neighbour), so the stacked tail interleaves the spaces per neighbour; returns the plan data and
    ``tail_maps[s]``: position in the stacked tail of every tail entry of space s."""
    rank = plans[0].rank
    own_off, _ = union_space_plan_offsets([p.n_owned for p in plans], [p.n_halo for p in plans])
    nb = sorted(set(int(r) for p in plans for r in p.neighbors))
    send_counts, recv_counts, send_idx = [], [], []
    tail_maps = [np.zeros(p.n_halo, dtype=np.int64) for p in plans]
    pos = 0
    for r in nb:
        sc = rc = 0
        for s, p in enumerate(plans):
            where = np.nonzero(p.neighbors == r)[0]
            if where.size == 0:
                continue
            k = int(where[0])
            s0 = int(p.send_counts[:k].sum()); s1 = s0 + int(p.send_counts[k])
            r0 = int(p.recv_counts[:k].sum()); r1 = r0 + int(p.recv_counts[k])
            send_idx.append(p.send_idx[s0:s1].astype(np.int64) + own_off[s])
            sc += s1 - s0
            tail_maps[s][r0:r1] = pos + np.arange(r1 - r0)
            pos += r1 - r0
            rc += r1 - r0
        send_counts.append(sc)
        recv_counts.append(rc)
    send_idx = np.concatenate(send_idx).astype(np.int32) if send_idx else np.zeros(0, np.int32)
    return dict(rank=rank, n_owned=int(own_off[-1]), n_halo=pos, neighbors=np.asarray(nb, dtype=np.int32),
                send_counts=np.asarray(send_counts, dtype=np.int64), send_idx=send_idx,
                recv_counts=np.asarray(recv_counts, dtype=np.int64)), tail_maps


# ------------------------------------------------------------------------------------------------
# x-slab partition of the structured Taylor-Hood Stokes blocks (synthetic.assemble_stokes(slab=...))
# ------------------------------------------------------------------------------------------------
def stokes_slab_spaces(cells, nslabs):
    """(velocity IndexSpace, pressure IndexSpace) of the x-slab row partition used by
    synthetic.assemble_stokes: slab s owns the vertex planes [(c_x+1)s/P, (c_x+1)(s+1)/P) and the midpoint
    plane right of each; velocity dofs are component-major, so a slab owns d global ranges."""
    d = len(cells)
    if nslabs > cells[0] + 1:
        raise ValueError("stokes_slab_spaces: %d slabs for %d vertex planes in x -- a rank would own no rows" % (nslabs, cells[0] + 1))
    dims2 = [2 * c + 1 for c in cells]
    dims1 = [c + 1 for c in cells]
    n2 = int(np.prod(dims2))
    plane2 = int(np.prod(dims2[1:]))
    plane1 = int(np.prod(dims1[1:]))
    ru, rp = [], []
    for s in range(nslabs):
        vx0 = (cells[0] + 1) * s // nslabs
        vx1 = (cells[0] + 1) * (s + 1) // nslabs
        px0, px1 = 2 * vx0, min(2 * vx1, dims2[0])
        ru.append([(px0 * plane2 + c * n2, px1 * plane2 + c * n2) for c in range(d)])
        rp.append([(vx0 * plane1, vx1 * plane1)])
    return IndexSpace(ru), IndexSpace(rp)


@dataclass
class LocalStokes:
    """One rank's share: CSR triples with LOCAL column numbering plus the halo plans."""
    rank: int
    nranks: int
    space_u: IndexSpace
    space_p: IndexSpace
    plan_u: Plan           # velocity halo (union of the columns of A and B rows owned here)
    plan_p: Plan           # pressure halo (columns of the owned B^T rows)
    A: object              # synthetic.CSR, (n_u_owned) x (n_u_owned + halo_u)
    B: object              # (n_p_owned) x (n_u_owned + halo_u)
    BT: object             # (n_u_owned) x (n_p_owned + halo_p)
    dA: torch.Tensor
    dM: torch.Tensor
    f: torch.Tensor
    g: torch.Tensor
    u0: torch.Tensor
    owned_u: torch.Tensor
    owned_p: torch.Tensor

    @property
    def n_u(self):
        return self.plan_u.n_owned

    @property
    def n_p(self):
        return self.plan_p.n_owned


def localize_stokes(S, rank, nranks, group=None, plans=None) -> LocalStokes:
    """Turn the slab assembly ``S`` (global column ids) of ``rank`` into local numbering + halo plans.
    Collective (two make_plan calls) unless ``plans = (plan_u, plan_p)`` is given."""
    import navier_stokes_solver_b200.synthetic as syn
    su, sp = stokes_slab_spaces(S.cells, nranks)
    assert su.n_owned(rank) == S.owned_u.numel() and sp.n_owned(rank) == S.owned_p.numel()
    dA = syn.jacobi_inverse(S.A, row_ids=S.owned_u, free=S.free_u)
    dM = syn.jacobi_inverse(S.Mp, row_ids=S.owned_p)
    if plans is None:
        plan_u = make_plan(su, rank, torch.cat([S.A.colidx, S.B.colidx]), group)
        plan_p = make_plan(sp, rank, S.BT.colidx, group)
    else:
        plan_u, plan_p = plans
    nxu = plan_u.n_owned + plan_u.n_halo
    nxp = plan_p.n_owned + plan_p.n_halo
    A = syn.CSR((S.A.shape[0], nxu), S.A.rowptr, plan_u.localize(su, S.A.colidx), S.A.values)
    B = syn.CSR((S.B.shape[0], nxu), S.B.rowptr, plan_u.localize(su, S.B.colidx), S.B.values)
    BT = syn.CSR((S.BT.shape[0], nxp), S.BT.rowptr, plan_p.localize(sp, S.BT.colidx), S.BT.values)
    return LocalStokes(rank, nranks, su, sp, plan_u, plan_p, A, B, BT, dA, dM, S.f, S.g, S.u0, S.owned_u, S.owned_p)


def local_saddle(L: LocalStokes):
    """Rows of K = [[A, B^T], [B, 0]] owned by the rank (run.py:45) as one CSR over the stacked local vector
    [u_owned | p_owned | halo], plus the stacked halo plan data (for nsb_halo_create)."""
    import navier_stokes_solver_b200.synthetic as syn
    plan, tails = stack_plans([L.plan_u, L.plan_p])
    dev = L.A.values.device
    n_u, n_p, n_own = L.n_u, L.n_p, plan["n_owned"]
    tu = torch.as_tensor(tails[0], device=dev)
    tp = torch.as_tensor(tails[1], device=dev)

    def map_u(c):
        c = c.to(torch.int64)
        return torch.where(c < n_u, c, n_own + tu[(c - n_u).clamp(min=0)] if tu.numel() else c)

    def map_p(c):
        c = c.to(torch.int64)
        return torch.where(c < n_p, c + n_u, n_own + tp[(c - n_p).clamp(min=0)] if tp.numel() else c + n_u)

    def rows_of(csr, off):
        cnt = csr.rowptr[1:] - csr.rowptr[:-1]
        return torch.repeat_interleave(torch.arange(csr.shape[0], device=dev) + off, cnt)

    rows = torch.cat([rows_of(L.A, 0), rows_of(L.BT, 0), rows_of(L.B, n_u)])
    cols = torch.cat([map_u(L.A.colidx), map_p(L.BT.colidx), map_u(L.B.colidx)])
    vals = torch.cat([L.A.values, L.BT.values, L.B.values])
    order = torch.argsort(rows, stable=True)
    rows, cols, vals = rows[order], cols[order], vals[order]
    rowptr = torch.zeros(n_own + 1, dtype=torch.int64, device=dev)
    rowptr[1:] = torch.cumsum(torch.bincount(rows, minlength=n_own), 0)
    K = syn.CSR((n_own, n_own + plan["n_halo"]), rowptr, cols.to(torch.int32), vals)
    return K, plan


# ------------------------------------------------------------------------------------------------
# any set of operators over one column index space (the statically condensed velocity operator: A, harmonic_extension,
# harmonic_extension_trans, inner_solve, inner_matrix -- solvers/bramblepasciak_new.py:8-21,84-103 -- and B beside them)
# ------------------------------------------------------------------------------------------------
def _as_csr(op, device):
    """synthetic.CSR (torch) of a synthetic.CSR or a scipy sparse matrix."""
    import navier_stokes_solver_b200.synthetic as syn
    if isinstance(op, syn.CSR):
        return op
    m = op.tocsr()
    m.sort_indices()
    return syn.CSR(tuple(m.shape), torch.as_tensor(m.indptr.astype(np.int64), device=device),
                   torch.as_tensor(m.indices.astype(np.int32), device=device), torch.as_tensor(m.data.astype(np.float64), device=device))


def take_rows(csr, row_ids: torch.Tensor):
    """Rows ``row_ids`` (int64, any order) of a CSR, in that order, column ids unchanged."""
    import navier_stokes_solver_b200.synthetic as syn
    row_ids = row_ids.to(torch.int64)
    start = csr.rowptr[row_ids]
    cnt = csr.rowptr[row_ids + 1] - start
    rowptr = torch.zeros(row_ids.numel() + 1, dtype=torch.int64, device=csr.rowptr.device)
    rowptr[1:] = torch.cumsum(cnt, 0)
    total = int(rowptr[-1])
    # position p of the output belongs to output row k = searchsorted(rowptr, p, right) - 1 and is entry start[k] + (p - rowptr[k])
    p = torch.arange(total, dtype=torch.int64, device=csr.rowptr.device)
    k = torch.repeat_interleave(torch.arange(row_ids.numel(), dtype=torch.int64, device=p.device), cnt)
    src = start[k] + (p - rowptr[k])
    return syn.CSR((int(row_ids.numel()), csr.shape[1]), rowptr, csr.colidx[src], csr.values[src])


@dataclass
class LocalOperators:
    """One rank's rows of several operators that share a column index space, renumbered over ONE halo plan."""
    plan: Plan
    ops: list            # synthetic.CSR, columns in [owned | halo] numbering; a compact operator has only the owned columns
    touches_halo: list   # per operator: some owned row references an off-rank column (then the plan must be attached to it)

    @property
    def nx(self):
        return self.plan.n_owned + self.plan.n_halo


def distribute_operators(ops, row_spaces, col_space: IndexSpace, rank: int, group=None, compact=None, plan=None) -> LocalOperators:
    """Row partition of operators over one column space with ONE shared halo plan: every vector of that space is then exchanged
    the same way whichever operator reads it (csrc/cond.cuh applies five of them in a chain).

    ops         global operators (synthetic.CSR with global column ids, or scipy sparse), or already this rank's rows of them
                (row count == row_spaces[k].n_owned(rank))
    row_spaces  the IndexSpace of each operator's rows (the velocity space for A, H, ...; the pressure space for B)
    compact     per operator: True = must reference owned columns only (element-local inner_solve / inner_matrix with elements kept
                inside a slab); it is returned with n_owned columns, gets no plan, and its input needs no exchange.  ValueError if
                it does touch another rank's columns.
    Collective over ``group`` (one make_plan) unless ``plan`` is given."""
    import navier_stokes_solver_b200.synthetic as syn
    compact = list(compact) if compact is not None else [False] * len(ops)
    dev = None
    for op in ops:
        if isinstance(op, syn.CSR):
            dev = op.values.device
            break
    dev = dev if dev is not None else torch.device("cpu")
    mine = []
    for op, rs in zip(ops, row_spaces):
        csr = _as_csr(op, dev)
        if csr.shape[0] == rs.n_owned(rank) and csr.shape[0] != rs.n_global():
            mine.append(csr)
        else:
            if csr.shape[0] != rs.n_global():
                raise ValueError("distribute_operators: operator has %d rows, its row space %d (owned here: %d)"
                                 % (csr.shape[0], rs.n_global(), rs.n_owned(rank)))
            mine.append(take_rows(csr, rs.owned_ids(rank, device=csr.rowptr.device)))
    if plan is None:
        cols = [m.colidx.to(torch.int64) for m, cp in zip(mine, compact) if not cp]
        needed = torch.cat(cols) if cols else torch.zeros(0, dtype=torch.int64, device=dev)
        plan = make_plan(col_space, rank, needed, group)
    nx = plan.n_owned + plan.n_halo
    out, touches = [], []
    for k, (m, cp) in enumerate(zip(mine, compact)):
        owner, _ = col_space.lookup(m.colidx.to(torch.int64)) if m.colidx.numel() else (torch.zeros(0, dtype=torch.int64), None)
        off = bool((owner != rank).any()) if m.colidx.numel() else False
        if cp and off:
            raise ValueError("distribute_operators: operator %d is declared compact (element-local) but its rows on rank %d reference "
                             "columns owned elsewhere" % (k, rank))
        if cp:
            _, loc = col_space.lookup(m.colidx.to(torch.int64)) if m.colidx.numel() else (None, torch.zeros(0, dtype=torch.int64))
            out.append(syn.CSR((m.shape[0], plan.n_owned), m.rowptr, loc.to(torch.int32), m.values))
        else:
            out.append(syn.CSR((m.shape[0], nx), m.rowptr, plan.localize(col_space, m.colidx), m.values))
        touches.append(off)
    return LocalOperators(plan, out, touches)


def balanced_row_space(rowptr, nranks: int) -> IndexSpace:
    """One contiguous range of rows per rank, balanced by stored nonzeros (SpMV work and HBM bytes follow nnz, not rows): the cut
    after rank r is the first row at which the running nnz reaches (r+1)/nranks of the total.  For operators exported from an
    unstructured mesh (the reference's Netgen meshes), where no slab structure exists."""
    rp = np.asarray(rowptr.cpu().numpy() if isinstance(rowptr, torch.Tensor) else rowptr, dtype=np.int64)
    n = len(rp) - 1
    if nranks > n:
        raise ValueError("balanced_row_space: %d ranks for %d rows -- a rank would own no rows" % (nranks, n))
    total = int(rp[-1])
    cuts = [0]
    for r in range(1, nranks):
        c = int(np.searchsorted(rp, total * r / nranks, side="left"))
        c = min(max(c, cuts[-1] + 1), n - (nranks - r))          # every rank keeps at least one row
        cuts.append(c)
    cuts.append(n)
    return IndexSpace([[(cuts[r], cuts[r + 1])] for r in range(nranks)])


def bandwidth_reducing_order(pattern) -> np.ndarray:
    """Reverse Cuthill-McKee permutation of a symmetric sparsity pattern (scipy): rows that couple end up close together, so contiguous
    row ranges cut few couplings.  Apply it to every operator of the space (``M[perm][:, perm]``) and to the vectors before
    ``balanced_row_space`` / ``distribute_operators``; a mesh numbered by a mesh generator needs it, the structured channel does not."""
    import scipy.sparse as sp
    from scipy.sparse.csgraph import reverse_cuthill_mckee
    P = sp.csr_matrix(pattern)
    P = (abs(P) + abs(P).T).tocsr()
    return np.asarray(reverse_cuthill_mckee(P, symmetric_mode=True), dtype=np.int64)


def localize_blocks(blocks, color_ptr, space: IndexSpace, rank: int):
    """This rank's share of a coloured block list (la.color_blocks of the GLOBAL pattern: ``blocks`` colour-major, ``color_ptr`` the
    block ranges per colour): the blocks whose dofs it owns, in local numbering, with a colour table of the SAME number of colours
    (a colour may be empty here) -- every rank walks all colours, the halo of the iterate is exchanged before each
    (csrc/blockpre.cu).  A block that straddles two ranks is refused: the partition must keep a facet's dofs together."""
    nb = len(blocks)
    sizes = np.array([len(b) for b in blocks], dtype=np.int64)
    ncol = len(color_ptr) - 1
    if nb == 0 or int(sizes.sum()) == 0:
        return [], np.zeros(ncol + 1, dtype=np.int64)
    dofs = np.concatenate([np.asarray(b, dtype=np.int64) for b in blocks])
    owner, loc = space.lookup(torch.as_tensor(dofs))            # one lookup for all blocks
    owner, loc = owner.numpy(), loc.numpy()
    ptr = np.concatenate([[0], np.cumsum(sizes)])
    bid = np.repeat(np.arange(nb), sizes)
    mine_cnt = np.bincount(bid, weights=(owner == rank).astype(np.float64), minlength=nb).astype(np.int64)
    cut = np.nonzero((mine_cnt > 0) & (mine_cnt < sizes))[0]
    if cut.size:
        b = int(cut[0])
        raise ValueError("localize_blocks: block %d has dofs on ranks %s; blocks must not be cut by the row partition"
                         % (b, sorted(set(owner[ptr[b]:ptr[b + 1]].tolist()))))
    keep = (mine_cnt == sizes) & (sizes > 0)
    out = [loc[ptr[b]:ptr[b + 1]].astype(np.int32) for b in np.nonzero(keep)[0]]
    kept_before = np.concatenate([[0], np.cumsum(keep)])         # blocks kept among the first k
    lptr = np.array([kept_before[int(color_ptr[c])] for c in range(ncol + 1)], dtype=np.int64)
    return out, lptr


@dataclass
class DistributedCondensed:
    """Device side of a row-partitioned statically condensed form: an object with the attributes of a condensed BilinearForm
    (what solvers.BramblePasciakCG / la.CondensedOperator.from_bilinear_form read) plus B, B^T and the plans that keep them alive."""
    mat: object
    harmonic_extension: object
    harmonic_extension_trans: object
    inner_solve: object
    inner_matrix: object
    B: object
    BT: object
    plan_u: Plan
    plan_p: Plan
    halos: tuple = ()
    condense: bool = True


def build_distributed_condensed(ctx, rank, space_u: IndexSpace, space_p: IndexSpace, mat, H, HT, inner_solve, inner_matrix, B, BT,
                                group=None, inner_is_local=True) -> DistributedCondensed:
    """Upload this rank's rows of the five operators of a condensed velocity form and of B / B^T, with one velocity halo plan shared by
    everything that reads a velocity vector and one pressure plan for B^T.  Collective (two make_plan calls + the plans' device set-up).
    ``inner_is_local``: inner_solve / inner_matrix couple dofs of one element only and elements are not split across ranks."""
    from .la import SparseMatrix
    Lu = distribute_operators([mat, H, HT, inner_solve, inner_matrix, B], [space_u] * 5 + [space_p], space_u, rank, group,
                              compact=[False, False, False, inner_is_local, inner_is_local, False])
    Lp = distribute_operators([BT], [space_u], space_p, rank, group)
    hu, hp = upload_halo(ctx, Lu.plan), upload_halo(ctx, Lp.plan)
    dev = [SparseMatrix.from_csr(m, ctx) for m in Lu.ops]
    for m, cp in zip(dev, [False, False, False, inner_is_local, inner_is_local, False]):
        if not cp:
            hu.attach(m)
    dBT = SparseMatrix.from_csr(Lp.ops[0], ctx)
    hp.attach(dBT)
    dev[5].set_transpose(dBT)
    return DistributedCondensed(dev[0], dev[1], dev[2], dev[3], dev[4], dev[5], dBT, Lu.plan, Lp.plan, (hu, hp))


# ------------------------------------------------------------------------------------------------
# device upload (needs libnsb200 + a row-partitioned Context)
# ------------------------------------------------------------------------------------------------
@dataclass
class DistributedStokes:
    A: object
    B: object
    BT: object
    dA: torch.Tensor
    dM: torch.Tensor
    f: torch.Tensor
    g: torch.Tensor
    u_start: torch.Tensor
    lanczos_start: torch.Tensor
    local: LocalStokes
    n_u_global: int
    n_p_global: int
    nnzA_global: int
    nnzB_global: int
    halos: tuple = ()


def upload_halo(ctx, plan):
    from .la import HaloPlan
    if isinstance(plan, Plan):
        return HaloPlan(ctx, plan.n_owned, plan.n_halo, plan.neighbors, plan.send_counts, plan.send_idx, plan.recv_counts)
    return HaloPlan(ctx, plan["n_owned"], plan["n_halo"], plan["neighbors"], plan["send_counts"], plan["send_idx"],
                    plan["recv_counts"])


def build_distributed_stokes(cells, lengths, ctx, rank, world, device="cuda", group=None, keep_local=False) -> DistributedStokes:
    """Assemble the rank's x-slab of the 3D/2D Taylor-Hood channel on its GPU, renumber, upload the blocks
    and attach the halo plans.  Collective over the default process group."""
    import torch.distributed as dist
    import navier_stokes_solver_b200.synthetic as syn
    from .la import SparseMatrix
    S = syn.assemble_stokes(cells, lengths, device=device, slab=rank, nslabs=world)
    L = localize_stokes(S, rank, world, group)
    del S
    if torch.cuda.is_available():
        torch.cuda.empty_cache()   # the library allocates with cudaMalloc: torch's cache still holds the assembly scratch (tens of GB)
    hu, hp = upload_halo(ctx, L.plan_u), upload_halo(ctx, L.plan_p)
    A = SparseMatrix.from_csr(L.A, ctx)
    B = SparseMatrix.from_csr(L.B, ctx)
    BT = SparseMatrix.from_csr(L.BT, ctx)
    hu.attach(A)
    hu.attach(B)
    hp.attach(BT)
    B.set_transpose(BT)
    counts = torch.tensor([L.A.nnz, L.B.nnz], dtype=torch.int64, device=L.A.values.device)
    if world > 1:
        dist.all_reduce(counts, group=group)
    u_start = torch.cat([L.u0, torch.zeros(L.n_p, dtype=torch.float64, device=L.u0.device)])
    lz = 1.0 + 0.5 * torch.sin(L.owned_u.to(torch.float64) + 1.0)
    out = DistributedStokes(A, B, BT, L.dA, L.dM, L.f, L.g, u_start, lz, L, L.space_u.n_global(), L.space_p.n_global(),
                            int(counts[0]), int(counts[1]), (hu, hp))
    if not keep_local:
        L.A = L.B = L.BT = None   # the CSR triples now live in the library (kept only for local_saddle)
    return out


def host_halo_exchange(plan, x_owned: torch.Tensor, group=None) -> torch.Tensor:
    """The halo exchange of dist.cu restated with torch.distributed point-to-point calls on host
    tensors: returns [x_owned | halo tail].  Used to check plans under gloo (tests) -- the solvers run
    the exchange on the device (nsb_halo_create / halo_exchange)."""
    import torch.distributed as dist
    p = plan if isinstance(plan, dict) else dict(n_owned=plan.n_owned, n_halo=plan.n_halo, neighbors=plan.neighbors,
                                                 send_counts=plan.send_counts, send_idx=plan.send_idx,
                                                 recv_counts=plan.recv_counts)
    out = torch.zeros(p["n_owned"] + p["n_halo"], dtype=x_owned.dtype)
    out[:p["n_owned"]] = x_owned
    send_idx = torch.as_tensor(np.asarray(p["send_idx"], dtype=np.int64))
    reqs, so, ro = [], 0, 0
    bufs = []
    for k, nb in enumerate(p["neighbors"]):
        sc, rc = int(p["send_counts"][k]), int(p["recv_counts"][k])
        if sc:
            buf = x_owned[send_idx[so:so + sc]].contiguous()
            bufs.append(buf)
            reqs.append(dist.isend(buf, int(nb), group=group))
        if rc:
            reqs.append(dist.irecv(out[p["n_owned"] + ro:p["n_owned"] + ro + rc], int(nb), group=group))
        so += sc
        ro += rc
    for r in reqs:
        r.wait()
    return out
