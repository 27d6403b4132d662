"""Synthetic operators of the SIMPLE-type time step (templates/NavierStokesSIMPLE_iterative.py:79-144) on the structured
Taylor-Hood channel, built with torch on the CPU or directly on the GPU, whole or one x-slab of a row partition.

This is an *input generator* for tests and benchmarks (like synthetic.py): the reference assembles these forms with NGSolve's
MCS H(div) element library, which is absent here.  The operators keep the reference's shapes and roles:

    a      = nu * vector Laplacian                          self.a      (:79-80)
    mstar  = M + dt * a                                     self.mstar  (:85-86)   solved by CGSolver(mstar, pre, precision=1e-4) (:93)
    Xproj  = [V2 | Q]; the mixed projection system [[-Ml, B^T], [B, 0]] with lumped velocity mass Ml, condensed onto Q:
             aproj = [[0, 0], [0, B Ml^-1 B^T]], harmonic_extension H = [[0, Ml^-1 B^T], [0, 0]], inner_solve = diag(-Ml^-1, 0)   (:118-131)
    bproj  = [0; B]  (Xproj.ndof x V.ndof)                  self.bproj  (:133-135)
    mapV   : V dof (c, node) lives at V2 dof ((c + 1) % d, node)        self.mapV   (:138-144; node-local, so slab-local)

(same construction as oracle.cases.simple_case, which is the scipy twin used by the small parity cases and by the golden fixture
of the reference's own DoTimeStep / Project method bodies).  B is taken with the Dirichlet velocity columns removed.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import synthetic as syn
from .synthetic import CSR, _coo_to_csr


def _free_velocity(gids, cells, outlet=True):
    """Free-dof mask of GLOBAL velocity ids (same rule as synthetic.assemble_stokes)."""
    d = len(cells)
    dims2 = [2 * c + 1 for c in cells]
    n2 = int(np.prod(dims2))
    rem = gids % n2
    on = torch.zeros_like(gids, dtype=torch.bool)
    for a in reversed(range(d)):
        ca = rem % dims2[a]
        rem = torch.div(rem, dims2[a], rounding_mode="floor")
        on |= ca == 0
        if not (a == 0 and outlet):
            on |= ca == dims2[a] - 1
    return ~on


def _rows_of(csr):
    cnt = csr.rowptr[1:] - csr.rowptr[:-1]
    return torch.repeat_interleave(torch.arange(csr.shape[0], device=csr.values.device), cnt)


def mass_diagonal(cells, lengths, device="cpu"):
    """Diagonal of the scalar P2 mass matrix at EVERY node of the refined lattice (element loop over the whole mesh; only the
    diagonal is accumulated, in a fixed order)."""
    d = len(cells)
    dev = torch.device(device)
    h = np.array([lengths[a] / cells[a] for a in range(d)])
    dims2 = tuple(2 * c + 1 for c in cells)
    n2 = int(np.prod(dims2))
    ranges = [torch.arange(0, cells[a], device=dev, dtype=torch.int64) for a in range(d)]
    cell = torch.stack(torch.meshgrid(*ranges, indexing="ij"), dim=-1).reshape(-1, d)
    rows, vals = [], []
    for sv in syn.kuhn_simplices(d):
        loc = syn.simplex_p2p1_matrices(sv * h)
        svt = torch.as_tensor(sv, device=dev)
        vtx = cell[:, None, :] + svt[None, :, :]
        mids = torch.stack([vtx[:, a] + vtx[:, b] for a, b in loc["edges"]], dim=1)
        dof2 = syn._lattice_index(torch.cat([2 * vtx, mids], dim=1), dims2)
        rows.append(dof2.reshape(-1))
        vals.append(torch.as_tensor(np.diag(loc["M2"]).copy(), device=dev)[None].expand(dof2.shape[0], -1).reshape(-1))
    r = torch.cat(rows)
    c = _coo_to_csr(r, torch.zeros_like(r), torch.cat(vals), (n2, 1))
    out = torch.zeros(n2, dtype=torch.float64, device=dev)
    out[(c.rowptr[1:] - c.rowptr[:-1]) > 0] = c.values
    return out


def schur_rows(Bown: CSR, Bext: CSR, ext_row_ids: torch.Tensor, dinv_cols: torch.Tensor, n_p: int, chunk_rows=8192) -> CSR:
    """Rows of  S = B D B^T  for the rows of ``Bown``; ``Bext`` holds every row of B that shares a column with them (row ids
    ``ext_row_ids``, global), ``dinv_cols`` the diagonal D over the global columns.  Expansion + sort + ordered segment sums in row
    chunks (deterministic: the same bits whatever the partition)."""
    dev = Bown.values.device
    n_u = Bown.shape[1]
    # transpose of Bext: for every global column the list of (ext row, value), sorted by column then row
    er = ext_row_ids[_rows_of(Bext)]
    ec = Bext.colidx.to(torch.int64)
    order = torch.argsort(ec * (n_p + 1) + er)
    tc, tr, tv = ec[order], er[order], Bext.values[order]
    tptr = torch.zeros(n_u + 1, dtype=torch.int64, device=dev)
    tptr[1:] = torch.cumsum(torch.bincount(tc, minlength=n_u), 0)
    del er, ec, order, tc
    parts = []
    nrows = Bown.shape[0]
    for r0 in range(0, nrows, chunk_rows):
        r1 = min(r0 + chunk_rows, nrows)
        p0, p1 = int(Bown.rowptr[r0]), int(Bown.rowptr[r1])
        j = Bown.colidx[p0:p1].to(torch.int64)
        v = Bown.values[p0:p1] * dinv_cols[j]
        q = torch.repeat_interleave(torch.arange(r0, r1, device=dev), Bown.rowptr[r0 + 1:r1 + 1] - Bown.rowptr[r0:r1])
        cnt = tptr[j + 1] - tptr[j]
        src = torch.repeat_interleave(torch.arange(j.numel(), device=dev), cnt)
        off = torch.arange(src.numel(), device=dev) - torch.repeat_interleave(torch.cumsum(cnt, 0) - cnt, cnt)
        pos = tptr[j][src] + off
        parts.append(_coo_to_csr(q[src] - r0, tr[pos], v[src] * tv[pos], (r1 - r0, n_p)))
    rowptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev)] +
                       [p.rowptr[1:] + off for p, off in zip(parts, np.concatenate([[0], np.cumsum([p.nnz for p in parts])[:-1]]).tolist())])
    return CSR((nrows, n_p), rowptr, torch.cat([p.colidx for p in parts]), torch.cat([p.values for p in parts]))


@dataclass
class SimpleSystem:
    """Operators of one rank (or of the whole mesh) with GLOBAL column ids.  Row sets: velocity-space operators have the owned
    velocity rows; Xproj-space operators have the owned [V2 | Q] rows (V2 rows first)."""
    cells: tuple
    lengths: tuple
    dt: float
    n_u: int               # global V.ndof (= V2.ndof)
    n_p: int               # global Q.ndof
    owned_u: torch.Tensor  # global ids of the owned velocity rows
    owned_p: torch.Tensor
    a: CSR                 # owned_u x n_u
    mstar: CSR             # owned_u x n_u
    dm: torch.Tensor       # Jacobi of mstar on the free dofs
    bproj: CSR             # owned Xproj rows x n_u
    aproj: CSR             # owned Xproj rows x (n_u + n_p)
    H: CSR
    HT: CSR
    inner: CSR
    dp: torch.Tensor       # Jacobi of aproj (zero on the V2 rows)
    map_v2: torch.Tensor   # per owned velocity row: GLOBAL Xproj id of its V2 twin
    f: torch.Tensor
    u0: torch.Tensor
    free_u: torch.Tensor
    Bf: CSR                # owned_p x n_u (for divergence checks)


def assemble_simple(cells, lengths, dt=1e-3, nu=1.0, device="cpu", slab=None, nslabs=1) -> SimpleSystem:
    d = len(cells)
    cells = tuple(int(c) for c in cells)
    dev = torch.device(device)
    S = syn.assemble_stokes(cells, lengths, nu=nu, device=device, slab=slab, nslabs=nslabs, with_mass2=True, parts=("A", "B", "BT"))
    n2, n_u, n_p = S.n2, S.n_u, S.n_p
    M2 = S.extras["M2"]
    assert torch.equal(M2.colidx, S.extras["K2"].colidx) and torch.equal(M2.rowptr, S.extras["K2"].rowptr)
    S.extras.pop("K2")
    mstar = CSR(S.A.shape, S.A.rowptr, S.A.colidx, M2.values.repeat(d) + dt * S.A.values)
    dm = syn.jacobi_inverse(mstar, row_ids=S.owned_u, free=S.free_u)
    Ml = mass_diagonal(cells, lengths, device)                                  # every node
    dinv_cols = (1.0 / Ml).repeat(d)                                            # D over the global velocity ids
    # B with the Dirichlet columns removed (values zeroed, pattern kept)
    Bf = CSR(S.B.shape, S.B.rowptr, S.B.colidx, S.B.values * _free_velocity(S.B.colidx.to(torch.int64), cells).to(torch.float64))
    # rows of B two vertex planes either side: every row that shares a velocity column with an owned row
    plane1 = int(np.prod([c + 1 for c in cells[1:]]))
    if slab is None:
        Bext, ext_ids = Bf, S.owned_p
    else:
        v0 = int(S.owned_p[0]) // plane1
        v1 = int(S.owned_p[-1]) // plane1 + 1
        E = syn.assemble_stokes(cells, lengths, nu=nu, device=device, vrange=(v0 - 2, v1 + 2), parts=("B",))
        Bext = CSR(E.B.shape, E.B.rowptr, E.B.colidx, E.B.values * _free_velocity(E.B.colidx.to(torch.int64), cells).to(torch.float64))
        ext_ids = E.owned_p
        del E
    Spp = schur_rows(Bf, Bext, ext_ids, dinv_cols, n_p)
    del Bext
    n_pl, n_ul = int(S.owned_p.numel()), int(S.owned_u.numel())
    nl = n_ul // d                                                              # owned nodes
    zero_rows = torch.zeros(n_ul, dtype=torch.int64, device=dev)

    def under_v2(csr, col_shift):
        """Rows of an operator that lives on the Q rows of Xproj: n_ul empty V2 rows on top."""
        return CSR((n_ul + n_pl, csr.shape[1] + col_shift), torch.cat([zero_rows, csr.rowptr]),
                   (csr.colidx.to(torch.int64) + col_shift).to(torch.int32), csr.values)

    bproj = under_v2(Bf, 0)
    aproj = under_v2(Spp, n_u)
    dS = syn.jacobi_inverse(Spp, row_ids=S.owned_p)
    dp = torch.cat([torch.zeros(n_ul, dtype=torch.float64, device=dev), torch.where(dS > 0, dS, torch.zeros_like(dS))])
    # H: V2 row ((c+1)%d, node) = Ml^-1 * free * (B^T row (c, node)); V2 rows are component-major like V rows
    bt_cnt = (S.BT.rowptr[1:] - S.BT.rowptr[:-1]).reshape(d, nl)
    scale = (dinv_cols[S.owned_u] * S.free_u.to(torch.float64)).reshape(d, nl)
    btv = S.BT.values * torch.repeat_interleave(scale.reshape(-1), bt_cnt.reshape(-1))
    rot = [(c - 1) % d for c in range(d)]                                       # V2 block c' is filled from V block (c' - 1) % d
    seg = [(int(S.BT.rowptr[c * nl]), int(S.BT.rowptr[(c + 1) * nl])) for c in range(d)]
    hcnt = torch.cat([bt_cnt[c] for c in rot])
    hptr = torch.zeros(n_ul + n_pl + 1, dtype=torch.int64, device=dev)
    hptr[1:n_ul + 1] = torch.cumsum(hcnt, 0)
    hptr[n_ul + 1:] = hptr[n_ul]
    H = CSR((n_ul + n_pl, n_u + n_p), hptr, torch.cat([(S.BT.colidx[a:b].to(torch.int64) + n_u).to(torch.int32) for a, b in (seg[c] for c in rot)]),
            torch.cat([btv[a:b] for a, b in (seg[c] for c in rot)]))
    # HT: Q row q = (B D)[q, :] with the columns moved to the V2 numbering (rotation), re-sorted
    cj = Bf.colidx.to(torch.int64)
    c2 = (cj + n2) % n_u
    HTq = _coo_to_csr(_rows_of(Bf), c2, Bf.values * dinv_cols[cj], (n_pl, n_u))
    HT = under_v2(HTq, 0)
    HT = CSR((n_ul + n_pl, n_u + n_p), HT.rowptr, HT.colidx, HT.values)
    # inner_solve = diag(-1 / Ml2) on the V2 rows: V2 dof k = twin of V dof j = (k - n2) % n_u, same node => same Ml
    own_v2 = S.owned_u                                                          # the owned V2 gids equal the owned V gids (node-local rotation)
    iptr = torch.cat([torch.arange(0, n_ul + 1, dtype=torch.int64, device=dev), torch.full((n_pl,), n_ul, dtype=torch.int64, device=dev)])
    inner = CSR((n_ul + n_pl, n_u + n_p), iptr, own_v2.to(torch.int32), -dinv_cols[own_v2])
    map_v2 = (S.owned_u + n2) % n_u
    return SimpleSystem(cells, tuple(lengths), float(dt), n_u, n_p, S.owned_u, S.owned_p, S.A, mstar, dm, bproj, aproj, H, HT, inner, dp,
                        map_v2, S.f, S.u0, S.free_u, Bf)


# ------------------------------------------------------------------------------------------------
# device upload: whole mesh on one GPU, or one slab of a row partition
# ------------------------------------------------------------------------------------------------
@dataclass
class DeviceSimple:
    stepper: object
    sys: SimpleSystem
    ops: dict
    halos: tuple
    n_x_global: int
    n_proj_global: int
    nnz: dict              # global nonzero counts per operator
    plan_u: object = None
    plan_x: object = None
    host: dict = None      # keep_host: numpy CSR triples (shape, rowptr, col, val) and vectors, for the CPU baseline
    parts: dict = None     # pm, pp (Jacobi handles), mapV, f: what a second stepper on the same operators needs

    def make_stepper(self, gfu, mstar_precision, mstar_maxsteps, proj_precision, proj_maxsteps):
        """Another time stepper on the same uploaded operators (other tolerances / step caps, own state vector)."""
        from .simple import ExtendedInverse, SimpleStepper
        from .solvers import CGSolver
        o, q, s = self.ops, self.parts, self.stepper
        invm = ExtendedInverse(CGSolver(o["mstar"], q["pm"], precision=mstar_precision, maxsteps=mstar_maxsteps))
        invp = ExtendedInverse(CGSolver(o["aproj"], q["pp"], precision=proj_precision, maxsteps=proj_maxsteps), o["H"], o["HT"], o["inner"])
        return SimpleStepper(o["a"], invm, o["bproj"], invp, q["mapV"], gfu, s.n_vel, s.p_range, s.timestep, f=q["f"], ctx=s.ctx)


def build_device_simple(cells, lengths, dt, ctx, rank=0, world=1, device="cuda", nu=1.0, mstar_precision=1e-4, mstar_maxsteps=200,
                        proj_precision=1e-8, proj_maxsteps=200, group=None, keep_host=False) -> DeviceSimple:
    """Assemble (the rank's slab of) the synthetic SIMPLE operators, renumber [owned | halo], upload, attach the halo plans and
    build the device time stepper.  Collective over the default process group when world > 1."""
    from . import partition as P
    from .la import DiagonalPreconditioner, SparseMatrix
    from .simple import ExtendedInverse, PermutationMatrix, SimpleStepper
    from .solvers import CGSolver
    T = assemble_simple(cells, lengths, dt, nu, device, slab=rank if world > 1 else None, nslabs=world)
    n_ul, n_pl = int(T.owned_u.numel()), int(T.owned_p.numel())
    dev = T.a.values.device
    halos = ()
    plan_u = plan_x = None
    if world > 1:
        su, sp = P.stokes_slab_spaces(cells, world)
        sx = P.IndexSpace([su.ranges[r] + [(a + T.n_u, b + T.n_u) for a, b in sp.ranges[r]] for r in range(world)])
        plan_u = P.make_plan(su, rank, torch.cat([T.a.colidx, T.bproj.colidx]), group)
        plan_x = P.make_plan(sx, rank, torch.cat([T.aproj.colidx, T.H.colidx, T.HT.colidx, T.inner.colidx]), group)
        nxu, nxx = plan_u.n_owned + plan_u.n_halo, plan_x.n_owned + plan_x.n_halo

        def loc(csr, plan, space, ncols):
            return CSR((csr.shape[0], ncols), csr.rowptr, plan.localize(space, csr.colidx), csr.values)
        for k in ("a", "mstar", "bproj"):
            setattr(T, k, loc(getattr(T, k), plan_u, su, nxu))
        for k in ("aproj", "H", "HT", "inner"):
            setattr(T, k, loc(getattr(T, k), plan_x, sx, nxx))
        map_loc = plan_x.localize(sx, T.map_v2)
        hu, hx = P.upload_halo(ctx, plan_u), P.upload_halo(ctx, plan_x)
        halos = (hu, hx)
    else:
        map_loc = T.map_v2.to(torch.int32)
    host = None
    if keep_host:
        def trip(m):
            return (tuple(m.shape), m.rowptr.cpu().numpy(), m.colidx.cpu().numpy(), m.values.cpu().numpy())
        host = {k: trip(getattr(T, k)) for k in ("a", "mstar", "bproj", "aproj", "H", "HT", "inner")}
        host.update(dm=T.dm.cpu().numpy(), dp=T.dp.cpu().numpy(), map=map_loc.cpu().numpy().astype(np.int32), f=T.f.cpu().numpy(),
                    u0=T.u0.cpu().numpy())
    ops = {}
    panels = [0, n_ul, n_ul + n_pl]
    for k in ("a", "mstar", "bproj", "aproj", "H", "HT", "inner"):
        m = SparseMatrix.from_csr(getattr(T, k), ctx)
        if k not in ("a", "mstar"):
            m.set_row_panels(panels)      # V2 rows and Q rows differ completely in length (several are empty)
        if world > 1:
            (halos[0] if k in ("a", "mstar", "bproj") else halos[1]).attach(m)
        ops[k] = m
    counts = torch.tensor([getattr(T, k).nnz for k in ("a", "mstar", "bproj", "aproj", "H", "HT", "inner")], dtype=torch.int64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(counts, group=group)
    nnz = dict(zip(("a", "mstar", "bproj", "aproj", "H", "HT", "inner"), counts.tolist()))
    pm = DiagonalPreconditioner.from_diagonal(T.dm, ctx)
    pp = DiagonalPreconditioner.from_diagonal(T.dp, ctx)
    invmstar = ExtendedInverse(CGSolver(ops["mstar"], pm, precision=mstar_precision, maxsteps=mstar_maxsteps))
    invproj = ExtendedInverse(CGSolver(ops["aproj"], pp, precision=proj_precision, maxsteps=proj_maxsteps), ops["H"], ops["HT"], ops["inner"])
    mapV = PermutationMatrix(int(ops["aproj"].width), map_loc.to(torch.int32), device=dev)
    gfu = T.u0.clone()
    stepper = SimpleStepper(ops["a"], invmstar, ops["bproj"], invproj, mapV, gfu, n_ul, (n_ul, n_ul + n_pl), dt, f=T.f, ctx=ctx)
    for k in ("a", "mstar", "bproj", "aproj", "H", "HT", "inner"):   # the CSR triples now live in the library
        setattr(T, k, None)
    return DeviceSimple(stepper, T, ops, halos, T.n_u, T.n_u + T.n_p, nnz, plan_u, plan_x, host, dict(pm=pm, pp=pp, mapV=mapV, f=T.f))
