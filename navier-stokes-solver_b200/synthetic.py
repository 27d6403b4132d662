"""Synthetic FEM assembler: Taylor-Hood P2/P1 Stokes blocks on structured simplicial meshes.

This is the *input generator* for tests and benchmarks, not part of the solve path.  The reference
assembles with NGSolve (absent here); its weak forms are restated on a structured
Kuhn triangulation of a box so that the same block shapes can be produced at any size, on the CPU
(small parity cases) or directly on the GPU (the ~20M-DOF case never touches host memory):

    a  = int  grad(u_c) . grad(v_c)           (run.py:77-78)    -> A  (n_u x n_u), component-major
    b  = int  div(u) q                        (run.py:80-81)    -> B  (n_p x n_u)
    mp = int  p q                             (run.py:83-84)    -> Mp (n_p x n_p)
    f  = int  (x - 0.5) v_1                   (run.py:92-94)
    inlet profile 1.5*4y(0.41-y)/0.41^2       (run.py:101-103)  2D
    inlet profile 16 y(0.41-y) z(0.41-z)/0.41^4   (templates/NavierStokesSIMPLE_test_3D.py:25)  3D
    heat: mass, stiffness                     (heat.py:43-52)

Layout facts used throughout:  every point of the refined lattice (2n_x+1) x (2n_y+1) [x (2n_z+1)] is
a P2 node (vertices, edge midpoints; the diagonal midpoints fall on cell/face centres), every
even-even[-even] point is a P1 node.  Node numbering has x slowest so that slabs along the channel
axis are contiguous node ranges; velocity dofs are component-major like NGSolve's
compound space FESpace([V, V]) (discretizations.py:9-10).
"""
from __future__ import annotations

import itertools
import math
from dataclasses import dataclass, field

import numpy as np
import torch


# ----------------------------------------------------------------------------------------------
# exact local matrices on one simplex, via  int lambda^alpha = vol * d! * alpha! / (d + |alpha|)!
# ----------------------------------------------------------------------------------------------
def _poly_mul(p, q):
    out = {}
    for a, ca in p.items():
        for b, cb in q.items():
            k = tuple(x + y for x, y in zip(a, b))
            out[k] = out.get(k, 0.0) + ca * cb
    return out


def _poly_int(p, d):
    s = 0.0
    for a, c in p.items():
        num = math.factorial(d)
        for x in a:
            num *= math.factorial(x)
        s += c * num / math.factorial(d + sum(a))
    return s  # times the simplex volume


def _p2_basis(d):
    """P2 Lagrange basis as polynomials in barycentric coordinates.
    Order: d+1 vertices, then edges (a,b) a<b in lexicographic order.  Returns (polys, edges)."""
    nv = d + 1

    def mono(*idx):
        e = [0] * nv
        for i in idx:
            e[i] += 1
        return tuple(e)

    one = {mono(k): 1.0 for k in range(nv)}  # 1 = sum lambda_k  (keeps everything homogeneous)
    polys = []
    for i in range(nv):
        lam = {mono(i): 1.0}
        # lambda_i (2 lambda_i - 1) = 2 lambda_i^2 - lambda_i * (sum_k lambda_k)
        p = {mono(i, i): 2.0}
        for k, c in _poly_mul(lam, one).items():
            p[k] = p.get(k, 0.0) - c
        polys.append(p)
    edges = [(a, b) for a in range(nv) for b in range(a + 1, nv)]
    for a, b in edges:
        polys.append({mono(a, b): 4.0})
    return polys, edges


def _poly_dlam(p, k):
    out = {}
    for a, c in p.items():
        if a[k] > 0:
            b = list(a)
            b[k] -= 1
            out[tuple(b)] = out.get(tuple(b), 0.0) + c * a[k]
    return out


def simplex_p2p1_matrices(verts: np.ndarray) -> dict:
    """Local P2/P1 matrices on the simplex with vertex coordinates ``verts`` ((d+1) x d).

    Returns dict with  K2 (P2 stiffness), M2 (P2 mass), K1, M1 (P1), Bc[c] ((d+1) x nP2: int psi_i d_c phi_j),
    T (  int lambda_k phi_j , (d+1) x nP2 ),  I2 ( int phi_j ), vol.
    """
    verts = np.asarray(verts, dtype=np.float64)
    d = verts.shape[1]
    nv = d + 1
    J = (verts[1:] - verts[0]).T  # d x d
    vol = abs(np.linalg.det(J)) / math.factorial(d)
    # grad lambda_k : rows of inverse of [1 x]^T system
    Mx = np.hstack([np.ones((nv, 1)), verts])  # nv x (d+1)
    G = np.linalg.inv(Mx)[1:, :].T  # nv x d ; G[k] = grad lambda_k
    phis, edges = _p2_basis(d)
    n2 = len(phis)
    lam = []
    for k in range(nv):
        e = [0] * nv
        e[k] = 1
        lam.append({tuple(e): 1.0})
    dphi = [[_poly_dlam(p, k) for k in range(nv)] for p in phis]  # d phi_j / d lambda_k
    GG = G @ G.T
    K2 = np.zeros((n2, n2))
    M2 = np.zeros((n2, n2))
    for i in range(n2):
        for j in range(i, n2):
            s = 0.0
            for k in range(nv):
                for l in range(nv):
                    if GG[k, l] != 0.0 and dphi[i][k] and dphi[j][l]:
                        s += GG[k, l] * _poly_int(_poly_mul(dphi[i][k], dphi[j][l]), d)
            K2[i, j] = K2[j, i] = s * vol
            M2[i, j] = M2[j, i] = _poly_int(_poly_mul(phis[i], phis[j]), d) * vol
    M1 = np.array([[_poly_int(_poly_mul(lam[i], lam[j]), d) for j in range(nv)] for i in range(nv)]) * vol
    K1 = GG * vol
    Bc = np.zeros((d, nv, n2))
    for c in range(d):
        for i in range(nv):
            for j in range(n2):
                s = 0.0
                for k in range(nv):
                    if G[k, c] != 0.0 and dphi[j][k]:
                        s += G[k, c] * _poly_int(_poly_mul(lam[i], dphi[j][k]), d)
                Bc[c, i, j] = s * vol
    T = np.array([[_poly_int(_poly_mul(lam[k], phis[j]), d) for j in range(n2)] for k in range(nv)]) * vol
    I2 = np.array([_poly_int(phis[j], d) for j in range(n2)]) * vol
    return dict(K2=K2, M2=M2, K1=K1, M1=M1, Bc=Bc, T=T, I2=I2, vol=vol, edges=edges)


def kuhn_simplices(d: int):
    """Kuhn (Freudenthal) split of the unit d-cube: one simplex per axis permutation,
    vertices 0, e_p0, e_p0+e_p1, ...  All share the main diagonal; the triangulation is conforming
    under translation."""
    out = []
    for perm in itertools.permutations(range(d)):
        v = np.zeros((d + 1, d), dtype=np.int64)
        for s, ax in enumerate(perm):
            v[s + 1] = v[s]
            v[s + 1, ax] += 1
        out.append(v)
    return out


# ----------------------------------------------------------------------------------------------
@dataclass
class CSR:
    """Plain CSR triple (torch tensors; int64 rowptr, int32 colidx, float64 values)."""
    shape: tuple
    rowptr: torch.Tensor
    colidx: torch.Tensor
    values: torch.Tensor

    @property
    def nnz(self):
        return int(self.values.numel())

    def to_scipy(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.values.cpu().numpy(), self.colidx.cpu().numpy().astype(np.int64),
                              self.rowptr.cpu().numpy()), shape=self.shape)


def _coo_to_csr(rows, cols, vals, shape, row_select=None) -> CSR:
    """Deterministic duplicate summation: stable sort by (row, col) key, then segmented sum in
    sorted order.  ``row_select`` (sorted int64 global row ids) keeps only those rows; the result
    has len(row_select) rows with *global* column indices."""
    nr, nc = shape
    if row_select is not None:
        pos = torch.searchsorted(row_select, rows)
        pos.clamp_(max=row_select.numel() - 1)
        keep = row_select[pos] == rows
        rows, cols, vals = pos[keep], cols[keep], vals[keep]
        nr = int(row_select.numel())
    key = rows * nc + cols
    key, order = torch.sort(key, stable=True)
    vals = vals[order]
    del order
    ukey, counts = torch.unique_consecutive(key, return_counts=True)
    del key
    # segmented sum in sorted order (each segment <= a few dozen entries): loop over the position
    # inside the segment so the floating-point order is fixed.
    starts = torch.cumsum(counts, 0) - counts
    out = vals[starts].clone()
    maxc = int(counts.max().item()) if counts.numel() else 0
    for t in range(1, maxc):
        m = counts > t
        out[m] += vals[starts[m] + t]
    urow = torch.div(ukey, nc, rounding_mode="floor")
    ucol = (ukey - urow * nc).to(torch.int32)
    rowptr = torch.zeros(nr + 1, dtype=torch.int64, device=vals.device)
    rowptr[1:] = torch.cumsum(torch.bincount(urow, minlength=nr), 0)
    return CSR((nr, nc), rowptr, ucol, out)


@dataclass
class StokesSystem:
    """Assembled blocks of one (sub)domain.  Row sets are the *owned* rows (all rows when the whole
    mesh is assembled); column indices are global."""
    dim: int
    cells: tuple
    lengths: tuple
    n2: int            # number of P2 nodes (global)
    n1: int            # number of P1 nodes (global)
    A: CSR             # owned velocity rows x n_u
    B: CSR             # owned pressure rows x n_u
    BT: CSR            # owned velocity rows x n_p
    Mp: CSR            # owned pressure rows x n_p
    f: torch.Tensor    # owned velocity rows
    g: torch.Tensor    # owned pressure rows
    u0: torch.Tensor   # owned velocity rows: Dirichlet data (inlet profile), zero elsewhere
    free_u: torch.Tensor  # bool, owned velocity rows: True where the dof is free
    owned_u: torch.Tensor  # int64 global ids of owned velocity rows (sorted)
    owned_p: torch.Tensor  # int64 global ids of owned pressure rows (sorted)
    extras: dict = field(default_factory=dict)

    @property
    def n_u(self):
        return self.dim * self.n2

    @property
    def n_p(self):
        return self.n1


def _lattice_index(idx, dims):
    """x slowest."""
    out = idx[..., 0]
    for a in range(1, len(dims)):
        out = out * dims[a] + idx[..., a]
    return out


def assemble_stokes(cells, lengths=None, nu=1.0, device="cpu", slab=None, nslabs=1,
                    outlet=True, with_mass2=False, vrange=None, parts=("A", "B", "BT", "Mp")) -> StokesSystem:
    """Taylor-Hood P2/P1 Stokes blocks on a box ``[0,L_0] x ...`` split into ``cells`` cubes, each
    cut into d! Kuhn simplices.

    Velocity Dirichlet data on every boundary face except the outlet face x = L_0 when ``outlet``
    (natural outflow, as in run.py:22-25 'wall|inlet|cyl' vs 'outlet'); the inlet x = 0 carries the
    parabolic profile of the reference scripts.

    ``slab``/``nslabs``: assemble only the rows owned by x-slab ``slab`` of ``nslabs`` (row partition
    of SURVEY.md 8(e)); columns stay global.  ``vrange = (v0, v1)``: the rows of the vertex planes [v0, v1) instead
    (an arbitrary slab, e.g. a rank's rows plus the neighbouring planes).  ``parts``: which blocks to build (the others
    are None).
    """
    d = len(cells)
    cells = tuple(int(c) for c in cells)
    if lengths is None:
        lengths = (1.0,) * d
    lengths = tuple(float(x) for x in lengths)
    h = np.array([lengths[a] / cells[a] for a in range(d)])
    dev = torch.device(device)
    dims2 = tuple(2 * c + 1 for c in cells)
    dims1 = tuple(c + 1 for c in cells)
    n2 = int(np.prod(dims2))
    n1 = int(np.prod(dims1))
    n_u = d * n2

    # ---- ownership: slabs along x of the refined lattice (vertex planes split evenly)
    def vertex_cut(s):
        return (cells[0] + 1) * s // nslabs  # first owned vertex plane of slab s

    if vrange is not None:
        vx0, vx1 = max(int(vrange[0]), 0), min(int(vrange[1]), cells[0] + 1)
        cx0, cx1 = max(vx0 - 1, 0), min(vx1, cells[0])
        slab = -1   # "not the whole mesh"
    elif slab is None:
        vx0, vx1 = 0, cells[0] + 1
        cx0, cx1 = 0, cells[0]
    else:
        vx0, vx1 = vertex_cut(slab), vertex_cut(slab + 1)
        cx0, cx1 = max(vx0 - 1, 0), min(vx1, cells[0])  # cells touching owned nodes
    # refined-lattice planes owned: vertex plane v and the midpoint plane right of it
    px0 = 2 * vx0
    px1 = min(2 * vx1, dims2[0])
    plane2 = int(np.prod(dims2[1:]))
    plane1 = int(np.prod(dims1[1:]))
    own_nodes2 = torch.arange(px0 * plane2, px1 * plane2, dtype=torch.int64, device=dev)
    owned_u = torch.cat([own_nodes2 + c * n2 for c in range(d)])
    owned_p = torch.arange(vx0 * plane1, vx1 * plane1, dtype=torch.int64, device=dev)

    # ---- cells of this slab
    ranges = [torch.arange(cx0, cx1, device=dev, dtype=torch.int64)] + \
             [torch.arange(0, cells[a], device=dev, dtype=torch.int64) for a in range(1, d)]
    cell = torch.stack(torch.meshgrid(*ranges, indexing="ij"), dim=-1).reshape(-1, d)  # ne x d
    ne = cell.shape[0]

    simp = kuhn_simplices(d)
    coo = {k: ([], [], []) for k in ("A1", "B", "BT", "Mp", "M2")}
    f = torch.zeros(n_u, dtype=torch.float64, device=dev) if slab is None else None
    f_rows, f_vals = [], []
    for sv in simp:
        loc = simplex_p2p1_matrices(sv * h)
        edges = loc["edges"]
        svt = torch.as_tensor(sv, device=dev)
        # global vertex coords (coarse lattice) and P2 node coords (refined lattice) of every element
        vtx = cell[:, None, :] + svt[None, :, :]                     # ne x (d+1) x d
        mids = torch.stack([vtx[:, a] + vtx[:, b] for a, b in edges], dim=1)  # refined coords
        nodes2 = torch.cat([2 * vtx, mids], dim=1)                    # ne x nP2 x d
        dof2 = _lattice_index(nodes2, dims2)                          # ne x nP2
        dof1 = _lattice_index(vtx, dims1)                             # ne x (d+1)
        nP2 = dof2.shape[1]
        K2 = torch.as_tensor(loc["K2"] * nu, device=dev)
        if "A" in parts or with_mass2:
            r = dof2[:, :, None].expand(ne, nP2, nP2).reshape(-1)
            c = dof2[:, None, :].expand(ne, nP2, nP2).reshape(-1)
        if "A" in parts:
            coo["A1"][0].append(r); coo["A1"][1].append(c)
            coo["A1"][2].append(K2[None].expand(ne, nP2, nP2).reshape(-1))
        if with_mass2:
            coo["M2"][0].append(r); coo["M2"][1].append(c)
            coo["M2"][2].append(torch.as_tensor(loc["M2"], device=dev)[None].expand(ne, nP2, nP2).reshape(-1))
        for comp in range(d if ("B" in parts or "BT" in parts) else 0):
            Bl = torch.as_tensor(loc["Bc"][comp], device=dev)        # (d+1) x nP2
            r1 = dof1[:, :, None].expand(ne, d + 1, nP2).reshape(-1)
            c1 = (dof2 + comp * n2)[:, None, :].expand(ne, d + 1, nP2).reshape(-1)
            v1 = Bl[None].expand(ne, d + 1, nP2).reshape(-1)
            if "B" in parts:
                coo["B"][0].append(r1); coo["B"][1].append(c1); coo["B"][2].append(v1)
            if "BT" in parts:
                coo["BT"][0].append(c1); coo["BT"][1].append(r1); coo["BT"][2].append(v1)
        if "Mp" in parts:
            M1 = torch.as_tensor(loc["M1"], device=dev)
            coo["Mp"][0].append(dof1[:, :, None].expand(ne, d + 1, d + 1).reshape(-1))
            coo["Mp"][1].append(dof1[:, None, :].expand(ne, d + 1, d + 1).reshape(-1))
            coo["Mp"][2].append(M1[None].expand(ne, d + 1, d + 1).reshape(-1))
        # f = int (x - 0.5) v_1 : x = sum_k x_k lambda_k   (component 1, run.py:92-94)
        T = torch.as_tensor(loc["T"], device=dev)                     # (d+1) x nP2
        xk = vtx[:, :, 0].to(torch.float64) * h[0] - 0.5              # ne x (d+1)
        f_rows.append((dof2 + 1 * n2).reshape(-1))
        f_vals.append((xk @ T).reshape(-1))

    def build(name, shape, rows_sel):
        r = torch.cat(coo[name][0]); c = torch.cat(coo[name][1]); v = torch.cat(coo[name][2])
        coo[name] = None
        return _coo_to_csr(r, c, v, shape, rows_sel)

    whole = slab is None
    extras = {}
    A = None
    if "A" in parts:
        A1 = build("A1", (n2, n2), None if whole else own_nodes2)
        # vector Laplacian: d copies of the scalar block on the diagonal (component-major)
        nr1 = A1.shape[0]
        counts = A1.rowptr[1:] - A1.rowptr[:-1]
        rowptr = torch.zeros(d * nr1 + 1, dtype=torch.int64, device=dev)
        rowptr[1:] = torch.cumsum(counts.repeat(d), 0)
        colidx = torch.cat([(A1.colidx.to(torch.int64) + comp * n2).to(torch.int32) for comp in range(d)])
        A = CSR((d * nr1, n_u), rowptr, colidx, A1.values.repeat(d))
        if with_mass2:
            extras["K2"] = A1
        del A1
    if with_mass2:
        extras["M2"] = build("M2", (n2, n2), None if whole else own_nodes2)
    B = build("B", (n1, n_u), None if whole else owned_p) if "B" in parts else None
    BT = build("BT", (n_u, n1), None if whole else owned_u) if "BT" in parts else None
    Mp = build("Mp", (n1, n1), None if whole else owned_p) if "Mp" in parts else None

    # load vector, deterministic accumulation: sort by row then segmented sum via the CSR helper
    fr = torch.cat(f_rows); fv = torch.cat(f_vals)
    fcsr = _coo_to_csr(fr, torch.zeros_like(fr), fv, (n_u, 1), None if whole else owned_u)
    fvec = torch.zeros(fcsr.shape[0], dtype=torch.float64, device=dev)
    nzrow = (fcsr.rowptr[1:] - fcsr.rowptr[:-1]) > 0
    fvec[nzrow] = fcsr.values

    # ---- boundary data on owned velocity rows
    node = owned_u % n2
    coords = []
    rem = node
    for a in reversed(range(d)):
        coords.append(rem % dims2[a])
        rem = torch.div(rem, dims2[a], rounding_mode="floor")
    coords = coords[::-1]  # refined lattice integer coords per axis
    on_bnd = torch.zeros_like(node, dtype=torch.bool)
    for a in range(d):
        on_bnd |= coords[a] == 0
        if a == 0 and outlet:
            # outlet face x = L: natural, except where it meets the walls
            continue
        on_bnd |= coords[a] == dims2[a] - 1
    free_u = ~on_bnd
    comp = torch.div(owned_u, n2, rounding_mode="floor")
    inlet = (coords[0] == 0) & (comp == 0)
    u0 = torch.zeros(owned_u.numel(), dtype=torch.float64, device=dev)
    y = coords[1].to(torch.float64) * (h[1] / 2)
    Ly = lengths[1]
    if d == 2:
        prof = 1.5 * 4.0 * y * (Ly - y) / (Ly * Ly)
    else:
        z = coords[2].to(torch.float64) * (h[2] / 2)
        Lz = lengths[2]
        prof = 16.0 * y * (Ly - y) * z * (Lz - z) / (Ly * Ly * Lz * Lz)
    u0[inlet] = prof[inlet]
    g = torch.zeros(owned_p.numel(), dtype=torch.float64, device=dev)
    return StokesSystem(d, cells, lengths, n2, n1, A, B, BT, Mp, fvec, g, u0, free_u, owned_u, owned_p, extras)


def jacobi_inverse(csr: CSR, row_ids: torch.Tensor | None = None, free: torch.Tensor | None = None) -> torch.Tensor:
    """1/diag restricted to free dofs (NGSolve's 'local' preconditioner, run.py:61-62).  ``row_ids``
    are the global ids of the rows held (defaults to 0..nrows-1)."""
    nr = csr.shape[0]
    dev = csr.values.device
    if row_ids is None:
        row_ids = torch.arange(nr, dtype=torch.int64, device=dev)
    rows = torch.repeat_interleave(torch.arange(nr, device=dev), csr.rowptr[1:] - csr.rowptr[:-1])
    hit = csr.colidx.to(torch.int64) == row_ids[rows]
    diag = torch.zeros(nr, dtype=torch.float64, device=dev)
    diag[rows[hit]] = csr.values[hit]
    inv = torch.where(diag != 0, 1.0 / diag, torch.zeros_like(diag))
    if free is not None:
        inv = torch.where(free, inv, torch.zeros_like(inv))
    return inv


def stokes_cells_for_dofs(target_dofs: float, aspect=(6, 1, 1)):
    """Cell counts for the 3D channel [0,2.5]x[0,0.41]^2 (templates/NavierStokesSIMPLE_test_3D.py:9-13)
    with about ``target_dofs`` Taylor-Hood unknowns (25 per vertex asymptotically)."""
    verts = target_dofs / 25.0
    base = (verts / float(np.prod(aspect))) ** (1.0 / 3.0)
    return tuple(max(2, int(round(base * a))) for a in aspect)
