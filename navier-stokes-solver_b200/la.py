"""Operator / vector protocol of the device solvers -- the host-side mirror of the slice of
NGSolve's ``la`` layer that the reference's solver call sites use:

    K   = BlockMatrix([[a, b.T], [b, None]])          run.py:45
    C   = BlockMatrix([[preA, None], [None, preS]])   run.py:46
    rhs = BlockVector([f, g]); sol = BlockVector([gfu, gfp])     run.py:47-48
    Preconditioner(m, 'local')                        run.py:62, stokes_hcurldiv.py:53

Operators are *handles to device-resident CSR* (uploaded once per assembly through the C ABI,
include/nsb200.h); vectors stay caller-owned host arrays (numpy, NGSolve vectors through their
numpy view) or device tensors (torch CUDA), and are passed to the solvers by pointer.  Anything that
cannot be turned into CSR / a diagonal is rejected with TypeError -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib

# ------------------------------------------------------------------------------------------------
# context
# ------------------------------------------------------------------------------------------------
_default_ctx = None


class Context:
    """One CUDA device + stream (+ NCCL communicator when row-partitioned)."""

    def __init__(self, device=None, rank=0, nranks=1, unique_id=None):
        lib = _lib.load()
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0")) if nranks > 1 else 0
        self.device, self.rank, self.nranks = int(device), int(rank), int(nranks)
        h = _lib.c_ctx()
        if nranks > 1:
            assert unique_id is not None and len(unique_id) == 128
            _lib.check(lib.nsb_ctx_create_dist(self.device, rank, nranks, unique_id, C.byref(h)), "nsb_ctx_create_dist")
        else:
            _lib.check(lib.nsb_ctx_create(self.device, C.byref(h)), "nsb_ctx_create")
        self.h = h
        self._lib = lib

    @classmethod
    def from_torch_distributed(cls, device=None):
        """Row-partitioned context for the calling rank of the default torch.distributed group: rank 0
        draws the NCCL id through the C ABI, the id travels by broadcast (gloo or nccl)."""
        import torch
        import torch.distributed as dist
        rank, n = dist.get_rank(), dist.get_world_size()
        if n == 1:
            return cls(device=device)
        lib = _lib.load()
        buf = C.create_string_buffer(128)
        if rank == 0:
            _lib.check(lib.nsb_nccl_unique_id(buf), "nsb_nccl_unique_id")
        t = torch.tensor(list(buf.raw), dtype=torch.uint8)
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().tolist())
        return cls(device=device, rank=rank, nranks=n, unique_id=uid)

    def sync(self):
        _lib.check(self._lib.nsb_ctx_sync(self.h), "nsb_ctx_sync")

    def launch_count(self):
        return int(self._lib.nsb_ctx_launch_count(self.h))

    def sync_stats(self, reset=False):
        """Diagnostic (library built with -DNSB_SYNC_STATS, else zeros): (nanoseconds, polls) the exchange kernels of
        this rank spent waiting on peer flags, by kind."""
        out = (C.c_uint64 * 16)()
        _lib.check(self._lib.nsb_ctx_sync_stats(self.h, out, int(bool(reset))), "nsb_ctx_sync_stats")
        names = ["allreduce", "push_wait_consumed", "spmv_wait_halo", "wait_copy"]
        return {n: {"ns": int(out[2 * i]), "count": int(out[2 * i + 1])} for i, n in enumerate(names)}

    def stream(self):
        """Raw ``cudaStream_t`` every kernel of this context is launched on (for CUDA-event timing from the
        caller: ``torch.cuda.ExternalStream(ctx.stream())``)."""
        return int(self._lib.nsb_ctx_stream(self.h) or 0)

    def set_graphs(self, enabled):
        _lib.check(self._lib.nsb_ctx_set_graphs(self.h, int(bool(enabled))), "nsb_ctx_set_graphs")

    def close(self):
        if getattr(self, "h", None):
            self._lib.nsb_ctx_destroy(self.h)
            self.h = None


def default_context() -> Context:
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context()
    return _default_ctx


def set_default_context(ctx):
    global _default_ctx
    _default_ctx = ctx


# ------------------------------------------------------------------------------------------------
# raw buffers
# ------------------------------------------------------------------------------------------------
def _is_torch(x):
    return type(x).__module__.startswith("torch") and hasattr(x, "data_ptr")


def buffer_of(x, dtype=np.float64):
    """(pointer, length, keepalive) of a flat contiguous array: numpy, torch (CPU or CUDA), or a
    vector of this module / NGSolve (through ``FV().NumPy()``)."""
    if isinstance(x, BaseVector):
        x = x._a
    elif not isinstance(x, np.ndarray) and not _is_torch(x) and hasattr(x, "FV"):
        x = x.FV().NumPy()      # NGSolve BaseVector: zero-copy numpy view
    if _is_torch(x):
        import torch
        want = {np.float64: torch.float64, np.int32: torch.int32, np.int64: torch.int64, np.uint8: torch.uint8}[dtype]
        if x.dtype != want or not x.is_contiguous():
            raise TypeError("tensor must be contiguous %s" % want)
        if x.is_cuda:
            # the library runs on its own stream: work torch queued on this tensor must have landed
            torch.cuda.current_stream(x.device).synchronize()
        return x.data_ptr(), x.numel(), x
    if isinstance(x, np.ndarray):
        if x.dtype != dtype or not x.flags.c_contiguous:
            raise TypeError("array must be C-contiguous %s (got %s)" % (np.dtype(dtype), x.dtype))
        return x.ctypes.data, x.size, x
    raise TypeError("cannot take a device-solver buffer from %r" % (type(x),))


# ------------------------------------------------------------------------------------------------
# vectors (host mirrors of NGSolve BaseVector / BlockVector)
# ------------------------------------------------------------------------------------------------
class BaseVector:
    def __init__(self, n_or_array):
        if isinstance(n_or_array, np.ndarray):
            assert n_or_array.dtype == np.float64
            self._a = n_or_array
        elif _is_torch(n_or_array):
            self._a = n_or_array
        else:
            self._a = np.zeros(int(n_or_array))

    def CreateVector(self):
        return BaseVector(len(self))

    def Copy(self):
        return BaseVector(self._a.clone() if _is_torch(self._a) else self._a.copy())

    def FV(self):
        return self

    def NumPy(self):
        return self._a

    def __len__(self):
        return int(self._a.numel()) if _is_torch(self._a) else len(self._a)

    @property
    def size(self):
        return len(self)

    @property
    def data(self):
        return self

    @data.setter
    def data(self, other):
        if other is self:
            return
        self._a[:] = other._a if isinstance(other, BaseVector) else other

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return BaseVector(self._a[idx])
        return float(self._a[idx])

    def __setitem__(self, idx, val):
        self._a[idx] = val._a if isinstance(val, BaseVector) else val


class BlockVector(BaseVector):
    """Components are views into one flat (u | p) array, as the C ABI expects; the component vectors
    passed in are re-homed onto that storage so the caller's grid-function vectors see the result
    (in-place ``sol`` semantics, run.py:48,107-109)."""

    def __init__(self, comps):
        comps = [c if isinstance(c, BaseVector) else BaseVector(np.ascontiguousarray(c, dtype=np.float64))
                 for c in comps]
        sizes = [len(c) for c in comps]
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(int)
        if any(_is_torch(c._a) for c in comps):
            import torch
            flat = torch.cat([c._a for c in comps])
        else:
            flat = np.concatenate([c._a for c in comps]) if comps else np.zeros(0)
        super().__init__(flat)
        self._offs = offs
        self._comps = comps
        for i, c in enumerate(comps):
            c._a = self._a[offs[i]:offs[i + 1]]

    def CreateVector(self):
        return BlockVector([BaseVector(int(self._offs[i + 1] - self._offs[i])) for i in range(len(self._comps))])

    @property
    def nblocks(self):
        return len(self._comps)

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return BaseVector(self._a[idx])
        return self._comps[idx]


def InnerProduct(a, b):
    return float((a._a * b._a).sum())


def Norm(a):
    return float((a._a * a._a).sum()) ** 0.5


# ------------------------------------------------------------------------------------------------
# operators
# ------------------------------------------------------------------------------------------------
class BaseMatrix:
    """Operator protocol subset: height/width, .T, Create*Vector, ``mat * vec`` (device SpMV)."""

    @property
    def height(self):
        raise NotImplementedError

    @property
    def width(self):
        raise NotImplementedError

    def CreateColVector(self):
        return BaseVector(self.height)

    def CreateRowVector(self):
        return BaseVector(self.width)

    def CreateVector(self):
        return self.CreateColVector()


class SparseMatrix(BaseMatrix):
    """Device-resident fp64 CSR (int32 column indices).  Stands where ``BilinearForm.mat`` stands at the
    reference call sites (run.py:105-106, stokes_hcurldiv.py:76)."""

    def __init__(self, rowptr, colidx, values, shape, ctx: Context | None = None):
        self.ctx = ctx or default_context()
        lib = self.ctx._lib
        rows, cols = int(shape[0]), int(shape[1])
        bits = 64
        if isinstance(rowptr, np.ndarray):
            if rowptr.dtype == np.int32:
                bits = 32
            elif rowptr.dtype != np.int64:
                rowptr = rowptr.astype(np.int64)
            rowptr = np.ascontiguousarray(rowptr)
            colidx = np.ascontiguousarray(colidx, dtype=np.int32)
            values = np.ascontiguousarray(values, dtype=np.float64)
        elif _is_torch(rowptr):
            import torch
            bits = 32 if rowptr.dtype == torch.int32 else 64
        rp, nrp, k1 = buffer_of(rowptr, np.int32 if bits == 32 else np.int64)
        ci, nnz, k2 = buffer_of(colidx, np.int32)
        va, nv, k3 = buffer_of(values, np.float64)
        if nrp != rows + 1 or nv != nnz:
            raise ValueError("inconsistent CSR arrays")
        h = _lib.c_mat()
        _lib.check(lib.nsb_csr_create(self.ctx.h, rows, cols, nnz, rp, bits, ci, va, C.byref(h)), "nsb_csr_create")
        self.h = h
        self.shape = (rows, cols)
        self.nnz = int(nnz)
        self._T = None
        self._halo = None

    @classmethod
    def _wrap(cls, h, ctx):
        self = cls.__new__(cls)
        self.ctx, self.h = ctx, h
        r, c_, z = C.c_int64(), C.c_int64(), C.c_int64()
        _lib.check(ctx._lib.nsb_mat_info(h, C.byref(r), C.byref(c_), C.byref(z)), "nsb_mat_info")
        self.shape, self.nnz, self._T, self._halo = (r.value, c_.value), z.value, None, None
        return self

    @classmethod
    def from_scipy(cls, m, ctx=None):
        m = m.tocsr()
        m.sort_indices()
        return cls(m.indptr, m.indices, m.data, m.shape, ctx)

    @classmethod
    def from_csr(cls, csr, ctx=None):
        """From the synthetic assembler's CSR triple (torch tensors on CPU or on the GPU)."""
        return cls(csr.rowptr, csr.colidx, csr.values, csr.shape, ctx)

    @classmethod
    def from_ngsolve(cls, mat, ctx=None):
        """From an assembled NGSolve SparseMatrix (``bf.mat``): one CSR export per assembly."""
        vals, cols, rowptr = mat.CSR()
        return cls(np.asarray(rowptr, dtype=np.int64), np.asarray(cols, dtype=np.int32),
                   np.asarray(vals, dtype=np.float64), (mat.height, mat.width), ctx)

    @property
    def height(self):
        return self.shape[0]

    @property
    def width(self):
        return self.shape[1]

    @property
    def T(self):
        return self.CreateTranspose()

    def CreateBlockSmoother(self, blocks, pattern=None, gauss_seidel=False):
        """``a.mat.CreateBlockSmoother(jacblocks)`` (templates/NavierStokesSIMPLE_iterative.py:373)."""
        return BlockSmoother(self, blocks, pattern, gauss_seidel)

    def CreateTranspose(self):
        """Explicit device transpose, built once and cached (solvers/bramblepasciak_new.py:198)."""
        if self._T is None:
            h = _lib.c_mat()
            _lib.check(self.ctx._lib.nsb_csr_transpose(self.ctx.h, self.h, C.byref(h)), "nsb_csr_transpose")
            self._T = SparseMatrix._wrap(h, self.ctx)
            self._T._T = self
        return self._T

    def set_transpose(self, t):
        """Register an already uploaded transpose (row-partitioned operators upload both)."""
        self._T, t._T = t, self

    def format(self):
        """Which SpMV kernel this operator runs: dict(rows_per_chunk, chunks, max_distinct, total_distinct,
        lanes_per_row, stored_nnz, column_runs); rows_per_chunk == 0 means the plain CSR-vector kernel, > 0 the SELL-LX kernel."""
        a, b, c_, t = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        d, e, f = C.c_int64(), C.c_int64(), C.c_int64()
        _lib.check(self.ctx._lib.nsb_mat_format(self.h, C.byref(a), C.byref(b), C.byref(c_), C.byref(d), C.byref(t),
                                                C.byref(e), C.byref(f)), "nsb_mat_format")
        return dict(rows_per_chunk=a.value, chunks=b.value, max_distinct=c_.value, total_distinct=d.value,
                    lanes_per_row=t.value, stored_nnz=e.value, column_runs=f.value)

    def set_row_panels(self, bounds):
        """Give row ranges [bounds[k], bounds[k+1]) their own SpMV storage shape (stacked saddle-point rows)."""
        b = np.ascontiguousarray(bounds, dtype=np.int64)
        _lib.check(self.ctx._lib.nsb_mat_set_row_panels(self.ctx.h, self.h, len(b), b.ctypes.data), "nsb_mat_set_row_panels")

    def diagonal(self):
        out = np.empty(self.height)
        _lib.check(self.ctx._lib.nsb_mat_diagonal(self.ctx.h, self.h, out.ctypes.data), "nsb_mat_diagonal")
        return out

    def __mul__(self, v):
        """``mat * vec`` evaluated on the device (not lazy)."""
        n_in = self._halo.n_owned if self._halo is not None else self.width
        px, nx, keep = buffer_of(v)
        if nx != n_in:
            raise ValueError("vector length %d != %d" % (nx, n_in))
        if _is_torch(keep):
            import torch
            out = torch.empty(self.height, dtype=torch.float64, device=keep.device)
        else:
            out = np.empty(self.height)
        py, _, _ = buffer_of(out)
        _lib.check(self.ctx._lib.nsb_spmv(self.ctx.h, self.h, px, py), "nsb_spmv")
        return BaseVector(out)

    def spmv_bench(self, warmup=20, reps=200):
        ms = C.c_double()
        _lib.check(self.ctx._lib.nsb_spmv_bench(self.ctx.h, self.h, warmup, reps, C.byref(ms)), "nsb_spmv_bench")
        return ms.value

    def spmv_bytes(self):
        """Algorithmic bytes of one SpMV: 12 nnz + 4 (r+1) + 8 c + 8 r   (SURVEY.md 8(d))."""
        r, c_ = self.shape
        return 12 * self.nnz + 4 * (r + 1) + 8 * c_ + 8 * r

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_mat_destroy(self.h)
        except Exception:
            pass


class CondensedOperator(BaseMatrix):
    """matA = (I - HT)(A + inner_matrix)(I - H) of a statically condensed BilinearForm
    (myAmatrix, solvers/bramblepasciak_new.py:84-103).  Holds references to the five device matrices.  Row-partitioned: this
    rank's rows of each, over one shared velocity halo plan (partition.build_distributed_condensed)."""

    def __init__(self, mat, harmonic_extension, harmonic_extension_trans, inner_solve, inner_matrix):
        parts = (mat, harmonic_extension, harmonic_extension_trans, inner_solve, inner_matrix)
        for m in parts:
            if not isinstance(m, SparseMatrix):
                raise TypeError("CondensedOperator needs device SparseMatrix parts (export the element matrices as CSR); "
                                "got %r" % (type(m),))
        self.parts = parts
        self.ctx = mat.ctx
        h = C.c_void_p()
        _lib.check(self.ctx._lib.nsb_cond_create(self.ctx.h, *[m.h for m in parts], C.byref(h)), "nsb_cond_create")
        self.h = h

    @classmethod
    def from_bilinear_form(cls, blfA):
        """From an object with the attributes of a condensed NGSolve BilinearForm (``.mat``, ``.harmonic_extension``,
        ``.harmonic_extension_trans``, ``.inner_solve``, ``.inner_matrix``) holding device matrices."""
        return cls(blfA.mat, blfA.harmonic_extension, blfA.harmonic_extension_trans, blfA.inner_solve, blfA.inner_matrix)

    @property
    def height(self):
        return self.parts[0].height

    @property
    def width(self):
        return self.parts[0].width

    def __mul__(self, v):
        px, nx, keep = buffer_of(v)
        if nx != self.height:      # row-partitioned: the owned entries; the chain exchanges the halo itself (csrc/cond.cuh)
            raise ValueError("vector length %d != %d" % (nx, self.height))
        out = keep.new_empty(self.height) if _is_torch(keep) else np.empty(self.height)
        py, _, _ = buffer_of(out)
        _lib.check(self.ctx._lib.nsb_cond_apply(self.ctx.h, self.h, px, py), "nsb_cond_apply")
        return BaseVector(out)

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_cond_destroy(self.h)
        except Exception:
            pass


class HaloPlan:
    """Device copy of a partition.Plan (send lists, receive counts) -- SURVEY.md 8(e)."""

    def __init__(self, ctx, n_owned, n_halo, neighbors, send_counts, send_idx, recv_counts):
        self.ctx = ctx
        self.n_owned, self.n_halo = int(n_owned), int(n_halo)
        nb = np.ascontiguousarray(neighbors, dtype=np.int32)
        sc = np.ascontiguousarray(send_counts, dtype=np.int64)
        si = np.ascontiguousarray(send_idx, dtype=np.int32)
        rc = np.ascontiguousarray(recv_counts, dtype=np.int64)
        h = _lib.c_halo()
        _lib.check(ctx._lib.nsb_halo_create(ctx.h, self.n_owned, self.n_halo, len(nb), nb.ctypes.data, sc.ctypes.data,
                                           si.ctypes.data, rc.ctypes.data, C.byref(h)), "nsb_halo_create")
        self.h = h

    def attach(self, mat: SparseMatrix):
        _lib.check(self.ctx._lib.nsb_mat_set_halo(mat.h, self.h), "nsb_mat_set_halo")
        mat._halo = self

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_halo_destroy(self.h)
        except Exception:
            pass


class DiagonalPreconditioner(BaseMatrix):
    """y = d .* x on the device: NGSolve's 'local' (Jacobi) preconditioner restricted to the free dofs."""

    def __init__(self, h, n, ctx):
        self.h, self.n, self.ctx = h, int(n), ctx

    @classmethod
    def from_diagonal(cls, d, ctx=None):
        ctx = ctx or default_context()
        p, n, keep = buffer_of(d if not isinstance(d, np.ndarray) else np.ascontiguousarray(d, dtype=np.float64))
        h = _lib.c_pre()
        _lib.check(ctx._lib.nsb_pre_create_diag(ctx.h, n, p, C.byref(h)), "nsb_pre_create_diag")
        return cls(h, n, ctx)

    @classmethod
    def jacobi(cls, mat: SparseMatrix, freedofs=None):
        mask_p = None
        keep = None
        if freedofs is not None:
            if _is_torch(freedofs):
                import torch
                keep = freedofs.to(torch.uint8).contiguous()
                mask_p = keep.data_ptr()
            else:
                keep = np.ascontiguousarray(np.asarray(freedofs, dtype=bool).astype(np.uint8))
                mask_p = keep.ctypes.data
            if len(keep) != mat.height:
                raise ValueError("freedofs length mismatch")
        h = _lib.c_pre()
        _lib.check(mat.ctx._lib.nsb_pre_create_jacobi(mat.ctx.h, mat.h, mask_p, C.byref(h)), "nsb_pre_create_jacobi")
        return cls(h, mat.height, mat.ctx)

    @property
    def height(self):
        return self.n

    @property
    def width(self):
        return self.n

    def __mul__(self, v):
        px, nx, keep = buffer_of(v)
        out = np.empty(self.n) if not _is_torch(keep) else keep.new_empty(self.n)
        py, _, _ = buffer_of(out)
        _lib.check(self.ctx._lib.nsb_pre_apply(self.ctx.h, self.h, px, py), "nsb_pre_apply")
        return BaseVector(out)

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_pre_destroy(self.h)
        except Exception:
            pass


def color_blocks(pattern, blocks):
    """Greedy colouring of dof blocks: two blocks get different colours when ``pattern`` (scipy sparse, the matrix
    the smoother belongs to) couples any of their dofs.  Returns (order, color_ptr): the blocks in colour-major order
    (stable inside a colour) and the block ranges per colour."""
    import scipy.sparse as sp
    nb = len(blocks)
    n = pattern.shape[0]
    sizes = np.array([len(b) for b in blocks], dtype=np.int64)
    dofs = np.concatenate([np.asarray(b, dtype=np.int64) for b in blocks]) if nb else np.zeros(0, np.int64)
    owner = np.repeat(np.arange(nb), sizes)
    if len(np.unique(dofs)) != len(dofs):
        raise ValueError("block smoother: blocks must be disjoint")
    P = sp.csr_matrix((np.ones(len(dofs)), (dofs, owner)), shape=(n, nb))
    pat = sp.csr_matrix(pattern, copy=True)
    pat.data[:] = 1.0
    G = (P.T @ pat @ P).tocsr()
    color = -np.ones(nb, dtype=np.int64)
    for b in range(nb):
        nbr = G.indices[G.indptr[b]:G.indptr[b + 1]]
        used = set(color[nbr[nbr != b]].tolist())
        k = 0
        while k in used:
            k += 1
        color[b] = k
    ncol = int(color.max()) + 1 if nb else 0
    order = np.argsort(color, kind="stable")
    color_ptr = np.concatenate([[0], np.cumsum(np.bincount(color, minlength=ncol))]).astype(np.int64)
    return order, color_ptr


class BlockSmoother(BaseMatrix):
    """``a.mat.CreateBlockSmoother(blocks)`` on the device (templates/NavierStokesSIMPLE_iterative.py:360-373).

    ``blocks``: list of disjoint dof lists (the facet blocks of :360-362).  With ``pattern`` (the scipy matrix, used for
    colouring only) the sweeps ``Smooth`` / ``SmoothBack`` are available; they equal the sequential block Gauss-Seidel
    sweep over ``self.blocks`` (the colour-major reordering of the input, kept for the caller).  As a preconditioner the
    object acts as ``jacobi * x`` (``gauss_seidel=False``, :383) or as ``y=0; Smooth(y,x); SmoothBack(y,x)``
    (``gauss_seidel=True``, MypreA.Mult :375-381 without the NGSolve-internal auxiliary-space term)."""

    def __init__(self, mat, blocks, pattern=None, gauss_seidel=False, color_ptr=None):
        if not isinstance(mat, SparseMatrix):
            raise TypeError("BlockSmoother needs a device SparseMatrix")
        self.ctx, self.mat, self.n = mat.ctx, mat, mat.height
        if color_ptr is not None:
            # row-partitioned (partition.localize_blocks): this rank's blocks in LOCAL numbering, already colour-major under a
            # colouring of the GLOBAL block graph; every rank passes every colour (possibly empty here), so no block is dropped
            blocks = [np.asarray(b, dtype=np.int32) for b in blocks]
            color_ptr = np.ascontiguousarray(color_ptr, dtype=np.int64)
            if any(len(b) == 0 for b in blocks) or color_ptr[0] != 0 or color_ptr[-1] != len(blocks):
                raise ValueError("BlockSmoother: colour table does not cover the (non-empty) blocks")
        else:
            if getattr(mat, "_halo", None) is not None:
                raise ValueError("a row-partitioned BlockSmoother needs the global colouring: pass the blocks and color_ptr of "
                                 "partition.localize_blocks")
            blocks = [np.asarray(b, dtype=np.int32) for b in blocks if len(b) > 0]
        if color_ptr is not None:
            pass
        elif pattern is not None:
            order, color_ptr = color_blocks(pattern, blocks)
            blocks = [blocks[i] for i in order]
        else:
            if gauss_seidel:
                raise ValueError("Gauss-Seidel sweeps need `pattern` for the colouring")
            color_ptr = np.zeros(1, dtype=np.int64)
        self.blocks = blocks
        self.color_ptr = color_ptr
        bptr = np.concatenate([[0], np.cumsum([len(b) for b in blocks])]).astype(np.int64)
        bd = np.concatenate(blocks).astype(np.int32) if blocks else np.zeros(1, np.int32)
        h = _lib.c_pre()
        ncol = len(color_ptr) - 1
        _lib.check(self.ctx._lib.nsb_pre_create_blocks(self.ctx.h, mat.h, len(blocks), bptr.ctypes.data, bd.ctypes.data, ncol,
                                                       color_ptr.ctypes.data if ncol else None, int(bool(gauss_seidel)),
                                                       C.byref(h)), "nsb_pre_create_blocks")
        self.h = h

    @property
    def height(self):
        return self.n

    @property
    def width(self):
        return self.n

    def __mul__(self, v):
        px, nx, keep = buffer_of(v)
        out = np.empty(self.n) if not _is_torch(keep) else keep.new_empty(self.n)
        py, _, _ = buffer_of(out)
        _lib.check(self.ctx._lib.nsb_pre_apply(self.ctx.h, self.h, px, py), "nsb_pre_apply")
        return BaseVector(out)

    def Smooth(self, y, x):
        """One forward block Gauss-Seidel sweep on ``y`` for the right-hand side ``x`` (:377)."""
        px, _, _ = buffer_of(x)
        py, _, _ = buffer_of(y)
        _lib.check(self.ctx._lib.nsb_pre_smooth(self.ctx.h, self.h, px, py, 0), "nsb_pre_smooth")

    def SmoothBack(self, y, x):
        """One backward sweep (:381)."""
        px, _, _ = buffer_of(x)
        py, _, _ = buffer_of(y)
        _lib.check(self.ctx._lib.nsb_pre_smooth(self.ctx.h, self.h, px, py, 1), "nsb_pre_smooth")

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_pre_destroy(self.h)
        except Exception:
            pass


class AuxiliarySpacePreconditioner(BaseMatrix):
    """``MypreA(space, a, jacblocks, GS)`` of templates/NavierStokesSIMPLE_iterative.py:364-391 on the device:

        GS=True   y = 0; jacobi.Smooth(y, x); t = x - a.mat*y; y += (transform @ coarse @ transform.T)*t; jacobi.SmoothBack(y, x)
        GS=False  y = ((transform @ coarse @ transform.T) + jacobi) * x

    ``smoother``: a BlockSmoother of ``a.mat`` (built with ``pattern`` when GS); ``transform``: device SparseMatrix
    (X2.ndof x H1.ndof); ``coarse``: any device preconditioner of the auxiliary H1 problem -- the reference puts
    component-wise h1amg there (:310-357), which is NGSolve-internal; a Jacobi / block smoother of the exported H1 matrix
    (or another composite) takes its place."""

    def __init__(self, smoother, transform, coarse, GS=True, transform_t=None):
        if not isinstance(smoother, BlockSmoother):
            raise TypeError("AuxiliarySpacePreconditioner needs a BlockSmoother")
        if not isinstance(transform, SparseMatrix):
            raise TypeError("AuxiliarySpacePreconditioner needs `transform` as a device SparseMatrix")
        if not hasattr(coarse, "h"):
            raise TypeError("the coarse preconditioner must be a device preconditioner (Jacobi, BlockSmoother, ...)")
        self.ctx, self.n = smoother.ctx, smoother.n
        self.smoother, self.transform, self.coarse = smoother, transform, coarse
        if transform_t is None and getattr(transform, "_halo", None) is not None:
            raise ValueError("row-partitioned: pass transform_t (this rank's rows of the global transpose, partition.distribute_operators)")
        self.transform_t = transform_t if transform_t is not None else transform.CreateTranspose()
        h = _lib.c_pre()
        _lib.check(self.ctx._lib.nsb_pre_create_auxspace(self.ctx.h, smoother.h, transform.h, self.transform_t.h, coarse.h,
                                                         int(bool(GS)), C.byref(h)), "nsb_pre_create_auxspace")
        self.h = h

    height = property(lambda self: self.n)
    width = property(lambda self: self.n)

    def __mul__(self, v):
        px, nx, keep = buffer_of(v)
        out = np.empty(self.n) if not _is_torch(keep) else keep.new_empty(self.n)
        py, _, _ = buffer_of(out)
        _lib.check(self.ctx._lib.nsb_pre_apply(self.ctx.h, self.h, px, py), "nsb_pre_apply")
        return BaseVector(out)

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.ctx.h:
                self.ctx._lib.nsb_pre_destroy(self.h)
        except Exception:
            pass


def Preconditioner(mat, type="local", freedofs=None, **kwargs):
    """``Preconditioner(bf, 'local')`` (run.py:62).  Only the Jacobi type can be built from CSR; BDDC,
    h1amg and friends are NGSolve-internal and are refused rather than emulated."""
    if type != "local":
        raise NotImplementedError("preconditioner type %r lives inside NGSolve and cannot be rebuilt from CSR; "
                                  "only 'local' (Jacobi) is available on the device" % (type,))
    if not isinstance(mat, SparseMatrix):
        raise TypeError("Preconditioner needs a device SparseMatrix")
    return DiagonalPreconditioner.jacobi(mat, freedofs)


class IdentityMatrix(BaseMatrix):
    def __init__(self, n=None):
        self.n = -1 if n is None else int(n)

    @property
    def height(self):
        return self.n

    @property
    def width(self):
        return self.n


class BlockMatrix(BaseMatrix):
    """2x2 block operator.  Recognised shapes: the saddle-point matrix [[A, B^T], [B, C|None]] (merged
    into one device CSR on first use) and the block-diagonal preconditioner [[preA, None], [None, preS]]."""

    def __init__(self, rows):
        if len(rows) != 2 or any(len(r) != 2 for r in rows):
            raise TypeError("device BlockMatrix supports 2x2 block structure only")
        self.rows = rows
        self._K = None
        self._P = None

    def _first(self, items, attr):
        for m in items:
            if m is not None and getattr(m, attr) >= 0:
                return getattr(m, attr)
        raise ValueError("block row/column without a sized block")

    @property
    def height(self):
        return sum(self._first(r, "height") for r in self.rows)

    @property
    def width(self):
        return sum(self._first([self.rows[0][j], self.rows[1][j]], "width") for j in range(2))

    def CreateColVector(self):
        return BlockVector([BaseVector(self._first(r, "height")) for r in self.rows])

    def is_saddle(self):
        (a, bt), (b, c) = self.rows
        return isinstance(a, SparseMatrix) and isinstance(b, SparseMatrix) and isinstance(bt, SparseMatrix) \
            and (c is None or isinstance(c, SparseMatrix))

    def saddle(self) -> SparseMatrix:
        if self._K is None:
            if not self.is_saddle():
                raise TypeError("BlockMatrix is not of the form [[A, B.T], [B, C]] with device SparseMatrix blocks")
            (a, bt), (b, c) = self.rows
            ctx = a.ctx
            h = _lib.c_mat()
            _lib.check(ctx._lib.nsb_saddle_create(ctx.h, a.h, b.h, bt.h, c.h if c is not None else None, C.byref(h)),
                       "nsb_saddle_create")
            self._K = SparseMatrix._wrap(h, ctx)
        return self._K

    def blockdiag_preconditioner(self):
        """[[pre0, None], [None, pre1]] (run.py:46) of device preconditioners: two Jacobi blocks merge into one diagonal, a block
        smoother / auxiliary-space composite in either block gives a block-by-block composite."""
        if self._P is None:
            (p0, z01), (z10, p1) = self.rows
            kinds = (DiagonalPreconditioner, BlockSmoother, AuxiliarySpacePreconditioner)
            if z01 is not None or z10 is not None or not isinstance(p0, kinds) or not isinstance(p1, kinds):
                raise TypeError("block preconditioner must be [[pre0, None], [None, pre1]] of device preconditioners (Jacobi, block "
                                "smoother, auxiliary-space composite); there is no CPU fallback for general operators")
            ctx = p0.ctx
            h = _lib.c_pre()
            _lib.check(ctx._lib.nsb_pre_create_blockdiag2(ctx.h, p0.h, p1.h, C.byref(h)), "nsb_pre_create_blockdiag2")
            self._P = DiagonalPreconditioner(h, p0.n + p1.n, ctx)
            self._P._blocks = (p0, p1)      # borrowed by the composite handle
        return self._P

    def __mul__(self, v):
        return self.saddle() * v
