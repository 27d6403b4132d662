"""Build and bind oracle/c/krylov_omp.c (gcc -O3 -fopenmp) -> oracle/c/libkrylov_omp.so.  TEST INFRASTRUCTURE:
only tests/, __graft_entry__ and bench.py's CPU legs use it."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "c", "krylov_omp.c")
LIB = os.path.join(HERE, "c", "libkrylov_omp.so")
_lib = None


def build():
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        cmd = ["gcc", "-O3", "-march=native", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-o", LIB, SRC, "-lm"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("gcc failed:\n" + res.stderr)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.ko_num_threads.restype = C.c_int
    return _lib


def _csr(m):
    m = m.tocsr()
    m.sort_indices()
    return (np.ascontiguousarray(m.indptr, dtype=np.int64), np.ascontiguousarray(m.indices, dtype=np.int32),
            np.ascontiguousarray(m.data, dtype=np.float64))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def num_threads():
    return int(load().ko_num_threads())


def set_threads(n=None):
    """Use ``n`` OpenMP threads (default: every core this process may run on), whatever OMP_NUM_THREADS the launcher
    exported (torch.distributed.run sets it to 1).  Returns the count now in force."""
    if n is None:
        try:
            n = len(os.sched_getaffinity(0))
        except AttributeError:
            n = os.cpu_count() or 1
    load().ko_set_threads(C.c_int(int(n)))
    return num_threads()


class Blocks:
    """CSR triples of A, B, B^T converted once (the conversion is not part of any timed region)."""

    def __init__(self, A, B, BT=None):
        self.n_u, self.n_p = A.shape[0], B.shape[0]
        self.A, self.B = _csr(A), _csr(B)
        self.BT = _csr(BT if BT is not None else B.T)

    def args(self):
        out = [C.c_int64(self.n_u), C.c_int64(self.n_p)]
        for t in (self.A, self.B, self.BT):
            out += [_p(t[0]), _p(t[1]), _p(t[2])]
        return out


def spmv(m, x):
    rp, ci, va = _csr(m)
    y = np.empty(m.shape[0])
    x = np.ascontiguousarray(x, dtype=np.float64)
    load().ko_spmv(C.c_int64(m.shape[0]), C.c_int64(m.shape[1]), _p(rp), _p(ci), _p(va), _p(x), _p(y))
    return y


def bpcg_v2(blocks: Blocks, f, g, dA, dM, k, sol, tol=1e-6, maxsteps=100, initialize=True, rel_err=True):
    """solvers/bramblepasciak_new.py:24-253 -> (it, history, err0, loop_seconds); sol mutated in place."""
    hist = np.zeros(maxsteps + 1)
    nh, err0, secs = C.c_int(), C.c_double(), C.c_double()
    f, g = np.ascontiguousarray(f, dtype=np.float64), np.ascontiguousarray(g, dtype=np.float64)
    dA, dM = np.ascontiguousarray(dA, dtype=np.float64), np.ascontiguousarray(dM, dtype=np.float64)
    lib = load()
    lib.ko_bpcg_v2.restype = C.c_int
    it = lib.ko_bpcg_v2(*blocks.args(), _p(dA), _p(dM), C.c_double(k), _p(f), _p(g), _p(sol), C.c_double(tol), C.c_int(maxsteps),
                        C.c_int(int(initialize)), C.c_int(int(rel_err)), _p(hist), C.byref(nh), C.byref(err0), C.byref(secs))
    return it, hist[:nh.value].copy(), err0.value, secs.value


def minres(blocks: Blocks, dC, rhs, sol, tol=1e-7, maxsteps=100, initialize=True):
    """minres.py:12-149 -> (errors, converged, loop_seconds); sol mutated in place."""
    errors = np.zeros(maxsteps + 2)
    conv, secs = C.c_int(), C.c_double()
    dC, rhs = np.ascontiguousarray(dC, dtype=np.float64), np.ascontiguousarray(rhs, dtype=np.float64)
    lib = load()
    lib.ko_minres.restype = C.c_int
    n = lib.ko_minres(*blocks.args(), _p(dC), _p(rhs), _p(sol), C.c_double(tol), C.c_int(maxsteps), C.c_int(int(initialize)),
                      _p(errors), C.byref(conv), C.byref(secs))
    return errors[:n + 1].copy(), bool(conv.value), secs.value


class _CsrT(C.Structure):
    _fields_ = [("rows", C.c_int64), ("cols", C.c_int64), ("rowptr", C.c_void_p), ("col", C.c_void_p), ("val", C.c_void_p)]


def _csr_t(m, keep):
    if m is None:
        return _CsrT(-1, -1, None, None, None)
    if isinstance(m, tuple):      # (shape, rowptr int64, col int32, val float64) already in the C layout
        shape, rp, ci, va = m
        t = (np.ascontiguousarray(rp, dtype=np.int64), np.ascontiguousarray(ci, dtype=np.int32), np.ascontiguousarray(va, dtype=np.float64))
        keep.append(t)
        return _CsrT(shape[0], shape[1], t[0].ctypes.data, t[1].ctypes.data, t[2].ctypes.data)
    t = _csr(m)
    keep.append(t)
    return _CsrT(m.shape[0], m.shape[1], t[0].ctypes.data, t[1].ctypes.data, t[2].ctypes.data)


def pcg(A, dinv, rhs, tol=1e-8, maxsteps=200):
    """NGSolve CGSolver from u = 0 with a diagonal preconditioner -> (u, iterations)."""
    keep = []
    a = _csr_t(A, keep)
    u = np.zeros(A.shape[0])
    dinv, rhs = np.ascontiguousarray(dinv, dtype=np.float64), np.ascontiguousarray(rhs, dtype=np.float64)
    lib = load()
    lib.ko_pcg.restype = C.c_int
    it = lib.ko_pcg(C.byref(a), _p(dinv), _p(rhs), _p(u), C.c_double(tol), C.c_int(maxsteps))
    return u, it


class SimpleOps:
    """The ten operators of DoTimeStep / Project converted once (a, mstar, mH, mHT, m_inner, bproj, aproj, pH, pHT,
    p_inner; the condensation parts may be None) plus the two Jacobi diagonals and the V -> Xproj index map."""

    def __init__(self, a, mstar, dm, bproj, aproj, dp, mapV, n_vel, p_range, mcond=None, pcond=None):
        self.keep = []
        mc = mcond or {}
        pc = pcond or {}
        mats = [a, mstar, mc.get("H"), mc.get("HT"), mc.get("inner_solve"), bproj, aproj, pc.get("H"), pc.get("HT"),
                pc.get("inner_solve")]
        self.ops = (_CsrT * 10)(*[_csr_t(m, self.keep) for m in mats])
        self.dm = np.ascontiguousarray(dm, dtype=np.float64)
        self.dp = np.ascontiguousarray(dp, dtype=np.float64)
        self.mapV = np.ascontiguousarray(mapV, dtype=np.int32)
        self.n_vel, self.p_range = int(n_vel), (int(p_range[0]), int(p_range[1]))


def simple_step(S: SimpleOps, gfu, dt, f=None, conv=None, prec_m=1e-4, max_m=200, prec_p=1e-8, max_p=200):
    """templates/NavierStokesSIMPLE_iterative.py:427-443 -> (gfup, (its_mstar, its_proj), seconds); gfu mutated in place."""
    lib = load()
    gfup = np.zeros(S.p_range[1] - S.p_range[0])
    its = (C.c_int * 2)()
    secs = C.c_double()
    f = None if f is None else np.ascontiguousarray(f, dtype=np.float64)
    conv = None if conv is None else np.ascontiguousarray(conv, dtype=np.float64)
    lib.ko_simple_step(S.ops, _p(S.dm), _p(S.dp), _p(S.mapV), C.c_int64(S.n_vel), C.c_int64(S.p_range[0]), C.c_int64(S.p_range[1]),
                       C.c_double(dt), C.c_double(prec_m), C.c_int(max_m), C.c_double(prec_p), C.c_int(max_p), _p(gfu),
                       _p(f) if f is not None else None, _p(conv) if conv is not None else None, _p(gfup), its, C.byref(secs))
    return gfup, (its[0], its[1]), secs.value


def orthonormalize(Q, tries=3):
    """orthonormalization.py:5-16 on an (m, n) C-contiguous array, in place."""
    m, n = Q.shape
    load().ko_orthonormalize(_p(Q), C.c_int64(n), C.c_int(m), C.c_int(tries))
    return Q
