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
