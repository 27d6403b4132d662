"""ctypes binding of libnsb200.so (include/nsb200.h).  The library is the only compute path: if it
is missing or no CUDA device is present the calls raise -- there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnsb200.so")

NSB_OK = 0
STATUS_NAMES = {0: "converged", 1: "maxsteps", 2: "abstol", 3: "zero_residual", 4: "breakdown"}
CONVERGED, MAXSTEPS, ABSTOL, ZERO_RESIDUAL, BREAKDOWN = 0, 1, 2, 3, 4

c_ctx = C.c_void_p
c_mat = C.c_void_p
c_pre = C.c_void_p
c_halo = C.c_void_p
P = C.POINTER

# name -> (restype, argtypes); mirrors include/nsb200.h declaration by declaration
SIGNATURES = {
    "nsb_version": (C.c_int, []),
    "nsb_last_error": (C.c_char_p, []),
    "nsb_ctx_create": (C.c_int, [C.c_int, P(c_ctx)]),
    "nsb_nccl_unique_id": (C.c_int, [C.c_char_p]),
    "nsb_ctx_create_dist": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_char_p, P(c_ctx)]),
    "nsb_ctx_destroy": (C.c_int, [c_ctx]),
    "nsb_ctx_sync": (C.c_int, [c_ctx]),
    "nsb_ctx_stream": (C.c_void_p, [c_ctx]),
    "nsb_ctx_launch_count": (C.c_int64, [c_ctx]),
    "nsb_ctx_sync_stats": (C.c_int, [c_ctx, C.POINTER(C.c_uint64), C.c_int]),
    "nsb_ctx_set_graphs": (C.c_int, [c_ctx, C.c_int]),
    "nsb_csr_create": (C.c_int, [c_ctx, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, P(c_mat)]),
    "nsb_csr_transpose": (C.c_int, [c_ctx, c_mat, P(c_mat)]),
    "nsb_saddle_create": (C.c_int, [c_ctx, c_mat, c_mat, c_mat, c_mat, P(c_mat)]),
    "nsb_mat_destroy": (C.c_int, [c_mat]),
    "nsb_mat_info": (C.c_int, [c_mat, P(C.c_int64), P(C.c_int64), P(C.c_int64)]),
    "nsb_mat_format": (C.c_int, [c_mat, P(C.c_int), P(C.c_int), P(C.c_int), P(C.c_int64), P(C.c_int), P(C.c_int64), P(C.c_int64)]),
    "nsb_mat_set_row_panels": (C.c_int, [c_ctx, c_mat, C.c_int, C.c_void_p]),
    "nsb_mat_diagonal": (C.c_int, [c_ctx, c_mat, C.c_void_p]),
    "nsb_spmv": (C.c_int, [c_ctx, c_mat, C.c_void_p, C.c_void_p]),
    "nsb_spmv_bench": (C.c_int, [c_ctx, c_mat, C.c_int, C.c_int, P(C.c_double)]),
    "nsb_halo_create": (C.c_int, [c_ctx, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, P(c_halo)]),
    "nsb_halo_destroy": (C.c_int, [c_halo]),
    "nsb_mat_set_halo": (C.c_int, [c_mat, c_halo]),
    "nsb_pre_create_diag": (C.c_int, [c_ctx, C.c_int64, C.c_void_p, P(c_pre)]),
    "nsb_pre_create_jacobi": (C.c_int, [c_ctx, c_mat, C.c_void_p, P(c_pre)]),
    "nsb_pre_create_blockdiag2": (C.c_int, [c_ctx, c_pre, c_pre, P(c_pre)]),
    "nsb_pre_create_blocks": (C.c_int, [c_ctx, c_mat, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, P(c_pre)]),
    "nsb_pre_smooth": (C.c_int, [c_ctx, c_pre, C.c_void_p, C.c_void_p, C.c_int]),
    "nsb_pre_create_auxspace": (C.c_int, [c_ctx, c_pre, c_mat, c_mat, c_pre, C.c_int, P(c_pre)]),
    "nsb_pre_destroy": (C.c_int, [c_pre]),
    "nsb_pre_apply": (C.c_int, [c_ctx, c_pre, C.c_void_p, C.c_void_p]),
    "nsb_minres": (C.c_int, [c_ctx, c_mat, c_pre, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p,
                             P(C.c_int), P(C.c_double), P(C.c_int)]),
    "nsb_bpcg": (C.c_int, [c_ctx, c_mat, c_mat, c_mat, c_pre, c_pre, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p,
                           C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_void_p, P(C.c_int), P(C.c_double), P(C.c_int)]),
    "nsb_cg": (C.c_int, [c_ctx, c_mat, c_pre, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p,
                         P(C.c_int), P(C.c_double), P(C.c_int)]),
    "nsb_lanczos_eigs": (C.c_int, [c_ctx, c_mat, c_pre, C.c_void_p, C.c_int, C.c_double, P(C.c_double), P(C.c_double), P(C.c_int)]),
    "nsb_orthonormalize": (C.c_int, [c_ctx, C.c_void_p, C.c_int64, C.c_int, C.c_int]),
    "nsb_orthonormalize_noalloc": (C.c_int, [c_ctx, C.c_void_p, C.c_int64, C.c_int, C.c_int]),
    "nsb_cond_create": (C.c_int, [c_ctx, c_mat, c_mat, c_mat, c_mat, c_mat, P(C.c_void_p)]),
    "nsb_cond_destroy": (C.c_int, [C.c_void_p]),
    "nsb_cond_apply": (C.c_int, [c_ctx, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nsb_bpcg_condensed": (C.c_int, [c_ctx, C.c_void_p, c_mat, c_mat, c_pre, c_pre, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_double, C.c_int, C.c_int, C.c_void_p, P(C.c_int), P(C.c_double), P(C.c_int)]),
    "nsb_lanczos_eigs_condensed": (C.c_int, [c_ctx, C.c_void_p, c_pre, C.c_void_p, C.c_int, C.c_double, P(C.c_double),
                                             P(C.c_double), P(C.c_int)]),
    "nsb_vec_axpby": (C.c_int, [c_ctx, C.c_void_p, C.c_double, C.c_void_p, C.c_double, C.c_int64]),
    "nsb_vec_gather_axpy": (C.c_int, [c_ctx, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int64]),
    "nsb_simple_create": (C.c_int, [c_ctx, c_mat, c_mat, c_pre, c_mat, c_mat, c_mat, c_mat, c_mat, c_pre, c_mat, c_mat, c_mat, C.c_void_p,
                                    C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_double, C.c_int, C.c_double, C.c_int, P(C.c_void_p)]),
    "nsb_simple_destroy": (C.c_int, [C.c_void_p]),
    "nsb_simple_step": (C.c_int, [c_ctx, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, P(C.c_int), P(C.c_double)]),
    "nsb_simple_project": (C.c_int, [c_ctx, C.c_void_p, C.c_void_p, C.c_void_p, P(C.c_int)]),
    "nsb_heat_step": (C.c_int, [c_ctx, c_mat, c_mat, c_mat, c_pre, C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_void_p,
                                C.c_void_p, C.c_double, C.c_int, P(C.c_int), C.c_void_p, C.c_void_p]),
}


class NsbError(RuntimeError):
    pass


_lib = None


def load():
    """Load libnsb200.so; raises if it has not been built (``__graft_entry__.build()``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NsbError("libnsb200.so is not built (%s missing): run `python __graft_entry__.py` / "
                       "navier_stokes_solver_b200.build.build(); there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != NSB_OK:
        msg = load().nsb_last_error()
        raise NsbError("%s failed (code %d): %s" % (what or "libnsb200 call", rc, msg.decode() if msg else ""))
