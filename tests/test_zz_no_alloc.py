"""nsb_orthonormalize_noalloc (no reference counterpart): the allocation-free variant
of nsb_orthonormalize -- scratch in a file-scope device array -- against the existing entry point, through the C ABI, on two calls
in a row (the array keeps the previous call's contents).  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _basis(torch, m, n, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    Q = torch.rand((m, n), dtype=torch.float64, device="cuda", generator=g) - 0.5
    if m > 3:
        Q[3] = 0.5 * Q[0] + 1e-3 * Q[3]      # nearly dependent column: the three sweeps matter
    return Q


def test_noalloc_orthonormalize_matches_existing_entry_point_twice():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import navier_stokes_solver_b200 as nsb
    from navier_stokes_solver_b200 import _lib
    ctx = nsb.default_context()
    lib = ctx._lib
    # two calls in a row, different shapes: the second one finds the first one's state in the shared scratch array
    for call, (m, n, seed) in enumerate([(5, 200003, 1), (8, 77777, 2), (5, 200003, 1)]):
        base = _basis(torch, m, n, seed)
        want, got = base.clone(), base.clone()
        torch.cuda.synchronize()
        _lib.check(lib.nsb_orthonormalize(ctx.h, want.data_ptr(), n, m, 3), "nsb_orthonormalize")
        free0 = torch.cuda.mem_get_info()[0]
        _lib.check(lib.nsb_orthonormalize_noalloc(ctx.h, got.data_ptr(), n, m, 3), "nsb_orthonormalize_noalloc")
        free1 = torch.cuda.mem_get_info()[0]
        assert free1 >= free0, "the no-alloc entry point took device memory"
        w, g_ = want.cpu().numpy(), got.cpu().numpy()
        np.testing.assert_allclose(g_, w, rtol=1e-10, atol=1e-12 * np.abs(w).max(), err_msg="call %d" % call)
        gram = g_ @ g_.T
        assert np.abs(gram - np.eye(m)).max() < 1e-11


def test_noalloc_orthonormalize_refuses_what_the_static_scratch_cannot_hold():
    import torch
    import navier_stokes_solver_b200 as nsb
    from navier_stokes_solver_b200 import _lib
    ctx = nsb.default_context()
    lib = ctx._lib
    Q = torch.zeros((9, 64), dtype=torch.float64, device="cuda")
    with pytest.raises(_lib.NsbError, match="exceeds the maximum"):
        _lib.check(lib.nsb_orthonormalize_noalloc(ctx.h, Q.data_ptr(), 64, 9, 3), "nsb_orthonormalize_noalloc")
    with pytest.raises(_lib.NsbError, match="exceeds the maximum length"):
        _lib.check(lib.nsb_orthonormalize_noalloc(ctx.h, Q.data_ptr(), (1 << 31) + 1, 2, 3), "nsb_orthonormalize_noalloc")
    host = np.zeros((2, 64))
    with pytest.raises(_lib.NsbError, match="device-resident"):
        _lib.check(lib.nsb_orthonormalize_noalloc(ctx.h, host.ctypes.data, 64, 2, 3), "nsb_orthonormalize_noalloc")
