"""`ldc_diagnostics_noalloc` (include/ldc_b200.h): the allocation-free variant of `ldc_diagnostics`, called through the
C-ABI with device pointers, against the existing entry point on the same state -- two calls in a row on different
states, so that the second call meets the first call's contents in the shared static scratch array.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12   # same kernels, same summation order: the entry points should agree to round-off


def _dev(t):
    return C.c_void_p(t.data_ptr())


def _case(ldc_b200, torch, N, B, seed):
    """Existing entry point on a stepped random-ish state; returns its result and the device operands of the new one."""
    rng = np.random.default_rng(seed)
    n, m = N + 1, N - 1
    with ldc_b200.BatchedSpectralLDC(N, batch=B) as s:
        s.set_cases(np.linspace(100.0, 400.0, B))
        s.set_state(0.1 * rng.standard_normal((B, n, n)), 0.1 * rng.standard_normal((B, n, n)),
                    0.1 * rng.standard_normal((B, m, m)))
        s.step(3)
        ref = s.diagnostics()
        u, v, _ = s.get_state()
        lv = s.level
        host = (u, v, lv.Dx, np.ascontiguousarray(np.asarray(lv.Dy).T), lv.W)
    dev = [torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda() for a in host]
    return ref, dev


def test_noalloc_matches_diagnostics_twice(built_lib):
    import torch
    from ldc_b200 import _lib

    ldc_b200 = built_lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    # second call: a smaller problem, with the first call's vorticity still behind it in the array
    for call, (N, B, seed) in enumerate(((32, 5, 11), (24, 3, 12))):
        ref, dev = _case(ldc_b200, torch, N, B, seed)
        out = torch.full((B, 3), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.ldc_diagnostics_noalloc(0, None, N, N, B, *[_dev(t) for t in dev], _dev(out))
        _lib.check(rc, "ldc_diagnostics_noalloc")
        torch.cuda.synchronize()
        if call > 0:   # the first call may still load the kernels' module; after that free device memory must not shrink
            assert torch.cuda.mem_get_info()[0] >= free_before
        got = out.cpu().numpy()
        assert np.isfinite(got).all()
        np.testing.assert_allclose(got, ref, rtol=RTOL, atol=0.0)


def test_noalloc_rejects_oversize_and_null(built_lib):
    import torch
    from ldc_b200 import _lib

    lib = _lib.load()
    t = torch.zeros(16, dtype=torch.float64, device="cuda")
    p = _dev(t)
    # 64 cavities at N = 128: 64 * 129 * 129 doubles > LDC_NOALLOC_MAX_DOUBLES (524288); refused before any pointer is used
    assert lib.ldc_diagnostics_noalloc(0, None, 128, 128, 64, p, p, p, p, p, p) == 1
    assert b"LDC_NOALLOC_MAX_DOUBLES" in lib.ldc_last_error()
    assert lib.ldc_diagnostics_noalloc(0, None, 16, 16, 1, p, p, None, p, p, p) == 1
    assert lib.ldc_diagnostics_noalloc(0, None, 1, 16, 1, p, p, p, p, p, p) == 1
