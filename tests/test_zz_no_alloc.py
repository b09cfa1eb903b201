"""b200hypre_multi_dot_noalloc (include/b200hypre.h): the multi-dot kernel of the GMRES path with its scratch (per-block partial
sums + arrival counter) in a file-scope device array, for callers that cannot call cudaMalloc.  It has no counterpart in the
reference; it is checked against the existing entry point for the same numbers, HYPRE_ParVectorInnerProd, through the C ABI.

The file is named so that pytest collects it LAST.  The calls below run back to back on the
same static array with different sizes -- a large grid first, then a small one that leaves most of the old partials in place, then
the large one again -- so a result that depended on what the previous call left behind would show.  Same kernel, same grid, same
summation order as the existing path: the results are normally bit-identical; the assertion uses 1e-12 relative (the value tests of
this suite use 1e-10 .. 1e-13)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12
DEVICE, HOST = 1, 0   # HYPRE_MemoryLocation


def _bind(b200):
    L = b200.lib()
    L.b200hypre_multi_dot_noalloc.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p]
    L.b200hypre_multi_dot_noalloc.restype = C.c_int
    L.HYPRE_ParVectorInnerProd.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    b200.check(L.HYPRE_Init(), "HYPRE_Init")
    return L


def _uploads_done():
    # cudaMemcpy from pageable host memory may return before the DMA has landed, and the library stream is non-blocking
    import torch
    torch.cuda.synchronize()


def _case(b200, L, n, k, seed):
    """k vectors of n entries: <V_j, x> by the existing entry point (one ParVector pair at a time) and by the new one (one call)"""
    rng = np.random.default_rng(seed)
    xh = rng.standard_normal(n)
    Vh = rng.standard_normal((k, n))
    ld = (n + 1) & ~1
    x = b200.IJVector(0, n - 1)
    x.set_values(xh)
    x.assemble()
    existing = np.zeros(k)
    for j in range(k):
        v = b200.IJVector(0, n - 1)
        v.set_values(Vh[j])
        v.assemble()
        prod = C.c_double(0.0)
        b200.check(L.HYPRE_ParVectorInnerProd(v.object(), x.object(), C.byref(prod)), "ParVectorInnerProd")
        existing[j] = prod.value
        v.destroy()
    # the caller's device buffers: the k vectors at stride ld, and the k results
    V = L.b200hypre_malloc(k * ld * 8, DEVICE)
    out = L.b200hypre_malloc(k * 8, DEVICE)
    assert V and out
    for j in range(k):
        L.b200hypre_memcpy(V + j * ld * 8, Vh[j].ctypes.data, n * 8, DEVICE, HOST)
    poison = np.full(k, np.nan)
    L.b200hypre_memcpy(out, poison.ctypes.data, k * 8, DEVICE, HOST)
    _uploads_done()
    b200.check(L.b200hypre_multi_dot_noalloc(V, ld, k, x.data_ptr(), n, out), "multi_dot_noalloc")
    b200.check(L.b200hypre_device_synchronize(), "device_synchronize")
    got = np.zeros(k)
    L.b200hypre_memcpy(got.ctypes.data, out, k * 8, HOST, DEVICE)
    L.b200hypre_free(V, DEVICE)
    L.b200hypre_free(out, DEVICE)
    x.destroy()
    scale = np.sqrt((Vh * Vh).sum(axis=1) * (xh * xh).sum())      # |<v, x>| <= |v| |x|: relative to the size of the terms
    assert np.all(np.isfinite(got)), got
    assert np.abs(got - existing).max() <= RTOL * scale.max(), (got, existing)
    assert np.abs(got - Vh @ xh).max() <= 1e-10 * scale.max()
    return got, existing


def test_noalloc_multi_dot_matches_inner_prod_on_calls_in_a_row(b200):
    L = _bind(b200)
    launches0 = L.b200hypre_launch_count()
    first, _ = _case(b200, L, 1_000_003, 11, seed=1)      # odd length, two passes (8 + 3 vectors), grid of 489 blocks
    _case(b200, L, 4_099, 3, seed=2)                      # 3 blocks: most of the partials of the first call stay in the array
    _case(b200, L, 2_500_000, 1, seed=3)                  # k = 1 is exactly what HYPRE_ParVectorInnerProd runs; full grid
    again, _ = _case(b200, L, 1_000_003, 11, seed=1)
    assert np.array_equal(first, again)                   # same inputs, same bits, whatever ran in between
    assert L.b200hypre_launch_count() > launches0         # kernels of the library ran (no host fallback)


def test_noalloc_multi_dot_rejects_what_the_static_scratch_cannot_serve(b200):
    L = _bind(b200)
    n = 1024
    buf = L.b200hypre_malloc(2 * n * 8, DEVICE)
    out = L.b200hypre_malloc(128 * 8, DEVICE)
    try:
        rc = L.b200hypre_multi_dot_noalloc(buf, n, 65, buf, n, out)           # more vectors than the stated maximum
        msg = L.b200hypre_last_error().decode()
        L.HYPRE_ClearAllErrors()
        assert rc != 0 and "64" in msg and "k = 65" in msg, (rc, msg)
        rc = L.b200hypre_multi_dot_noalloc(None, n, 1, buf, n, out)
        L.HYPRE_ClearAllErrors()
        assert rc != 0
        rc = L.b200hypre_multi_dot_noalloc(buf + 8, n, 1, buf, n - 2, out)    # misaligned vector
        L.HYPRE_ClearAllErrors()
        assert rc != 0
        b200.check(L.b200hypre_multi_dot_noalloc(buf, n, 1, buf, 0, out), "n = 0")   # empty vectors: out[0] = 0
        b200.check(L.b200hypre_device_synchronize())
        got = np.full(1, np.nan)
        L.b200hypre_memcpy(got.ctypes.data, out, 8, HOST, DEVICE)
        assert got[0] == 0.0
    finally:
        L.b200hypre_free(buf, DEVICE)
        L.b200hypre_free(out, DEVICE)
