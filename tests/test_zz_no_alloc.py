"""gk_truncnorm_draw_static: the allocation-free variant of gk_truncnorm_draw (static device scratch, include/gk.h) against
the allocating entry point, through the C ABI, on two calls in a row (the second smaller, so stale contents of the shared
array would show), plus the size limit.  Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(ctx, fn, mu, edge, u):
    from graphistician_b200._lib import ptr
    out = np.full(mu.shape, np.nan)
    rc = fn(ctx.h, ptr(mu), ptr(edge), ptr(u), ptr(out), mu.size)
    return rc, out


def test_static_scratch_entry_point_matches_the_allocating_one():
    from graphistician_b200 import _lib
    ctx = _lib.default_context()
    lib = ctx.lib
    nmax = int(lib.gk_truncnorm_static_max())
    assert nmax == 65536
    rng = np.random.RandomState(77)
    for n in (nmax, 1000, 1):                                   # full array, then smaller calls over its stale contents
        mu = np.ascontiguousarray(3.0 * rng.randn(n))
        edge = np.ascontiguousarray((rng.rand(n) < 0.5).astype(np.uint8))
        u = np.ascontiguousarray(rng.rand(n))
        rc0, ref = _call(ctx, lib.gk_truncnorm_draw, mu, edge, u)
        rc1, got = _call(ctx, lib.gk_truncnorm_draw_static, mu, edge, u)
        assert rc0 == 0 and rc1 == 0, lib.gk_last_error(ctx.h).decode()
        assert np.all(np.isfinite(got))
        assert np.allclose(got, ref, rtol=1e-12, atol=0.0) and np.array_equal(got, ref)     # same kernel: bit-identical
    # beyond the static scratch: refused with a clear error, nothing written
    n = nmax + 1
    mu, edge, u = np.zeros(n), np.zeros(n, np.uint8), np.full(n, 0.5)
    rc, out = _call(ctx, lib.gk_truncnorm_draw_static, mu, edge, u)
    assert rc == -1 and b"exceeds the static scratch" in lib.gk_last_error(ctx.h)
    assert np.all(np.isnan(out))
