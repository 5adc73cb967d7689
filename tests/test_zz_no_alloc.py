"""dls_sum_static (include/dls_b200.h): the allocation-free variant of dls_sum.  It has no counterpart in the reference,
so its oracle is the existing entry point: same inputs through both,
on two calls in a row that share the file-scope scratch array, within the reduction tolerance of the parity tests
(REL_REDUCTION, tests/test_gpu_parity.py:15) -- and in fact bit for bit on a B200, where the two grids are equal.

Collected last on purpose.  Uses nothing but the C ABI through ctypes."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_REDUCTION = 1e-5


@pytest.fixture()
def env():
    from deeplearning.scala_b200 import _native as N
    ctx = C.c_void_p()
    N.check(N.lib.dls_ctx_create(C.byref(N.CtxOpts()), C.byref(ctx)))
    held = []
    yield N, ctx, held
    for h in held:
        N.lib.dls_buf_release(h)
    N.lib.dls_ctx_destroy(ctx)


def upload(N, ctx, held, x):
    h = C.c_void_p()
    N.check(N.lib.dls_buf_from_host(ctx, x.ctypes.data_as(C.c_void_p), x.size, C.byref(h), None))
    held.append(h)
    return h


def read1(N, ctx, h):
    got = np.zeros(1, np.float32)
    N.check(N.lib.dls_buf_read(ctx, h, 0, got.ctypes.data_as(C.c_void_p), 1, None, 0, None))   # blocking
    return got[0]


def sum_alloc(N, ctx, h, n):
    out = C.c_void_p()
    N.check(N.lib.dls_sum(ctx, h, n, None, 0, C.byref(out), None))
    v = read1(N, ctx, out)
    N.lib.dls_buf_release(out)
    return v


def driver_allocs(N, ctx):
    st = N.Stats()
    N.check(N.lib.dls_ctx_stats(ctx, C.byref(st)))
    return st.driver_allocs, st.bytes_in_use, st.bytes_cached, st.kernel_launches


def test_static_sum_matches_the_allocating_sum_twice_in_a_row(env):
    N, ctx, held = env
    rng = np.random.RandomState(77)
    # ragged tail, more CTAs than one wave, then a SMALLER input: stale partials of the first call must not leak in
    xs = [rng.standard_normal(8 * 1000 * 1000 + 3).astype(np.float32), rng.standard_normal(4099).astype(np.float32) + 1.0]
    bufs = [upload(N, ctx, held, x) for x in xs]
    out = C.c_void_p()
    N.check(N.lib.dls_buf_alloc(ctx, 1, C.byref(out)))
    held.append(out)
    want = [sum_alloc(N, ctx, h, x.size) for h, x in zip(bufs, xs)]
    N.check(N.lib.dls_ctx_sync(ctx))
    before = driver_allocs(N, ctx)
    N.check(N.lib.dls_sum_static(ctx, bufs[0], xs[0].size, out))      # two calls in a row, nothing in between
    N.check(N.lib.dls_sum_static(ctx, bufs[1], xs[1].size, out))
    after = driver_allocs(N, ctx)
    assert after[:3] == before[:3], "the allocator was touched: %r -> %r" % (before, after)
    assert after[3] == before[3] + 2
    second = read1(N, ctx, out)
    N.check(N.lib.dls_sum_static(ctx, bufs[0], xs[0].size, out))
    first = read1(N, ctx, out)
    for got, ref, x in ((first, want[0], xs[0]), (second, want[1], xs[1])):
        scale = max(1.0, float(np.abs(x).sum()))
        assert abs(float(got) - float(ref)) <= REL_REDUCTION * scale
        assert abs(float(got) - float(x.astype(np.float64).sum())) <= REL_REDUCTION * scale
        assert np.float32(got).tobytes() == np.float32(ref).tobytes()   # equal grids on a B200: the same bits
    # prefix, empty and one-element inputs
    for n in (0, 1, 5, 4096, 1000003):
        N.check(N.lib.dls_sum_static(ctx, bufs[0], n, out))
        got = read1(N, ctx, out)
        assert np.float32(got).tobytes() == np.float32(sum_alloc(N, ctx, bufs[0], n)).tobytes(), n


def test_static_sum_rejects_what_it_cannot_serve(env):
    N, ctx, held = env
    x = np.ones(16, np.float32)
    h = upload(N, ctx, held, x)
    out = C.c_void_p()
    N.check(N.lib.dls_buf_alloc(ctx, 1, C.byref(out)))
    held.append(out)
    with pytest.raises(ValueError, match="DLS_SUM_STATIC_MAX_FLOATS"):
        N.check(N.lib.dls_sum_static(ctx, h, N.DLS_SUM_STATIC_MAX_FLOATS + 1, out))
    with pytest.raises(ValueError, match="exceeds the input buffer"):
        N.check(N.lib.dls_sum_static(ctx, h, 17, out))
    with pytest.raises(ValueError, match="alias"):
        N.check(N.lib.dls_sum_static(ctx, h, 16, h))
    with pytest.raises(ValueError, match="NULL"):
        N.check(N.lib.dls_sum_static(ctx, h, 16, None))
    N.check(N.lib.dls_sum_static(ctx, h, 16, out))     # and the scratch still works after the refusals
    assert read1(N, ctx, out) == np.float32(16.0)
