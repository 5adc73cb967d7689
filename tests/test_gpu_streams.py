"""Multi-stream contexts (`streams_per_device > 1`, the reference's numberOfCommandQueuesPerDevice, OpenCL:425): launches
go round robin over the pool and are ordered by events.  The in-place entry points (dls_sgd_update, dls_optimizer_step,
dls_buf_write, dls_buf_copy) have no counterpart in the reference -- it never writes a buffer twice -- and must wait for
every stream that may still be reading the buffer (write-after-read), and a staged tensor-core operand that is reused
from another stream must wait for the staging kernel (read-after-write).  Stress: long readers on one stream, an
in-place writer issued right behind them on another."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def env():
    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200 import _native as N
    be = B.Backend(streams_per_device=4)
    be.ctx
    yield B, N, be
    be.close()


def test_in_place_writers_wait_for_readers_on_other_streams(env):
    B, N, be = env
    n = 1 << 24
    rng = np.random.RandomState(0)
    w0 = rng.standard_normal(n).astype(np.float32)
    g = rng.standard_normal(n).astype(np.float32)
    for trial in range(6):
        w = be.tensor(w0).do_cache()
        gt = be.tensor(g).do_cache()
        be.synchronize()
        # several long readers of w on round-robin streams (tanh chains keep them busy), then the in-place update
        readers = [(be.tanh(be.tanh(w * float(k + 1))) + w).do_cache() for k in range(3)]
        N.check(N.lib.dls_sgd_update(be.ctx, w._buffer().handle, gt._buffer().handle, n, C.c_float(0.5), None, 0, None))
        after = w.flat_array()
        assert (after == (w0 - np.float32(0.5) * g).astype(np.float32)).all()
        for k, r in enumerate(readers):
            want = np.tanh(np.tanh((w0 * np.float32(k + 1)).astype(np.float64))).astype(np.float32) + w0
            got = r.flat_array()
            assert np.abs(got - want).max() <= 1e-5, "reader %d of trial %d saw the in-place update" % (k, trial)


def test_buffer_write_and_copy_wait_for_readers(env):
    B, N, be = env
    n = 1 << 23
    a0 = np.random.RandomState(1).standard_normal(n).astype(np.float32)
    fresh = np.random.RandomState(2).standard_normal(n).astype(np.float32)
    for trial in range(4):
        a = be.tensor(a0).do_cache()
        src = be.tensor(fresh).do_cache()
        be.synchronize()
        readers = [(be.tanh(be.tanh(a * float(k + 2))) * 1.0).do_cache() for k in range(3)]
        if trial % 2:
            be.write(a._buffer(), fresh)
        else:
            be.copy(a._buffer(), src._buffer(), n)
        assert (a.flat_array() == fresh).all()
        for k, r in enumerate(readers):
            want = np.tanh(np.tanh((a0 * np.float32(k + 2)).astype(np.float64))).astype(np.float32)
            assert np.abs(r.flat_array() - want).max() <= 1e-5


def test_staged_operands_are_ordered_across_streams(env):
    """The same activation feeds several contractions launched on different streams: the bf16 shadow staged by the first
    must be complete before the second reads it."""
    B, N, be = env
    rng = np.random.RandomState(3)
    a = rng.standard_normal((1024, 2048)).astype(np.float32)
    bs = [rng.standard_normal((2048, 256)).astype(np.float32) for _ in range(4)]
    from oracle import tensors_oracle as O
    for trial in range(3):
        at = be.tensor(a).do_cache()
        outs = [be.matmul(at, be.tensor(b), mode="bf16") for b in bs]
        bufs = [o.do_cache() for o in outs]
        for b, o in zip(bs, bufs):
            want = O.round_operand(a, "bf16").astype(np.float64) @ O.round_operand(b, "bf16").astype(np.float64)
            got = o.flat_array().reshape(1024, 256)
            assert np.abs(got - want).max() <= 2e-6 * (np.abs(a).astype(np.float64) @ np.abs(b).astype(np.float64)).max()
