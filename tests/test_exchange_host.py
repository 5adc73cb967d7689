"""Host-side routing of the data-parallel exchange (harness.ExchangeSgd), without a GPU: which gradients leave over
NCCL, which stay for the one-launch peer kernel, and what that launch is told.  The device side is covered by
tests/test_gpu_peer.py (two GPUs)."""
import ctypes as C
from types import SimpleNamespace

import pytest


class FakeBuffer:
    def __init__(self, handle):
        self.handle = handle


class FakeTensor:
    """Just enough of tensors.Tensor for ExchangeSgd: shape / size, broadcast, do_cache, _buffer."""
    def __init__(self, shape, handle):
        self.shape = list(shape)
        self.size = 1
        for s in shape:
            self.size *= s
        self._buf = FakeBuffer(handle)
        self.cached = False

    def broadcast(self, shape):
        assert list(shape) == self.shape
        return self

    def do_cache(self):
        self.cached = True
        return self

    def _buffer(self):
        return self._buf


class FakeExchange:
    peer = True
    peer_bytes = 1 << 20

    def __init__(self):
        self._over_nccl = set()
        self.sent = []
        self.backend = SimpleNamespace(ctx=C.c_void_p(1234))

    def allreduce(self, tensor):
        cached = tensor.do_cache()
        self.sent.append(cached)
        return cached


@pytest.fixture()
def H():
    import deeplearning.scala_b200.harness as H
    return H


def test_large_gradients_leave_over_nccl_small_ones_wait_for_the_peer_launch(H, monkeypatch):
    from deeplearning.scala_b200 import _native as N
    ex = FakeExchange()
    opt = H.ExchangeSgd(ex, 0.125)
    shapes = [(256, 10), (10,), (8192, 256), (256,), (3, 3, 64, 128)]          # fc2, b, fc1 (8 MiB), b, conv3
    weights = [SimpleNamespace(data=FakeTensor(s, 100 + i)) for i, s in enumerate(shapes)]
    pending = [(w, opt.gradient(w, FakeTensor(s, 200 + i))) for i, (w, s) in enumerate(zip(weights, shapes))]
    assert [t._buffer().handle for t in ex.sent] == [202]                      # only the 8 MiB gradient
    assert all(g.cached for _, g in pending)                                   # every gradient is materialised

    calls = []

    def fake_step(ctx, weights_, grads_, counts_, flags_, n, lr):
        calls.append(([weights_[i] for i in range(n)], [grads_[i] for i in range(n)], [counts_[i] for i in range(n)],
                      [flags_[i] for i in range(n)], lr))
        return 0
    monkeypatch.setattr(N, "lib", SimpleNamespace(dls_peer_sgd_step=fake_step))
    opt.step_many(pending)
    assert len(calls) == 1                                                     # the whole end of the step is one call
    w, g, counts, flags, lr = calls[0]
    assert w == [100, 101, 102, 103, 104] and g == [200, 201, 202, 203, 204]   # arrival order, identical on every rank
    assert counts == [2560, 10, 8192 * 256, 256, 3 * 3 * 64 * 128]
    assert flags == [1, 1, 0, 1, 1]                                            # already reduced by NCCL -> local update only
    assert lr == 0.125
    assert not ex._over_nccl                                                   # the next step starts clean


def test_plugins_hand_all_deferred_updates_to_step_many_at_once():
    from deeplearning.scala_b200.plugins import Plugins
    seen = []

    class Opt:
        def begin_step(self, backend=None):
            pass

        def step_many(self, pending):
            seen.append(list(pending))

        def step(self, weight, delta):                                        # must not be used when step_many exists
            raise AssertionError("per-weight step called")
    P = Plugins(optimizer=Opt(), backend=object())
    order = []
    P.before_flush.append(lambda: order.append("exchange.flush"))
    P.deferred_updates.extend([("w0", "g0"), ("w1", "g1")])
    P.flush_updates()
    assert order == ["exchange.flush"]                                         # leftover NCCL bucket goes out first
    assert seen == [[("w0", "g0"), ("w1", "g1")]] and P.deferred_updates == []
    P.flush_updates()                                                          # nothing pending: no hooks, no launch
    assert order == ["exchange.flush"] and len(seen) == 1


def test_reduce_routes_by_size_and_flush_issues_nccl_before_the_peer_launch(H):
    """The route stateful optimizers take: `Exchange.reduce` per gradient, `Exchange.flush` at the end of the step."""
    log = []

    def many(ctx, handles, counts, n):
        log.append(("nccl", [handles[i] for i in range(n)], [counts[i] for i in range(n)]))
        return 0

    def peer(ctx, weights, grads, counts, flags, n, lr):
        log.append(("peer", weights, [grads[i] for i in range(n)], [counts[i] for i in range(n)], [flags[i] for i in range(n)], lr))
        return 0
    ex = object.__new__(H.Exchange)                                            # no communicator: only the routing is under test
    ex.backend = SimpleNamespace(ctx=C.c_void_p(1))
    ex._C, ex._N = C, SimpleNamespace(lib=SimpleNamespace(dls_allreduce_sum_many=many, dls_peer_sgd_step=peer), check=lambda rc: None)
    ex._pending, ex._pending_bytes, ex._held, ex._over_nccl = [], 0, [], set()
    ex.peer, ex.peer_reduce, ex.peer_bytes, ex.bucket_bytes = True, True, 1 << 20, 1 << 20
    small0, big, small1 = FakeTensor((10,), 1), FakeTensor((8192, 256), 2), FakeTensor((3, 3, 64, 128), 3)
    assert ex.reduce(small0) is small0 and log == []                           # held, nothing launched
    assert ex.reduce(big) is big
    assert log == [("nccl", [2], [8192 * 256])]                                # a full bucket leaves at once ...
    assert ex.reduce(small1) is small1 and len(log) == 1                       # ... without dragging the held ones along
    ex.flush()
    assert log[1:] == [("peer", None, [1, 3], [10, 3 * 3 * 64 * 128], [1, 1], 0.0)]
    ex.flush()
    assert len(log) == 2                                                       # nothing pending
    # without the peer path everything is bucketed for NCCL, small gradients share one launch at flush
    ex.peer = False
    del log[:]
    ex.reduce(small0), ex.reduce(small1)
    assert log == []
    ex.flush()
    assert log == [("nccl", [1, 3], [10, 3 * 3 * 64 * 128])]
