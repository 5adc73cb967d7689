"""the differentiable operations that complete `layers.py` -- divide, min, max, abs, sqrt, sum, permute,
broadcast (derivative rules of FloatLayers.scala:133-274, per element) -- on the GPU against oracle/models_oracle.py
(`ops_step`, pinned by central finite differences in tests/test_models_oracle.py)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

LR = 0.05


@pytest.fixture(scope="module")
def env():
    from oracle import models_oracle as MO
    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200.plugins import Plugins
    from deeplearning.scala_b200.layers import Layers
    return MO, B, Plugins, Layers


def ulps(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, dtype=np.float32).view(np.int32).astype(np.int64)
    return np.abs(a - b).max()


def inputs(seed, n=37, m=53):
    rng = np.random.RandomState(seed)
    a = (rng.rand(n, m).astype(np.float32) + 0.5) * np.where(rng.rand(n, m) < 0.5, -1, 1).astype(np.float32)
    b = rng.rand(n, m).astype(np.float32) + 0.7
    c = (rng.rand(n).astype(np.float32) - 0.5) * 2
    # a few exact ties and zeros: min / max send a tie to operand 1, abs passes 0 through with factor +1
    a[0, 0] = c[0]
    b[1, 1] = max(a[1, 1], c[1])
    a[2, 2] = 0.0
    return a, b, c


@pytest.mark.parametrize("cumulative", [False, True], ids=["TensorLayers", "CumulativeTensorLayers"])
def test_every_remaining_operation_in_one_training_step(env, cumulative):
    MO, B, Plugins, Layers = env
    a, b, c = inputs(1)
    P = Plugins(cumulative=cumulative, optimizer_delta=lambda w, d: d * LR)
    L = Layers(P)
    wa, wb, wc = P.weight(B.tensor(a)), P.weight(B.tensor(b)), P.weight(B.tensor(c))
    q = L.divide(wa, wb)
    r = L.sqrt(L.abs(q) + 0.5)
    m = L.max(wa, L.broadcast(wc, list(a.shape)))
    k = L.min(wb, m)
    p = L.permute(r * k, [1, 0])
    got_p = p.predict().map(lambda t: t.flat_array().reshape(t.shape)).run()
    loss_layer = L.sum(p * p)
    got_loss = float(loss_layer.train().map(lambda t: t.flat_array()[0]).run())
    P.flush_updates()
    assert not P.handled_exceptions, P.handled_exceptions
    want_loss, na, nb, nc, want_p = MO.ops_step(a, b, c, LR)
    assert got_p.shape == want_p.shape and ulps(got_p, want_p) <= 2          # sqrt and / are correctly rounded, one rounding per node
    assert abs(got_loss - want_loss) <= 1e-5 * abs(want_loss)
    for name, w, want, old in (("a", wa, na, a), ("b", wb, nb, b), ("c", wc, nc, c)):
        got = w.data.flat_array().reshape(old.shape)
        g_got, g_want = (old - got).astype(np.float64), (old - want).astype(np.float64)
        assert np.isfinite(got).all()
        err = np.abs(g_got - g_want).max() / (np.abs(g_want).max() + 1e-30)
        assert err <= 1e-5, "gradient of %s: %g" % (name, err)


def test_sum_layer_hands_the_scalar_delta_to_every_element(env):
    MO, B, Plugins, Layers = env
    P = Plugins(optimizer_delta=lambda w, d: d * LR)
    L = Layers(P)
    x = np.random.RandomState(2).standard_normal((5, 7, 3)).astype(np.float32)
    w = P.weight(B.tensor(x))
    total = L.sum(w).train().map(lambda t: float(t.flat_array()[0])).run()
    P.flush_updates()
    assert abs(total - float(x.astype(np.float64).sum())) <= 1e-5 * np.abs(x).sum()
    assert (w.data.flat_array().reshape(x.shape) == (x - np.float32(LR))).all()


def test_permute_and_broadcast_layers_are_bit_exact_views(env):
    MO, B, Plugins, Layers = env
    P = Plugins(optimizer_delta=lambda w, d: d * LR)
    L = Layers(P)
    rng = np.random.RandomState(3)
    x = rng.standard_normal((4, 5, 6)).astype(np.float32)
    v = rng.standard_normal(4).astype(np.float32)
    seed = rng.standard_normal((6, 4, 5)).astype(np.float32)
    wx, wv = P.weight(B.tensor(x)), P.weight(B.tensor(v))
    out = L.permute(wx, [2, 0, 1]) * B.tensor(seed) + L.permute(L.broadcast(wv, [4, 5, 6]), [2, 0, 1])
    got = out.train().map(lambda t: t.flat_array().reshape(t.shape)).run()
    P.flush_updates()
    want = x.transpose(2, 0, 1) * seed + MO.left_broadcast(v, (4, 5, 6)).transpose(2, 0, 1)
    assert (got.view(np.uint32) == want.astype(np.float32).view(np.uint32)).all()
    # d loss / d x = seed through the inverse permutation; d loss / d v = the number of elements each v[i] was copied to
    assert (wx.data.flat_array().reshape(x.shape) == x - (seed.transpose(1, 2, 0) * np.float32(LR))).all()
    assert np.allclose(wv.data.flat_array(), v - np.float32(30 * LR), rtol=1e-6)
