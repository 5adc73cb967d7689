"""The three convolution passes of csrc/conv.cu (implicit GEMM on the tcgen05 tensor cores through tap-shifted 4-D TMA
boxes; direct kernels for a first layer with Cin <= 4), through the C ABI, against the plain-numpy convolution of
oracle/models_oracle.py -- itself pinned on the reference's convolution known-answers (TensorsSpec:144-249, golden
vectors G17 in tests/test_oracle_golden.py) -- evaluated on operands rounded the way the mode rounds them.

Identical inputs on both sides, so only the fp32 accumulation order differs and the bound is tight:
    |got - want| <= 4e-6 * (the same convolution of |x|, |w| / |dz|)          per element
which is what proves that every tap, every image border (the TMA zero fill IS the padding), every tile and every split of
the pixel contraction addresses the right elements.  Shapes: the three layers of the benched C3 at full size, ragged
batches (partial last tile), wide images, odd channel counts (padded operand channels), and the in-place read of a
down-sampled source.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from oracle import models_oracle as MO
    from oracle import tensors_oracle as O
    import deeplearning.scala_b200.tensors as B
    return MO, O, B


@pytest.fixture(scope="module", params=["bf16", "tf32"])
def backend(request, env):
    B = env[2]
    be = B.Backend(gemm_mode=request.param)
    be.ctx
    yield request.param, be
    be.close()


def make(batch, h, w, cin, cout, seed):
    rng = np.random.RandomState(seed)
    x = rng.standard_normal((batch, h, w, cin)).astype(np.float32)
    wt = (rng.standard_normal((3, 3, cin, cout)) * 0.2).astype(np.float32)
    bias = rng.standard_normal(cout).astype(np.float32)
    dz = (rng.standard_normal((batch, h, w, cout)) * (rng.rand(batch, h, w, cout) < 0.5)).astype(np.float32)
    return x, wt, bias, dz


# batch, h, w, cin, cout
SHAPES = [
    (256, 32, 32, 3, 32),      # C3 conv1 (direct kernels)
    (256, 16, 16, 32, 64),     # C3 conv2: 2 tiles per image, channels padded 32 -> 64 in bf16
    (256, 8, 8, 64, 128),      # C3 conv3: 2 images per tile
    (3, 8, 8, 64, 128),        # ragged batch: the last tile holds one image + zero fill
    (5, 32, 32, 64, 64),       # 8 tiles per image
    (2, 16, 16, 40, 24),       # channel counts that are not multiples of anything
    (7, 4, 4, 32, 32),         # 8 images per tile, ragged
    (2, 64, 64, 32, 16),       # two rows per tile
    (4, 16, 16, 1, 8),         # direct kernels, one input channel
    (2, 12, 20, 4, 32),        # direct kernels, odd spatial size
]
IDS = ["%dx%dx%dx%d-%d" % s for s in SHAPES]


def supported(be, which, shape):
    _, h, w, cin, cout = shape
    return be.conv3x3_supported(which, h, w, cin, cout)


@pytest.mark.parametrize("shape", SHAPES, ids=IDS)
@pytest.mark.parametrize("activation", ["none", "relu"])
def test_forward(env, backend, shape, activation):
    MO, O, B = env
    mode, be = backend
    if not supported(be, "forward", shape):
        pytest.skip("no dedicated forward kernel for this shape in %s (composed from views + matmul instead)" % mode)
    x, wt, bias, _ = make(*shape, seed=sum(shape))
    r = lambda a: O.round_operand(a, mode)
    got = be.conv3x3_forward(be.tensor(x), be.tensor(wt), be.tensor(bias), activation).flat_array().reshape(shape[:3] + (shape[4],))
    want = MO.conv_forward(x, wt, bias, r)
    scale = MO.conv_forward(np.abs(x), np.abs(wt), np.abs(bias), r) + 1.0
    if activation == "relu":
        want = np.maximum(want, 0.0)
    err = np.abs(got.astype(np.float64) - want) / scale
    assert err.max() <= 4e-6, "worst element %r: %g" % (np.unravel_index(err.argmax(), err.shape), err.max())


@pytest.mark.parametrize("shape", SHAPES, ids=IDS)
def test_backward_data(env, backend, shape):
    MO, O, B = env
    mode, be = backend
    if not supported(be, "backward_data", shape):
        pytest.skip("no dedicated backward-data kernel for this shape in %s" % mode)
    x, wt, _, dz = make(*shape, seed=sum(shape) + 1)
    r = lambda a: O.round_operand(a, mode)
    got = be.conv3x3_backward_data(be.tensor(dz), be.tensor(wt)).flat_array().reshape(x.shape)
    want, _, _ = MO.conv_backward(x, wt, dz, r)
    scale, _, _ = MO.conv_backward(np.abs(x), np.abs(wt), np.abs(dz), r)
    err = np.abs(got.astype(np.float64) - want.astype(np.float64)) / (scale.astype(np.float64) + 1.0)
    assert err.max() <= 4e-6, "worst element %r: %g" % (np.unravel_index(err.argmax(), err.shape), err.max())


@pytest.mark.parametrize("shape", SHAPES, ids=IDS)
def test_backward_weights(env, backend, shape):
    MO, O, B = env
    mode, be = backend
    if not supported(be, "backward_weights", shape):
        pytest.skip("no dedicated weight-gradient kernel for this shape in %s" % mode)
    x, wt, _, dz = make(*shape, seed=sum(shape) + 2)
    r = lambda a: O.round_operand(a, mode)
    got = be.conv3x3_backward_weights(be.tensor(x), be.tensor(dz)).flat_array().reshape(wt.shape)
    _, want, _ = MO.conv_backward(x, wt, dz, r, need_dx=False)
    _, scale, _ = MO.conv_backward(np.abs(x), np.abs(wt), np.abs(dz), r, need_dx=False)
    err = np.abs(got.astype(np.float64) - want.astype(np.float64)) / (scale.astype(np.float64) + 1.0)
    assert err.max() <= 4e-6, "worst element %r: %g" % (np.unravel_index(err.argmax(), err.shape), err.max())


def test_downsampled_source_is_read_in_place(env, backend):
    """conv(x.scale(half)) with x materialised: the kernels sample x at stride 2 themselves (no gather launch), forward
    and weight gradient, and give what the convolution of the materialised half-size tensor gives."""
    MO, O, B = env
    mode, be = backend
    batch, h, w, cin, cout = 6, 16, 16, 32, 64
    rng = np.random.RandomState(11)
    full = rng.standard_normal((batch, 2 * h, 2 * w, cin)).astype(np.float32)
    wt = (rng.standard_normal((3, 3, cin, cout)) * 0.2).astype(np.float32)
    dz = rng.standard_normal((batch, h, w, cout)).astype(np.float32)
    half = np.ascontiguousarray(full[:, ::2, ::2, :])
    xt = be.tensor(full).do_cache()
    view = xt.scale([batch, h, w, cin])
    wd, dzd = be.tensor(wt).do_cache(), be.tensor(dz).do_cache()
    be.tensor(half).do_cache().flat_array()
    launches0 = be.stats()["kernel_launches"]
    y = be.conv3x3_forward(view, wd).flat_array()
    fwd_launches = be.stats()["kernel_launches"] - launches0
    y_ref = be.conv3x3_forward(be.tensor(half), be.tensor(wt)).flat_array()
    assert (y.view(np.uint32) == y_ref.view(np.uint32)).all()
    if be.conv3x3_supported("backward_weights", h, w, cin, cout):
        dw = be.conv3x3_backward_weights(view, dzd).flat_array()
        dw_ref = be.conv3x3_backward_weights(be.tensor(half), be.tensor(dz)).flat_array()
        assert (dw.view(np.uint32) == dw_ref.view(np.uint32)).all()
    assert fwd_launches <= 3                   # stage the sampled operand, pack the weights, the convolution
    r = lambda a: O.round_operand(a, mode)
    want = MO.conv_forward(half, wt, np.zeros(cout, np.float32), r)
    scale = MO.conv_forward(np.abs(half), np.abs(wt), np.zeros(cout, np.float32), r) + 1.0
    assert (np.abs(y.reshape(want.shape) - want) / scale).max() <= 4e-6


def test_reference_convolution_known_answer(env, backend):
    """TensorsSpec:144-249: the reference's own convolution test -- input 1x? ... reproduced through the layer API:
    conv2d_3x3 of the golden case must equal the oracle's `convolute` composition (which is pinned on the reference's
    expected output, tests/test_oracle_golden.py G17)."""
    MO, O, B = env
    mode, be = backend
    from deeplearning.scala_b200.plugins import Plugins
    from deeplearning.scala_b200.layers import Layers
    rng = np.random.RandomState(5)
    x = rng.standard_normal((2, 8, 8, 32)).astype(np.float32)
    wt = (rng.standard_normal((3, 3, 32, 16)) * 0.2).astype(np.float32)
    bias = rng.standard_normal(16).astype(np.float32)
    P = Plugins(backend=be)
    L = Layers(P)
    # the forward value is pinned only while the Do is open (TL:57-68): read it inside
    out = L.conv2d_3x3(be.tensor(x), P.weight(be.tensor(wt)), P.weight(be.tensor(bias)), "relu").predict().map(
        lambda t: t.flat_array()).run()
    r = lambda a: O.round_operand(a, mode)
    want = np.maximum(MO.conv_forward(x, wt, bias, r), 0.0)
    scale = MO.conv_forward(np.abs(x), np.abs(wt), np.abs(bias), r) + 1.0
    assert (np.abs(out.reshape(want.shape) - want) / scale).max() <= 4e-6


SAMPLED = [(256, 32, 32, 3, 32, 2), (256, 16, 16, 32, 64, 2), (3, 16, 16, 32, 64, 2), (5, 12, 20, 4, 32, 2), (2, 16, 16, 1, 8, 4),
           (6, 8, 8, 64, 128, 2)]


@pytest.mark.parametrize("shape", SAMPLED, ids=["%dx%dx%dx%d-%d-s%d" % s for s in SAMPLED])
def test_sampled_delta_equals_the_zero_filled_delta(env, backend, shape):
    """`dz_stride` = s: the delta that a down-sampling after the layer hands back arrives at its sampled size
    [B, H/s, W/s, Cout]; both backward passes must give what they give for the full-size delta that is zero except at
    (s*y, s*x) -- against the oracle on that zero-filled delta, and (tensor-core kernels: same staged operand) bit for
    bit against the dense call."""
    MO, O, B = env
    mode, be = backend
    batch, h, w, cin, cout, s = shape
    rng = np.random.RandomState(sum(shape))
    x = rng.standard_normal((batch, h, w, cin)).astype(np.float32)
    wt = (rng.standard_normal((3, 3, cin, cout)) * 0.2).astype(np.float32)
    dzs = rng.standard_normal((batch, h // s, w // s, cout)).astype(np.float32)
    dz = np.zeros((batch, h, w, cout), dtype=np.float32)
    dz[:, ::s, ::s, :] = dzs
    r = lambda a: O.round_operand(a, mode)
    want_dx, want_dw, _ = MO.conv_backward(x, wt, dz, r)
    scale_dx, scale_dw, _ = MO.conv_backward(np.abs(x), np.abs(wt), np.abs(dz), r)
    ran = 0
    if be.conv3x3_supported("backward_weights", h, w, cin, cout):
        got = be.conv3x3_backward_weights(be.tensor(x), be.tensor(dzs), dz_stride=s).flat_array().reshape(wt.shape)
        err = np.abs(got.astype(np.float64) - want_dw.astype(np.float64)) / (scale_dw.astype(np.float64) + 1.0)
        assert err.max() <= 4e-6, "dW worst element %r: %g" % (np.unravel_index(err.argmax(), err.shape), err.max())
        ran += 1
    if be.conv3x3_supported("backward_data", h, w, cin, cout):
        got = be.conv3x3_backward_data(be.tensor(dzs), be.tensor(wt), dz_stride=s).flat_array().reshape(x.shape)
        err = np.abs(got.astype(np.float64) - want_dx.astype(np.float64)) / (scale_dx.astype(np.float64) + 1.0)
        assert err.max() <= 4e-6, "dX worst element %r: %g" % (np.unravel_index(err.argmax(), err.shape), err.max())
        dense = be.conv3x3_backward_data(be.tensor(dz), be.tensor(wt)).flat_array().reshape(x.shape)
        assert (got.view(np.uint32) == dense.view(np.uint32)).all()
        ran += 1
    if not ran:
        pytest.skip("no dedicated backward kernel for this shape in %s" % mode)


def test_downsample_after_a_convolution_folds_into_the_layer(env, backend):
    """`Layers.downsample(conv2d_3x3(...))` is one layer: same value as the composition, and one training step moves the
    weights exactly as the composition's does (loss = sum of the down-sampled outputs)."""
    MO, O, B = env
    mode, be = backend
    from deeplearning.scala_b200.plugins import Plugins
    from deeplearning.scala_b200.layers import Layers
    rng = np.random.RandomState(17)
    x = rng.standard_normal((4, 16, 16, 3)).astype(np.float32)
    wt = (rng.standard_normal((3, 3, 3, 8)) * 0.3).astype(np.float32)
    bias = (rng.standard_normal(8) * 0.1).astype(np.float32)
    seed = rng.standard_normal((4, 8, 8, 8)).astype(np.float32)
    results = []
    for fold in (True, False):
        P = Plugins(backend=be, optimizer_delta=lambda w_, d: d * 0.1)
        L = Layers(P)
        w_, b_ = P.weight(be.tensor(wt)), P.weight(be.tensor(bias))
        conv = L.conv2d_3x3(be.tensor(x), w_, b_, "relu")
        if not fold:
            del conv._conv_spec
        out = L.downsample(conv, 2) * be.tensor(seed)
        val = out.train().map(lambda t: t.flat_array()).run()
        P.flush_updates()
        assert not P.handled_exceptions, P.handled_exceptions
        results.append((val, w_.data.flat_array(), b_.data.flat_array()))
    (v1, w1, b1), (v2, w2, b2) = results
    assert (v1.view(np.uint32) == v2.view(np.uint32)).all()
    gw1, gw2 = (wt.reshape(-1) - w1).astype(np.float64), (wt.reshape(-1) - w2).astype(np.float64)
    assert np.abs(gw1 - gw2).max() <= 1e-5 * np.abs(gw2).max()
    assert np.abs((bias - b1) - (bias - b2)).max() <= 1e-5 * np.abs(bias - b2).max()
    assert np.abs(gw2).max() > 0


@pytest.mark.parametrize("case", [(256, 16, 16, 32, 64, 2, 2), (6, 8, 8, 64, 128, 1, 2), (5, 16, 16, 32, 64, 2, 1), (3, 8, 8, 64, 32, 1, 1)],
                         ids=["c3-conv2", "c3-conv3", "sampled-dz-dense-gate", "dense"])
@pytest.mark.parametrize("gate_mode", ["relu", "tanh"])
def test_backward_data_with_the_derivative_of_the_layer_below(env, backend, case, gate_mode):
    """dls_conv3x3_backward_data_gated: dx * act'(y[b, gs*y, gs*x, c]) in the epilogue equals, bit for bit, the plain pass
    followed by the elementwise tree on the `scale`d view of y."""
    MO, O, B = env
    mode, be = backend
    batch, h, w, cin, cout, dzs, gs = case
    if not be.conv3x3_supported("backward_data", h, w, cin, cout):
        pytest.skip("no dedicated backward-data kernel for this shape in %s" % mode)
    rng = np.random.RandomState(sum(case))
    wt = (rng.standard_normal((3, 3, cin, cout)) * 0.2).astype(np.float32)
    dz = rng.standard_normal((batch, h // dzs, w // dzs, cout)).astype(np.float32)
    y = rng.standard_normal((batch, h * gs, w * gs, cin)).astype(np.float32)
    y = np.maximum(y, 0) if gate_mode == "relu" else np.tanh(y)
    dzt, wtt, yt = be.tensor(dz).do_cache(), be.tensor(wt).do_cache(), be.tensor(y).do_cache()
    got = be.conv3x3_backward_data(dzt, wtt, dz_stride=dzs, gate=yt, gate_mode=gate_mode, gate_stride=gs).flat_array()
    plain = be.conv3x3_backward_data(dzt, wtt, dz_stride=dzs).do_cache()
    ys = yt.scale([batch, h, w, cin]) if gs > 1 else yt
    if gate_mode == "tanh":
        want = (plain * (1.0 - ys * ys)).flat_array()
    else:
        want = (plain * B.minimum(ys * 3.4028234663852886e38, be.scalar(1.0))).flat_array()
    assert (got.view(np.uint32) == want.view(np.uint32)).all()


@pytest.mark.parametrize("shape", [(256, 32, 32, 3, 32, 2), (5, 12, 20, 4, 32, 1), (4, 16, 16, 1, 8, 2)],
                         ids=["c3-conv1-sampled", "dense", "one-channel-sampled"])
def test_first_layer_weight_and_bias_gradient_in_one_pass(env, backend, shape):
    """dls_conv3x3_backward_weights_bias: the first-layer kernel returns dW and the bias gradient from one pass over the
    delta; dW as from the plain entry, db = the fp32 column sums of the unrounded delta (1e-5 relative, the reductions bar)."""
    MO, O, B = env
    mode, be = backend
    batch, h, w, cin, cout, s = shape
    rng = np.random.RandomState(sum(shape))
    x = rng.standard_normal((batch, h, w, cin)).astype(np.float32)
    dzs = rng.standard_normal((batch, h // s, w // s, cout)).astype(np.float32)
    xt, dt = be.tensor(x).do_cache(), be.tensor(dzs).do_cache()
    n0 = be.stats()["kernel_launches"]
    dw, db = be.conv3x3_backward_weights_bias(xt, dt, dz_stride=s)
    got_db = db.flat_array()
    got_dw = dw.flat_array()
    assert be.stats()["kernel_launches"] - n0 == 2                      # the pass + its fixed-order reduce, once for both
    ref_dw = be.conv3x3_backward_weights(xt, dt, dz_stride=s).flat_array()
    assert (got_dw.view(np.uint32) == ref_dw.view(np.uint32)).all()
    want_db = dzs.astype(np.float64).sum(axis=(0, 1, 2))
    scale = np.abs(dzs).astype(np.float64).sum(axis=(0, 1, 2))
    assert (np.abs(got_db - want_db) / scale).max() <= 1e-6
