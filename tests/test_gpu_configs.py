"""Config-level parity at the BENCHED sizes (BASELINE.json configs C3 and C4, built by the same `bench_train.build` the
bench times) against oracle/models_oracle.py, plus the tensor-core contraction at C4's full GEMM shape.

The small-size model tests (tests/test_gpu_models.py) cannot reach what only the full sizes exercise: the convolution
weight gradients contract over K = 262 144 / 65 536 / 16 384 pixels (split-K), conv2 / conv3 run many waves of tiles, the
down-sampling backward gathers cross image boundaries thousands of times, and an 8192 x 4096 x 4096 product covers 1024
output tiles (7 waves, grouped raster wrap-around).

Bars (same as tests/test_gpu_models.py): fp32 mode 1e-5 relative; tensor-core modes 2e-4 against the oracle evaluated
with the same operand rounding (bf16 RNE / tf32 RNA), i.e. only the accumulation order differs.
"""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

LR = 0.01


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


@pytest.fixture(scope="module")
def env():
    from oracle import models_oracle as MO
    from oracle import tensors_oracle as O
    import deeplearning.scala_b200.tensors as B
    import bench_train
    return MO, O, B, bench_train


def read_pair(pair):
    return pair[0].data.flat_array().reshape(pair[0].data.shape), pair[1].data.flat_array()


@pytest.mark.parametrize("mode,tol", [("fp32", 1e-5), ("tf32", 2e-4), ("bf16", 2e-4)])
@pytest.mark.parametrize("compiled", [False, True], ids=["eager", "captured"])
def test_c3_step_at_the_benched_size(env, mode, tol, compiled):
    """ONE step of the benched C3 (batch 256, 32x32x3, 3-32-64-128, fc 8192-256-10): loss, every bias gradient and every
    entry of every weight gradient against `cnn_step`."""
    MO, O, B, bench_train = env
    if compiled and mode == "fp32":
        pytest.skip("the captured step is covered in the two tensor-core modes (fp32 contractions are the parity mode)")
    be = B.Backend(gemm_mode=mode)
    try:
        model, x, t, batch, _, _ = bench_train.build("c3", be, learning_rate=LR)
        assert batch == 256 and x.shape == (256, 32, 32, 3)
        before = [read_pair(p) for p in model.convs] + [read_pair(model.fc1), read_pair(model.fc2)]
        rounding = None if mode == "fp32" else (lambda a: O.round_operand(a, mode))
        want_loss, nconvs, nfc1, nfc2 = MO.cnn_step(before[:3], before[3], before[4], x, t, LR, rounding)
        want = nconvs + [nfc1, nfc2]
        if compiled:
            # the recording runs one eager step first (weights move), so replay from restored weights
            xd, td = be.tensor(x).do_cache(), be.tensor(t).do_cache()
            step = model.compile(xd, td)
            for (w, b), (w0, b0) in zip(model.convs + [model.fc1, model.fc2], before):
                be.write(w.data._buffer(), w0)
                be.write(b.data._buffer(), b0)
            got_loss = step()
        else:
            got_loss = model.step(be.tensor(x), be.tensor(t))
        got = [read_pair(p) for p in model.convs] + [read_pair(model.fc1), read_pair(model.fc2)]
        assert abs(got_loss - want_loss) <= tol * abs(want_loss), (got_loss, want_loss)
        names = ["conv1", "conv2", "conv3", "fc1", "fc2"]
        for name, (gw, gb), (rw, rb), (ow, ob) in zip(names, got, want, before):
            assert np.isfinite(gw).all() and np.isfinite(gb).all()
            assert rel(ow - gw, ow - rw) <= tol, "weight gradient of %s: %g" % (name, rel(ow - gw, ow - rw))
            assert rel(ob - gb, ob - rb) <= tol, "bias gradient of %s: %g" % (name, rel(ob - gb, ob - rb))
            # no stray element: the worst entry stays within 50 x tol of the gradient's rms (catches one bad tile that
            # a Frobenius norm over 2 M entries would hide)
            g_got, g_ref = (ow - gw).astype(np.float64), (ow - rw).astype(np.float64)
            rms = np.sqrt((g_ref ** 2).mean()) + 1e-30
            assert np.abs(g_got - g_ref).max() <= 50 * tol * rms + 4 * np.spacing(np.abs(ow).max()), name
    finally:
        be.close()


@pytest.mark.parametrize("mode", ["tf32", "bf16"])
def test_c4_model_at_width_4096_on_a_256_sample_slice(env, mode):
    """the 4 x 4096 tanh MLP of C4 (no bias, loss = sum of outputs) at its real width, batch cut to 256 so the
    float64 oracle finishes in seconds."""
    MO, O, B, _ = env
    import deeplearning.scala_b200.harness as H
    be = B.Backend(gemm_mode=mode)
    try:
        model = H.Mlp((4096,) * 5, bias=False, last_activation="tanh", backend=be, learning_rate=LR)
        x = np.random.RandomState(9).standard_normal((4096, 256)).astype(np.float32)
        before = [(w.data.flat_array().reshape(w.data.shape), None) for w, _ in model.layers]
        got_loss = model.step(be.tensor(x), None)
        want_loss, want, _ = MO.mlp_step(before, x, None, LR, "tanh", lambda a: O.round_operand(a, mode))
        got = [w.data.flat_array().reshape(w.data.shape) for w, _ in model.layers]
        assert abs(got_loss - want_loss) <= 2e-4 * abs(want_loss) + 1e-3, (got_loss, want_loss)
        for i, (gw, (rw, _), (ow, _)) in enumerate(zip(got, want, before)):
            assert rel(ow - gw, ow - rw) <= 2e-4, "weight gradient of layer %d: %g" % (i, rel(ow - gw, ow - rw))
    finally:
        be.close()


@pytest.mark.parametrize("mode", ["bf16", "tf32"])
@pytest.mark.parametrize("shape", [(8192, 4096, 4096, False, False), (4096, 8192, 4096, False, False), (4096, 4096, 8192, False, True),
                                   (4096, 4096, 8192, True, False)],
                         ids=["8192x4096x4096", "c4-forward", "c4-dW", "c4-dX"])
def test_contraction_at_c4_shape_on_sampled_rows(env, mode, shape):
    """The benched GEMM shapes: one random row of EVERY 128-row tile (so every output tile of every wave is touched) against
    the operand-rounded float64 product; bound 2e-6 * sum |a||b| as in tests/test_gpu_gemm.py."""
    MO, O, B, _ = env
    m, n, k, ta, tb = shape
    rng = np.random.RandomState(m + n + k)
    a = rng.standard_normal((k, m) if ta else (m, k)).astype(np.float32)
    b = rng.standard_normal((n, k) if tb else (k, n)).astype(np.float32)
    got = B.matmul(B.tensor(a), B.tensor(b), transpose_a=ta, transpose_b=tb, mode=mode).flat_array().reshape(m, n)
    rows = np.array([t * 128 + rng.randint(0, min(128, m - t * 128)) for t in range((m + 127) // 128)])
    ar = O.round_operand(a, mode).astype(np.float64)
    br = O.round_operand(b, mode).astype(np.float64)
    A = (ar.T if ta else ar)[rows]
    Bm = br.T if tb else br
    want = A @ Bm
    scale = np.abs(A) @ np.abs(Bm) + 1.0
    err = np.abs(got[rows].astype(np.float64) - want) / scale
    assert err.max() <= 2e-6, "row %d col %d: %g" % (rows[err.argmax() // n], err.argmax() % n, err.max())
    assert np.isfinite(got).all()
    # and a cheap whole-matrix property: column sums of the product = (column sums of op(A)) . op(B)
    colsum_want = (ar.T if ta else ar).sum(axis=0) @ Bm
    colsum_scale = np.abs(ar.T if ta else ar).sum(axis=0) @ np.abs(Bm) + 1.0
    assert (np.abs(got.astype(np.float64).sum(axis=0) - colsum_want) / colsum_scale).max() <= 2e-6
