"""Optimizer plugins (deeplearning/scala_b200/optimizers.py, dls_optimizer_step) against oracle/optim_oracle.py.

CPU part: the oracle itself against closed forms and float64 references.  GPU part: the fused kernels are bit-exact
with the oracle (same operation order, one fp32 rounding per operation), the stateless `delta` overrides of
CumulativeFloatLayersSpec:18-47 stay one fused launch per weight, and a captured step with a stateful optimizer
reproduces the eager steps bit for bit (hyper-parameters are re-published before every replay)."""
import numpy as np
import pytest

from oracle import optim_oracle as OO

F = np.float32


def rnd(seed, *shape, scale=1.0):
    return (np.random.RandomState(seed).standard_normal(shape) * scale).astype(F)


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


# ---- oracle (CPU) ----------------------------------------------------------------------------------------------------------

def test_oracle_stateless_rules_follow_the_reference_spec():
    w, d = rnd(1, 5, 7), rnd(2, 5, 7)
    assert (OO.learning_rate(d, w, 0.03) == (d * F(0.03))).all()                    # CumulativeFloatLayersSpec:26-32
    assert (OO.l2(d, w, 0.1) == (d + w * F(0.1)).astype(F)).all()                   # :42-46
    assert (OO.l1(d, w, 0.1) == (d + np.sign(w) * F(0.1)).astype(F)).all()          # :34-40


def test_oracle_momentum_matches_the_closed_form():
    w0, g = rnd(3, 64), rnd(4, 64)
    h = OO.momentum_hyper(0.1, 0.5)
    w, v = w0.copy(), np.zeros_like(w0)
    for _ in range(3):                                  # constant gradient: v_k = g (1 + mu + mu^2 ...)
        w, v = OO.momentum_step(w, v, g, h)
    want_v = g.astype(np.float64) * (1 + 0.5 + 0.25)
    want_w = w0.astype(np.float64) - 0.1 * g.astype(np.float64) * (1 + 1.5 + 1.75)
    assert np.allclose(v, want_v, rtol=1e-6) and np.allclose(w, want_w, rtol=1e-5, atol=1e-6)
    wn, vn = OO.momentum_step(w0, np.zeros_like(w0), g, OO.momentum_hyper(0.1, 0.5, nesterov=True))
    assert np.allclose(wn, w0 - 0.1 * (g + 0.5 * g), rtol=1e-6)
    wd, _ = OO.momentum_step(w0, np.zeros_like(w0), g, OO.momentum_hyper(0.1, 0.0, weight_decay=0.01))
    assert np.allclose(wd, w0 - 0.1 * (g + 0.01 * w0), rtol=1e-6)


def test_oracle_adam_matches_a_float64_reference():
    w0 = rnd(5, 257)
    w, m, v = w0.copy(), np.zeros_like(w0), np.zeros_like(w0)
    w64, m64, v64 = w0.astype(np.float64), np.zeros(257), np.zeros(257)
    lr, b1, b2, eps = 1e-2, 0.9, 0.999, 1e-8
    for t in range(1, 6):
        g = rnd(10 + t, 257)
        w, m, v = OO.adam_step(w, m, v, g, OO.adam_hyper(lr, b1, b2, eps, 0.0, t))
        m64 = b1 * m64 + (1 - b1) * g
        v64 = b2 * v64 + (1 - b2) * g.astype(np.float64) ** 2
        w64 = w64 - lr * (m64 / (1 - b1 ** t)) / (np.sqrt(v64 / (1 - b2 ** t)) + eps)
    assert np.allclose(w, w64, rtol=2e-5, atol=2e-6)
    first, _, _ = OO.adam_step(w0, np.zeros_like(w0), np.zeros_like(w0), rnd(11, 257), OO.adam_hyper(lr, b1, b2, eps, 0.0, 1))
    assert np.allclose(np.abs(first - w0), lr, rtol=1e-3)          # the first Adam step has magnitude lr


# ---- product (GPU) ---------------------------------------------------------------------------------------------------------

@pytest.fixture(scope="module")
def gpu():
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    import deeplearning.scala_b200.optimizers as Opt
    from deeplearning.scala_b200.plugins import Plugins
    be = B.Backend()
    be.ctx
    yield B, H, Opt, Plugins, be
    be.close()


def _weight_step(P, w, grad_tensor):
    """One `train` of the layer `w * g` whose delta w.r.t. w is exactly g (the seed delta is the scalar 1)."""
    (w * grad_tensor).train().run()
    P.flush_updates()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 3, 1024, 4099])
@pytest.mark.parametrize("kind", ["momentum", "nesterov_wd", "adam", "adam_wd"])
def test_fused_optimizer_kernels_are_bit_exact_with_the_oracle(gpu, kind, n):
    B, H, Opt, Plugins, be = gpu
    if kind == "momentum":
        opt, hyper = Opt.Momentum(0.05, 0.9), lambda t: OO.momentum_hyper(0.05, 0.9)
    elif kind == "nesterov_wd":
        opt, hyper = Opt.Momentum(0.05, 0.8, 1e-2, True), lambda t: OO.momentum_hyper(0.05, 0.8, 1e-2, True)
    elif kind == "adam":
        opt, hyper = Opt.Adam(1e-2), lambda t: OO.adam_hyper(1e-2, 0.9, 0.999, 1e-8, 0.0, t)
    else:
        opt, hyper = Opt.Adam(3e-3, 0.8, 0.99, 1e-6, 0.1), lambda t: OO.adam_hyper(3e-3, 0.8, 0.99, 1e-6, 0.1, t)
    P = Plugins(optimizer=opt, backend=be)
    w0 = rnd(20, n)
    w = P.weight(be.tensor(w0))
    want_w, s1, s2 = w0.copy(), np.zeros(n, F), np.zeros(n, F)
    for t in range(1, 5):
        g = rnd(30 + t, n)
        _weight_step(P, w, be.tensor(g))
        if kind.startswith("adam"):
            want_w, s1, s2 = OO.adam_step(want_w, s1, s2, g, hyper(t))
        else:
            want_w, s1 = OO.momentum_step(want_w, s1, g, hyper(t))
        got = w.data.flat_array()
        assert (bits(got) == bits(want_w)).all(), "step %d: %d elements differ" % (t, (bits(got) != bits(want_w)).sum())
    state = opt.state_arrays(w)
    assert (bits(state[0]) == bits(s1)).all()
    if kind.startswith("adam"):
        assert (bits(state[1]) == bits(s2)).all()


@pytest.mark.gpu
def test_stateless_rules_stay_one_fused_launch_and_match_the_oracle(gpu):
    B, H, Opt, Plugins, be = gpu
    w0, g = rnd(40, 33, 17), rnd(41, 33, 17)
    w0[0, :5] = 0.0                                                     # signum(0) == 0
    rule = Opt.chain(Opt.l2(0.01), Opt.l1(0.003), Opt.learning_rate(0.1))
    P = Plugins(optimizer_delta=rule, backend=be)
    w = P.weight(be.tensor(w0))
    gt = be.tensor(g).do_cache()
    w.data.flat_array()
    launches0 = be.stats()["kernel_launches"]
    (w * gt).train().run()
    assert be.stats()["kernel_launches"] - launches0 == 1               # data - delta: regularisers + lr + step fused
    d = OO.learning_rate(OO.l1(OO.l2(g, w0, 0.01), w0, 0.003), w0, 0.1)
    want = (w0 - d).astype(F)
    got = w.data.flat_array().reshape(33, 17)
    assert (bits(got) == bits(want)).all()


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["momentum", "adam"])
def test_compiled_step_with_a_stateful_optimizer_equals_eager(gpu, which):
    B, H, Opt, Plugins, _ = gpu

    def run(compiled):
        be = B.Backend(gemm_mode="fp32")
        try:
            opt = Opt.Momentum(1e-3, 0.9) if which == "momentum" else Opt.Adam(1e-3)
            model = H.Mlp((20, 16, 6), bias=True, last_activation="none", backend=be, optimizer=opt)
            rng = np.random.RandomState(3)
            x = rng.rand(20, 12).astype(F)
            t = H.one_hot(rng.randint(0, 6, size=12), 6, feature_major=True)
            xd, td = be.tensor(x).do_cache(), be.tensor(t).do_cache()
            losses = []
            if compiled:
                step = model.compile(xd, td)
                for _ in range(4):
                    losses.append(step())
            else:
                model.step(xd, td)
                for _ in range(4):
                    losses.append(model.step(xd, td))
            return losses, [p.copy() for p in model.parameters()], opt.t
        finally:
            be.close()
    want_l, want_p, want_t = run(False)
    got_l, got_p, got_t = run(True)
    assert got_t == want_t == 5
    assert got_l == want_l and len(set(got_l)) > 1
    for g, w in zip(got_p, want_p):
        assert (bits(g) == bits(w)).all()


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["momentum", "adam"])
def test_in_place_rules_see_the_pre_update_weights_of_other_weights(gpu, which):
    """`w1 * w2` under an in-place rule: the gradient of w2 is `d * w1.data`, a LAZY expression over w1's live buffer
    when the updates are flushed.  The reference swaps buffers (TW:63-72), so every delta reads pre-update values; here
    all pending deltas must be materialised before the first in-place step, and the data `train` yields must be the
    forward value of this step, not a recomputation from the updated weights."""
    B, H, Opt, Plugins, be = gpu
    n = 1000
    a0, b0 = rnd(50, n), rnd(51, n)
    if which == "momentum":
        opt, step, hyper = Opt.Momentum(0.05, 0.9), OO.momentum_step, (lambda t: OO.momentum_hyper(0.05, 0.9))
    else:
        opt, hyper = Opt.Adam(1e-2), (lambda t: OO.adam_hyper(1e-2, 0.9, 0.999, 1e-8, 0.0, t))
    P = Plugins(optimizer=opt, backend=be)
    w1, w2 = P.weight(be.tensor(a0)), P.weight(be.tensor(b0))
    a, b = a0.copy(), b0.copy()
    sa = [np.zeros(n, F), np.zeros(n, F)]
    sb = [np.zeros(n, F), np.zeros(n, F)]
    for t in range(1, 4):
        out = (w1 * w2).train().run().flat_array()
        P.flush_updates()
        assert (bits(out) == bits(a * b)).all(), "step %d: train() yielded data computed from updated weights" % t
        ga, gb = b.copy(), a.copy()                     # d(w1*w2)/dw1 = w2, /dw2 = w1, both PRE-update (seed delta 1.0)
        if which == "momentum":
            a, sa[0] = OO.momentum_step(a, sa[0], ga, hyper(t))
            b, sb[0] = OO.momentum_step(b, sb[0], gb, hyper(t))
        else:
            a, sa[0], sa[1] = OO.adam_step(a, sa[0], sa[1], ga, hyper(t))
            b, sb[0], sb[1] = OO.adam_step(b, sb[0], sb[1], gb, hyper(t))
        assert (bits(w1.data.flat_array()) == bits(a)).all(), "w1 after step %d" % t
        assert (bits(w2.data.flat_array()) == bits(b)).all(), "w2 after step %d" % t
