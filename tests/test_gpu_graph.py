"""Captured steps (csrc/graph.cu, Backend.capture, harness.CompiledStep): a recorded step replayed N times must leave
exactly the bits the same N eager steps leave -- the same kernels run in the same order on the same values -- and the
recording's arena must stay out of the eager allocator's reach."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    from oracle import tensors_oracle as O
    return B, H, O


def rnd(seed, *shape):
    return np.random.RandomState(seed).standard_normal(shape).astype(np.float32)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_recorded_expression_replays_on_refreshed_inputs(env):
    B, _, O = env
    be = B.Backend()
    try:
        a0, b0, c0 = rnd(1, 96, 256), rnd(2, 96, 256), rnd(3, 96)
        a, b, c = be.tensor(a0).do_cache(), be.tensor(b0).do_cache(), be.tensor(c0).do_cache()
        expr = lambda M, x, y, z: M.tanh(x * y + z)                      # noqa: E731
        expr(be, a, b, c).flat_array()                                   # eager once: the plan exists
        launches0 = be.stats()["kernel_launches"]
        with be.capture() as step:
            y = expr(be, a, b, c).do_cache()
            s = y.sum().do_cache()
        assert be.stats()["kernel_launches"] == launches0 + 2             # counted while recording ...
        info = step.info()
        assert info["kernel_nodes"] == 2 and info["arena_bytes"] >= y.size * 4
        step.launch()
        assert be.stats()["kernel_launches"] == launches0 + 4             # ... and per replay
        want = np.asarray(expr(O, O.tensor(a0), O.tensor(b0), O.tensor(c0)).flat_array(), dtype=np.float32)
        got = y.flat_array()
        assert np.abs(bits(got).astype(np.int64) - bits(want).astype(np.int64)).max() <= 2
        # eager work between replays must not be handed the recording's buffers
        junk = [(be.tensor(rnd(10 + i, 96, 256)) * 2.0).do_cache() for i in range(8)]
        ptrs = {t._buffer().device_ptr for t in junk}
        assert y._buffer().device_ptr not in ptrs and s._buffer().device_ptr not in ptrs
        # new input values, same recording
        a1 = rnd(4, 96, 256)
        be.write(a._buffer(), a1)
        step.launch()
        want1 = np.asarray(expr(O, O.tensor(a1), O.tensor(b0), O.tensor(c0)).flat_array(), dtype=np.float32)
        got1 = y.flat_array()
        assert np.abs(bits(got1).astype(np.int64) - bits(want1).astype(np.int64)).max() <= 2
        assert abs(float(s.flat_array()[0]) - float(want1.astype(np.float64).sum())) <= 1e-5 * np.abs(want1).sum()
        assert all(np.isfinite(t.flat_array()).all() for t in junk)
    finally:
        be.close()


def test_host_transfers_are_rejected_while_recording_and_the_context_survives(env):
    B, _, _ = env
    be = B.Backend()
    try:
        a = be.tensor(rnd(5, 64, 64)).do_cache()
        (a + a).flat_array()
        with pytest.raises(ValueError, match="captured step"):
            with be.capture():
                (a + a).flat_array()                                     # D2H inside a recording
        with pytest.raises(ValueError, match="captured step"):
            with be.capture():
                be.tensor(rnd(6, 8)).do_cache()                          # H2D inside a recording
        got = (a + a).flat_array()                                       # the aborted captures left the context usable
        assert (bits(got) == bits(a.flat_array() * 2)).all()
        with pytest.raises(ValueError, match="no capture is open"):
            import ctypes
            from deeplearning.scala_b200 import _native as N
            N.check(N.lib.dls_graph_end(be.ctx, ctypes.byref(ctypes.c_void_p())))
    finally:
        be.close()


def test_multi_stream_contexts_refuse_to_record(env):
    B, _, _ = env
    from deeplearning.scala_b200._native import NativeError
    be = B.Backend(streams_per_device=2)
    try:
        with pytest.raises(NativeError):
            with be.capture():
                pass
    finally:
        be.close()


def _train(env, kind, mode, compiled, steps):
    B, H, _ = env
    be = B.Backend(gemm_mode=mode)
    try:
        rng = np.random.RandomState(11)
        if kind == "mlp":
            model = H.Mlp((784, 500, 10), bias=True, last_activation="none", backend=be, learning_rate=0.01 / 128)
            x = rng.rand(784, 128).astype(np.float32)
            t = H.one_hot(rng.randint(0, 10, size=128), 10, feature_major=True)
        else:
            model = H.Cnn(image=16, channels=(3, 8, 16, 32), hidden=32, classes=10, backend=be, learning_rate=1e-3)
            x = rng.rand(8, 16, 16, 3).astype(np.float32)
            t = H.one_hot(rng.randint(0, 10, size=8), 10, feature_major=False)
        xs = [x, x[::-1].copy() if kind == "cnn" else x[:, ::-1].copy(), x * 0.5]
        xd, td = be.tensor(xs[0]).do_cache(), be.tensor(t).do_cache()
        losses = []
        if compiled:
            step = model.compile(xd, td)                                 # runs step 0 eagerly, then records
            for i in range(1, steps):
                losses.append(step(xs[i % 3], t))
            launches = step.graph.info()["kernel_nodes"]
        else:
            losses.append(model.step(xd, td))
            for i in range(1, steps):
                losses.append(model.step(be.tensor(xs[i % 3]), be.tensor(t)))
            losses = losses[1:]
            launches = None
        return losses, [p.copy() for p in model.parameters()], launches
    finally:
        be.close()


@pytest.mark.parametrize("kind,mode", [("mlp", "fp32"), ("mlp", "bf16"), ("cnn", "bf16"), ("cnn", "tf32")])
def test_compiled_training_step_is_bit_identical_to_eager_steps(env, kind, mode):
    want_losses, want_params, _ = _train(env, kind, mode, False, 5)
    got_losses, got_params, launches = _train(env, kind, mode, True, 5)
    assert launches and launches >= 10
    assert got_losses == want_losses
    for g, w in zip(got_params, want_params):
        assert g.shape == w.shape and (bits(g) == bits(w)).all()
    assert len({float(v) for v in got_losses}) > 1                       # the weights really moved between replays


def test_a_step_that_outlives_its_backend_is_harmless(env):
    """The context takes the CUDA side of live recordings with it (NCCL would otherwise wait for them forever in
    ncclCommDestroy); the handle stays valid for an error return and for its release."""
    B, _, _ = env
    from deeplearning.scala_b200 import _native as N
    be = B.Backend()
    a = be.tensor(rnd(1, 64, 64)).do_cache()
    with be.capture() as step:
        y = (a * 2.0).do_cache()
    step.launch()
    assert np.isfinite(y.flat_array()).all()
    handle, step.handle = step.handle, None
    del a, y
    be.close()
    assert N.lib.dls_graph_launch(handle) != 0
    assert "destroyed" in N.last_error()
    N.lib.dls_graph_release(handle)


@pytest.mark.parametrize("compiled", [False, True])
def test_fused_single_gpu_update_equals_the_per_weight_swaps(env, monkeypatch, compiled):
    """The default single-GPU update of the harness (DLS_FUSED_SGD=0 turns it off): every gradient materialised, ONE
    in-place multi-tensor launch (dls_peer_sgd_step without exchange) instead of one `data - delta * lr` launch + swap
    (+ copy-back when captured) per weight.  Same arithmetic, so the same bits; fewer launches."""
    B, H, _ = env

    def run(fused):
        monkeypatch.setenv("DLS_FUSED_SGD", "1" if fused else "0")
        be = B.Backend(gemm_mode="fp32")
        try:
            model = H.Mlp((20, 16, 6), bias=True, last_activation="none", backend=be, learning_rate=1e-2)
            rng = np.random.RandomState(3)
            x = be.tensor(rng.rand(20, 12).astype(np.float32)).do_cache()
            t = be.tensor(H.one_hot(rng.randint(0, 6, size=12), 6, feature_major=True)).do_cache()
            if compiled:
                step = model.compile(x, t)
                launches0 = be.stats()["kernel_launches"]
                losses = [step() for _ in range(4)]
                per_step = (be.stats()["kernel_launches"] - launches0) / 4
                del step
            else:
                model.step(x, t)
                launches0 = be.stats()["kernel_launches"]
                losses = [model.step(x, t) for _ in range(4)]
                per_step = (be.stats()["kernel_launches"] - launches0) / 4
            return losses, [p.copy() for p in model.parameters()], per_step
        finally:
            be.close()
    want_l, want_p, want_n = run(False)
    got_l, got_p, got_n = run(True)
    assert got_l == want_l and len(set(got_l)) > 1
    for g, w in zip(got_p, want_p):
        assert (bits(g) == bits(w)).all()
    assert got_n < want_n
