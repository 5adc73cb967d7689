"""One SGD step of the BASELINE.json model configs through the product stack (TensorWeights + TensorLayers + the custom
layers of layers.py, i.e. tcgen05 GEMM / gather / reduction / fused elementwise kernels behind the C ABI) against the
plain-numpy restatement of the reference's autodiff rules in oracle/models_oracle.py.

Bars: fp32 mode -- loss and every updated parameter within 1e-5 relative (reductions bar of north_star);
tensor-core modes -- gradients within 2e-4 of the oracle evaluated with the same operand rounding (only the
accumulation order differs), and within the stated 1e-3 (TF32) / 4e-3 (BF16) of the fp32 oracle.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LR = 0.01


@pytest.fixture(scope="module")
def env():
    from oracle import models_oracle as MO
    from oracle import tensors_oracle as O
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    B.default_backend().ctx
    return MO, O, B, H


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def make_backend(B, mode):
    be = B.Backend(gemm_mode=mode)
    be.ctx
    return be


def mlp_case(env, mode, sizes, batch, bias, last_activation, with_target, steps=2):
    MO, O, B, H = env
    be = make_backend(B, mode)
    try:
        model = H.Mlp(sizes, bias=bias, last_activation=last_activation, backend=be, learning_rate=LR)
        rng = np.random.RandomState(7)
        x = rng.rand(sizes[0], batch).astype(np.float32)
        target = H.one_hot(rng.randint(0, sizes[-1], size=batch), sizes[-1], feature_major=True) if with_target else None
        ref = [(w.data.flat_array().reshape(w.data.shape), None if b is None else b.data.flat_array()) for w, b in model.layers]
        rounding = None if mode == "fp32" else (lambda a: O.round_operand(a, mode))
        out = []
        for _ in range(steps):
            got_loss = model.step(be.tensor(x), None if target is None else be.tensor(target))
            before = ref
            want_loss, ref, _ = MO.mlp_step(ref, x, target, LR, last_activation, rounding)
            _, ref32, _ = MO.mlp_step(before, x, target, LR, last_activation, None)
            got = [(w.data.flat_array().reshape(w.data.shape), None if b is None else b.data.flat_array()) for w, b in model.layers]
            out.append((got_loss, want_loss, before, got, ref, ref32))
            ref = got            # continue from the product's weights so steps do not compound differences
        return out
    finally:
        be.close()


@pytest.mark.parametrize("sizes,batch", [((20, 16, 6), 12), ((784, 500, 10), 128)], ids=["tiny", "c1"])
def test_mlp_step_fp32(env, sizes, batch):
    for got_loss, want_loss, before, got, ref, _ in mlp_case(env, "fp32", sizes, batch, True, "none", True):
        assert abs(got_loss - want_loss) <= 1e-5 * abs(want_loss)
        for (gw, gb), (rw, rb), (ow, ob) in zip(got, ref, before):
            assert rel(ow - gw, ow - rw) <= 1e-5          # the applied update lr * gradient
            assert rel(ob - gb, ob - rb) <= 1e-5
            assert rel(gw, rw) <= 1e-6


@pytest.mark.parametrize("mode,stated", [("tf32", 1e-3), ("bf16", 4e-3)])
def test_c1_mlp_step_on_tensor_cores(env, mode, stated):
    for got_loss, want_loss, before, got, ref, ref32 in mlp_case(env, mode, (784, 500, 10), 128, True, "none", True):
        assert abs(got_loss - want_loss) <= 2e-4 * abs(want_loss)
        for (gw, gb), (rw, rb), (fw, fb), (ow, ob) in zip(got, ref, ref32, before):
            assert rel(ow - gw, ow - rw) <= 2e-4
            assert rel(ob - gb, ob - rb) <= 2e-4
            assert rel(ow - gw, ow - fw) <= 3 * stated    # gradients chain three rounded contractions
            assert rel(gw, fw) <= stated


def test_wide_mlp_sum_loss_bf16(env):
    """C4's structure (no bias, tanh everywhere, loss = sum of outputs) at a size the oracle finishes quickly."""
    for got_loss, want_loss, before, got, ref, _ in mlp_case(env, "bf16", (256, 256, 256), 384, False, "tanh", False, steps=1):
        assert abs(got_loss - want_loss) <= 2e-4 * abs(want_loss) + 1e-3
        for (gw, _), (rw, _), (ow, _) in zip(got, ref, before):
            assert rel(ow - gw, ow - rw) <= 2e-4


def cnn_case(env, mode, image, channels, hidden, batch):
    MO, O, B, H = env
    be = make_backend(B, mode)
    try:
        model = H.Cnn(image=image, channels=channels, hidden=hidden, classes=10, backend=be, learning_rate=LR)
        rng = np.random.RandomState(21)
        x = rng.rand(batch, image, image, channels[0]).astype(np.float32)
        target = H.one_hot(rng.randint(0, 10, size=batch), 10, feature_major=False)
        read = lambda pair: (pair[0].data.flat_array().reshape(pair[0].data.shape), pair[1].data.flat_array())
        convs, fc1, fc2 = [read(p) for p in model.convs], read(model.fc1), read(model.fc2)
        rounding = None if mode == "fp32" else (lambda a: O.round_operand(a, mode))
        got_loss = model.step(be.tensor(x), be.tensor(target))
        want_loss, nconvs, nfc1, nfc2 = MO.cnn_step(convs, fc1, fc2, x, target, LR, rounding)
        got = [read(p) for p in model.convs] + [read(model.fc1), read(model.fc2)]
        return got_loss, want_loss, convs + [fc1, fc2], got, nconvs + [nfc1, nfc2]
    finally:
        be.close()


@pytest.mark.parametrize("mode,tol", [("fp32", 1e-5), ("tf32", 2e-4), ("bf16", 2e-4)])
def test_cnn_step(env, mode, tol):
    got_loss, want_loss, before, got, ref = cnn_case(env, mode, 8, (3, 4, 6, 8), 16, 4)
    assert abs(got_loss - want_loss) <= tol * abs(want_loss)
    for i, ((gw, gb), (rw, rb), (ow, ob)) in enumerate(zip(got, ref, before)):
        assert rel(ow - gw, ow - rw) <= tol, "weight gradient of layer %d" % i
        assert rel(ob - gb, ob - rb) <= tol, "bias gradient of layer %d" % i


def _dp_rank(rank, world, uid_queue, out_queue, kind):
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    be = B.Backend(device=rank, gemm_mode="fp32")
    be.ctx
    if rank == 0:
        uid = H.Exchange.unique_id()
        for _ in range(world - 1):
            uid_queue.put(uid)
    else:
        uid = uid_queue.get(timeout=120)
    ex = H.Exchange(be, rank, world, uid)
    ex.verify()
    rng = np.random.RandomState(5)
    if kind == "mlp":
        model = H.Mlp((24, 16, 6), bias=True, backend=be, learning_rate=LR, exchange=ex)
        x = rng.rand(24, 16).astype(np.float32)
        t = H.one_hot(rng.randint(0, 6, size=16), 6, feature_major=True)
        axis = 1
        pairs = model.layers
    else:                                # the C5 model (C3's CNN) at a size the oracle steps in milliseconds
        model = H.Cnn(image=8, channels=(3, 4, 6, 8), hidden=16, classes=10, backend=be, learning_rate=LR, exchange=ex)
        x = rng.rand(8, 8, 8, 3).astype(np.float32)
        t = H.one_hot(rng.randint(0, 10, size=8), 10, feature_major=False)
        axis = 0
        pairs = model.convs + [model.fc1, model.fc2]
    read = lambda: [(w.data.flat_array().reshape(w.data.shape), b.data.flat_array()) for w, b in pairs]
    before = read()
    loss = model.step(be.tensor(H.shard(x, axis, rank, world)), be.tensor(H.shard(t, axis, rank, world)))
    after = read()
    out_queue.put((rank, loss, before, after, x, t))
    be.close()


@pytest.mark.parametrize("kind", ["mlp", "cnn"])
def test_data_parallel_step_equals_the_full_batch_step(env, kind):
    """C5's exchange on two B200s: replicas initialised from the same counter-based RNG, batch sharded, every weight
    gradient SUM-all-reduced (NCCL for the large ones, the NVLink peer-memory launch for the small ones) before the
    update  ==  the oracle's full-batch step; `cnn` is the C5 model at a small size."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (ranks must not share a device)")
    import torch.multiprocessing as mp
    MO = env[0]
    ctx = mp.get_context("spawn")
    uid_q, out_q = ctx.Queue(), ctx.Queue()
    procs = [ctx.Process(target=_dp_rank, args=(r, 2, uid_q, out_q, kind)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted([out_q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        if p.is_alive():
            p.kill()
            p.join(timeout=10)
            pytest.fail("a rank process did not exit after delivering its result")
    (_, loss0, before, after0, x, t), (_, loss1, _, after1, _, _) = results
    if kind == "mlp":
        want_loss, want, _ = MO.mlp_step(before, x, t, LR)
    else:
        want_loss, nconvs, nfc1, nfc2 = MO.cnn_step(before[:3], before[3], before[4], x, t, LR)
        want = nconvs + [nfc1, nfc2]
    assert abs((loss0 + loss1) - want_loss) <= 1e-5 * abs(want_loss)
    for (w0, b0), (w1, b1), (rw, rb), (ow, ob) in zip(after0, after1, want, before):
        assert (w0 == w1).all() and (b0 == b1).all()            # replicas stay bit-identical
        assert rel(ow - w0, ow - rw) <= 1e-5
        assert rel(ob - b0, ob - rb) <= 1e-5


@pytest.mark.parametrize("mode", ["tf32", "bf16"])
def test_update_in_the_gradient_contraction_equals_the_separate_update(env, mode):
    """Long contractions (batch >= 1024) apply `w - lr * dw` in the weight-gradient GEMM's epilogue on a single GPU: two
    steps must leave exactly the weights the materialise-then-update path leaves."""
    MO, O, B, H = env
    sizes, batch = (96, 160, 64, 32), 1024
    x = np.random.RandomState(21).standard_normal((sizes[0], batch)).astype(np.float32)
    results = []
    for fused in (True, False):
        old = H.ExchangeSgd.fuse_contractions
        H.ExchangeSgd.fuse_contractions = fused
        be = B.Backend(gemm_mode=mode)
        try:
            model = H.Mlp(sizes, bias=True, last_activation="tanh", backend=be, learning_rate=0.01)
            n0 = be.stats()["kernel_launches"]
            losses = [model.step(be.tensor(x), None) for _ in range(2)]
            launches = be.stats()["kernel_launches"] - n0
            results.append((losses, [p.copy() for p in model.parameters()], launches))
        finally:
            H.ExchangeSgd.fuse_contractions = old
            be.close()
    (l1, p1, n1), (l2, p2, n2) = results
    assert l1 == l2
    for u, v in zip(p1, p2):
        assert (u.view(np.uint32) == v.view(np.uint32)).all()
    assert n1 < n2 if mode == "bf16" else n1 <= n2     # bf16: the weights' staging passes are gone as well
