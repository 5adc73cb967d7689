"""Weight checkpoints (deeplearning/scala_b200/checkpoint.py): the file format on the CPU, resume-equals-uninterrupted on
the GPU (weights, Adam state and step count), restore into the buffers a captured step was recorded on."""
import struct

import numpy as np
import pytest

F = np.float32


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


@pytest.fixture()
def ck():
    import deeplearning.scala_b200.checkpoint as ck
    return ck


def test_round_trip_keeps_bits_shapes_and_order(ck, tmp_path):
    rng = np.random.RandomState(0)
    arrays = {"w0": rng.standard_normal((500, 784)).astype(F), "w1": rng.standard_normal(500).astype(F),
              "scalar": np.array(3.5, dtype=F), "empty": np.zeros((0, 4), dtype=F), "odd": rng.standard_normal(17).astype(F),
              "special": np.array([np.nan, np.inf, -np.inf, -0.0, 1e-45], dtype=F)}
    path = str(tmp_path / "a.dlsw")
    n = ck.save_arrays(path, arrays)
    assert n % 64 == 0 or n > 0
    back = ck.load_arrays(path)
    assert list(back) == list(arrays)
    for k in arrays:
        assert back[k].shape == arrays[k].shape and back[k].dtype == F
        assert (bits(back[k]) == bits(arrays[k])).all()
    raw = open(path, "rb").read()
    assert raw[:4] == b"DLSW" and struct.unpack_from("<II", raw, 4) == (1, len(arrays))


def test_malformed_files_are_rejected(ck, tmp_path):
    path = str(tmp_path / "b.dlsw")
    ck.save_arrays(path, {"w0": np.arange(100, dtype=F)})
    raw = open(path, "rb").read()
    for name, blob in (("magic", b"XXXX" + raw[4:]), ("version", raw[:4] + struct.pack("<I", 9) + raw[8:]),
                       ("truncated-table", raw[:24]), ("truncated-arena", raw[:-200])):
        bad = str(tmp_path / (name + ".dlsw"))
        open(bad, "wb").write(blob)
        with pytest.raises(ValueError):
            ck.load_arrays(bad)
    with pytest.raises(ValueError, match="float32"):
        ck.save_arrays(path, {"w0": np.arange(4, dtype=np.float64)})


def _model(B, H, Opt, optimizer):
    be = B.Backend(gemm_mode="fp32")
    opt = {"sgd": None, "adam": Opt.Adam(1e-3), "momentum": Opt.Momentum(1e-3, 0.9)}[optimizer]
    model = H.Mlp((20, 16, 6), bias=True, last_activation="none", backend=be, learning_rate=1e-3, optimizer=opt)
    rng = np.random.RandomState(3)
    x = rng.rand(20, 12).astype(F)
    t = H.one_hot(rng.randint(0, 6, size=12), 6, feature_major=True)
    return be, model, be.tensor(x).do_cache(), be.tensor(t).do_cache()


@pytest.mark.gpu
@pytest.mark.parametrize("optimizer", ["sgd", "adam", "momentum"])
def test_resume_equals_the_uninterrupted_run(ck, tmp_path, optimizer):
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    import deeplearning.scala_b200.optimizers as Opt
    path = str(tmp_path / "run.dlsw")
    be, model, x, t = _model(B, H, Opt, optimizer)
    try:
        for _ in range(2):
            model.step(x, t)
        ck.save(path, model)
        want_losses = [model.step(x, t) for _ in range(3)]
        want = [p.copy() for p in model.parameters()]
    finally:
        be.close()
    be, model, x, t = _model(B, H, Opt, optimizer)         # a fresh process would look like this: new init, new buffers
    try:
        ck.restore(path, model)
        got_losses = [model.step(x, t) for _ in range(3)]
        got = [p.copy() for p in model.parameters()]
    finally:
        be.close()
    assert got_losses == want_losses
    for g, w in zip(got, want):
        assert (bits(g) == bits(w)).all()


@pytest.mark.gpu
def test_restore_into_a_recorded_step_and_shape_mismatch(ck, tmp_path):
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    import deeplearning.scala_b200.optimizers as Opt
    path = str(tmp_path / "run.dlsw")
    be, model, x, t = _model(B, H, Opt, "sgd")
    try:
        step = model.compile(x, t)                            # one eager step + the recording
        ck.save(path, model)
        first = [step() for _ in range(3)]
        ck.restore(path, model)                               # back to the saved weights, written into the same buffers
        again = [step() for _ in range(3)]
        assert again == first
        other = H.Mlp((20, 8, 6), bias=True, last_activation="none", backend=be, learning_rate=1e-3)
        with pytest.raises(ValueError, match="shape"):
            ck.restore(path, other)
    finally:
        be.close()
