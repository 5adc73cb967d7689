import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B
import deeplearning.scala_b200.harness as H
B.default_backend().ctx
rng = np.random.RandomState(0)
for mode in ("fp32", "bf16", "tf32"):
    be = B.Backend(gemm_mode=mode)
    m = H.Mlp((256, 256, 256), bias=False, last_activation="tanh", backend=be, learning_rate=0.01)
    x = be.tensor(rng.rand(256, 384).astype(np.float32))
    w0 = [w_.data.flat_array().copy() for w_, _ in m.layers]
    loss = m.step(x)
    print(mode, "loss", loss, "exceptions", m.P.handled_exceptions)
    for (w_, _), old in zip(m.layers, w0):
        print("  moved", np.abs(w_.data.flat_array() - old).max())
    be.close()
