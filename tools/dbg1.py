import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B
import deeplearning.scala_b200.harness as H
be = B.default_backend()
rng = np.random.RandomState(0)
w = B.tensor(rng.randn(256, 256).astype(np.float32)).do_cache()
a = B.tensor(rng.randn(256, 64).astype(np.float32))
b = B.tensor(rng.randn(64, 256).astype(np.float32))
d = B.matmul(a, b, mode="fp32")
expr = w - d * 0.01
print(be.describe_plan([expr.closure()], expr.shape))
new = expr.do_cache()
W, D, NW = w.flat_array(), d.flat_array(), new.flat_array()
print("update err", np.abs(NW - (W - D * np.float32(0.01))).max(), "moved", np.abs(NW - W).max())
# the model path
m = H.Mlp((256, 256, 256), bias=False, last_activation="tanh", backend=be, learning_rate=0.01)
x = B.tensor(rng.rand(256, 384).astype(np.float32))
w0 = [w_.data.flat_array().copy() for w_, _ in m.layers]
ids0 = [id(w_.data) for w_, _ in m.layers]
loss = m.step(x)
print("loss", loss, "exceptions", m.P.handled_exceptions)
for (w_, _), old, i0 in zip(m.layers, w0, ids0):
    print("moved", np.abs(w_.data.flat_array() - old).max(), "same object", id(w_.data) == i0)
