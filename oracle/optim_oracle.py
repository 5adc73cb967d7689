"""TEST INFRASTRUCTURE -- CPU oracle for the optimizer rules behind `TensorOptimizerApi.delta` (never imported by the
product package; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use oracle/).

What the reference pins:
  * the extension point itself: `delta` is overridden by optimizer plugins, the default is the identity
    (plugins-Weights Weights.scala:46-51, plugins-TensorWeights TensorWeights.scala:155-161);
  * learning rate, L1 and L2 as `delta` overrides -- for scalars only, in a test
    (plugins-CumulativeFloatLayers/src/test/.../CumulativeFloatLayersSpec.scala:18-47):
        FixedLearningRate: delta = super.delta * learningRate
        L1Regularization:  delta = super.delta + signum(weight.data) * l1Regularization
        L2Regularization:  delta = super.delta + weight.data * l2Regularization
    `learning_rate`, `l1`, `l2` below restate those three for tensors (elementwise fp32, one rounding per operation).
What it does not pin: momentum SGD and Adam exist only as README links (README.md:70).  PARITY UNPINNED for those two:
the rules below are the published algorithms (Polyak / Nesterov momentum; Kingma & Ba 2015, algorithm 1, with the
L2-style weight decay of its reference code), written with exactly the operation order of
deeplearning/scala_b200/csrc/aux_kernels.cu `momentum_one` / `adam_one`, one fp32 rounding per operation, no FMA.
"""
import numpy as np

F = np.float32


def learning_rate(delta, weight, lr):
    return (np.asarray(delta, F) * F(lr)).astype(F)


def l2(delta, weight, coefficient):
    return (np.asarray(delta, F) + np.asarray(weight, F) * F(coefficient)).astype(F)


def l1(delta, weight, coefficient):
    return (np.asarray(delta, F) + np.sign(np.asarray(weight, F)).astype(F) * F(coefficient)).astype(F)


def momentum_hyper(lr, mu, weight_decay=0.0, nesterov=False):
    return np.array([lr, mu, weight_decay, 1.0 if nesterov else 0.0, 0.0, 0.0], dtype=F)


def adam_hyper(lr, beta1, beta2, eps, weight_decay, t):
    """Bias corrections are folded on the host in double precision, then rounded once to fp32."""
    return np.array([lr / (1.0 - beta1 ** t), beta1, beta2, eps, weight_decay, 1.0 / (1.0 - beta2 ** t)], dtype=F)


def momentum_step(w, v, g, hyper):
    """-> (w', v')"""
    lr, mu, wd, nesterov = (F(hyper[0]), F(hyper[1]), F(hyper[2]), hyper[3] != 0)
    w, v, g = np.asarray(w, F), np.asarray(v, F), np.asarray(g, F)
    if wd != 0:
        g = (g + (wd * w).astype(F)).astype(F)
    v = ((mu * v).astype(F) + g).astype(F)
    step = (g + (mu * v).astype(F)).astype(F) if nesterov else v
    w = (w - (lr * step).astype(F)).astype(F)
    return w, v


def adam_step(w, m, v, g, hyper):
    """-> (w', m', v')"""
    lr_c, b1, b2, eps, wd, inv_c2 = (F(h) for h in hyper)
    w, m, v, g = (np.asarray(a, F) for a in (w, m, v, g))
    one = F(1.0)
    if wd != 0:
        g = (g + (wd * w).astype(F)).astype(F)
    m = ((b1 * m).astype(F) + ((one - b1) * g).astype(F)).astype(F)
    v = ((b2 * v).astype(F) + (((one - b2) * g).astype(F) * g).astype(F)).astype(F)
    denom = (np.sqrt((v * inv_c2).astype(F)).astype(F) + eps).astype(F)
    w = (w - ((lr_c * m).astype(F) / denom).astype(F)).astype(F)
    return w, m, v
