"""Optimizer plugins for tensor weights, behind the reference's extension point `TensorOptimizerApi.delta`
(plugins-TensorWeights TensorWeights.scala:155-161; default = identity, plugins-Weights Weights.scala:46-51).

Two families:

* stateless `delta` overrides, stacked like the reference stacks mixins (`abstract override def delta = super.delta ...`,
  CumulativeFloatLayersSpec.scala:18-47): `learning_rate`, `l2`, `l1`, combined with `chain`.  They stay lazy tensor
  expressions, so the whole `data - delta` tree of TW:64 -- backward expression, regularisers, learning rate, SGD step --
  is still ONE fused launch per weight;
* stateful rules (`Momentum`, `Adam`): the gradient is materialised, then weight, state and gradient are updated in place
  by one kernel (`dls_optimizer_step`, csrc/aux_kernels.cu).  In-place means the swap of TW:63-72 is replaced, which is
  only safe after every launch that reads the old weight has been issued: `Plugins(optimizer=...)` postpones weight
  updates to the end of the step (`Plugins.flush_updates`, called by `TensorLayer.train`).
"""
from __future__ import annotations

from typing import Callable, Dict

import numpy as np

from . import _native as N
from . import tensors as T

_BIG = 3.4028234663852886e38


# -- stateless delta overrides (CumulativeFloatLayersSpec:18-47 for tensors) ---------------------------------------------

def learning_rate(lr: float) -> Callable:
    """FixedLearningRate: delta = super.delta * learningRate."""
    return lambda weight, delta: delta * float(lr)


def l2(coefficient: float) -> Callable:
    """L2Regularization: delta = super.delta + weight.data * l2Regularization."""
    return lambda weight, delta: delta + weight.data * float(coefficient)


def l1(coefficient: float) -> Callable:
    """L1Regularization: delta = super.delta + signum(weight.data) * l1Regularization.  The tensor API has no signum;
    clamp(w * MAX_FLOAT, -1, 1) equals it for every finite w (0 stays 0, the product saturates to +-inf otherwise)."""
    def rule(weight, delta):
        be = weight.data.backend
        sign = be.max(be.min(weight.data * _BIG, be.scalar(1.0)), be.scalar(-1.0))
        return delta + sign * float(coefficient)
    return rule


def chain(*rules: Callable) -> Callable:
    """Mixin linearisation: `chain(l2(c), learning_rate(lr))` == `with L2Regularization with FixedLearningRate`
    (the right-most rule is the outermost override, so it sees the others' result as `super.delta`)."""
    def rule(weight, delta):
        for r in rules:
            delta = r(weight, delta)
        return delta
    return rule


# -- stateful, fused ----------------------------------------------------------------------------------------------------

class _Fused:
    kind = -1
    n_state = 1

    def __init__(self):
        self.t = 0
        self._state: Dict[int, list] = {}
        self._hyper = None
        self._backend = None

    def hyper(self) -> np.ndarray:
        raise NotImplementedError

    def begin_step(self, backend=None):
        """Advance the step counter and publish this step's hyper-parameters to the device (24 bytes).  Called once per
        training step -- by `TensorLayer.train` eagerly, by `CompiledStep.launch` before each replay."""
        backend = backend or self._backend
        self.t += 1
        if backend is None:
            return
        self._backend = backend
        values = self.hyper()
        if self._hyper is None:
            self._hyper = backend.upload(values)
        else:
            backend.write(self._hyper, values)

    def state_of(self, weight):
        st = self._state.get(id(weight))
        if st is None:
            be = weight.data.backend
            st = [be.fill(0.0, weight.data.shape).do_cache() for _ in range(self.n_state)]
            self._state[id(weight)] = st
        return st

    def step(self, weight, delta):
        """weight.data, the state and the materialised delta go through one in-place kernel."""
        be = weight.data.backend
        if self._hyper is None:
            self.t = max(self.t, 1)
            self._backend = be
            self._hyper = be.upload(self.hyper())
        grad = delta.broadcast(weight.data.shape).do_cache() if list(delta.shape) != list(weight.data.shape) else delta.do_cache()
        st = self.state_of(weight)
        s2 = st[1]._buffer().handle if self.n_state > 1 else None
        N.check(N.lib.dls_optimizer_step(be.ctx, self.kind, weight.data._buffer().handle, st[0]._buffer().handle, s2,
                                         grad._buffer().handle, weight.data.size, self._hyper.handle))

    def state_arrays(self, weight):
        return [s.flat_array().reshape(weight.data.shape) for s in self.state_of(weight)]


class Momentum(_Fused):
    """v = mu v + (g + wd w);  w -= lr (nesterov ? g + mu v : v)."""
    kind = N.DLS_OPT_MOMENTUM

    def __init__(self, lr: float, momentum: float = 0.9, weight_decay: float = 0.0, nesterov: bool = False):
        super().__init__()
        self.lr, self.momentum, self.weight_decay, self.nesterov = lr, momentum, weight_decay, nesterov

    def hyper(self):
        return np.array([self.lr, self.momentum, self.weight_decay, 1.0 if self.nesterov else 0.0, 0.0, 0.0], dtype=np.float32)


class Adam(_Fused):
    """Kingma & Ba 2015, algorithm 1; weight decay enters the gradient (L2 form); bias corrections folded on the host."""
    kind = N.DLS_OPT_ADAM
    n_state = 2

    def __init__(self, lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.0):
        super().__init__()
        self.lr, self.beta1, self.beta2, self.eps, self.weight_decay = lr, beta1, beta2, eps, weight_decay

    def hyper(self):
        t = max(self.t, 1)
        return np.array([self.lr / (1.0 - self.beta1 ** t), self.beta1, self.beta2, self.eps, self.weight_decay,
                         1.0 / (1.0 - self.beta2 ** t)], dtype=np.float32)
