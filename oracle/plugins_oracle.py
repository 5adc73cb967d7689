"""CPU ORACLE (test infrastructure, NOT product code) for the DeepLearning.scala tensor plugins.

Restates, on top of `oracle/tensors_oracle.py`:
  * plugins-TensorLayers      (`TL`  = plugins-TensorLayers/src/main/scala/.../TensorLayers.scala)
  * plugins-TensorWeights     (`TW`  = plugins-TensorWeights/src/main/scala/.../TensorWeights.scala)
  * plugins-CumulativeTensorLayers (`CTL` = plugins-CumulativeTensorLayers/src/main/scala/.../CumulativeTensorLayers.scala)
  * plugins-TensorLiterals    (`TLit`)
The RAII.scala `Do` monad (asynchronous + scoped resource) is modelled synchronously: a `Do`
is a thunk `acquire() -> (value, release)`; `.shared` is reference counted exactly like
`Do.shared`, which is what makes CumulativeTensorLayers flush once on the last release.

Pinned by G14 (TLSpec:19-47), G15 (CTLSpec:66-127), G6 (CTLSpec:129-208).
Only tests/, smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

from . import tensors_oracle as T


# ---------------------------------------------------------------------------------------------
# Minimal synchronous model of com.thoughtworks.raii.asynchronous.Do
# ---------------------------------------------------------------------------------------------


class Do:
    def __init__(self, acquire: Callable[[], Tuple[object, Callable[[], None]]]):
        self._acquire = acquire

    def acquire(self):
        return self._acquire()

    @staticmethod
    def now(value) -> "Do":
        return Do(lambda: (value, lambda: None))

    def map(self, f) -> "Do":
        def acq():
            v, rel = self.acquire()
            return f(v), rel
        return Do(acq)

    def flat_map(self, f) -> "Do":
        def acq():
            v, rel = self.acquire()
            v2, rel2 = f(v).acquire()

            def release():
                rel2()
                rel()
            return v2, release
        return Do(acq)

    def shared(self) -> "Do":
        state = {"count": 0, "value": None, "release": None}

        def acq():
            if state["count"] == 0:
                state["value"], state["release"] = self.acquire()
            state["count"] += 1

            def release():
                state["count"] -= 1
                if state["count"] == 0:
                    rel, state["release"], state["value"] = state["release"], None, None
                    rel()
            return state["value"], release
        return Do(acq)

    def run(self):
        """Do.garbageCollected(...).run: acquire, take the value, release."""
        v, rel = self.acquire()
        rel()
        return v


def sequence(dos: Sequence[Do]) -> Do:
    def acq():
        vals, rels = [], []
        for d in dos:
            v, r = d.acquire()
            vals.append(v)
            rels.append(r)

        def release():
            for r in reversed(rels):
                r()
        return vals, release
    return Do(acq)


class Tape:
    """DeepLearning.scala:18 -- Tape(data, backward: Do[Delta] => UnitContinuation[Unit])."""

    def __init__(self, data, backward):
        self.data, self.backward = data, backward


# ---------------------------------------------------------------------------------------------
# TensorLayers
# ---------------------------------------------------------------------------------------------


def broadcast_then_sum(tensor: T.Tensor, shape) -> T.Tensor:
    """TL:110-121."""
    shape = list(shape)
    if len(tensor.shape) < len(shape):
        tensor = tensor.broadcast([tensor.shape[i] if i < len(tensor.shape) else shape[i] for i in range(len(shape))])
    return sum_as(tensor, shape)


def sum_as(tensor: T.Tensor, shape) -> T.Tensor:
    """TL:124-162 (host-unrolled `split(j).map(loop).reduce(_ + _)`, left fold)."""
    shape = list(shape)
    if len(shape) > len(tensor.shape):
        raise ValueError("Cannot sum %r to %r" % (tensor.shape, shape))

    def loop(t, i, j):
        if j < len(t.shape):
            if i >= len(shape):
                parts = [loop(s, i + 1, j) for s in t.split(j)]
            elif t.shape[i] != 1 and shape[i] == 1:
                # TL:139-148: the permute result is discarded in the reference; kept verbatim (defective branch).
                parts = [loop(s, i + 1, j + 1) for s in t.split(j)]
            elif t.shape[j] == shape[i]:
                return loop(t, i + 1, j + 1)
            else:
                raise ValueError("Cannot sum %r to %r" % (tensor.shape, shape))
            acc = parts[0]
            for p in parts[1:]:
                acc = acc + p
            return acc
        return t

    out = loop(tensor, 0, 0)
    assert list(out.shape) == shape, "ensuring(shape) failed (TL:159-161)"
    return out


class Differentiable:
    def forward(self) -> Do:
        raise NotImplementedError

    # operator sugar (plugins-Operators)
    def __add__(self, o): return self.plugins.plus(self, o)
    def __mul__(self, o): return self.plugins.times(self, o)


class Plugins:
    """One mixin stack (`Factory[TensorLayers with TensorWeights ...].newInstance()`).
    `cumulative=True` mixes in CumulativeTensorLayers.  `optimizer_delta(weight, delta)` is the
    TensorOptimizerApi.delta override point (TW:155-161, Weights.scala:46-51)."""

    def __init__(self, cumulative: bool = False, optimizer_delta=None):
        self.cumulative = cumulative
        self.optimizer_delta = optimizer_delta or (lambda weight, delta: delta)
        self.handled_exceptions: List[BaseException] = []

    # -- Do helpers ------------------------------------------------------------------------
    @staticmethod
    def do_now(value) -> Do:
        return Do.now(value)

    @staticmethod
    def do(acquire) -> Do:
        return Do(acquire)

    # -- lifting ---------------------------------------------------------------------------
    def lift(self, x) -> Differentiable:
        if isinstance(x, Differentiable):
            return x
        return Literal(self, T.as_tensor(x))

    def layer(self, raw_forward: Do) -> "TensorLayer":
        """TensorLayer.apply (TL:98-108)."""
        return TensorLayer(self, raw_forward)

    def weight(self, initial: T.Tensor) -> "TensorWeight":
        """TensorWeight.allocate (TW:139-143)."""
        return TensorWeight(self, initial.do_cache())

    # -- operators ---------------------------------------------------------------------------
    def plus(self, a, b) -> "TensorLayer":
        """`Tensor+Tensor` TL:291-314."""
        a, b = self.lift(a), self.lift(b)

        def build(tapes):
            t0, t1 = tapes
            s0, s1 = t0.data.shape, t1.data.shape
            out = t0.data + t1.data

            def backward(do_delta: Do):
                t0.backward(do_delta.map(lambda d: broadcast_then_sum(d, s0)))
                t1.backward(do_delta.map(lambda d: broadcast_then_sum(d, s1)))
            return Tape(out, backward)
        return self.layer(sequence([a.forward(), b.forward()]).map(build))

    def times(self, a, b) -> "TensorLayer":
        """`Tensor*Tensor` TL:316-340."""
        a, b = self.lift(a), self.lift(b)

        def build(tapes):
            t0, t1 = tapes
            s0, s1 = t0.data.shape, t1.data.shape
            d0, d1 = t0.data, t1.data
            out = d0 * d1

            def backward(do_delta: Do):
                t0.backward(do_delta.map(lambda d: broadcast_then_sum(d * d1, s0)))
                t1.backward(do_delta.map(lambda d: broadcast_then_sum(d * d0, s1)))
            return Tape(out, backward)
        return self.layer(sequence([a.forward(), b.forward()]).map(build))

    def translate(self, a, offset) -> "TensorLayer":
        """TL:225-238."""
        a = self.lift(a)
        offset = [float(o) for o in offset]

        def build(tape):
            shape0 = tape.data.shape
            out = tape.data.translate(offset)

            def backward(do_delta: Do):
                tape.backward(do_delta.map(lambda d: d.broadcast(shape0).translate([-o for o in offset])))
            return Tape(out, backward)
        return self.layer(a.forward().map(build))

    def join(self, operands, dimension=None) -> "TensorLayer":
        """TL:164-215."""
        operands = [self.lift(o) for o in operands]

        def build(tapes):
            out = T.join([t.data for t in tapes], dimension)
            out_shape = out.shape
            dim = len(out_shape) - 1 if dimension is None else dimension

            def backward(do_delta: Do):
                for i, tape in enumerate(tapes):
                    tape.backward(do_delta.map(lambda d, i=i: d.broadcast(out_shape).split(dim)[i]))
            return Tape(out, backward)
        return self.layer(sequence([o.forward() for o in operands]).map(build))

    def split(self, a, dimension) -> Do:
        """TL:240-287: Do[Seq[TensorLayer]]; slice deltas are collected and re-joined on release."""
        a = self.lift(a)

        def bind(tape):
            def acq():
                slices = tape.data.split(dimension)
                do_zero = Do.now(T.scalar(0.0))
                delta_slices: List[Do] = [do_zero] * len(slices)

                def release():
                    dos = list(delta_slices)
                    for i in range(len(delta_slices)):
                        delta_slices[i] = do_zero
                    tape.backward(sequence(dos).map(lambda ts: T.join(ts, dimension)))

                def make(i):
                    slice_shape = slices[i].shape

                    def slice_backward(do_delta: Do):
                        prev = delta_slices[i]
                        if prev is None:                        # TL:272-273 (never taken: slots start at doZero)
                            delta_slices[i] = do_delta.map(lambda d: d.broadcast(slice_shape))
                        else:
                            delta_slices[i] = sequence([prev, do_delta]).map(lambda ab: ab[0] + ab[1])
                    return self.layer(Do.now(Tape(slices[i], slice_backward)))
                return [make(i) for i in range(len(slices))], release
            return Do(acq).shared()
        return a.forward().flat_map(bind)


class Literal(Differentiable):
    """TensorLiterals (TLit:13-27): no-op backward."""

    def __init__(self, plugins, tensor):
        self.plugins, self.tensor = plugins, tensor

    def forward(self) -> Do:
        return Do.now(Tape(self.tensor, lambda do_delta: None))


class TensorLayer(Differentiable):
    """TensorLayerApi (TL:57-82) [+ CumulativeTensorLayers.TensorLayerApi CTL:30-83]."""

    def __init__(self, plugins: Plugins, raw_forward: Do):
        self.plugins, self.raw_forward = plugins, raw_forward
        self._shared_accumulator = None

    def forward(self) -> Do:
        if not self.plugins.cumulative:
            return self.raw_forward                                  # TL:68
        if self._shared_accumulator is None:                         # CTL:55-63
            def bind(tape):
                def acq():
                    zero = Do.now(T.scalar(0.0))
                    acc = {"delta": zero, "zero": zero}

                    def release():                                   # CTL:39-52
                        delta, acc["delta"] = acc["delta"], None
                        if delta is not acc["zero"]:
                            tape.backward(delta)
                    return (tape, acc), release
                return Do(acq)
            self._shared_accumulator = self.raw_forward.flat_map(bind).shared()

        def wrap(pair):
            tape, acc = pair

            def cumulative_backward(do_delta: Do):                   # CTL:67-79
                cur = acc["delta"]
                if cur is None:
                    raise RuntimeError("Cannot perform Tape.backward after the Tape is released")
                # CTL:72 `case `zero`` compares a Do with a Tensor and never matches => always `cur + delta`
                acc["delta"] = sequence([cur, do_delta]).map(lambda ab: ab[0] + ab[1])
            return Tape(tape.data, cumulative_backward)
        return self._shared_accumulator.map(wrap)

    def train(self) -> Do:
        """TL:70-76: forward, backward(scalar 1), yield forward data."""
        def bind(tape):
            tape.backward(Do.now(T.scalar(1.0)))
            return Do.now(tape.data)
        return self.forward().flat_map(bind)

    def predict(self) -> Do:
        return self.forward().map(lambda tape: tape.data)


class TensorWeight(Differentiable):
    """TW:27-74, :139-153."""

    def __init__(self, plugins: Plugins, data: T.CachedTensor):
        self.plugins, self.data = plugins, data

    def forward(self) -> Do:
        def acq():
            data = self.data.do_cache()                              # TW:31

            def backward(do_delta: Do):
                try:
                    self._backward(do_delta)
                except Exception as e:                               # TW:36-39 -> handleException
                    self.plugins.handled_exceptions.append(e)
            return Tape(data, backward), (lambda: None)
        return Do(acq)

    def _backward(self, original_delta: Do):
        """TW:50-74: delta = optimizer.delta; data = (data - delta).doCache."""
        delta = original_delta.map(lambda d: self.plugins.optimizer_delta(self, d)).run()
        self.data = (self.data - delta).do_cache()
