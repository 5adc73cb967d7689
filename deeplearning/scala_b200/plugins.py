"""Host mirror of the DeepLearning.scala tensor autodiff plugins on top of `deeplearning.scala_b200.tensors`.

  TL   = plugins-TensorLayers/src/main/scala/com/thoughtworks/deeplearning/plugins/TensorLayers.scala
  TW   = plugins-TensorWeights/src/main/scala/com/thoughtworks/deeplearning/plugins/TensorWeights.scala
  CTL  = plugins-CumulativeTensorLayers/src/main/scala/com/thoughtworks/deeplearning/plugins/CumulativeTensorLayers.scala
  TLit = plugins-TensorLiterals/src/main/scala/com/thoughtworks/deeplearning/plugins/TensorLiterals.scala

`Plugins(cumulative=..., optimizer_delta=...)` is one mixin stack (`Factory[TensorLayers with TensorWeights with
[Cumulative]TensorLayers ...].newInstance()`), exposing the same operations under the same names: `+`, `*`, `translate`,
`split`, `join`, `TensorLayer(forward)`, `train`, `predict`, `TensorWeight.allocate` (here `weight`).  All tensor work
stays lazy; the only device launches are the ones the reference performs -- `doCache` of each weight's
`data - delta` tree (TW:64), i.e. forward recompute + backward + SGD step fused into a single kernel per weight.
"""
from __future__ import annotations

from typing import Callable, List, Optional

from . import tensors as T
from .do import Do, sequence


class Tape:
    """DeepLearning.scala:18 -- Tape(data, backward)."""
    __slots__ = ("data", "backward")

    def __init__(self, data, backward):
        self.data, self.backward = data, backward


def sum_as(tensor, shape):
    """TL:124-162: reduce `tensor` to `shape` by summing away trailing dims with host-unrolled `split(j).reduce(_ + _)`
    (a left fold).  The `shape(i) == 1` branch reproduces the reference's defect (its permute is discarded, TL:139-148)."""
    shape = list(shape)
    if len(shape) > len(tensor.shape):
        raise ValueError("Cannot sum %r to %r" % (tensor.shape, shape))

    def fold(parts):
        acc = parts[0]
        for p in parts[1:]:
            acc = acc + p
        return acc

    def loop(t, i, j):
        if j >= len(t.shape):
            return t
        if i >= len(shape):
            return fold([loop(s, i + 1, j) for s in t.split(j)])
        if t.shape[i] != 1 and shape[i] == 1:
            return fold([loop(s, i + 1, j + 1) for s in t.split(j)])
        if t.shape[j] == shape[i]:
            return loop(t, i + 1, j + 1)
        raise ValueError("Cannot sum %r to %r" % (tensor.shape, shape))

    out = loop(tensor, 0, 0)
    if list(out.shape) != shape:
        raise AssertionError("sumAs produced %r, expected %r (TL:159-161)" % (out.shape, shape))
    return out


def broadcast_then_sum(tensor, shape):
    """TL:110-121: pad the delta's shape on the right up to the target rank, then sumAs."""
    shape = list(shape)
    have = list(tensor.shape)
    if len(have) < len(shape):
        tensor = tensor.broadcast(have + shape[len(have):])
    return sum_as(tensor, shape)


def _may_read_weights(data, pending) -> bool:
    """Does materialising the lazy tensor `data` later read a buffer that an in-place update in `pending` is about to
    overwrite?  Leaves that are pinned buffers are compared by identity; a leaf whose buffer does not exist yet is of
    unknown provenance and counts as a reader."""
    if not pending or isinstance(data, T.CachedTensor):
        return False
    from .trees import serialize
    live = {id(w.data._buf) for w, _ in pending if isinstance(w.data, T.CachedTensor)}
    _, params = serialize([data.closure()], data.shape)
    for p in params:
        if isinstance(p, T.CachedTensor):
            if id(p._buf) in live:
                return True
        elif not (isinstance(p, T.UploadTensor) and p._buf is not None):
            return True
    return False


class Differentiable:
    plugins: "Plugins"

    def forward(self) -> Do:
        raise NotImplementedError

    def __add__(self, other): return self.plugins.plus(self, other)
    def __radd__(self, other): return self.plugins.plus(other, self)
    def __mul__(self, other): return self.plugins.times(self, other)
    def __rmul__(self, other): return self.plugins.times(other, self)


class TensorLiteral(Differentiable):
    """TLit:13-27: a raw Tensor in the graph; backward is a no-op."""

    def __init__(self, plugins, tensor):
        self.plugins, self.tensor = plugins, tensor

    def forward(self) -> Do:
        return Do.now(Tape(self.tensor, lambda do_delta: None))


class TensorLayer(Differentiable):
    """TensorLayerApi (TL:57-82); with `plugins.cumulative` also CumulativeTensorLayers.TensorLayerApi (CTL:30-83)."""

    def __init__(self, plugins: "Plugins", raw_forward: Do):
        self.plugins, self.raw_forward = plugins, raw_forward
        self._accumulated: Optional[Do] = None

    def forward(self) -> Do:
        if not self.plugins.cumulative:
            return self.raw_forward                                          # TL:66-68
        if self._accumulated is None:                                        # CTL:55-63
            self._accumulated = self.raw_forward.flat_map(self._accumulator).shared()
        return self._accumulated.map(self._cumulative_tape)

    def _accumulator(self, tape: Tape) -> Do:
        def acquire():
            zero = Do.now(self.plugins.scalar(0.0))
            state = {"delta": zero, "zero": zero}

            def flush():                                                     # Accumulator.release, CTL:39-52
                delta, state["delta"] = state["delta"], None
                if delta is not state["zero"]:
                    tape.backward(delta)
            return (tape, state), flush
        return Do(acquire)

    @staticmethod
    def _cumulative_tape(pair) -> Tape:
        tape, state = pair

        def backward(do_delta: Do):                                          # CTL:67-79
            current = state["delta"]
            if current is None:
                raise RuntimeError("Cannot perform Tape.backward after the Tape is released")
            # CTL:72 compares a Do with a Tensor, so the `zero` case never matches: every delta is `0.0f + delta`
            state["delta"] = sequence([current, do_delta]).map(lambda pair: pair[0] + pair[1])
        return Tape(tape.data, backward)

    def train(self) -> Do:
        """TL:70-76: forward, backward with the scalar delta 1.0f (TL:55), yield the forward data."""
        def step(tape):
            tape.backward(Do.now(self.plugins.scalar(1.0)))
            data = tape.data
            if self.plugins.optimizer is not None and _may_read_weights(data, self.plugins.deferred_updates):
                # an in-place rule is about to overwrite the weights a lazy forward expression still reads: pin the
                # forward value first, so `train` yields the data of THIS step like the reference (TL:70-76)
                data = data.do_cache()
            if not self.plugins.cumulative and self.plugins.flush_in_train:   # cumulative layers flush on release: the
                self.plugins.flush_updates()                                 # caller flushes after running the Do
            return Do.now(data)

        def begin():
            self.plugins.begin_step()
            return None, (lambda: None)
        return Do(begin).flat_map(lambda _: self.forward()).flat_map(step)

    def predict(self) -> Do:
        """TL:78-80."""
        return self.forward().map(lambda tape: tape.data)


# returned by an optimizer hook that has already applied the update (harness.ExchangeSgd: the weight-gradient contraction
# wrote `data - delta * learningRate` in its epilogue), so there is nothing left to swap or defer
APPLIED = object()


class TensorWeight(Differentiable):
    """TensorWeightApi (TW:27-74) + allocate (TW:139-153)."""

    def __init__(self, plugins: "Plugins", data):
        self.plugins, self.data = plugins, data
        self.uses, self.uses_step = 0, -1

    def forward(self) -> Do:
        def acquire():
            # TW:31 `data.doCache`: the tape holds its OWN reference to the weight buffer (clRetainMemObject in the
            # reference), so a backward pass that swaps the weight does not pull the buffer from under other consumers.
            # It is dropped with the tape object rather than at the Do's release: the reference's own specs read the
            # result of `train` after running it (CumulativeTensorLayersSpec:96-127).
            current = self.data
            data = T.CachedTensor(current.backend, current._buffer(), list(current.shape), current.padding)
            step = self.plugins.step_id                                      # how many tapes of this step read the weight
            self.uses = self.uses + 1 if self.uses_step == step else 1
            self.uses_step = step

            def backward(do_delta: Do):
                try:
                    self.backward(do_delta)
                except Exception as exc:                                     # TW:36-39 -> handleException
                    self.plugins.handle_exception(exc)
            return Tape(data, backward), (lambda: None)
        return Do(acquire)

    def backward(self, original_delta: Do):
        """TW:50-74: delta = optimizer.delta; newData = (data - delta).doCache; release the old buffer; swap."""
        delta = original_delta.map(lambda d: self.plugins.optimizer_delta(self, d)).run()
        if delta is APPLIED:                                                 # the optimizer hook updated the weight itself
            return
        if self.plugins.deferred_updates is not None:
            # data-parallel training: the optimizer hook has put an all-reduce of this gradient in flight on the exchange
            # stream; applying `data - delta` now would make the compute stream wait for it.  The swap is postponed to
            # `flush_updates` (end of the step), so the collective overlaps the rest of the backward pass.  Forward
            # values were computed from the tapes' own references, so the result is the same.
            self.plugins.deferred_updates.append((self, delta))
            return
        self.apply(delta)

    def apply(self, delta):
        old = self.data
        self.data = (old - delta).do_cache()
        old.release()                                                        # TW:65-72; tapes keep their own reference


class Plugins:
    def __init__(self, cumulative: bool = False, optimizer_delta: Optional[Callable] = None, backend=None,
                 defer_updates: bool = False, optimizer=None):
        self.cumulative = cumulative
        self.step_id = 0
        # False: `train` leaves the deferred weight updates to the caller's `flush_updates()`, so whatever the caller maps
        # over the forward data (the loss reduction) is issued BEFORE the end-of-step exchange and update instead of behind it
        self.flush_in_train = True
        # a stateful optimizer (optimizers.Momentum / Adam) updates weights in place, which is only safe at the end of
        # the step: it implies deferred updates
        self.optimizer = optimizer
        self.deferred_updates: Optional[list] = [] if (defer_updates or optimizer is not None) else None
        self.before_flush: List[Callable] = []                                    # e.g. the gradient exchange's flush
        self.recording = False                                                    # inside Backend.capture (harness.CompiledStep)
        self.backend = backend                                                    # None = the `gpu` singleton
        self.optimizer_delta = optimizer_delta or (lambda weight, delta: delta)   # Weights.scala:46-51 default
        self.handled_exceptions: List[BaseException] = []

    def flush_updates(self):
        """Apply the weight swaps postponed by `defer_updates`, in the order their deltas arrived."""
        pending = self.deferred_updates
        if not pending:
            return
        for hook in self.before_flush:
            hook()
        if hasattr(self.optimizer, "step_many"):                             # every update of the step in one launch
            self.optimizer.step_many(list(pending))
            del pending[:]
        elif self.optimizer is not None:
            # in-place rules: every still-lazy delta is materialised BEFORE the first weight is overwritten -- a delta may
            # read another weight's live buffer (`w1 * w2`: the gradient of w2 is d * w1.data), and the reference's
            # swap (TW:63-72) leaves the tapes the pre-update values
            pending[:] = [(w, d.broadcast(w.data.shape).do_cache()) for w, d in pending]
        while pending:
            weight, delta = pending.pop(0)
            if self.optimizer is not None:
                self.optimizer.step(weight, delta)
            else:
                weight.apply(delta)

    def begin_step(self):
        """Once per training step: a stateful optimizer advances its step counter and publishes this step's
        hyper-parameters to the device.  Skipped while a step is being recorded (host writes cannot be captured; the
        replays call it themselves)."""
        self.step_id += 1
        if self.optimizer is not None and not self.recording:
            self.optimizer.begin_step(self.backend or T.default_backend())

    def handle_exception(self, exc):                                         # Differentiables.scala:10-14
        self.handled_exceptions.append(exc)

    def scalar(self, value: float):
        return (self.backend or T.default_backend()).scalar(value)

    # -- Do helpers used by hand-written layers --------------------------------------------------------------------
    @staticmethod
    def do_now(value) -> Do:
        return Do.now(value)

    @staticmethod
    def do(acquire) -> Do:
        return Do(acquire)

    # -- constructors ----------------------------------------------------------------------------------------------
    def lift(self, x) -> Differentiable:
        return x if isinstance(x, Differentiable) else TensorLiteral(self, x if isinstance(x, T.Tensor) else self.scalar(float(x)))

    def layer(self, raw_forward: Do) -> TensorLayer:
        """TensorLayer.apply (TL:98-108): the extension point for custom differentiable operations."""
        return TensorLayer(self, raw_forward)

    def weight(self, initial) -> TensorWeight:
        """TensorWeight.allocate (TW:139-143)."""
        return TensorWeight(self, initial.do_cache())

    # -- differentiable operators ----------------------------------------------------------------------------------
    def _binary(self, a, b, data_of, deltas_of) -> TensorLayer:
        a, b = self.lift(a), self.lift(b)

        def build(tapes):
            t0, t1 = tapes
            d0, d1 = t0.data, t1.data
            s0, s1 = list(d0.shape), list(d1.shape)

            def backward(do_delta: Do):
                t0.backward(do_delta.map(lambda d: broadcast_then_sum(deltas_of(d, d0, d1)[0], s0)))
                t1.backward(do_delta.map(lambda d: broadcast_then_sum(deltas_of(d, d0, d1)[1], s1)))
            return Tape(data_of(d0, d1), backward)
        return self.layer(sequence([a.forward(), b.forward()]).map(build))

    def plus(self, a, b) -> TensorLayer:
        """`Tensor+Tensor` (TL:291-314)."""
        return self._binary(a, b, lambda x, y: x + y, lambda d, x, y: (d, d))

    def times(self, a, b) -> TensorLayer:
        """`Tensor*Tensor` (TL:316-340)."""
        return self._binary(a, b, lambda x, y: x * y, lambda d, x, y: (d * y, d * x))

    def translate(self, a, offset) -> TensorLayer:
        """TL:225-238: backward translates the (broadcast) delta back by -offset."""
        a = self.lift(a)
        offset = [float(o) for o in offset]

        def build(tape):
            shape0 = list(tape.data.shape)

            def backward(do_delta: Do):
                tape.backward(do_delta.map(lambda d: d.broadcast(shape0).translate([-o for o in offset])))
            return Tape(tape.data.translate(offset), backward)
        return self.layer(a.forward().map(build))

    def join(self, operands, dimension=None) -> TensorLayer:
        """TL:164-215: operand i receives `delta.broadcast(outShape).split(dim)(i)`."""
        operands = [self.lift(o) for o in operands]

        def build(tapes):
            out = T.join([t.data for t in tapes], dimension)
            out_shape = list(out.shape)
            dim = len(out_shape) - 1 if dimension is None else dimension

            def backward(do_delta: Do):
                for i, tape in enumerate(tapes):
                    tape.backward(do_delta.map(lambda d, i=i: d.broadcast(out_shape).split(dim)[i]))
            return Tape(out, backward)
        return self.layer(sequence([o.forward() for o in operands]).map(build))

    def split(self, a, dimension) -> Do:
        """TL:240-287: Do[Seq[TensorLayer]]; slice deltas accumulate per slot and are re-joined and sent upstream once,
        when the shared Do is released."""
        a = self.lift(a)

        def bind(tape):
            def acquire():
                slices = tape.data.split(dimension)
                zero = Do.now(self.scalar(0.0))
                slots: List[Do] = [zero] * len(slices)

                def flush():
                    pending = list(slots)
                    for i in range(len(slots)):
                        slots[i] = zero
                    tape.backward(sequence(pending).map(lambda ts: T.join(ts, dimension)))

                def slice_layer(i):
                    def backward(do_delta: Do):
                        slots[i] = sequence([slots[i], do_delta]).map(lambda pair: pair[0] + pair[1])
                    return self.layer(Do.now(Tape(slices[i], backward)))
                return [slice_layer(i) for i in range(len(slices))], flush
            return Do(acquire).shared()
        return a.forward().flat_map(bind)
