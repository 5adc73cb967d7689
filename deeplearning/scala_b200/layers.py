"""First-class differentiable tensor operations on top of `TensorLayer(forward)` (TL:98-108).

The reference ships only `+`, `*`, `translate`, `split` and `join` for tensors (TL:164-340); every real model adds its
own operations through the `TensorLayer.apply` extension point, writing the forward tensor expression and the backward
`Do[Tensor] => Unit` by hand, exactly like plugins-FloatLayers does for scalars (FloatLayers.scala:58-274: `-`, `/`,
`exp`, `log`, `abs`, `min`, `max`, ...).  This module is that set of custom layers for the new backend (SURVEY.md
section 8(f).2): the elementwise ones mirror FloatLayers case by case, and the contractions (`matmul`, `linear`,
`conv2d_3x3`) emit `Tensor.matmul` -- the tcgen05 GEMM -- directly instead of the host-unrolled
`split(j).reduce(_ + _)` chains of TensorsSpec:337-393 / :144-210.

Conventions (all inherited from the reference):
  * `train` seeds the backward pass with the scalar 1.0f (TL:55, :72): the implicit loss is the SUM of the outputs;
  * broadcasting is left-aligned (Tensors:208-222), so bias vectors live on the leading dimension (`bias_mode="row"`,
    feature-major activations `[features, batch]`) or are reshaped to `[1, n]` (`bias_mode="col"`);
  * a layer's forward value is pinned with `doCache` while its tape is alive and released with the `Do`.
"""
from __future__ import annotations

from typing import Optional, Sequence

from . import tensors as T
from .do import Do, sequence
from .plugins import Plugins, Tape, TensorLayer, TensorWeight, broadcast_then_sum

_BIG = 3.4028234663852886e38


class Layers:
    def __init__(self, plugins: Plugins):
        self.P = plugins

    # -- helpers ---------------------------------------------------------------------------------------------------
    def _pinned(self, inputs: Sequence, build, view=None) -> TensorLayer:
        """`build(tapes) -> (data, backward[, owned])`; `data` is materialised and pinned (doCache) until the Do is
        released.  `view(cached, tapes)` optionally re-shapes the pinned buffer into the tape's data (same buffer).
        `owned` lists further pinned buffers (im2col matrix, dz, ...) that are released with the Do, so device memory is
        returned to the allocator at a deterministic point instead of whenever the host collector finds the closures
        (with plain TensorLayers every backward has run by then; CumulativeTensorLayers flush later, so there the
        buffers are left to the collector)."""
        P = self.P
        inputs = [P.lift(i) for i in inputs]

        def bind(tapes):
            def acquire():
                built = build(tapes)
                data, backward = built[0], built[1]
                owned = built[2] if len(built) > 2 else []
                cached = data.do_cache()
                tape = Tape(cached if view is None else view(cached, tapes), backward(cached))

                def release():
                    # CumulativeTensorLayers flush their accumulators on release, innermost first: a weight's lazy delta
                    # (`d * 0.5 / y`) is materialised after this layer's Do has closed, so there the pinned value and the
                    # owned buffers are left to the host collector instead of being freed at this point
                    if not P.cumulative:
                        cached.release()
                        while owned:
                            owned.pop().release()
                return tape, release
            return Do(acquire)
        return P.layer(sequence([i.forward() for i in inputs]).flat_map(bind))

    def _lazy(self, inputs: Sequence, build) -> TensorLayer:
        """Same, but the forward value stays a lazy expression (fused into whatever consumes it)."""
        P = self.P
        inputs = [P.lift(i) for i in inputs]

        def make(tapes):
            data, backward = build(tapes)
            return Tape(data, backward(data))
        return P.layer(sequence([i.forward() for i in inputs]).map(make))

    # -- elementwise (mirrors FloatLayers.scala:58-274 on tensors) -----------------------------------------------------
    def minus(self, a, b) -> TensorLayer:
        def build(tapes):
            t0, t1 = tapes
            s0, s1 = list(t0.data.shape), list(t1.data.shape)

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: broadcast_then_sum(d, s0)))
                    t1.backward(do_delta.map(lambda d: broadcast_then_sum(-d, s1)))
                return run
            return t0.data - t1.data, backward
        return self._lazy([a, b], build)

    def negate(self, a) -> TensorLayer:
        return self._unary(a, lambda x: -x, lambda d, x, y: -d)

    def tanh(self, a) -> TensorLayer:
        return self._unary(a, T.tanh, lambda d, x, y: d * (1.0 - y * y), pin=True)

    def exp(self, a) -> TensorLayer:
        return self._unary(a, T.exp, lambda d, x, y: d * y, pin=True)

    def log(self, a) -> TensorLayer:
        return self._unary(a, T.log, lambda d, x, y: d / x)

    def relu(self, a) -> TensorLayer:
        """`Tensor.max(x, fill 0)`; the derivative 1[y > 0] is `min(y * FLT_MAX, 1)` (the algebra has no comparison)."""
        return self._unary(a, lambda x: T.maximum(x, x.backend.scalar(0.0)), lambda d, x, y: d * relu_mask(y), pin=True)

    def abs(self, a) -> TensorLayer:
        """FloatLayers.scala:239-256 per element: negative inputs negate value and delta, everything else (0, +0, NaN)
        passes through -- the factor is 1 - 2 * 1[x < 0]."""
        return self._unary(a, T.abs_, lambda d, x, y: d * (1.0 - positive_mask(-x) * 2.0))

    def sqrt(self, a) -> TensorLayer:
        """FloatLayers.scala:258-272: delta * 0.5 / sqrt(x)."""
        return self._unary(a, T.sqrt, lambda d, x, y: d * 0.5 / y, pin=True)

    def divide(self, a, b) -> TensorLayer:
        """FloatLayers.scala:133-148: delta_a = delta / b, delta_b = -delta * a / (b * b), each summed back to its
        operand's shape like `+` and `*` (TL:110-162)."""
        def build(tapes):
            t0, t1 = tapes
            A, B = t0.data, t1.data
            s0, s1 = list(A.shape), list(B.shape)

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: broadcast_then_sum(d / B, s0)))
                    t1.backward(do_delta.map(lambda d: broadcast_then_sum(-d * A / (B * B), s1)))
                return run
            return A / B, backward
        return self._lazy([a, b], build)

    def min(self, a, b) -> TensorLayer:
        """FloatLayers.scala:150-165 per element: the delta goes to operand 0 where a < b, to operand 1 everywhere else
        (ties and NaN included, as `if (data0 < data1) tape0 else tape1` does)."""
        return self._select(a, b, T.minimum, lambda A, B: positive_mask(B - A))

    def max(self, a, b) -> TensorLayer:
        """FloatLayers.scala:167-182 per element: operand 0 where a > b, operand 1 everywhere else."""
        return self._select(a, b, T.maximum, lambda A, B: positive_mask(A - B))

    def _select(self, a, b, pick, first_mask) -> TensorLayer:
        def build(tapes):
            t0, t1 = tapes
            A, B = t0.data, t1.data
            s0, s1 = list(A.shape), list(B.shape)

            def backward(_):
                def run(do_delta: Do):
                    m = first_mask(A, B)
                    t0.backward(do_delta.map(lambda d: broadcast_then_sum(d * m, s0)))
                    t1.backward(do_delta.map(lambda d: broadcast_then_sum(d * (1.0 - m), s1)))
                return run
            return pick(A, B), backward
        return self._lazy([a, b], build)

    def sum(self, a) -> TensorLayer:
        """`Tensor.sum` (Tensors:673-766) as a layer: one scalar out, and every element receives the scalar delta
        (d out / d x_i = 1), i.e. the delta broadcast to the operand's shape."""
        def build(tapes):
            (t0,) = tapes
            shape0 = list(t0.data.shape)

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: d.broadcast(shape0)))
                return run
            return t0.data.sum(), backward
        return self._pinned([a], build)

    def permute(self, a, dimensions) -> TensorLayer:
        """`Tensor.permute` (Tensors:1008-1024) as a layer; the delta travels through the inverse permutation."""
        dimensions = [int(d) for d in dimensions]
        inverse = [dimensions.index(i) for i in range(len(dimensions))]

        def build(tapes):
            (t0,) = tapes
            out_shape = [t0.data.shape[d] for d in dimensions]

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: d.broadcast(out_shape).permute(inverse)))
                return run
            return t0.data.permute(dimensions), backward
        return self._lazy([a], build)

    def broadcast(self, a, new_shape) -> TensorLayer:
        """`Tensor.broadcast` (Tensors:893-945, left-aligned) as a layer; the delta is summed back over the broadcast
        dimensions (`sumAs`, TL:124-162)."""
        new_shape = [int(s) for s in new_shape]

        def build(tapes):
            (t0,) = tapes
            shape0 = list(t0.data.shape)

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: broadcast_then_sum(d.broadcast(new_shape), shape0)))
                return run
            return t0.data.broadcast(new_shape), backward
        return self._lazy([a], build)

    def _unary(self, a, forward, delta, pin: bool = False) -> TensorLayer:
        def build(tapes):
            (t0,) = tapes
            x = t0.data

            def backward(y):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: delta(d, x, y)))
                return run
            return forward(x), backward
        return (self._pinned if pin else self._lazy)([a], build)

    def squared_error(self, out, target) -> TensorLayer:
        """(out - target)^2 elementwise; under `train` the loss is its sum.  `target` gets no gradient."""
        def build(tapes):
            t0, t1 = tapes
            diff = t0.data - t1.data

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: d * (diff * 2.0)))
                return run
            return diff * diff, backward
        return self._lazy([out, target], build)

    # -- views -----------------------------------------------------------------------------------------------------
    def reshape(self, a, new_shape) -> TensorLayer:
        new_shape = [int(s) for s in new_shape]

        def build(tapes):
            (t0,) = tapes
            shape0 = list(t0.data.shape)

            def back(d):
                r = d.broadcast(new_shape).reshape(shape0)
                spec = getattr(d, "_ungated", None)
                if spec is not None and list(d.shape) == new_shape:   # lets the layer below fuse its activation derivative
                    r._ungated_flat = dict(spec=spec, source=d)      # into the contraction that produces this delta
                return r

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(back))
                return run
            return t0.data.reshape(new_shape), backward
        return self._lazy([a], build)

    def downsample(self, a, factor: int = 2) -> TensorLayer:
        """Nearest-neighbour 1/factor resampling of the two middle dims of an NHWC tensor by `Tensor.scale`
        (Tensors:950-965): out[b, y, x, c] = in[b, factor*y, factor*x, c].  Backward routes each delta to the sample it
        was read from (a `scale` back up, masked to the sampled positions)."""
        spec = getattr(a, "_conv_spec", None)
        if spec is not None:                                         # conv -> down-sample: one layer, sampled delta (conv2d_3x3)
            return self.conv2d_3x3(spec[0], spec[1], spec[2], spec[3], downsample=factor)

        def build(tapes):
            (t0,) = tapes
            b, h, w, c = t0.data.shape
            out_shape = [b, h // factor, w // factor, c]
            mask = sample_mask(t0.data.backend, h, w, factor)

            def backward(_):
                def run(do_delta: Do):
                    t0.backward(do_delta.map(lambda d: d.broadcast(out_shape).scale([b, h, w, c]) * mask))
                return run
            return t0.data.scale(out_shape), backward
        return self._lazy([a], build)

    # -- contractions ------------------------------------------------------------------------------------------------
    def matmul(self, a, b, transpose_a: bool = False, transpose_b: bool = False) -> TensorLayer:
        return self.linear(a, b, transpose_a=transpose_a, transpose_b=transpose_b)

    def linear(self, a, b, bias=None, bias_mode: str = "none", activation: str = "none", transpose_a: bool = False,
               transpose_b: bool = False) -> TensorLayer:
        """y = act(op(a) op(b) + bias) as ONE tcgen05 GEMM with a fused epilogue; backward is two more GEMMs
        (delta_a = dz op(b)^T, delta_b = op(a)^T dz) and one row / column reduction for the bias.
        Feature-major dense layer: linear(W, x, bias, "row", "tanh"); batch-major: linear(x, W, bias, "col", ...)."""
        if activation not in ("none", "tanh", "relu"):
            raise ValueError("linear: unknown activation %r" % (activation,))
        inputs = [a, b] + ([bias] if bias_mode != "none" else [])

        def build(tapes):
            ta, tb = tapes[0], tapes[1]
            A, B = ta.data, tb.data
            bias_t = tapes[2].data if bias_mode != "none" else None
            # the value feeds the next contraction (and this layer's weight gradient): its operand copy comes with it
            y = T.matmul(A, B, transpose_a=transpose_a, transpose_b=transpose_b, bias=bias_t, bias_mode=bias_mode,
                         activation=activation, emit_operand_copy=True)
            out_shape = list(y.shape)
            owned = []

            def backward(y_cached):
                def run(do_delta: Do):
                    memo = {}

                    def dz_of(d):                                   # shared by the three consumers below
                        if "dz" not in memo:
                            spec = getattr(d, "_ungated", None)
                            if (activation != "none" and spec is not None and list(d.shape) == out_shape
                                    and (d._weak is None or d._weak() is None)):
                                # `d` is the still unevaluated product of the layer above (its delta_a): issue it with
                                # this layer's activation derivative in its epilogue -- `d * (1 - y * y)` /
                                # `d * 1[y > 0]` node for node -- instead of materialising d and launching the tree
                                d = T.matmul(spec["a"], spec["b"], transpose_a=spec["transpose_a"], transpose_b=spec["transpose_b"],
                                             mode=spec["mode"], gate=y_cached, gate_mode=activation, emit_operand_copy=True)
                            else:
                                d = d.broadcast(out_shape)
                                if activation == "tanh":
                                    d = d * (1.0 - y_cached * y_cached)
                                elif activation == "relu":
                                    d = d * relu_mask(y_cached)
                            memo["dz"] = d.do_cache()
                            owned.append(memo["dz"])
                        return memo["dz"]

                    def delta_a(d):
                        dz = dz_of(d)
                        if transpose_a:                             # a is stored [K, M]: delta = op(b) dz^T
                            return T.matmul(B, dz, transpose_a=transpose_b, transpose_b=True)
                        return T.matmul(dz, B, transpose_b=not transpose_b)

                    def delta_b(d):
                        dz = dz_of(d)
                        if transpose_b:                             # b is stored [N, K]: delta = dz^T op(a)
                            return T.matmul(dz, A, transpose_a=True, transpose_b=transpose_a)
                        return T.matmul(A, dz, transpose_a=not transpose_a)

                    # weights first: their gradients are complete the moment dz exists, and under data parallelism the
                    # large ones leave for the all-reduce right away and travel under the rest of the backward pass --
                    # descending into the layer below first would put every weight gradient at the END of the step
                    if bias_mode != "none":
                        axis = 1 if bias_mode == "row" else 0
                        tapes[2].backward(do_delta.map(lambda d: T.sum_axis(dz_of(d), axis)))
                    order = [(ta, delta_a, a), (tb, delta_b, b)]
                    # ... unless the optimizer applies an update the moment it receives the gradient (single GPU: the
                    # weight-gradient contraction writes `w - lr * dw` in place): then the pass that still reads the old
                    # weight -- this layer's backward-data -- has to be issued first
                    last = getattr(self.P.optimizer, "applies_in_backward", False)
                    order.sort(key=lambda e: (1 if last else 0) if isinstance(e[2], TensorWeight) else (0 if last else 1))
                    for tape, delta, _ in order:
                        tape.backward(do_delta.map(delta))
                return run
            return y, backward, owned
        return self._pinned(inputs, build)

    def conv2d_3x3(self, x, w, bias=None, activation: str = "none", downsample: int = 1) -> TensorLayer:
        """'Same' 3x3 convolution of an NHWC tensor with weights [3, 3, Cin, Cout] (the layout and the index convention
        of TensorsSpec:144-210: out[b,y,x,f] = bias[f] + sum x[b, y-(ky-1), x-(kx-1), c] * w[ky,kx,c,f], zero padding from
        the bounds check of `translate`).
        In the tensor-core modes each of the three passes has a kernel of its own (csrc/conv.cu: implicit GEMM whose
        operand tiles are tap-shifted TMA boxes of the activation, direct kernels for a first layer with Cin <= 4), and a
        preceding `downsample` view is read in place.  Where no such kernel applies (fp32 parity mode, odd shapes) the
        pass is composed the way the reference composes it: the nine translated views written by ONE multi-root gather
        launch (`Tensor.join`), then out = col[BHW, Cin*9] . wmat[Cin*9, Cout].
        `downsample` = s > 1 is the layer followed by `Layers.downsample(., s)` (which calls this): the value is the same
        `scale` view of the pinned full-size output, and the delta that comes back -- zero except at the sampled
        positions -- stays at its sampled size [B, H/s, W/s, Cout] all the way into the backward kernels
        (`dz_stride`), instead of being scaled back up, masked and written out at full size first."""
        offsets = [(ky - 1, kx - 1) for ky in range(3) for kx in range(3)]
        inputs = [x, w] + ([bias] if bias is not None else [])
        s = int(downsample)

        def build(tapes):
            tx, tw = tapes[0], tapes[1]
            X, W = tx.data, tw.data
            be = X.backend
            bsz, h, wd, cin = X.shape
            cout = W.shape[3]
            m = bsz * h * wd
            if s > 1 and (h % s or wd % s):
                raise ValueError("conv2d_3x3: downsample %d does not divide the %d x %d image" % (s, h, wd))
            full, half = [bsz, h, wd, cout], [bsz, h // s, wd // s, cout]
            bias_t = tapes[2].data if bias is not None else None
            owned = []
            lazy = {}

            def col():                                               # the im2col matrix, only where a composed pass needs it
                if "col" not in lazy:
                    lazy["col"] = T.join([X.translate([0, oy, ox, 0]) for (oy, ox) in offsets]).reshape([m, cin * 9]).do_cache()
                    owned.append(lazy["col"])
                return lazy["col"]

            def wmat():
                if "wmat" not in lazy:
                    lazy["wmat"] = W.permute([2, 0, 1, 3]).reshape([cin * 9, cout]).do_cache()
                    owned.append(lazy["wmat"])
                return lazy["wmat"]

            if be.conv3x3_supported("forward", h, wd, cin, cout):
                y = be.conv3x3_forward(X, W, bias_t, activation, out_stride=s)
            else:
                y = T.matmul(col(), wmat(), bias=bias_t, bias_mode="col" if bias is not None else "none",
                             activation=activation).reshape(full)

            def backward(y_cached):
                def run(do_delta: Do):
                    memo = {}

                    def dz_sampled(d):                               # [B, H/s, W/s, Cout]: the delta where it can be non-zero
                        if "dzs" not in memo:
                            fused = None
                            if activation != "none" and list(d.shape) == half:
                                fused = fuse_activation_derivative(d, y_cached, activation, s)
                            if fused is not None:
                                # `d` was the still unevaluated backward-data pass / delta contraction of the layer above:
                                # it is issued with `d * act'(y)` -- y read at the sampled positions -- in its epilogue
                                d = fused
                            else:
                                d = d.broadcast(half)
                                ys = y_cached.scale(half) if s > 1 else y_cached
                                if activation == "tanh":
                                    d = d * (1.0 - ys * ys)
                                elif activation == "relu":
                                    d = d * relu_mask(ys)
                            memo["dzs"] = d.do_cache()
                            owned.append(memo["dzs"])
                        return memo["dzs"]

                    def dz_of(d):                                    # [B, H, W, Cout], for the passes composed from views
                        if s == 1:
                            return dz_sampled(d)
                        if "dz" not in memo:
                            memo["dz"] = (dz_sampled(d).scale(full) * sample_mask(be, h, wd, s)).do_cache()
                            owned.append(memo["dz"])
                        return memo["dz"]

                    def delta_x(d):
                        if be.conv3x3_supported("backward_data", h, wd, cin, cout):
                            return be.conv3x3_backward_data(dz_sampled(d), W, dz_stride=s)
                        dcol = T.matmul(dz_of(d).reshape([m, cout]), wmat(), transpose_b=True).reshape([bsz, h, wd, cin, 9])
                        parts = [p.translate([0, -oy, -ox, 0]) for p, (oy, ox) in zip(dcol.split(4), offsets)]
                        acc = parts[0]
                        for p in parts[1:]:
                            acc = acc + p
                        return acc

                    # first layer (Cin <= 4): one kernel pass over the delta gives the weight AND the bias gradient
                    both_ok = bias is not None and cin <= 4 and cout <= 32 and wd <= 512 and \
                        be.conv3x3_supported("backward_weights", h, wd, cin, cout)

                    def dw_db(d):
                        if "dw_db" not in memo:
                            memo["dw_db"] = be.conv3x3_backward_weights_bias(X, dz_sampled(d), dz_stride=s)
                        return memo["dw_db"]

                    def delta_w(d):
                        if both_ok:
                            return dw_db(d)[0]
                        if be.conv3x3_supported("backward_weights", h, wd, cin, cout):
                            return be.conv3x3_backward_weights(X, dz_sampled(d), dz_stride=s)
                        dw = T.matmul(col(), dz_of(d).reshape([m, cout]), transpose_a=True)     # [Cin*9, Cout]
                        return dw.reshape([cin, 3, 3, cout]).permute([1, 2, 0, 3])

                    # weights before the input, for the same reason as in `linear`
                    if bias is not None:                             # the zeros of a sampled delta add nothing
                        tapes[2].backward(do_delta.map(lambda d: dw_db(d)[1] if both_ok else
                                                       T.sum_axis(dz_sampled(d).reshape([m // (s * s), cout]), 0)))
                    tw.backward(do_delta.map(delta_w))
                    tx.backward(do_delta.map(delta_x))
                return run
            return y, backward, owned

        view = (lambda cached, tapes: cached.scale([cached.shape[0], cached.shape[1] // s, cached.shape[2] // s, cached.shape[3]])) \
            if s > 1 else None
        layer = self._pinned(inputs, build, view)
        if s == 1:
            layer._conv_spec = (x, w, bias, activation)              # lets a following `downsample` fold itself into this layer
        return layer


def _unevaluated(t) -> bool:
    return getattr(t, "_weak", None) is None or t._weak() is None


def fuse_activation_derivative(d, y, activation: str, y_stride: int):
    """`d * act'(y)` for a delta `d` that is still an unevaluated product of the layer above, issued as ONE kernel with the
    derivative in that product's epilogue; None when `d` is anything else.  `y` is the pinned full-size output of the layer
    this delta belongs to, read at stride `y_stride` (the `scale` view between the two layers)."""
    be = y.backend
    conv = getattr(d, "_ungated_conv", None)
    if conv is not None and _unevaluated(d):
        return be.conv3x3_backward_data(conv["dz"], conv["w"], mode=conv["mode"], dz_stride=conv["dz_stride"], gate=y,
                                        gate_mode=activation, gate_stride=y_stride)
    flat = getattr(d, "_ungated_flat", None)                         # a contraction seen through a `reshape`
    if flat is not None and y_stride == 1 and _unevaluated(flat["source"]) and _unevaluated(d):
        spec = flat["spec"]
        return T.matmul(spec["a"], spec["b"], transpose_a=spec["transpose_a"], transpose_b=spec["transpose_b"], mode=spec["mode"],
                        gate=y, gate_mode=activation, emit_operand_copy=True).reshape(list(d.shape))
    return None


def relu_mask(y):
    """1[y > 0] for y = max(x, 0) >= 0, spelled with the operators the algebra has (Tensors:546-558)."""
    return T.minimum(y * _BIG, y.backend.scalar(1.0))


def positive_mask(t):
    """1[t > 0] for any finite or infinite t (0 for t <= 0, -0 and NaN-free inputs): clamp at 0, then scale twice so even
    the smallest subnormal (1.4e-45 * FLT_MAX = 4.8e-7, * FLT_MAX again = 1.6e32) saturates at 1."""
    zero, one = t.backend.scalar(0.0), t.backend.scalar(1.0)
    return T.minimum(T.minimum(T.maximum(t, zero) * _BIG, one) * _BIG, one)


_MASKS = {}


def sample_mask(backend, h: int, w: int, factor: int):
    """[1, h, w, 1] tensor that is 1 where both coordinates are multiples of `factor` (broadcasts over batch and channel)."""
    import numpy as np
    key = (id(backend), h, w, factor)
    if key not in _MASKS:
        m = np.zeros((1, h, w, 1), dtype=np.float32)
        m[0, ::factor, ::factor, 0] = 1.0
        _MASKS[key] = backend.tensor(m).do_cache()
    return _MASKS[key]
