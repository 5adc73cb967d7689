"""The training loops of BASELINE.json's configs on the new backend: C1 (MLP 784-500-10), C3 (CIFAR-shaped CNN),
C4 (wide MLP 4 x 4096) and C5 (C3 under data parallelism), written the way a DeepLearning.scala user writes them --
weights are `TensorWeight`s (TW:139-153), layers are `TensorLayer`s (TL:98-108), one step is `loss.train` (TL:70-76),
the learning rate is an optimizer plugin overriding `delta` (the pattern of CumulativeFloatLayersSpec:26-32).
Shapes, seeds and layouts follow SURVEY.md section 8(d).  Synthetic data only.
"""
from __future__ import annotations

import os
import sys
from typing import List, Optional, Sequence

import numpy as np

from . import tensors as T
from .layers import Layers
from .plugins import APPLIED, Plugins


class Exchange:
    """Gradient exchange of data-parallel training (new functionality): SUM all-reduce of one
    materialised gradient buffer over NCCL / NVLink, in place.  The reference's implicit loss is a sum over the batch,
    so N shards summed equal the full-batch gradient (no averaging)."""

    def __init__(self, backend: T.Backend, rank: int, world_size: int, unique_id: bytes):
        import ctypes as C
        from . import _native as N
        self.backend, self.rank, self.world_size = backend, rank, world_size
        self._N, self._C = N, C
        # the communicator is capped at DLS_EXCHANGE_CTAS (default 8) CTAs by dls_comm_init (ncclCommInitRankConfig): the
        # gradient buckets are a few MB and travel under the rest of the backward pass, which keeps the other SMs
        # NCCL reads NCCL_PROTO when it tunes a NEW communicator, so DLS_EXCHANGE_PROTO can pin the protocol of this one only
        # (the variable is restored right after).  Default: NCCL's own choice -- measured at 8 GPUs on the C5 step (ms):
        # NCCL's choice (LL) 0.268, Simple 0.269, LL128 0.289.
        proto, had = os.environ.get("DLS_EXCHANGE_PROTO", ""), os.environ.get("NCCL_PROTO")
        if had is None and proto:
            os.environ["NCCL_PROTO"] = proto
        try:
            N.check(N.lib.dls_comm_init(backend.ctx, unique_id, rank, world_size))
        finally:
            if had is None and proto:
                os.environ.pop("NCCL_PROTO", None)
        self._pending, self._pending_bytes = [], 0
        self._held = []                         # small gradients waiting for the one-launch peer sum at `flush`
        self._over_nccl = set()                 # ids of the gradient buffers this step sent through NCCL
        # gradients below `peer_bytes` skip NCCL: they are summed and applied by the one-launch end-of-step kernel over
        # NVLink peer memory (csrc/peer.cu).  Needs cudaIpc between the ranks (one node); otherwise NCCL carries everything.
        self.peer = False
        if world_size > 1 and self.peer_bytes > 0:
            rc = N.lib.dls_peer_init(backend.ctx, self.peer_staging_floats)
            if rc == 0:
                self.peer = True
            else:
                sys.stderr.write("dls exchange: peer-memory path unavailable (%s); NCCL carries every gradient\n" % N.last_error())

    def verify(self, n: int = 300001) -> None:
        """Known-answer check of the peer path (run once after construction by the bench and the tests): rank r
        contributes (i % 251) * (r + 1), every rank must read back (i % 251) * W (W + 1) / 2 -- exact in fp32 -- so a
        segment that is not mapped, or a rank that is left out of the sum, fails loudly before any training step."""
        if not self.peer:
            return
        C, N, be = self._C, self._N, self.backend
        base = (np.arange(n) % 251).astype(np.float32)
        buf = be.upload(base * np.float32(self.rank + 1))
        handles, counts, flags = (C.c_void_p * 1)(buf.handle), (C.c_size_t * 1)(n), (C.c_int * 1)(1)
        N.check(N.lib.dls_peer_sgd_step(be.ctx, None, handles, counts, flags, 1, 0.0))
        got = be.download(buf, n)
        want = base * np.float32(self.world_size * (self.world_size + 1) // 2)
        if not (got == want).all():
            raise RuntimeError("peer exchange self-test failed on rank %d: %d of %d sums are wrong"
                               % (self.rank, int((got != want).sum()), n))

    @staticmethod
    def unique_id() -> bytes:
        import ctypes as C
        from . import _native as N
        buf = C.create_string_buffer(128)
        N.check(N.lib.dls_comm_unique_id(buf))
        return buf.raw

    # gradients are collected until this much is pending, then exchanged by ONE grouped launch (DLS_BUCKET_BYTES overrides)
    bucket_bytes = int(os.environ.get("DLS_BUCKET_BYTES", 1 << 20))
    # one-shot peer exchange below this size (every rank reads world-1 remote copies: latency optimal, bandwidth
    # wasteful, so only for what is small); DLS_PEER_BYTES=0 turns the path off
    peer_bytes = int(os.environ.get("DLS_PEER_BYTES", 1 << 20))
    peer_staging_floats = 1 << 21
    # `reduce` (the route of stateful optimizers) sends its small gradients through the peer sum as well
    # (tests/test_gpu_peer.py::test_stateful_optimizer_takes_its_small_gradients_from_the_peer_sum); DLS_PEER_REDUCE=0
    # keeps them on NCCL
    peer_reduce = os.environ.get("DLS_PEER_REDUCE", "1") != "0"

    def allreduce(self, tensor: T.Tensor) -> T.Tensor:
        """Materialise the local partial gradient and queue it; the bucket goes out once `bucket_bytes` are pending or
        at `flush` (end of the step).  Small gradients (biases, early convolutions) then share one NCCL launch instead
        of paying its latency each, and the big ones leave as soon as they exist and overlap the rest of backward."""
        cached = tensor.do_cache()
        self._pending.append(cached)
        self._pending_bytes += cached.size * 4
        if self._pending_bytes >= self.bucket_bytes:
            self._flush_bucket()
        return cached

    def reduce(self, tensor: T.Tensor) -> T.Tensor:
        """SUM over the ranks, by whichever route fits the size: NCCL now (`allreduce`) for `peer_bytes` and more, else
        materialise and hold for ONE peer-memory launch at `flush` (dls_peer_sgd_step without weights: the sums are
        written back into the gradient buffers, which a stateful optimizer then consumes)."""
        if not (self.peer and self.peer_reduce) or tensor.size * 4 >= self.peer_bytes:
            return self.allreduce(tensor)
        cached = tensor.do_cache()
        self._held.append(cached)
        return cached

    def _flush_bucket(self):
        C, N = self._C, self._N
        pending, self._pending, self._pending_bytes = self._pending, [], 0
        if pending:
            bufs = [t._buffer() for t in pending]
            handles = (C.c_void_p * len(bufs))(*[b.handle for b in bufs])
            counts = (C.c_size_t * len(bufs))(*[t.size for t in pending])
            N.check(N.lib.dls_allreduce_sum_many(self.backend.ctx, handles, counts, len(bufs)))

    def flush(self):
        """End of the step: the leftover NCCL bucket goes out first (asynchronously, on the exchange stream), then the
        held small gradients are summed by one peer-memory launch on the compute stream.  Same order on every rank."""
        self._flush_bucket()
        C, N = self._C, self._N
        held, self._held = self._held, []
        if held:
            handles = (C.c_void_p * len(held))(*[t._buffer().handle for t in held])
            counts = (C.c_size_t * len(held))(*[t.size for t in held])
            flags = (C.c_int * len(held))(*([1] * len(held)))
            N.check(N.lib.dls_peer_sgd_step(self.backend.ctx, None, handles, counts, flags, len(held), 0.0))


class ExchangeSgd:
    """Plain SGD of a data-parallel job with the end of the step fused (takes the optimizer slot of `Plugins`): large
    gradients leave over NCCL as soon as they exist (`Exchange.allreduce`, overlapping backward), small ones are only
    materialised; `step_many` then issues ONE launch that sums the small ones over NVLink peer memory and updates every
    weight in place, w = w - lr * sum (dls_peer_sgd_step).  Same arithmetic as `data - delta * learningRate`."""
    t = 0

    def __init__(self, exchange: Optional["Exchange"], learning_rate: float, backend: Optional[T.Backend] = None):
        """`exchange=None`: single GPU -- every gradient is materialised and one multi-tensor launch updates all weights in
        place (no per-weight update launch, no swap, so a captured step needs no copy-back either)."""
        self.exchange, self.learning_rate = exchange, float(learning_rate)
        self.backend = exchange.backend if exchange is not None else (backend or T.default_backend())

    # single GPU: a layer orders its backward-data pass before its weight gradient, because `gradient` may apply the update
    # in place right away (layers.linear reads this flag)
    fuse_contractions = os.environ.get("DLS_FUSED_DW_SGD", "1") != "0"

    # the epilogue reads the weight tile with per-lane loads; under a long contraction that hides behind the next tile's main
    # loop (C4: K = 8192), under a short one (C1 / C3: K = batch = 128 / 256) it is exposed and the one-launch multi-tensor
    # update at full HBM bandwidth is faster (measured: C1 0.059 -> 0.072 ms with the epilogue update)
    fuse_min_k = int(os.environ.get("DLS_FUSED_DW_SGD_MIN_K", 1024))

    @staticmethod
    def _contraction_length(spec) -> int:
        a = spec["a"]
        return a.shape[0] if spec["transpose_a"] else a.shape[1]

    @property
    def applies_in_backward(self) -> bool:
        return self.exchange is None and self.fuse_contractions

    def gradient(self, weight, delta: T.Tensor) -> T.Tensor:
        ex = self.exchange
        spec = getattr(delta, "_ungated", None)
        if (self.applies_in_backward and not weight.plugins.cumulative and spec is not None and list(delta.shape) == list(weight.data.shape)
                and (delta._weak is None or delta._weak() is None) and weight.uses == 1 and isinstance(weight.data, T.CachedTensor)
                and self.backend.fuses_sgd_into_contractions(spec["mode"]) and self._contraction_length(spec) >= self.fuse_min_k):
            # the gradient is a still unevaluated contraction and this weight was read once this step: issue the
            # contraction with `data - delta * learningRate` (TW:63-72) as its epilogue, in place -- the gradient never
            # exists in HBM, and the new weight's bf16 operand copy comes out of the same kernel
            self.backend.matmul_sgd(spec, weight.data, self.learning_rate)
            return APPLIED
        d = delta.broadcast(weight.data.shape)
        if ex is not None and d.size * 4 >= ex.peer_bytes:
            cached = ex.allreduce(d)
            ex._over_nccl.add(id(cached._buffer()))
            return cached
        return d.do_cache()

    def begin_step(self, backend=None):
        pass

    def state_of(self, weight):
        return []

    def state_arrays(self, weight):
        return []

    def step_many(self, pending):
        import ctypes as C
        from . import _native as N
        ex = self.exchange
        n = len(pending)
        if n == 0:
            return
        wbufs = [w.data._buffer() for w, _ in pending]
        gbufs = [g._buffer() for _, g in pending]
        weights = (C.c_void_p * n)(*[b.handle for b in wbufs])
        grads = (C.c_void_p * n)(*[b.handle for b in gbufs])
        counts = (C.c_size_t * n)(*[w.data.size for w, _ in pending])
        flags = None
        if ex is not None:
            flags = (C.c_int * n)(*[0 if id(b) in ex._over_nccl else 1 for b in gbufs])
            ex._over_nccl.clear()
        N.check(N.lib.dls_peer_sgd_step(self.backend.ctx, weights, grads, counts, flags, n, self.learning_rate))


class Trainer:
    """One mixin stack + its weights.  `optimizer_delta` = learning_rate * (all-reduced) delta (TW:155-161)."""

    def __init__(self, backend: Optional[T.Backend] = None, learning_rate: float = 0.01, exchange=None, optimizer=None):
        """`optimizer`: None = plain SGD through the `delta` hook (`delta * learning_rate`, one fused launch per weight),
        or a stateful rule from `optimizers` (Momentum / Adam; `learning_rate` is then ignored)."""
        self.backend = backend or T.default_backend()
        self.learning_rate = float(learning_rate)
        self.exchange = exchange

        fused = None
        if exchange is not None and optimizer is None and exchange.peer:
            optimizer = fused = ExchangeSgd(exchange, self.learning_rate)
        elif exchange is None and optimizer is None and os.environ.get("DLS_FUSED_SGD", "1") != "0":
            # the single-GPU step with all updates in ONE in-place launch (no per-weight `data - delta` launch, no swap,
            # no copy-back in a captured step); DLS_FUSED_SGD=0 keeps the reference-literal per-weight swaps of TW:63-72
            optimizer = fused = ExchangeSgd(None, self.learning_rate, self.backend)

        def delta(weight, d):
            if fused is not None:
                return fused.gradient(weight, d)
            if self.exchange is not None:
                d = self.exchange.reduce(d.broadcast(weight.data.shape))
            return d if optimizer is not None else d * self.learning_rate
        self.P = Plugins(cumulative=False, optimizer_delta=delta, backend=self.backend, defer_updates=exchange is not None,
                         optimizer=optimizer)
        # `run` / `CompiledStep` flush explicitly after the loss reduction has been issued: the scalar loss does not wait
        # behind the gradient exchange and the update (5-9 us off the tail of a step)
        # (only with the fused SGD, whose gradients are materialised the moment they arrive: the lazy deltas of the other
        # optimizers read layer buffers that are released with the Do, so those still flush inside `train`)
        self.P.flush_in_train = fused is None
        if exchange is not None:
            self.P.before_flush.append(exchange.flush)
        self.L = Layers(self.P)
        self.weights: List = []

    def weight(self, shape: Sequence[int], seed: int, scale: float = 0.01):
        """`TensorWeight.allocate(Tensor.randomNormal(shape, seed) * scale)`: counter-based RNG, so every rank of a
        data-parallel job initialises identical replicas without a broadcast."""
        # Box-Muller in the reference does not guard u1 == 0 (Tensors:409-425): work-item 61 ^ seed draws hash 0 and
        # writes inf / NaN.  A seed with bit 30 set moves that work-item beyond every tensor used here.
        w = self.P.weight(self.backend.random_normal(list(shape), seed=seed | 0x40000000) * scale)
        self.weights.append(w)
        return w

    def run(self, loss_layer) -> float:
        """`loss.train` then the scalar loss read back (4 bytes D2H)."""
        loss = loss_layer.train().map(lambda data: data.sum().do_cache()).run()
        self.P.flush_updates()
        out = float(loss.flat_array()[0])
        if self.P.handled_exceptions:
            raise self.P.handled_exceptions.pop(0)
        return out

    def loss_layer(self, x: T.Tensor, target: Optional[T.Tensor] = None):
        raise NotImplementedError

    def step(self, x: T.Tensor, target: Optional[T.Tensor] = None) -> float:
        return self.run(self.loss_layer(x, target))

    def compile(self, x: T.Tensor, target: Optional[T.Tensor] = None) -> "CompiledStep":
        """Record one whole training step on the fixed input buffers `x` / `target` (see CompiledStep)."""
        return CompiledStep(self, x, target)

    def parameters(self) -> List[np.ndarray]:
        return [w.data.flat_array().reshape(w.data.shape) for w in self.weights]


class CompiledStep:
    """One training step (forward, backward, every weight update, the scalar loss) recorded into a CUDA graph
    (`Backend.capture`, csrc/graph.cu) and replayed with one driver call per step.  The user-level code is unchanged:
    the same `loss.train` runs once while recording.  The reference swaps each weight for a freshly allocated buffer
    (TW:63-72); inside the recording the new value is copied back into the weight's home buffer at the end of the step,
    so every replay reads the weights the previous one wrote.  `x` / `target` are device tensors whose buffers stay the
    step's inputs: refresh them with `write_inputs` between replays."""

    def __init__(self, trainer: "Trainer", x: T.Tensor, target: Optional[T.Tensor]):
        be = trainer.backend
        self.trainer, self.backend = trainer, be
        self.x = x.do_cache()
        self.target = target.do_cache() if target is not None else None
        trainer.step(self.x, self.target)                    # eager once: every plan is built, operands are staged
        homes = [(w, w.data._buffer(), list(w.data.shape), w.data.padding) for w in trainer.weights]
        trainer.P.recording = True
        try:
            with be.capture() as graph:
                loss = trainer.loss_layer(self.x, self.target).train().map(lambda data: data.sum().do_cache()).run()
                trainer.P.flush_updates()
                self._copy_back(homes)
        finally:
            trainer.P.recording = False
        if trainer.P.handled_exceptions:
            raise trainer.P.handled_exceptions.pop(0)
        self.graph, self._loss = graph, loss
        self._pipe = None

    # -- pipelined replays: the next batch travels while this step computes -------------------------------------------
    def submit(self, x_host: np.ndarray, target_host: Optional[np.ndarray] = None) -> Optional[float]:
        """One replay on a HOST batch, pipelined: the batch is uploaded on the context's copy stream into staging buffers
        (overlapping the replay submitted before), copied device-to-device into the step's input buffers, the step is
        replayed and its scalar loss is read back asynchronously.  Returns the loss of the PREVIOUS submitted step
        (None for the first): the host only ever waits for a step that has a successor queued behind it, so PCIe,
        the device and the host overlap.  `x_host` / `target_host` should be page-locked and must stay unchanged until
        the next `submit` (or `drain`) returns.  Same arithmetic as `__call__`; `drain()` yields the last loss."""
        import ctypes as C
        from . import _native as N
        be = self.backend
        if self._pipe is None:
            host = C.POINTER(C.c_float)()
            N.check(N.lib.dls_host_alloc(2, C.byref(host)))
            sx = be.fill(0.0, self.x.shape).do_cache()
            stt = be.fill(0.0, self.target.shape).do_cache() if self.target is not None else None
            self._pipe = {"x": sx, "t": stt, "free": None, "pending": None, "slot": 0, "host": host}
        p = self._pipe
        deps = [p["free"]] if p["free"] is not None else []
        uploads = [be.upload_async(p["x"]._buffer(), x_host, deps)]
        if target_host is not None:
            uploads.append(be.upload_async(p["t"]._buffer(), target_host, deps))
        for ev in uploads:
            be.stream_wait(ev)
        be.copy(self.x._buffer(), p["x"]._buffer(), self.x.size)
        if target_host is not None:
            be.copy(self.target._buffer(), p["t"]._buffer(), self.target.size)
        p["free"] = be.record_event()                        # the staging buffers may be overwritten from here on
        self.launch()
        previous, slot = p["pending"], p["slot"]
        p["slot"] ^= 1
        done = be.read_async(self._loss._buffer(), C.cast(C.addressof(p["host"].contents) + 4 * slot, C.c_void_p), 1)
        p["pending"] = (done, slot)
        if previous is None:
            return None
        previous[0].wait()
        return float(p["host"][previous[1]])

    def drain(self) -> Optional[float]:
        """Wait for the last submitted step and return its loss."""
        p = self._pipe
        if p is None or p["pending"] is None:
            return None
        done, slot = p["pending"]
        p["pending"] = None
        done.wait()
        return float(p["host"][slot])

    def __del__(self):
        p, self._pipe = getattr(self, "_pipe", None), None
        if p is not None and p.get("host"):
            from . import _native as N
            if N is not None:
                N.lib.dls_host_free(p["host"])

    def _copy_back(self, homes):
        be = self.backend
        for w, home, shape, padding in homes:
            if w.data._buffer() is not home:                 # swapped (TW:63-72): the next replay must read the new value
                be.copy(home, w.data._buffer(), w.data.size)
                w.data.release()
                w.data = T.CachedTensor(be, home, shape, padding)

    def write_inputs(self, x_host: np.ndarray, target_host: Optional[np.ndarray] = None):
        self.backend.write(self.x._buffer(), x_host)
        if target_host is not None:
            self.backend.write(self.target._buffer(), target_host)

    def launch(self):
        self.trainer.P.begin_step()                          # stateful optimizers: this replay's hyper-parameters
        self.graph.launch()

    def loss(self) -> float:
        """Scalar loss of the most recent replay (4-byte D2H, waits for the replay)."""
        return float(self._loss.flat_array()[0])

    def __call__(self, x_host: Optional[np.ndarray] = None, target_host: Optional[np.ndarray] = None) -> float:
        if x_host is not None:
            self.write_inputs(x_host, target_host)
        self.launch()
        return self.loss()


class Mlp(Trainer):
    """Feature-major MLP: h_{l+1} = act(W_l h_l + b_l), activations `[features, batch]` (left-aligned broadcast makes the
    bias a per-row vector).  C1: sizes (784, 500, 10), tanh hidden, linear output, loss sum((out - onehot)^2).
    C4: sizes (4096,)*5, no bias, tanh everywhere, loss sum(out)."""

    def __init__(self, sizes: Sequence[int], bias: bool = True, last_activation: str = "none", seed0: int = 11, **kw):
        super().__init__(**kw)
        self.sizes = list(sizes)
        self.last_activation = last_activation
        self.layers = []
        seed = seed0
        for fan_in, fan_out in zip(self.sizes[:-1], self.sizes[1:]):
            w = self.weight([fan_out, fan_in], seed)
            b = self.weight([fan_out], seed + 100) if bias else None
            seed += 1
            self.layers.append((w, b))

    def forward(self, x: T.Tensor):
        h = self.P.lift(x)
        for i, (w, b) in enumerate(self.layers):
            act = "tanh" if i + 1 < len(self.layers) else self.last_activation
            h = self.L.linear(w, h, b, "row" if b is not None else "none", act)
        return h

    def loss_layer(self, x: T.Tensor, target: Optional[T.Tensor] = None):
        out = self.forward(x)
        return self.L.squared_error(out, target) if target is not None else out


class Cnn(Trainer):
    """C3: NHWC input, three 'same' 3x3 convolutions (ReLU) with 2x nearest-neighbour down-sampling between them,
    flatten, FC -> 256 (ReLU) -> 10, loss sum((out - onehot)^2)."""

    def __init__(self, image: int = 32, channels: Sequence[int] = (3, 32, 64, 128), hidden: int = 256, classes: int = 10,
                 seed0: int = 51, **kw):
        super().__init__(**kw)
        self.image, self.channels = image, list(channels)
        self.convs = []
        seed = seed0
        for cin, cout in zip(self.channels[:-1], self.channels[1:]):
            self.convs.append((self.weight([3, 3, cin, cout], seed, scale=0.05), self.weight([cout], seed + 100)))
            seed += 1
        side = image // (2 ** (len(self.convs) - 1))
        self.flat = side * side * self.channels[-1]
        self.fc1 = (self.weight([self.flat, hidden], seed), self.weight([hidden], seed + 100))
        self.fc2 = (self.weight([hidden, classes], seed + 1), self.weight([classes], seed + 101))

    def forward(self, x: T.Tensor):
        L = self.L
        h = self.P.lift(x)
        batch = x.shape[0]
        for i, (w, b) in enumerate(self.convs):
            h = L.conv2d_3x3(h, w, b, "relu")
            if i + 1 < len(self.convs):
                h = L.downsample(h, 2)
        h = L.reshape(h, [batch, self.flat])
        h = L.linear(h, self.fc1[0], self.fc1[1], "col", "relu")
        return L.linear(h, self.fc2[0], self.fc2[1], "col", "none")

    def loss_layer(self, x: T.Tensor, target: Optional[T.Tensor] = None):
        return self.L.squared_error(self.forward(x), target)


def one_hot(labels: np.ndarray, classes: int, feature_major: bool) -> np.ndarray:
    eye = np.eye(classes, dtype=np.float32)[labels]
    return np.ascontiguousarray(eye.T if feature_major else eye)


def shard(array: np.ndarray, axis: int, rank: int, world_size: int) -> np.ndarray:
    """Batch shard of one rank: contiguous split of `axis` (the batch axis) into world_size equal parts."""
    n = array.shape[axis]
    if n % world_size:
        raise ValueError("batch %d is not divisible by %d ranks" % (n, world_size))
    per = n // world_size
    index = [slice(None)] * array.ndim
    index[axis] = slice(rank * per, (rank + 1) * per)
    return np.ascontiguousarray(array[tuple(index)])
