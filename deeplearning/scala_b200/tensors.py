"""The lazy `Tensor` algebra of Compute.scala's `Tensors` trait on top of libdlsb200 (CUDA, sm_100a).

Host-side mirror of `trait Tensors` (Tensors = Compute.scala/Tensors/src/main/scala/com/thoughtworks/compute/Tensors.scala):
same public surface, same laziness, same class-dependent view semantics (the checkpoint rule, Tensors:978-1003), but
every slow action goes through the C ABI (`dls_launch_tree`, `dls_sum`, `dls_random*`, ...) instead of generating and
compiling OpenCL C.  `Backend` is one `Factory[... with Tensors].newInstance(...)`; the module-level functions are the
`gpu` singleton (gpu.scala:15-27).  There is no CPU path in this module.
"""
from __future__ import annotations

import ctypes as C
import threading
import weakref
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N
from . import ndat
from ._format import format_tensor
from .trees import Binary, Ident, Literal, Unary, Xform, serialize

F32 = np.float32


# ----------------------------------------------------------------------------------------------------------------------
# device handles
# ----------------------------------------------------------------------------------------------------------------------


class DeviceBuffer:
    """Owns one reference to a `dls_buf` (OpenCL.DeviceBuffer, OpenCL.scala:633-717)."""
    __slots__ = ("backend", "handle", "length", "__weakref__")

    def __init__(self, backend, handle, length):
        self.backend, self.handle, self.length = backend, handle, length

    def __del__(self):
        h, self.handle = self.handle, None
        if h and N is not None and self.backend._ctx:
            N.lib.dls_buf_release(h)

    @property
    def device_ptr(self) -> int:
        return N.lib.dls_buf_device_ptr(self.handle)


class Event:
    """Owns one reference to a `dls_event` (OpenCL.Event, OpenCL.scala:565-612)."""
    __slots__ = ("handle",)

    def __init__(self, handle):
        self.handle = handle

    def wait(self):
        N.check(N.lib.dls_event_wait(self.handle))

    def __del__(self):
        h, self.handle = self.handle, None
        if h and N is not None:
            N.lib.dls_event_release(h)


def elapsed_ms(start: Event, end: Event) -> float:
    ms = C.c_float()
    N.check(N.lib.dls_event_elapsed_ms(start.handle, end.handle, C.byref(ms)))
    return ms.value


class CapturedStep:
    """Owns one reference to a `dls_graph`: everything the backend launched inside `with backend.capture() as step:`
    (csrc/graph.cu).  `step.launch()` replays it with one driver call."""
    __slots__ = ("backend", "handle")

    def __init__(self, backend):
        self.backend, self.handle = backend, None

    def launch(self):
        if not self.handle:
            raise RuntimeError("the capture did not complete")
        N.check(N.lib.dls_graph_launch(self.handle))

    def info(self) -> dict:
        k, n, b = C.c_uint64(), C.c_uint64(), C.c_uint64()
        N.check(N.lib.dls_graph_info(self.handle, C.byref(k), C.byref(n), C.byref(b)))
        return {"kernel_nodes": k.value, "nodes": n.value, "arena_bytes": b.value}

    def __del__(self):
        h, self.handle = self.handle, None
        if h and N is not None and self.backend._ctx:
            N.lib.dls_graph_release(h)


class _Capture:
    def __init__(self, backend):
        self.backend, self.step = backend, CapturedStep(backend)

    def __enter__(self) -> CapturedStep:
        N.check(N.lib.dls_graph_begin(self.backend.ctx))
        return self.step

    def __exit__(self, exc_type, exc, tb):
        if exc_type is not None:
            N.lib.dls_graph_abort(self.backend.ctx)
            return False
        h = C.c_void_p()
        N.check(N.lib.dls_graph_end(self.backend.ctx, C.byref(h)))
        self.step.handle = h
        return False


class Backend:
    """One instance of the `Tensors` trait: context, stream pool, allocator, plan cache (Tensors:199, OpenCL:1139)."""

    def __init__(self, device: int = 0, streams_per_device: int = 1, rng_hash: str = "wang", gemm_mode: str = "fp32",
                 allocator_cap_bytes: int = 0, fuse_reductions: bool = True):
        self._opts = N.CtxOpts(device=device, streams_per_device=streams_per_device,
                               rng_hash={"wang": N.DLS_HASH_WANG, "lcg": N.DLS_HASH_LCG}[rng_hash],
                               gemm_mode={"fp32": N.DLS_GEMM_FP32, "tf32": N.DLS_GEMM_TF32, "bf16": N.DLS_GEMM_BF16}[gemm_mode],
                               allocator_cap_bytes=allocator_cap_bytes, plan_cache_capacity=0, reserved=0)
        self.fuse_reductions = fuse_reductions
        self.gemm_mode = gemm_mode
        self._ctx = None
        self._lock = threading.Lock()

    # -- lifetime (lazy like OpenCL.context, OpenCL:1376-1393; close == monadicClose, OpenCL:1335) ---------------------
    @property
    def ctx(self):
        if self._ctx is None:
            with self._lock:
                if self._ctx is None:
                    h = C.c_void_p()
                    N.check(N.lib.dls_ctx_create(C.byref(self._opts), C.byref(h)))
                    self._ctx = h
        return self._ctx

    def close(self):
        ctx, self._ctx = self._ctx, None
        if ctx:
            N.lib.dls_ctx_destroy(ctx)

    def synchronize(self):
        N.check(N.lib.dls_ctx_sync(self.ctx))

    def stats(self) -> dict:
        s = N.Stats()
        N.check(N.lib.dls_ctx_stats(self.ctx, C.byref(s)))
        return {k: int(getattr(s, k)) for k, _ in N.Stats._fields_}

    def record_event(self, stream: int = 0) -> Event:
        h = C.c_void_p()
        N.check(N.lib.dls_event_record(self.ctx, stream, C.byref(h)))
        return Event(h)

    def stream_handle(self, index: int = 0) -> int:
        return N.lib.dls_ctx_stream(self.ctx, index)

    # -- raw device operations ---------------------------------------------------------------------------------------
    def upload(self, flat: np.ndarray) -> DeviceBuffer:
        flat = np.ascontiguousarray(flat, dtype=F32).reshape(-1)
        h = C.c_void_p()
        N.check(N.lib.dls_buf_from_host(self.ctx, flat.ctypes.data_as(C.c_void_p), flat.size, C.byref(h), None))
        # the source array may be freed by the caller: the copy above is enqueued from pageable memory, which the CUDA
        # runtime stages synchronously, so no extra wait is needed here
        return DeviceBuffer(self, h, flat.size)

    def capture(self) -> _Capture:
        """`with backend.capture() as step:` records every launch made inside instead of executing it (one CUDA graph);
        `step.launch()` replays the recording.  Inputs must already be on the device; nothing may be read back inside."""
        return _Capture(self)

    def copy(self, dst: DeviceBuffer, src: DeviceBuffer, n: Optional[int] = None):
        """Stream-ordered device-to-device copy (recordable)."""
        N.check(N.lib.dls_buf_copy(self.ctx, dst.handle, src.handle, min(dst.length, src.length) if n is None else n))

    def write(self, buf: DeviceBuffer, flat: np.ndarray, offset: int = 0):
        """Refresh an existing device buffer from host memory (asynchronous when `flat` is pinned)."""
        flat = np.ascontiguousarray(flat, dtype=F32).reshape(-1)
        N.check(N.lib.dls_buf_write(self.ctx, buf.handle, offset, flat.ctypes.data_as(C.c_void_p), flat.size, None))

    def upload_async(self, buf: DeviceBuffer, flat: np.ndarray, deps: Sequence[Event] = (), offset: int = 0) -> Event:
        """Host -> device on the context's COPY stream (`dls_buf_upload_async`): overlaps whatever the compute stream
        runs.  `flat` should live in page-locked memory (`dls_host_alloc`) and must not change until the returned event
        completes; consumers are ordered with `stream_wait` on that event."""
        flat = np.ascontiguousarray(flat, dtype=F32).reshape(-1)
        arr = (C.c_void_p * max(1, len(deps)))(*[d.handle for d in deps])
        h = C.c_void_p()
        N.check(N.lib.dls_buf_upload_async(self.ctx, buf.handle, offset, flat.ctypes.data_as(C.c_void_p), flat.size, arr,
                                           len(deps), C.byref(h)))
        return Event(h)

    def stream_wait(self, event: Event, stream: int = 0):
        """Everything queued on pool stream `stream` after this call waits for `event`."""
        N.check(N.lib.dls_stream_wait_event(self.ctx, stream, event.handle))

    def read_async(self, buf: DeviceBuffer, host_ptr, n: int, offset: int = 0) -> Event:
        """Non-blocking read-back into (page-locked) host memory; the data is valid once the returned event completes."""
        h = C.c_void_p()
        N.check(N.lib.dls_buf_read(self.ctx, buf.handle, offset, host_ptr, n, None, 0, C.byref(h)))
        return Event(h)

    def download(self, buf: DeviceBuffer, n: Optional[int] = None, offset: int = 0) -> np.ndarray:
        n = buf.length - offset if n is None else n
        out = np.empty(n, dtype=F32)
        N.check(N.lib.dls_buf_read(self.ctx, buf.handle, offset, out.ctypes.data_as(C.c_void_p), n, None, 0, None))
        return out

    def launch(self, roots, out_shape, reduce_sum: bool = False) -> DeviceBuffer:
        """enqueueClosure (Tensors:1163-1264): serialise, materialise every parameter tensor, launch."""
        blob, params = serialize(roots, out_shape)
        bufs = [p._buffer() for p in params]                         # Tensors:1208-1212
        for b in bufs:
            if b.backend is not self:                                # buffers of another context: no stream ordering
                raise ValueError("a tensor of another Backend instance was used in this expression")
        arr = (C.c_void_p * max(1, len(bufs)))(*[b.handle for b in bufs])
        out = C.c_void_p()
        flags = N.DLS_LAUNCH_REDUCE_SUM if reduce_sum else N.DLS_LAUNCH_DEFAULT
        N.check(N.lib.dls_launch_tree(self.ctx, blob, len(blob), arr, len(bufs), None, 0, flags, C.byref(out), None))
        n = 1 if reduce_sum else int(np.prod(out_shape, dtype=np.int64)) * len(roots)
        return DeviceBuffer(self, out, n)

    def describe_plan(self, roots, out_shape) -> str:
        blob, _ = serialize(roots, out_shape)
        text = C.create_string_buffer(4096)
        N.check(N.lib.dls_plan_describe(None, blob, len(blob), text, len(text)))
        return text.value.decode()

    def plan_cache_contains(self, tensor: "Tensor") -> bool:
        """`kernelCache.getIfPresent(tensor.getClosure)` (TensorsSpec:48-52)."""
        blob, _ = serialize([tensor.closure()], tensor.shape)
        found = C.c_int()
        N.check(N.lib.dls_plan_cache_contains(self.ctx, blob, len(blob), C.byref(found)))
        return bool(found.value)

    # -- object Tensor (Tensors:395-600) -------------------------------------------------------------------------------
    def tensor(self, elements, padding: float = 0.0) -> "NonInlineTensor":
        """Tensor.apply (Tensors:445-463); ragged input raises ValueError (TensorBuilder.shape, Tensors:78-91)."""
        if isinstance(elements, np.ndarray):
            data = np.ascontiguousarray(elements, dtype=F32)
            shape = list(data.shape)
        else:
            shape = _nested_shape(elements)
            flat: List[float] = []
            _nested_flatten(elements, flat)
            data = np.array(flat, dtype=F32)
        return UploadTensor(self, data.reshape(-1), shape, padding)

    def scalar(self, value, padding: float = 0.0) -> "FillTensor":
        return FillTensor(self, value, [], padding)

    def fill(self, value, shape, padding: float = 0.0) -> "FillTensor":
        return FillTensor(self, value, list(shape), padding)

    def random(self, shape, seed: int, padding: float = 0.0) -> "NonInlineTensor":
        """Tensor.random (Tensors:479-497)."""
        shape = [int(s) for s in shape]
        n = int(np.prod(shape, dtype=np.int64))

        def produce():
            h = C.c_void_p()
            N.check(N.lib.dls_random(self.ctx, n, seed & 0xFFFFFFFF, C.byref(h), None))
            return DeviceBuffer(self, h, n)
        return ProducedTensor(self, produce, shape, padding)

    def random_normal(self, shape, seed: int, padding: float = 0.0) -> "NonInlineTensor":
        """Tensor.randomNormal (Tensors:500-524)."""
        shape = [int(s) for s in shape]
        n = int(np.prod(shape, dtype=np.int64))

        def produce():
            h = C.c_void_p()
            N.check(N.lib.dls_random_normal(self.ctx, n, seed & 0xFFFFFFFF, C.byref(h), None))
            return DeviceBuffer(self, h, n)
        return ProducedTensor(self, produce, shape, padding)

    def join(self, tensors: Sequence["Tensor"], dimension: Optional[int] = None) -> "Tensor":
        """Tensor.join (Tensors:560-598): every element is evaluated at the HEAD tensor's shape and written as the new
        last dimension (Concatenate, KB:632-652); a `dimension` other than the last is a permute of that (Tensors:560-575)."""
        tensors = list(tensors)
        head = tensors[0]
        shape = list(head.shape) + [len(tensors)]
        roots = [t.closure() for t in tensors]
        joined = ProducedTensor(self, lambda: self.launch(roots, head.shape), shape, head.padding)
        if dimension is None or dimension == len(shape) - 1:
            return joined
        n = len(shape)
        perm = [i if i < dimension else (n - 1 if i == dimension else i - 1) for i in range(n)]
        return joined.permute(perm)

    def matmul(self, a: "Tensor", b: "Tensor", transpose_a: bool = False, transpose_b: bool = False, bias=None,
               bias_mode: str = "none", activation: str = "none", mode: Optional[str] = None, gate: Optional["Tensor"] = None,
               gate_mode: str = "none", emit_operand_copy: bool = False) -> "NonInlineTensor":
        """out = act(op(a) op(b) + bias) as ONE contraction kernel (`dls_matmul_ex`): what the reference can only spell
        as a host-unrolled `split(1) ... reduce(_ + _)` chain (TensorsSpec:337-344, :371-383).  `mode` None = this
        backend's gemm_mode; "fp32" is the CUDA-core kernel (same left-fold rounding as the chain), "tf32" / "bf16" run
        on the tcgen05 tensor cores with fp32 accumulation.
        `gate` / `gate_mode` ("tanh" | "relu"): the product is multiplied by the activation derivative of the [m, n]
        tensor `gate` in the same kernel (`delta * (1 - y * y)` / `delta * 1[y > 0]`, dls_matmul_gated).
        `emit_operand_copy`: the kernel also leaves the result in the mode's operand type on the output buffer, for a
        contraction that consumes it next."""
        if len(a.shape) != 2 or len(b.shape) != 2:
            raise ValueError("matmul: operands must be matrices, got %r and %r" % (a.shape, b.shape))
        m, k = (a.shape[1], a.shape[0]) if transpose_a else (a.shape[0], a.shape[1])
        n, k2 = (b.shape[0], b.shape[1]) if transpose_b else (b.shape[1], b.shape[0])
        if k != k2:
            raise ValueError("matmul: inner dimensions differ (%d vs %d)" % (k, k2))
        gm = {"none": 0, "tanh": 1, "relu": 2}[gate_mode]
        if gm and (gate is None or gate.size != m * n):
            raise ValueError("matmul: gate must have %d elements" % (m * n))
        opts = N.GemmOpts(mode={None: -1, "fp32": N.DLS_GEMM_FP32, "tf32": N.DLS_GEMM_TF32, "bf16": N.DLS_GEMM_BF16}[mode],
                          bias_mode={"none": 0, "row": 1, "col": 2}[bias_mode],
                          activation={"none": 0, "tanh": 1, "relu": 2}[activation], flags=(1 if emit_operand_copy else 0) | (gm << 8))
        if opts.bias_mode and (bias is None or bias.size != (m if bias_mode == "row" else n)):
            raise ValueError("matmul: bias must have %d elements" % (m if bias_mode == "row" else n))

        def produce():
            ba, bb = a._buffer(), b._buffer()
            bbias = bias._buffer() if opts.bias_mode else None
            bgate = gate._buffer() if gm else None
            h = C.c_void_p()
            N.check(N.lib.dls_matmul_gated(self.ctx, ba.handle, bb.handle, bbias.handle if bbias else None,
                                           bgate.handle if bgate else None, m, n, k, int(transpose_a), int(transpose_b),
                                           C.byref(opts), None, 0, C.byref(h), None))
            return DeviceBuffer(self, h, m * n)
        out = ProducedTensor(self, produce, [m, n], a.padding)
        if not opts.bias_mode and activation == "none" and not gm:
            # a consumer that multiplies this product by an activation derivative may re-issue it with the gate fused
            # (layers.linear backward) as long as nobody has produced it yet
            out._ungated = dict(a=a, b=b, transpose_a=transpose_a, transpose_b=transpose_b, mode=mode)
        return out

    def matmul_sgd(self, spec: dict, weight: "CachedTensor", learning_rate: float) -> None:
        """weight -= learning_rate * (op(a) op(b)) in place, in the contraction's epilogue (`dls_matmul_sgd`); `spec` is the
        `_ungated` record of the unevaluated product that would have been the gradient."""
        a, b = spec["a"], spec["b"]
        ta, tb = spec["transpose_a"], spec["transpose_b"]
        m, k = (a.shape[1], a.shape[0]) if ta else (a.shape[0], a.shape[1])
        n = b.shape[0] if tb else b.shape[1]
        N.check(N.lib.dls_matmul_sgd(self.ctx, a._buffer().handle, b._buffer().handle, weight._buffer().handle, m, n, k, int(ta), int(tb),
                                     float(learning_rate), self._MODES[spec["mode"]]))

    def fuses_sgd_into_contractions(self, mode: Optional[str] = None) -> bool:
        return (mode or self.gemm_mode) in ("tf32", "bf16")

    # -- 'same' 3x3 convolution without an im2col matrix (csrc/conv.cu; TensorsSpec:144-210 spells it with nine translates) --
    _MODES = {None: -1, "fp32": N.DLS_GEMM_FP32, "tf32": N.DLS_GEMM_TF32, "bf16": N.DLS_GEMM_BF16}

    def conv3x3_supported(self, which: str, h: int, w: int, cin: int, cout: int, mode: Optional[str] = None) -> bool:
        """Does libdlsb200 have a dedicated kernel for this pass ("forward" / "backward_data" / "backward_weights") at this
        shape in this contraction mode?  If not, compose the convolution from `translate`, `join` and `matmul`."""
        p = {"forward": N.DLS_CONV_FORWARD, "backward_data": N.DLS_CONV_BACKWARD_DATA, "backward_weights": N.DLS_CONV_BACKWARD_WEIGHTS}[which]
        return bool(N.lib.dls_conv3x3_supported(self.ctx, p, h, w, cin, cout, self._MODES[mode]))

    @staticmethod
    def _sampled_source(x: "Tensor"):
        """(tensor to read, stride): a nearest-neighbour `scale` down by an integer factor of the two middle dims of a
        MATERIALISED NHWC tensor is read in place by the convolution kernels instead of being gathered first."""
        if isinstance(x, TransformedTensor) and len(x.shape) == 4 and isinstance(x.checkpoint, NonInlineTensor):
            src = x.checkpoint.shape
            m = x.matrix
            if m.shape == (4, 5) and src[0] == x.shape[0] and src[3] == x.shape[3] and x.shape[1] and src[1] % x.shape[1] == 0:
                s = src[1] // x.shape[1]
                want = np.zeros((4, 5))
                want[0, 0] = want[3, 3] = 1.0
                want[1, 1] = want[2, 2] = float(s)
                if s > 1 and src[2] == x.shape[2] * s and np.array_equal(m, want):
                    return x.checkpoint, s
        return x, 1

    def conv3x3_forward(self, x: "Tensor", w: "Tensor", bias: Optional["Tensor"] = None, activation: str = "none",
                        mode: Optional[str] = None, out_stride: int = 0) -> "NonInlineTensor":
        """out[b,y,x,f] = act(bias[f] + sum x[b, y-(ky-1), x-(kx-1), c] w[ky,kx,c,f]) for NHWC `x` and `w` [3,3,Cin,Cout]."""
        batch, h, wd, cin = x.shape
        cout = w.shape[3]
        src, stride = self._sampled_source(x)
        # out_stride: the value is consumed next by a contraction / convolution (1) or through a `scale` down-sampling by s
        opts = N.ConvOpts(mode=self._MODES[mode], activation={"none": 0, "tanh": 1, "relu": 2}[activation], in_stride=stride,
                          dz_stride=out_stride)

        def produce():
            bx, bw = src._buffer(), w._buffer()
            bb = bias._buffer() if bias is not None else None
            hd = C.c_void_p()
            N.check(N.lib.dls_conv3x3_forward(self.ctx, bx.handle, bw.handle, bb.handle if bb else None, batch, h, wd, cin, cout,
                                              C.byref(opts), None, 0, C.byref(hd), None))
            return DeviceBuffer(self, hd, batch * h * wd * cout)
        return ProducedTensor(self, produce, [batch, h, wd, cout], x.padding)

    def conv3x3_backward_data(self, dz: "Tensor", w: "Tensor", mode: Optional[str] = None, dz_stride: int = 1,
                              gate: Optional["Tensor"] = None, gate_mode: str = "none", gate_stride: int = 1) -> "NonInlineTensor":
        """dx[b,y,x,c] = sum dz[b, y+(ky-1), x+(kx-1), f] w[ky,kx,c,f].  `dz_stride` = s > 1: `dz` is the SAMPLED delta
        [batch, h/s, w/s, cout] that a down-sampling after the layer hands back; it stands for the full-size delta that is
        zero everywhere except at (s*y, s*x), which is never materialised in fp32.
        `gate` / `gate_mode` / `gate_stride`: the result is multiplied by the activation derivative of `gate`
        [batch, h*gs, w*gs, cin] read at (gs*y, gs*x) in the same kernel -- the delta of the layer below."""
        batch, hs, ws, cout = dz.shape
        h, wd = hs * dz_stride, ws * dz_stride
        cin = w.shape[2]
        gm = {"none": 0, "tanh": 1, "relu": 2}[gate_mode]
        if gm and (gate is None or gate.size != batch * h * gate_stride * wd * gate_stride * cin):
            raise ValueError("conv3x3_backward_data: gate must be [batch, h*gs, w*gs, cin]")
        opts = N.ConvOpts(mode=self._MODES[mode], activation=0, in_stride=1, dz_stride=dz_stride)

        def produce():
            bd, bw = dz._buffer(), w._buffer()
            bg = gate._buffer() if gm else None
            hd = C.c_void_p()
            N.check(N.lib.dls_conv3x3_backward_data_gated(self.ctx, bd.handle, bw.handle, bg.handle if bg else None, gm, gate_stride, batch,
                                                          h, wd, cin, cout, C.byref(opts), None, 0, C.byref(hd), None))
            return DeviceBuffer(self, hd, batch * h * wd * cin)
        out = ProducedTensor(self, produce, [batch, h, wd, cin], dz.padding)
        if not gm:                                    # a consumer may re-issue this pass with its activation derivative fused
            out._ungated_conv = dict(dz=dz, w=w, mode=mode, dz_stride=dz_stride)
        return out

    def conv3x3_backward_weights(self, x: "Tensor", dz: "Tensor", mode: Optional[str] = None, dz_stride: int = 1) -> "NonInlineTensor":
        """dw[ky,kx,c,f] = sum over b,y,x of x[b, y-(ky-1), x-(kx-1), c] dz[b,y,x,f]; `dz_stride` as in backward_data."""
        batch, h, wd, cin = x.shape
        cout = dz.shape[3]
        src, stride = self._sampled_source(x)
        opts = N.ConvOpts(mode=self._MODES[mode], activation=0, in_stride=stride, dz_stride=dz_stride)

        def produce():
            bx, bd = src._buffer(), dz._buffer()
            hd = C.c_void_p()
            N.check(N.lib.dls_conv3x3_backward_weights(self.ctx, bx.handle, bd.handle, batch, h, wd, cin, cout, C.byref(opts), None, 0,
                                                       C.byref(hd), None))
            return DeviceBuffer(self, hd, 9 * cin * cout)
        return ProducedTensor(self, produce, [3, 3, cin, cout], x.padding)

    def conv3x3_backward_weights_bias(self, x: "Tensor", dz: "Tensor", mode: Optional[str] = None, dz_stride: int = 1):
        """(dw, db) from ONE pass over dz: the first-layer kernel (Cin <= 4) sums the bias gradient on the way
        (`dls_conv3x3_backward_weights_bias`); whichever of the two tensors is needed first runs the kernel."""
        batch, h, wd, cin = x.shape
        cout = dz.shape[3]
        src, stride = self._sampled_source(x)
        opts = N.ConvOpts(mode=self._MODES[mode], activation=0, in_stride=stride, dz_stride=dz_stride)
        made = {}

        def run():
            if "dw" not in made:
                bx, bd = src._buffer(), dz._buffer()
                hd, hb = C.c_void_p(), C.c_void_p()
                N.check(N.lib.dls_conv3x3_backward_weights_bias(self.ctx, bx.handle, bd.handle, batch, h, wd, cin, cout, C.byref(opts),
                                                                None, 0, C.byref(hd), C.byref(hb), None))
                made["dw"], made["db"] = DeviceBuffer(self, hd, 9 * cin * cout), DeviceBuffer(self, hb, cout)
            return made
        dw = ProducedTensor(self, lambda: run()["dw"], [3, 3, cin, cout], x.padding)
        db = ProducedTensor(self, lambda: run()["db"], [cout], x.padding)
        return dw, db

    def sum_axis(self, t: "Tensor", axis: int) -> "NonInlineTensor":
        """Sum away one dimension: what TensorLayers.sumAs spells as `split(axis).reduce(_ + _)` (TL:124-162), as one
        row / column reduction kernel (`dls_sum_axis`) instead of a host-unrolled chain of `shape[axis]` additions."""
        shape = list(t.shape)
        if not 0 <= axis < len(shape):
            raise ValueError("sum_axis: axis %d out of range for shape %r" % (axis, shape))
        outer = int(np.prod(shape[:axis], dtype=np.int64))
        k = shape[axis]
        inner = int(np.prod(shape[axis + 1:], dtype=np.int64))

        def produce():
            buf = t._buffer()
            h = C.c_void_p()
            N.check(N.lib.dls_sum_axis(self.ctx, buf.handle, outer, k, inner, None, 0, C.byref(h), None))
            return DeviceBuffer(self, h, outer * inner)
        return ProducedTensor(self, produce, shape[:axis] + shape[axis + 1:], t.padding)

    def abs(self, t): return t._derived(Unary("abs", t.closure()))
    def sqrt(self, t): return t._derived(Unary("sqrt", t.closure()))
    def tanh(self, t): return t._derived(Unary("tanh", t.closure()))
    def exp(self, t): return t._derived(Unary("exp", t.closure()))
    def log(self, t): return t._derived(Unary("log", t.closure()))
    def min(self, a, b): return a._binary("min", b)
    def max(self, a, b): return a._binary("max", b)


def _is_seq(a):
    return isinstance(a, (list, tuple)) or (isinstance(a, np.ndarray) and a.ndim > 0)


def _nested_shape(a):
    if _is_seq(a):
        if len(a) == 0:
            return [0]
        subs = [_nested_shape(x) for x in a]
        for s in subs[1:]:
            if s != subs[0]:
                raise ValueError("ragged nested sequence: %r vs %r" % (subs[0], s))
        return [len(a)] + subs[0]
    return []


def _nested_flatten(a, out):
    if _is_seq(a):
        for x in a:
            _nested_flatten(x, out)
    else:
        out.append(float(a))


def auto_broadcast_shape(s1, s2):
    """Tensors:208-222 -- dims are aligned from the LEFT (unlike numpy)."""
    out = []
    for i in range(max(len(s1), len(s2))):
        if i >= len(s1) or s1[i] == 1:
            out.append(s2[i])
        elif i >= len(s2) or s2[i] == 1:
            out.append(s1[i])
        elif s1[i] == s2[i]:
            out.append(s1[i])
        else:
            raise ValueError("Failed to automatically broadcast between shape %r and %r" % (list(s1), list(s2)))
    return out


# ----------------------------------------------------------------------------------------------------------------------
# sealed trait Tensor (Tensors:636-1133) and its five classes (Tensors:602, :1266-1312)
# ----------------------------------------------------------------------------------------------------------------------


class Tensor:
    backend: Backend
    shape: List[int]
    padding: float

    # -- closure / buffer ------------------------------------------------------------------------------------------
    def closure(self):
        raise NotImplementedError

    def _buffer(self) -> DeviceBuffer:
        """doBuffer: the row-major device buffer of this tensor."""
        raise NotImplementedError

    @property
    def size(self) -> int:
        return int(np.prod(self.shape, dtype=np.int64))

    # -- slow actions ----------------------------------------------------------------------------------------------
    def flat_array(self) -> np.ndarray:
        """flatArray (Tensors:1099-1118): materialise, then a blocking read-back."""
        return self.backend.download(self._buffer(), self.size)

    def to_string(self) -> str:
        """toString (Tensors:776-811)."""
        return format_tensor(self.shape, self.flat_array())

    __str__ = to_string

    def do_cache(self) -> "CachedTensor":
        """doCache (Tensors:642-666): pin the materialised buffer."""
        return CachedTensor(self.backend, self._buffer(), list(self.shape), self.padding)

    def non_inline(self) -> "NonInlineTensor":
        """nonInline (Tensors:671, :1277-1282)."""
        return ProducedTensor(self.backend, self._buffer, list(self.shape), self.padding)

    def sum(self) -> "NonInlineTensor":
        """sum (Tensors:771) = reduce(PlusPrograms) (Tensors:673-766).  The reference always materialises the receiver
        first (Tensors:678); with `fuse_reductions` an inline receiver is reduced inside the interpreter kernel instead
        (8 B/element instead of 16 for tanh(a*b+c).sum)."""
        be = self.backend

        def produce():
            if be.fuse_reductions and isinstance(self, InlineTensor) and self._live() is None:
                return be.launch([self.closure()], self.shape, reduce_sum=True)
            buf = self._buffer()
            h = C.c_void_p()
            N.check(N.lib.dls_sum(be.ctx, buf.handle, self.size, None, 0, C.byref(h), None))
            return DeviceBuffer(be, h, 1)
        return ProducedTensor(be, produce, [], self.padding)

    # -- views (Tensors:816-1074) ----------------------------------------------------------------------------------
    def _transform(self, new_shape, matrix1) -> "TransformedTensor":
        """Tensors:978-1003, the checkpoint rule: a view of a view composes against the ORIGINAL checkpoint; a view of
        anything else makes the receiver the checkpoint (it is materialised when the consumer runs)."""
        matrix1 = np.asarray(matrix1, dtype=np.float64)
        if isinstance(self, TransformedTensor):
            m = ndat.pre_concatenate(matrix1, self.matrix)
            return TransformedTensor(self.backend, self.checkpoint, m, list(new_shape), self.padding)
        return TransformedTensor(self.backend, self, matrix1, list(new_shape), self.padding)

    def broadcast(self, new_shape) -> "Tensor":
        """Tensors:816-855 (left aligned; a FillTensor stays a FillTensor)."""
        new_shape = [int(s) for s in new_shape]
        if new_shape == list(self.shape):
            return self
        if isinstance(self, FillTensor):
            return FillTensor(self.backend, self.value, new_shape, self.padding)
        n, n_new = len(self.shape), len(new_shape)
        if n_new < n:
            raise ValueError("Cannot broadcast %r to %r" % (self.shape, new_shape))
        m = np.zeros((n, n_new + 1), dtype=np.float64)
        for i in range(n):
            if self.shape[i] == new_shape[i]:
                m[i, i] = 1.0
            elif self.shape[i] != 1:
                raise ValueError("Cannot broadcast %r to %r" % (self.shape, new_shape))
        return self._transform(new_shape, m)

    def reshape(self, new_shape) -> "NonInlineTensor":
        """Tensors:879-888: same buffer, row-major reinterpretation; materialises the receiver."""
        new_shape = [int(s) for s in new_shape]
        if int(np.prod(new_shape, dtype=np.int64)) != self.size:
            raise ValueError("Cannot reshape %r to %r" % (self.shape, new_shape))
        return ProducedTensor(self.backend, self._buffer, new_shape, self.padding)

    def scale(self, new_shape) -> "TransformedTensor":
        """Tensors:950-965: nearest-neighbour resample, coefficient shape_i / newShape_i."""
        new_shape = [int(s) for s in new_shape]
        n = len(self.shape)
        if len(new_shape) != n:
            raise ValueError("scale: rank mismatch")
        m = np.zeros((n, n + 1), dtype=np.float64)
        for i in range(n):
            m[i, i] = float(self.shape[i]) / float(new_shape[i])
        return self._transform(new_shape, m)

    def translate(self, offset, new_shape=None) -> "TransformedTensor":
        """Tensors:970-976: out[g] = in[g - offset], out-of-range reads give the checkpoint's padding."""
        if len(offset) != len(self.shape):
            raise ValueError("translate: offset rank mismatch")
        new_shape = self.shape if new_shape is None else new_shape
        return self._transform(new_shape, ndat.translate([-float(o) for o in offset]))

    def permute(self, dimensions) -> "TransformedTensor":
        """Tensors:1008-1025."""
        n = len(self.shape)
        if len(dimensions) != n:
            raise ValueError("permute: rank mismatch")
        new_shape = [0] * n
        m = np.zeros((n, n + 1), dtype=np.float64)
        for new_d, old_d in enumerate(dimensions):
            new_shape[new_d] = self.shape[old_d]
            m[old_d, new_d] = 1.0
        return self._transform(new_shape, m)

    def transpose(self) -> "TransformedTensor":
        """Tensors:1030."""
        return self.permute(list(range(len(self.shape) - 1, -1, -1)))

    def split(self, dimension: int) -> List["TransformedTensor"]:
        """Tensors:1035-1074: slices along `dimension`, each of rank - 1."""
        n = len(self.shape)
        new_shape = list(self.shape[:dimension]) + list(self.shape[dimension + 1:])
        base = np.zeros((n, n), dtype=np.float64)
        for i in range(dimension):
            base[i, i] = 1.0
        for i in range(dimension + 1, n):
            base[i, i - 1] = 1.0
        out = []
        for index in range(self.shape[dimension]):
            m = base.copy()
            m[dimension, n - 1] = float(index)
            out.append(self._transform(new_shape, m))
        return out

    # -- elementwise operators (Tensors:857-863, :893-945) ------------------------------------------------------------
    def _derived(self, node) -> "InlineTensor":
        return InlineTensor(self.backend, node, list(self.shape), self.padding)

    def _binary(self, op, rhs) -> "InlineTensor":
        rhs = rhs if isinstance(rhs, Tensor) else self.backend.scalar(float(rhs))
        shape = auto_broadcast_shape(self.shape, rhs.shape)
        lhs, rhs = self.broadcast(shape), rhs.broadcast(shape)
        out = lhs._derived(Binary(op, lhs.closure(), rhs.closure()))
        if isinstance(lhs, FillTensor) and not isinstance(rhs, FillTensor):
            out.backend = rhs.backend          # a literal belongs to no device: launch where the data lives
        return out

    def __add__(self, o): return self._binary("add", o)
    def __sub__(self, o): return self._binary("sub", o)
    def __mul__(self, o): return self._binary("mul", o)
    def __truediv__(self, o): return self._binary("div", o)
    def __mod__(self, o): return self._binary("mod", o)
    def __neg__(self): return self._derived(Unary("neg", self.closure()))
    def __pos__(self): return self
    # host sugar: a Python number on the left is `Tensor.scalar(number)` (a FillTensor) combined with the same operator
    def __radd__(self, o): return self.backend.scalar(float(o))._binary("add", self)
    def __rsub__(self, o): return self.backend.scalar(float(o))._binary("sub", self)
    def __rmul__(self, o): return self.backend.scalar(float(o))._binary("mul", self)
    def __rtruediv__(self, o): return self.backend.scalar(float(o))._binary("div", self)


class InlineTensor(Tensor):
    """Tensors:1272-1283.  `doBuffer = enqueueClosure(closure, shape).shared`: the buffer is shared while somebody holds
    it and recomputed afterwards (Compute.scala README:157-166) -- modelled with a weak reference."""

    def __init__(self, backend, node, shape, padding):
        self.backend, self._node, self.shape, self.padding = backend, node, list(shape), float(padding)
        self._weak = None

    def closure(self):
        return self._node

    def _live(self):
        return self._weak() if self._weak is not None else None

    def _buffer(self):
        buf = self._live()
        if buf is None:
            buf = self.backend.launch([self.closure()], self.shape)
            self._weak = weakref.ref(buf)
        return buf

    def _see_through_closure(self):
        """The expression a view of this tensor may be pushed into (trees.see_through), or None when the buffer
        already exists (reading it is cheaper than recomputing)."""
        return self.closure() if self._live() is None else None


class FillTensor(InlineTensor):
    """Tensors:1266-1269."""

    def __init__(self, backend, value, shape, padding=0.0):
        self.value = F32(value)
        super().__init__(backend, Literal(value), shape, padding)


class TransformedTensor(InlineTensor):
    """Tensors:1285-1300."""

    def __init__(self, backend, checkpoint, matrix, shape, padding):
        rows = len(checkpoint.shape)
        self.checkpoint = checkpoint
        self.matrix = np.asarray(matrix, dtype=np.float64).reshape(rows, -1) if rows else np.zeros((0, len(shape) + 1))
        super().__init__(backend, None, shape, padding)

    def closure(self):
        if self._node is None:
            self._node = Xform(self.checkpoint, self.matrix)
        return self._node

    def _see_through_closure(self):
        return None          # a view of a view already composes against the original checkpoint (Tensors:978-1003)


class NonInlineTensor(Tensor):
    """Tensors:1303-1312: closure = Extract(ArrayParameter(id = this))."""
    _ident = None

    def closure(self):
        # one Ident per tensor while any tree still uses it (so `t + t` shares the leaf); held weakly, because a strong
        # tensor <-> Ident cycle would keep device buffers alive until the host's cycle collector runs
        ident = self._ident() if self._ident is not None else None
        if ident is None:
            ident = Ident(self)
            self._ident = weakref.ref(ident)
        return ident


class ProducedTensor(NonInlineTensor):
    """A NonInlineTensor whose buffer comes from a device producer (`.shared` semantics, see InlineTensor)."""

    def __init__(self, backend, produce, shape, padding):
        self.backend, self._produce, self.shape, self.padding = backend, produce, list(shape), float(padding)
        self._weak = None

    def _buffer(self):
        buf = self._weak() if self._weak is not None else None
        if buf is None:
            buf = self._produce()
            self._weak = weakref.ref(buf)
        return buf


class UploadTensor(NonInlineTensor):
    """Tensor.apply (Tensors:445-463): host data, uploaded on first use and then kept on the device."""

    def __init__(self, backend, flat, shape, padding):
        self.backend, self._host, self.shape, self.padding = backend, flat, list(shape), float(padding)
        self._buf = None

    def _buffer(self):
        if self._buf is None:
            self._buf = self.backend.upload(self._host)
            self._host = None
        return self._buf


class CachedTensor(NonInlineTensor):
    """Tensors:602: holds its buffer until released."""

    def __init__(self, backend, buf, shape, padding):
        self.backend, self._buf, self.shape, self.padding = backend, buf, list(shape), float(padding)

    def _buffer(self):
        if self._buf is None:
            raise RuntimeError("CachedTensor used after release")
        return self._buf

    def release(self):
        """monadicClose of the cached buffer (TW:65-72 releases the old weight buffer this way)."""
        self._buf = None


# ----------------------------------------------------------------------------------------------------------------------
# the `gpu` singleton (gpu.scala:15-27): module-level API bound to a lazily created default Backend
# ----------------------------------------------------------------------------------------------------------------------

_default: Optional[Backend] = None


def default_backend() -> Backend:
    global _default
    if _default is None:
        _default = Backend()
    return _default


def set_default_backend(backend: Optional[Backend]) -> None:
    global _default
    _default = backend


def tensor(elements, padding: float = 0.0): return default_backend().tensor(elements, padding)
def scalar(value, padding: float = 0.0): return default_backend().scalar(value, padding)
def fill(value, shape, padding: float = 0.0): return default_backend().fill(value, shape, padding)
def random(shape, seed, padding: float = 0.0): return default_backend().random(shape, seed, padding)
def random_normal(shape, seed, padding: float = 0.0): return default_backend().random_normal(shape, seed, padding)
def join(tensors, dimension=None): return tensors[0].backend.join(tensors, dimension)
def matmul(a, b, **kw): return a.backend.matmul(a, b, **kw)
def sum_axis(t, axis): return t.backend.sum_axis(t, axis)
def abs_(t): return t.backend.abs(t)
def sqrt(t): return t.backend.sqrt(t)
def tanh(t): return t.backend.tanh(t)
def exp(t): return t.backend.exp(t)
def log(t): return t.backend.log(t)
def minimum(a, b): return a.backend.min(a, b)
def maximum(a, b): return a.backend.max(a, b)
def plan_cache_contains(t): return t.backend.plan_cache_contains(t)
def as_tensor(x): return x if isinstance(x, Tensor) else scalar(float(x))
