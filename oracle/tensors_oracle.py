"""CPU ORACLE (test infrastructure, NOT product code).

A numpy restatement of the Compute.scala lazy `Tensor` semantics that the
DeepLearning.scala tensor plugins sit on.  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s cpu_baseline / `--impl reference` leg may import this module; the
product package `deeplearning.scala_b200` never does.

Parity status: PINNED against the reference's own golden vectors (G1-G17 of
SURVEY.md section 8c, transcribed in tests/golden/reference_vectors.json and
checked by tests/test_oracle_golden.py).  The reference itself cannot run in the
build container (no JVM, no OpenCL ICD), so there is no oracle/_ref build.

Arithmetic that lives outside /root/reference: `tanh exp log sqrt fmin fmax fabs`
and float `+ - * /` are whatever the OpenCL driver's compiler + libm produce
(unpinned; CI used AMD APP SDK 3.0, `.travis.yml:21`) under
`-cl-unsafe-math-optimizations` (`cpu.scala:105`).  This oracle defines each
operator as the IEEE-754 fp32 operation numpy performs (one rounding per node,
no FMA contraction), `%` as C `fmodf`.

Every function cites the reference file:line it follows (paths relative to
/root/reference/Compute.scala unless prefixed).
"""
from __future__ import annotations

import decimal
import math
from typing import List, Optional, Sequence

import numpy as np

F32 = np.float32

# --------------------------------------------------------------------------------------
# NDimensionalAffineTransform  (NDimensionalAffineTransform/src/main/scala/.../NDimensionalAffineTransform.scala)
# --------------------------------------------------------------------------------------


def ndat_translate(offset: Sequence[float]) -> np.ndarray:
    """NDAT:13-26 -- identity with last column = offset; shape (n, n+1)."""
    n = len(offset)
    m = np.zeros((n, n + 1), dtype=np.float64)
    for i in range(n):
        m[i, i] = 1.0
        m[i, n] = offset[i]
    return m


def _concatenate(m01: np.ndarray, m12: np.ndarray) -> np.ndarray:
    """NDAT:48-95.  m01 maps space0 -> space1 (rows = len1, cols = len0+1);
    m12 maps space1 -> space2 (rows = len2, cols = len1+1); result maps 0 -> 2.
    The accumulation order (index1 ascending, offset seeded first) is the reference's."""
    len1, c0 = m01.shape
    len0 = c0 - 1
    len2 = m12.shape[0]
    assert m12.shape[1] == len1 + 1
    out = np.zeros((len2, len0 + 1), dtype=np.float64)
    for i2 in range(len2):
        for i0 in range(len0):
            acc = 0.0
            for i1 in range(len1):
                acc = acc + m12[i2, i1] * m01[i1, i0]
            out[i2, i0] = acc
        acc = m12[i2, len1]
        for i1 in range(len1):
            acc = acc + m12[i2, i1] * m01[i1, len0]
        out[i2, len0] = acc
    return out


def ndat_pre_concatenate(matrix01: np.ndarray, matrix12: np.ndarray) -> np.ndarray:
    """NDAT:29-35 (`preConcatenate(matrix01, matrix12, length0)`)."""
    return _concatenate(matrix01, matrix12)


def ndat_concatenate(matrix12: np.ndarray, matrix01: np.ndarray) -> np.ndarray:
    """NDAT:38-46 (`concatenate(matrix12, matrix01, length2)`)."""
    return _concatenate(matrix01, matrix12)


def ndat_transform(matrix: np.ndarray, source: Sequence[float]) -> np.ndarray:
    """NDAT:97-127."""
    out = np.zeros(matrix.shape[0], dtype=np.float64)
    n = len(source)
    for y in range(matrix.shape[0]):
        acc = matrix[y, n]
        for x in range(n):
            acc = acc + matrix[y, x] * source[x]
        out[y] = acc
    return out


# --------------------------------------------------------------------------------------
# Literal formatting of matrix coefficients (OpenCLKernelBuilder.scala:14-32)
# --------------------------------------------------------------------------------------


def cl_matrix_literal(value: float):
    """KB:14-32 -- java.text.DecimalFormat, <=3 fraction digits, HALF_EVEN on the exact
    binary value, no grouping.  Returns (python int) when the printed literal has no
    fraction part (C int literal => int arithmetic) else (python float)."""
    q = decimal.Decimal(float(value)).quantize(decimal.Decimal("0.001"), rounding=decimal.ROUND_HALF_EVEN)
    if q == q.to_integral_value():
        return int(q)
    return float(q)


# --------------------------------------------------------------------------------------
# RNG  (Tensors.scala:98-117, :398-443)
# --------------------------------------------------------------------------------------


def wang_hash(v: np.ndarray) -> np.ndarray:
    """Tensors:107-116, uint32 wrap-around."""
    v = v.astype(np.uint32)
    with np.errstate(over="ignore"):
        v = (v ^ np.uint32(61)) ^ (v >> np.uint32(16))
        v = v * np.uint32(9)
        v = v ^ (v << np.uint32(4))
        v = v * np.uint32(0x27D4EB2D)
        v = v ^ (v >> np.uint32(15))
    return v


def lcg_hash(v: np.ndarray) -> np.ndarray:
    """Tensors:99-103."""
    with np.errstate(over="ignore"):
        return (np.uint32(1103515245) * v.astype(np.uint32) + np.uint32(12345)).astype(np.uint32)


def xorshift(seed: np.ndarray) -> np.ndarray:
    """Tensors:403-407."""
    seed = seed.astype(np.uint32)
    t1 = seed ^ (seed << np.uint32(13))
    t2 = t1 ^ (t1 >> np.uint32(17))
    return t2 ^ (t2 << np.uint32(5))


def random_uniform_flat(n: int, seed: int, hash_fn=wang_hash) -> np.ndarray:
    """Tensors:436-439: buffer[i] = hash(i ^ seed) / 4294967296.0f (uint->float RNE, then fp32 divide)."""
    i = np.arange(n, dtype=np.uint32)
    h = hash_fn(i ^ np.uint32(seed & 0xFFFFFFFF))
    return h.astype(F32) / F32(4294967296.0)


def random_normal_flat(n: int, seed: int, hash_fn=wang_hash) -> np.ndarray:
    """Tensors:409-425 Box-Muller; work-item i writes 2i, 2i+1; odd sizes pad by one (:508-518)."""
    padded = n + (n & 1)
    i = np.arange(padded // 2, dtype=np.uint32)
    r1 = hash_fn(i ^ np.uint32(seed & 0xFFFFFFFF))
    r2 = xorshift(r1)
    u1 = r1.astype(F32) / F32(4294967296.0)
    u2 = r2.astype(F32) / F32(4294967296.0)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.sqrt(F32(-2.0) * np.log(u1)).astype(F32)
        theta = (F32(2.0) * F32(math.pi)) * u2
        z0 = (r * np.cos(theta).astype(F32)).astype(F32)
        z1 = (r * np.sin(theta).astype(F32)).astype(F32)
    out = np.empty(padded, dtype=F32)
    out[0::2] = z0
    out[1::2] = z1
    return out[:n]


# --------------------------------------------------------------------------------------
# Expression tree (Trees.scala node kinds; evaluation follows OpenCLKernelBuilder.scala)
# --------------------------------------------------------------------------------------


class Node:
    __slots__ = ()


class Literal(Node):
    __slots__ = ("value",)

    def __init__(self, value):
        self.value = F32(value)


class Ident(Node):
    """Extract(ArrayParameter(id=tensor)) -- Tensors:1126-1132, KB:435-455."""
    __slots__ = ("tensor",)

    def __init__(self, tensor):
        self.tensor = tensor


class Xform(Node):
    """Extract(Transform(ArrayParameter(id=tensor), matrix)) -- Tensors:1296-1298, KB:348-411."""
    __slots__ = ("tensor", "matrix")

    def __init__(self, tensor, matrix):
        self.tensor = tensor
        self.matrix = matrix


class Unary(Node):
    __slots__ = ("op", "a")

    def __init__(self, op, a):
        self.op, self.a = op, a


class Binary(Node):
    __slots__ = ("op", "a", "b")

    def __init__(self, op, a, b):
        self.op, self.a, self.b = op, a, b


def _fmod(a, b):
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.fmod(a, b).astype(F32)


def _fmin(a, b):
    """C fmin (KB:271-280) with -0 < +0 (IEEE 754-2019 minimum); the reference leaves the zero-sign case to its libm."""
    out = np.fmin(a, b).astype(F32)
    both_zero = (a == 0) & (b == 0)
    return np.where(both_zero & (np.signbit(a) | np.signbit(b)), F32(-0.0), out).astype(F32)


def _fmax(a, b):
    out = np.fmax(a, b).astype(F32)
    both_zero = (a == 0) & (b == 0)
    return np.where(both_zero & ~(np.signbit(a) & np.signbit(b)), F32(0.0), out).astype(F32)


_UNARY = {
    "neg": lambda a: (-a).astype(F32),
    "abs": lambda a: np.abs(a).astype(F32),
    "sqrt": lambda a: np.sqrt(a).astype(F32),
    "tanh": lambda a: np.tanh(a.astype(np.float64)).astype(F32),   # correctly-rounded fp32 of the fp64 result
    "exp": lambda a: np.exp(a.astype(np.float64)).astype(F32),
    "log": lambda a: np.log(a.astype(np.float64)).astype(F32),
}
_BINARY = {
    "add": lambda a, b: (a + b).astype(F32),
    "sub": lambda a, b: (a - b).astype(F32),
    "mul": lambda a, b: (a * b).astype(F32),
    "div": lambda a, b: (a / b).astype(F32),
    "mod": _fmod,
    "min": _fmin,
    "max": _fmax,
}


def _affine_index(row: np.ndarray, grids: List[np.ndarray]):
    """KB:363-390 for one source row.  Returns (index array | None, checked: bool)."""
    ncols = row.shape[0]
    acc = None
    for x in range(ncols):
        scale = row[x]
        if scale == 0.0:
            continue
        if x < ncols - 1:
            if scale == 1.0:
                term = grids[x]
            else:
                term = grids[x] * cl_matrix_literal(scale)
        else:
            lit = cl_matrix_literal(scale)
            term = np.int64(lit) if isinstance(lit, int) else np.float64(lit)
        acc = term if acc is None else acc + term          # C left-to-right, int/double promotion
    if acc is None:
        return None, False                                  # `const int index = 0;` no bounds (KB:383-384)
    acc = np.asarray(acc)
    if acc.dtype.kind == "f":
        acc = np.trunc(acc).astype(np.int64)               # (int)(double) truncates toward zero (KB:386)
    else:
        acc = acc.astype(np.int64)
    return acc, True


def evaluate(node: Node, launch_shape: Sequence[int], memo: Optional[dict] = None) -> np.ndarray:
    """Evaluate one closure tree for every coordinate of `launch_shape`
    (one OpenCL work-item per output element, KB:135-221).  Each tree node is
    computed once (ExportContext identity CSE, Trees.scala:220)."""
    if memo is None:
        memo = {}
    key = id(node)
    if key in memo:
        return memo[key]
    S = tuple(int(s) for s in launch_shape)
    r = len(S)
    if isinstance(node, Literal):
        out = np.full(S, node.value, dtype=F32)
    elif isinstance(node, Ident):
        t = node.tensor
        buf = t.materialise()
        if len(t.shape) == 0:
            # rank-0 parameter: no `[__global_id]`, no bounds (KB:436-451 with empty shape)
            out = np.full(S, buf.reshape(()), dtype=F32)
        else:
            if len(t.shape) != r:
                raise ValueError("identity leaf rank != launch rank (reference would fail to compile)")
            out = np.full(S, F32(t.padding), dtype=F32)
            sl = tuple(slice(0, min(a, b)) for a, b in zip(S, t.shape))
            out[sl] = buf[sl]                               # only the upper bound is checked (KB:440-442)
    elif isinstance(node, Xform):
        t = node.tensor
        buf = t.materialise()
        rows = len(t.shape)
        if rows == 0:
            out = np.full(S, buf.reshape(()), dtype=F32)    # KB:351-356: `*p`, never checked
        else:
            m = np.asarray(node.matrix, dtype=np.float64).reshape(rows, -1)
            if m.shape[1] != r + 1:
                raise ValueError("matrix columns != launch rank + 1")
            grids = [g.astype(np.int64) for g in np.indices(S)] if r > 0 else []
            ok = np.ones(S, dtype=bool)
            idx = []
            for y in range(rows):
                iy, checked = _affine_index(m[y], grids)
                if not checked:
                    idx.append(np.zeros(S, dtype=np.int64))
                    continue
                iy = np.broadcast_to(iy, S)
                ok &= (iy >= 0) & (iy < t.shape[y])
                idx.append(iy)
            safe = tuple(np.where(ok, i, 0) for i in idx)
            out = np.where(ok, buf[safe], F32(t.padding)).astype(F32)
    elif isinstance(node, Unary):
        with np.errstate(all="ignore"):
            out = _UNARY[node.op](evaluate(node.a, S, memo))
    elif isinstance(node, Binary):
        a = evaluate(node.a, S, memo)
        b = evaluate(node.b, S, memo)
        with np.errstate(all="ignore"):
            out = _BINARY[node.op](a, b)
    else:
        raise TypeError(node)
    memo[key] = out
    return out


# --------------------------------------------------------------------------------------
# Reductions (Tensors.scala:308-393, :673-766)
# --------------------------------------------------------------------------------------


def sum_cpu_order(flat: np.ndarray) -> F32:
    """Tensors:313-351 with global size 1: 16 strided lanes, hi/lo folds, then the tail."""
    flat = np.ascontiguousarray(flat, dtype=F32).ravel()
    n = flat.size
    vlen = n // 16
    if vlen >= 1:
        body = flat[: 16 * vlen].reshape(vlen, 16)
        acc = body[0].copy()
        for k in range(1, vlen):
            acc = (acc + body[k]).astype(F32)
        f8 = (acc[8:] + acc[:8]).astype(F32)               # accumulator.hi + accumulator.lo
        f4 = (f8[4:] + f8[:4]).astype(F32)
        f2 = (f4[2:] + f4[:2]).astype(F32)
        total = F32(f2[0] + f2[1])
        for v in flat[16 * vlen:]:
            total = F32(total + v)
        return total
    total = F32(0.0)
    for v in flat:
        total = F32(total + v)
    return total


def sum_gpu_order(flat: np.ndarray, work_group_size: int = 256, max_compute_units: int = 148) -> F32:
    """Tensors:358-392 + driver :697-754 (two-stage parallel reduce; local size must be a
    power of two >= 2, Appendix B)."""
    flat = np.ascontiguousarray(flat, dtype=F32).ravel()

    def stage(buf, local, groups):
        gsz = local * groups
        res = np.zeros(groups, dtype=F32)
        for g in range(groups):
            scratch = np.zeros(local, dtype=F32)
            for l in range(local):
                acc = F32(0.0)
                for v in buf[g * local + l:: gsz]:
                    acc = F32(acc + v)
                scratch[l] = acc
            off = local // 2
            while off > 1:
                scratch[:off] = (scratch[:off] + scratch[off: 2 * off]).astype(F32)
                off //= 2
            res[g] = F32(scratch[0] + scratch[1]) if local > 1 else scratch[0]
        return res

    n = flat.size
    local = min(n, work_group_size)
    groups = min(n // local + (1 if n % local else 0), max_compute_units)
    if groups == 1:
        return stage(flat, local, 1)[0]
    partial = stage(flat, local, groups)
    return stage(partial, min(groups, work_group_size), 1)[0]


# --------------------------------------------------------------------------------------
# Tensor classes (Tensors.scala:636-1133, :1266-1312)
# --------------------------------------------------------------------------------------


def auto_broadcast_shape(s1, s2):
    """Tensors:208-222 -- LEFT-aligned."""
    out = []
    for i in range(max(len(s1), len(s2))):
        if i >= len(s1) or s1[i] == 1:
            out.append(s2[i])
        elif i >= len(s2) or s2[i] == 1:
            out.append(s1[i])
        elif s1[i] == s2[i]:
            out.append(s1[i])
        else:
            raise ValueError("Failed to automatically broadcast between shape %r and %r" % (s1, s2))
    return out


def _nested_shape(a):
    """TensorBuilder.shape (Tensors:78-91): ragged input -> IllegalArgumentException."""
    if isinstance(a, (list, tuple, np.ndarray)) and not (isinstance(a, np.ndarray) and a.ndim == 0):
        if len(a) == 0:
            return [0]
        subs = [_nested_shape(x) for x in a]
        for s in subs[1:]:
            if s != subs[0]:
                raise ValueError("ragged nested sequence")
        return [len(a)] + subs[0]
    return []


def _nested_flatten(a, out):
    if isinstance(a, (list, tuple, np.ndarray)) and not (isinstance(a, np.ndarray) and a.ndim == 0):
        for x in a:
            _nested_flatten(x, out)
    else:
        out.append(float(a))


class Tensor:
    """sealed trait Tensor (Tensors:636)."""
    shape: List[int]
    padding: float

    # -- closure / buffer -------------------------------------------------------------
    def closure(self) -> Node:
        raise NotImplementedError

    def materialise(self) -> np.ndarray:
        """doBuffer: the row-major buffer of `shape` (memoised == `.shared`)."""
        raise NotImplementedError

    # -- slow actions -----------------------------------------------------------------
    def flat_array(self) -> np.ndarray:
        """Tensors:1099-1118."""
        return np.ascontiguousarray(self.materialise(), dtype=F32).reshape(-1)[: int(np.prod(self.shape, dtype=np.int64))]

    def do_cache(self) -> "CachedTensor":
        """Tensors:642-666."""
        return CachedTensor(self.materialise(), list(self.shape), self.padding)

    def non_inline(self) -> "NonInlineTensor":
        return BufferTensor(lambda: self.materialise(), list(self.shape), self.padding)

    def sum(self, order: str = "cpu") -> "NonInlineTensor":
        """Tensors:771 -> reduce(PlusPrograms) :673-766; scalar NonInlineTensor, padding inherited."""
        fn = sum_cpu_order if order == "cpu" else sum_gpu_order
        return BufferTensor(lambda: np.array(fn(self.flat_array()), dtype=F32).reshape(()), [], self.padding)

    def to_string(self) -> str:
        """Tensors:776-811 with Scala Float.toString element formatting."""
        return format_tensor(self.shape, self.flat_array())

    __str__ = to_string

    # -- views ------------------------------------------------------------------------
    def _transform(self, new_shape, matrix1) -> "TransformedTensor":
        """Tensors:978-1003: the checkpoint rule."""
        if isinstance(self, TransformedTensor):
            m = ndat_pre_concatenate(np.asarray(matrix1, dtype=np.float64), self.matrix)
            return TransformedTensor(self.checkpoint, m, list(new_shape), self.padding)
        return TransformedTensor(self, np.asarray(matrix1, dtype=np.float64), list(new_shape), None)

    def broadcast(self, new_shape) -> "Tensor":
        """Tensors:816-855."""
        new_shape = list(new_shape)
        if new_shape == list(self.shape):
            return self
        if isinstance(self, FillTensor):
            return FillTensor(self.value, new_shape, self.padding)
        n_new, n = len(new_shape), len(self.shape)
        m = np.zeros((n, n_new + 1), dtype=np.float64)
        for i in range(n):
            if self.shape[i] == new_shape[i]:      # IndexError if n_new < n, like the reference's AIOOBE
                m[i, i] = 1.0
            elif self.shape[i] == 1:
                pass
            else:
                raise ValueError("Cannot broadcast %r to %r" % (self.shape, new_shape))
        return self._transform(new_shape, m)

    def reshape(self, new_shape) -> "NonInlineTensor":
        """Tensors:879-888: same buffer, new shape (materialises the receiver)."""
        new_shape = list(new_shape)
        if int(np.prod(new_shape, dtype=np.int64)) != int(np.prod(self.shape, dtype=np.int64)):
            raise ValueError("reshape size mismatch")
        return BufferTensor(lambda: self.flat_array().reshape(new_shape), new_shape, self.padding)

    def scale(self, new_shape) -> "TransformedTensor":
        """Tensors:950-965."""
        n = len(new_shape)
        if n != len(self.shape):
            raise ValueError
        m = np.zeros((n, n + 1), dtype=np.float64)
        for i in range(n):
            m[i, i] = float(self.shape[i]) / new_shape[i]
        return self._transform(new_shape, m)

    def translate(self, offset, new_shape=None) -> "TransformedTensor":
        """Tensors:970-976."""
        if len(offset) != len(self.shape):
            raise ValueError
        if new_shape is None:
            new_shape = self.shape
        return self._transform(new_shape, ndat_translate([-float(o) for o in offset]))

    def permute(self, dimensions) -> "TransformedTensor":
        """Tensors:1008-1025."""
        n = len(self.shape)
        if len(dimensions) != n:
            raise ValueError
        new_shape = [0] * n
        m = np.zeros((n, n + 1), dtype=np.float64)
        for new_d in range(n):
            old_d = dimensions[new_d]
            new_shape[new_d] = self.shape[old_d]
            m[old_d, new_d] = 1.0
        return self._transform(new_shape, m)

    def transpose(self) -> "TransformedTensor":
        """Tensors:1030."""
        return self.permute(list(reversed(range(len(self.shape)))))

    def split(self, dimension) -> List["TransformedTensor"]:
        """Tensors:1035-1074."""
        n = len(self.shape)
        new_shape = list(self.shape[:dimension]) + list(self.shape[dimension + 1:])
        out = []
        for index in range(self.shape[dimension]):
            m = np.zeros((n, n), dtype=np.float64)
            for i in range(dimension):
                m[i, i] = 1.0
            m[dimension, n - 1] = float(index)
            for i in range(dimension + 1, n):
                m[i, i - 1] = 1.0
            out.append(self._transform(new_shape, m))
        return out

    # -- elementwise operators (Tensors:857-863, :893-945, :526-558) ---------------------
    def _derived(self, node) -> "InlineTensor":
        return InlineTensor(node, list(self.shape), self.padding)

    def _binary(self, op, rhs) -> "InlineTensor":
        rhs = as_tensor(rhs)
        shape = auto_broadcast_shape(self.shape, rhs.shape)
        l, r = self.broadcast(shape), rhs.broadcast(shape)
        return l._derived(Binary(op, l.closure(), r.closure()))

    def __add__(self, o): return self._binary("add", o)
    def __sub__(self, o): return self._binary("sub", o)
    def __mul__(self, o): return self._binary("mul", o)
    def __truediv__(self, o): return self._binary("div", o)
    def __mod__(self, o): return self._binary("mod", o)
    def __neg__(self): return self._derived(Unary("neg", self.closure()))
    def __pos__(self): return self


def as_tensor(x) -> Tensor:
    return x if isinstance(x, Tensor) else scalar(float(x))


class InlineTensor(Tensor):
    """Tensors:1272-1283."""

    def __init__(self, node, shape, padding):
        self._node, self.shape, self.padding = node, list(shape), float(padding)
        self._buf = None

    def closure(self):
        return self._node

    def materialise(self):
        if self._buf is None:
            self._buf = evaluate(self.closure(), self.shape)
        return self._buf


class FillTensor(InlineTensor):
    """Tensors:1266-1269."""

    def __init__(self, value, shape, padding=0.0):
        self.value = F32(value)
        super().__init__(Literal(value), shape, padding)


class TransformedTensor(InlineTensor):
    """Tensors:1285-1300."""

    def __init__(self, checkpoint, matrix, shape, padding):
        self.checkpoint = checkpoint
        self.matrix = np.asarray(matrix, dtype=np.float64).reshape(len(checkpoint.shape), -1) \
            if len(checkpoint.shape) else np.zeros((0, len(shape) + 1))
        pad = checkpoint.padding if padding is None else padding
        Tensor.__init__(self)
        self._node = None
        self.shape, self.padding, self._buf = list(shape), float(pad), None

    def closure(self):
        if self._node is None:
            self._node = Xform(self.checkpoint, self.matrix)
        return self._node


class NonInlineTensor(Tensor):
    """Tensors:1303-1312."""
    _ident = None

    def closure(self):
        if self._ident is None:
            self._ident = Ident(self)
        return self._ident


class BufferTensor(NonInlineTensor):
    def __init__(self, thunk, shape, padding):
        self._thunk, self.shape, self.padding, self._buf = thunk, list(shape), float(padding), None

    def materialise(self):
        if self._buf is None:
            self._buf = np.asarray(self._thunk(), dtype=F32).reshape(self.shape)
        return self._buf


class CachedTensor(BufferTensor):
    """Tensors:602."""

    def __init__(self, buf, shape, padding):
        super().__init__(lambda: buf, shape, padding)


# -- object Tensor (Tensors:395-600) ------------------------------------------------------


def tensor(elements, padding: float = 0.0) -> NonInlineTensor:
    """Tensor.apply (Tensors:445-463)."""
    shape = _nested_shape(elements)
    flat: List[float] = []
    _nested_flatten(elements, flat)
    data = np.array(flat, dtype=F32).reshape(shape)
    return BufferTensor(lambda: data, shape, padding)


def scalar(value, padding=0.0) -> FillTensor:
    return FillTensor(value, [], padding)


def fill(value, shape, padding=0.0) -> FillTensor:
    return FillTensor(value, list(shape), padding)


def random(shape, seed, padding=0.0, hash_fn=wang_hash) -> NonInlineTensor:
    n = int(np.prod(shape, dtype=np.int64))
    return BufferTensor(lambda: random_uniform_flat(n, seed, hash_fn).reshape(shape), list(shape), padding)


def random_normal(shape, seed, padding=0.0, hash_fn=wang_hash) -> NonInlineTensor:
    n = int(np.prod(shape, dtype=np.int64))
    return BufferTensor(lambda: random_normal_flat(n, seed, hash_fn).reshape(shape), list(shape), padding)


def _unary_fn(op):
    def f(t: Tensor) -> InlineTensor:
        return t._derived(Unary(op, t.closure()))
    return f


abs_ = _unary_fn("abs")
sqrt = _unary_fn("sqrt")
tanh = _unary_fn("tanh")
exp = _unary_fn("exp")
log = _unary_fn("log")


def minimum(a: Tensor, b: Tensor) -> InlineTensor:
    return a._binary("min", b)


def maximum(a: Tensor, b: Tensor) -> InlineTensor:
    return a._binary("max", b)


def round_operand(x: np.ndarray, kind: Optional[str]) -> np.ndarray:
    """What the tensor cores see of an fp32 operand: bf16 = round to nearest even to 8 significand bits (the staging
    kernel's cvt.rn.bf16.f32); tf32 = round to nearest, ties away, to 11 significand bits (the staging kernel's
    cvt.rna.tf32.f32; the tensor core itself ignores the low 13 mantissa bits, which are zero after staging)."""
    x = np.ascontiguousarray(x, dtype=F32)
    if kind is None or kind == "fp32":
        return x
    bits = x.view(np.uint32).astype(np.uint64)
    if kind == "bf16":
        rounded = (bits + 0x7FFF + ((bits >> 16) & 1)) & 0xFFFF0000
        out = rounded.astype(np.uint32).view(F32)
        return np.where(np.isnan(x), x, out).astype(F32)
    if kind == "tf32":
        out = ((bits + 0x1000) & 0xFFFFE000).astype(np.uint32).view(F32)
        return np.where(np.isfinite(x), out, x).astype(F32)
    raise ValueError(kind)


def matmul(a: Tensor, b: Tensor, transpose_a: bool = False, transpose_b: bool = False, bias: Optional[Tensor] = None,
           bias_mode: str = "none", activation: str = "none", operands: Optional[str] = None) -> NonInlineTensor:
    """The contraction the reference spells as a host-unrolled chain (TensorsSpec:337-344): out[i,k] =
    ((a[i,0]*b[0,k] + a[i,1]*b[1,k]) + a[i,2]*b[2,k]) + ..., one fp32 rounding per node, j ascending (Scala `reduce` on
    an IndexedSeq is a left fold).  `operands` optionally rounds the inputs the way a tensor-core mode reads them."""
    def thunk():
        A = a.materialise().reshape(a.shape).astype(F32)
        B = b.materialise().reshape(b.shape).astype(F32)
        if transpose_a:
            A = A.T
        if transpose_b:
            B = B.T
        A, B = round_operand(A, operands), round_operand(B, operands)
        m, k = A.shape
        n = B.shape[1]
        acc = np.zeros((m, n), dtype=F32)
        for j in range(k):
            term = (A[:, j:j + 1] * B[j:j + 1, :]).astype(F32)
            acc = term if j == 0 else (acc + term).astype(F32)
        if bias_mode == "row":
            acc = (acc + bias.materialise().astype(F32).reshape(-1, 1)).astype(F32)
        elif bias_mode == "col":
            acc = (acc + bias.materialise().astype(F32).reshape(1, -1)).astype(F32)
        if activation == "tanh":
            acc = np.tanh(acc.astype(np.float64)).astype(F32)
        elif activation == "relu":
            acc = _fmax(acc, np.zeros_like(acc))
        return acc.reshape(-1)
    m = a.shape[1] if transpose_a else a.shape[0]
    n = b.shape[0] if transpose_b else b.shape[1]
    return BufferTensor(thunk, [m, n], a.padding)


def sum_axis(t: Tensor, axis: int) -> NonInlineTensor:
    """`t.split(axis).reduce(_ + _)` (TL:124-162): left fold over the slices, one fp32 rounding per addition."""
    shape = list(t.shape)

    def thunk():
        a = t.materialise().reshape(shape).astype(F32)
        a = np.moveaxis(a, axis, 0)
        acc = a[0].copy() if a.shape[0] else np.zeros(a.shape[1:], dtype=F32)
        for j in range(1, a.shape[0]):
            acc = (acc + a[j]).astype(F32)
        return acc.reshape(-1)
    return BufferTensor(thunk, shape[:axis] + shape[axis + 1:], t.padding)


def join(tensors: Sequence[Tensor], dimension: Optional[int] = None) -> Tensor:
    """Tensors:560-598: evaluate every element at the HEAD tensor's shape, write them as the new last
    dimension (KB:632-652), then optionally permute it to `dimension`."""
    tensors = list(tensors)
    head = tensors[0]
    shape = list(head.shape) + [len(tensors)]

    def thunk():
        memo: dict = {}
        cols = [evaluate(t.closure(), head.shape, memo) for t in tensors]
        return np.stack(cols, axis=-1)

    joined = BufferTensor(thunk, shape, head.padding)
    if dimension is None or len(shape) - 1 == dimension:
        return joined
    n = len(shape)
    perm = [i if i < dimension else (n - 1 if i == dimension else i - 1) for i in range(n)]
    return joined.permute(perm)


# -- formatting ---------------------------------------------------------------------------


def java_float_to_string(x) -> str:
    """Scala/Java Float.toString: shortest digits that round-trip, >=1 fraction digit, scientific
    outside [1e-3, 1e7)."""
    x = F32(x)
    if np.isnan(x):
        return "NaN"
    if np.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0:
        return "-0.0" if np.signbit(x) else "0.0"
    ax = abs(float(x))
    if 1e-3 <= ax < 1e7:
        s = np.format_float_positional(x, unique=True, trim="0")
        if s.endswith("."):
            s += "0"
        return s
    s = np.format_float_scientific(x, unique=True, trim="0", exp_digits=1)
    mant, exp = s.split("e")
    if "." not in mant:
        mant += ".0"
    elif mant.endswith("."):
        mant += "0"
    return "%sE%d" % (mant, int(exp))


def format_tensor(shape, flat) -> str:
    shape = list(shape)
    if not shape:
        if len(flat) != 1:
            raise ValueError("shape() does not match the data size")
        return java_float_to_string(flat[0])
    if len(shape) == 1:
        if shape[0] != len(flat):
            raise ValueError
        return "[" + ",".join(java_float_to_string(v) for v in flat) + "]"
    group = len(flat) // shape[0]
    return "[" + ",".join(format_tensor(shape[1:], flat[i * group:(i + 1) * group]) for i in range(shape[0])) + "]"
