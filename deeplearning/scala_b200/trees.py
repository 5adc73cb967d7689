"""Closure trees (mirror of Compute.scala `Trees`, Trees.scala:303-1050) and their wire format.

A tree is what a lazy Tensor *is*; `serialize` produces the blob `dls_launch_tree` takes (layout documented in
csrc/plan.h) together with the parameter tensors in first-visit order -- the same alpha-normalised numbering the
reference obtains with `alphaConversion` (Trees.scala:266-268, :807-822), so two structurally equal trees give equal
blobs and share one cached plan (Trees.scala:68-178).
"""
from __future__ import annotations

import struct
from array import array

import numpy as np

W_LITERAL, W_IDENT, W_XFORM = 1, 2, 3
UNARY = {"neg": 10, "abs": 11, "sqrt": 12, "tanh": 13, "exp": 14, "log": 15}
BINARY = {"add": 20, "sub": 21, "mul": 22, "div": 23, "mod": 24, "min": 25, "max": 26}
_MAGIC, _VERSION = 0x54534C44, 1


class Node:
    __slots__ = ()


class Literal(Node):
    """FloatLiteral (Trees.scala:309-330)."""
    __slots__ = ("bits",)

    def __init__(self, value):
        self.bits = int(np.float32(value).view(np.uint32))


class Ident(Node):
    """Extract(ArrayParameter(id = tensor)) (Tensors.scala:1126-1132)."""
    __slots__ = ("tensor", "__weakref__")

    def __init__(self, tensor):
        self.tensor = tensor


class Xform(Node):
    """Extract(Transform(ArrayParameter(id = checkpoint), matrix)) (Tensors.scala:1296-1298)."""
    __slots__ = ("tensor", "matrix")

    def __init__(self, tensor, matrix):
        self.tensor, self.matrix = tensor, matrix


class Unary(Node):
    __slots__ = ("op", "a")

    def __init__(self, op, a):
        self.op, self.a = op, a


class Binary(Node):
    __slots__ = ("op", "a", "b")

    def __init__(self, op, a, b):
        self.op, self.a, self.b = op, a, b


def _is_integral(m) -> bool:
    return bool(np.all(m == np.rint(m))) and bool(np.all(np.abs(m) < 2 ** 31))


def _injective_in_bounds(matrix, out_shape, shape) -> bool:
    """True when `matrix` (rows = dims of the viewed tensor, cols = launch dims + 1) has integer entries, reads every
    launch coordinate exactly once with coefficient 1 (split / permute / transpose, not broadcast or scale) and cannot
    leave [0, shape) anywhere in the launch box -- the views Appendix A.4 of SURVEY.md allows to be pushed through an
    elementwise expression without changing a single result."""
    rows, cols = matrix.shape
    if cols != len(out_shape) + 1 or rows != len(shape) or not _is_integral(matrix):
        return False
    used = [0] * (cols - 1)
    for y in range(rows):
        lo = hi = int(matrix[y, cols - 1])
        for x in range(cols - 1):
            c = int(matrix[y, x])
            if c == 0:
                continue
            if c != 1:
                return False
            used[x] += 1
            hi += max(int(out_shape[x]) - 1, 0)
        if lo < 0 or hi >= int(shape[y]):
            return False
    return all(u == 1 or int(out_shape[x]) == 1 for x, u in enumerate(used))


def _compose(inner, outer):
    """Affine composition: `outer` maps launch coordinates to the viewed tensor's coordinates, `inner` maps those to a
    leaf's coordinates (NDimensionalAffineTransform.preConcatenate, NDAT:38-46)."""
    rows, mid = inner.shape[0], inner.shape[1] - 1
    out = np.zeros((rows, outer.shape[1]), dtype=np.float64)
    if rows:
        out[:, :] = inner[:, :mid] @ outer
        out[:, -1] += inner[:, mid]
    return out


def _push_view(node, matrix, shape, memo):
    """The closure `node` (evaluated at `shape`) re-indexed through `matrix`; None when a leaf cannot be re-indexed
    exactly (non-integer inner matrices would round differently once composed, KB:14-32)."""
    key = id(node)
    if key in memo:
        return memo[key]
    cls = node.__class__
    if cls is Literal:
        out = node
    elif cls is Ident:
        out = Xform(node.tensor, matrix) if list(node.tensor.shape) == list(shape) else None
    elif cls is Xform:
        rows = len(node.tensor.shape)
        inner = np.asarray(node.matrix, dtype=np.float64).reshape(rows, -1) if rows else np.zeros((0, matrix.shape[0] + 1))
        out = Xform(node.tensor, _compose(inner, matrix)) if _is_integral(inner) else None
    elif cls is Unary:
        a = _push_view(node.a, matrix, shape, memo)
        out = Unary(node.op, a) if a is not None else None
    else:
        a = _push_view(node.a, matrix, shape, memo)
        b = _push_view(node.b, matrix, shape, memo) if a is not None else None
        out = Binary(node.op, a, b) if a is not None and b is not None else None
    memo[key] = out
    return out


def see_through(node, out_shape, cache):
    """An `Xform` whose checkpoint is a not-yet-materialised elementwise expression and whose view is an in-bounds
    split / permute: the reference materialises the checkpoint first (Tensors.scala:978-1003 -- 200 MB for the unrolled
    matmul of TensorsSpec:337-344 at MNIST size); here the view is pushed through the expression instead.  Every pushed
    index is provably in range, so no bounds check and no padding can differ."""
    getter = getattr(node.tensor, "_see_through_closure", None)
    if getter is None:
        return None
    closure = getter()
    if closure is None:
        return None
    rows = len(node.tensor.shape)
    if rows == 0:
        return None                      # a broadcast scalar: nothing to gain, and it is never bounds-checked (KB:351-356)
    matrix = np.asarray(node.matrix, dtype=np.float64).reshape(rows, -1)
    if not _injective_in_bounds(matrix, out_shape, node.tensor.shape):
        return None
    memo = cache.setdefault((id(closure), matrix.tobytes()), {})
    return _push_view(closure, matrix, node.tensor.shape, memo)


def expand_views(roots, out_shape):
    """Roots with every see-through view pushed into its checkpoint's expression (iterative, sharing-preserving)."""
    done, cache = {}, {}

    def rewrite(root):
        stack = [root]
        while stack:
            node = stack[-1]
            key = id(node)
            if key in done:
                stack.pop()
                continue
            cls = node.__class__
            if cls is Literal or cls is Ident:
                done[key] = node
            elif cls is Xform:
                pushed = see_through(node, out_shape, cache)
                if pushed is None:
                    done[key] = node
                elif id(pushed) in done:
                    done[key] = done[id(pushed)]
                else:
                    # the pushed expression may itself contain views of inline tensors: rewrite it, then alias
                    if ("v", key) not in done:
                        done[("v", key)] = pushed
                        stack.append(pushed)
                        continue
                    done[key] = done[id(pushed)]
            elif cls is Unary:
                if id(node.a) not in done:
                    stack.append(node.a)
                    continue
                a = done[id(node.a)]
                done[key] = node if a is node.a else Unary(node.op, a)
            else:
                if id(node.a) not in done:
                    stack.append(node.a)
                    continue
                if id(node.b) not in done:
                    stack.append(node.b)
                    continue
                a, b = done[id(node.a)], done[id(node.b)]
                done[key] = node if (a is node.a and b is node.b) else Binary(node.op, a, b)
            if key in done:
                stack.pop()
        return done[id(root)]
    return [rewrite(r) for r in roots], cache


def serialize(roots, out_shape):
    """Returns (blob: bytes, params: list of tensors in first-visit order)."""
    roots, _keep_alive = expand_views(roots, out_shape)
    index = {}
    params, pindex = [], {}
    body = array("I")
    n_nodes = 0

    def param_of(t):
        k = id(t)
        i = pindex.get(k)
        if i is None:
            i = pindex[k] = len(params)
            params.append(t)
        return i

    for root in roots:
        stack = [(root, False)]
        while stack:
            node, expanded = stack.pop()
            key = id(node)
            if key in index:
                continue
            cls = node.__class__
            if cls is Literal:
                body.extend((W_LITERAL, node.bits))
            elif cls is Ident:
                body.extend((W_IDENT, param_of(node.tensor)))
            elif cls is Xform:
                m = np.ascontiguousarray(node.matrix, dtype=np.float64)
                rows = len(node.tensor.shape)
                cols = m.size // rows if rows else len(out_shape) + 1
                body.extend((W_XFORM, param_of(node.tensor), rows, cols))
                if rows:
                    body.frombytes(m.tobytes())
            elif cls is Unary:
                ia = index.get(id(node.a))
                if ia is None:
                    stack.append((node, True))
                    stack.append((node.a, False))
                    continue
                body.extend((UNARY[node.op], ia))
            else:
                ia, ib = index.get(id(node.a)), index.get(id(node.b))
                if ia is None or ib is None:
                    stack.append((node, True))
                    if ib is None:
                        stack.append((node.b, False))
                    if ia is None:
                        stack.append((node.a, False))     # popped first: left operand visited first
                    continue
                body.extend((BINARY[node.op], ia, ib))
            index[key] = n_nodes
            n_nodes += 1

    head = array("I", (_MAGIC, _VERSION, len(params), n_nodes, len(roots), len(out_shape)))
    head.extend(int(s) for s in out_shape)
    for t in params:
        head.append(len(t.shape))
        head.extend(int(s) for s in t.shape)
        head.frombytes(struct.pack("<f", float(t.padding)))
    tail = array("I", (index[id(r)] for r in roots))
    return head.tobytes() + body.tobytes() + tail.tobytes(), params
