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


def serialize(roots, out_shape):
    """Returns (blob: bytes, params: list of tensors in first-visit order)."""
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
