"""N-dimensional affine transforms used by every Tensor view (host side, pure numpy).

Mirror of Compute.scala's NDimensionalAffineTransform (NDAT =
Compute.scala/NDimensionalAffineTransform/src/main/scala/com/thoughtworks/compute/NDimensionalAffineTransform.scala).
A matrix maps view coordinates to checkpoint coordinates: rows = rank(checkpoint), cols = rank(view) + 1, the last
column is the offset (Tensors.scala:1289-1293).
"""
import numpy as np


_TRANSLATE = {}
_COMPOSE = {}


def translate(offset):
    """NDAT:13-26 -- identity plus an offset column.  (Matrices are immutable values: results are memoised, a training
    step rebuilds the same few hundred views every iteration.)"""
    key = tuple(float(o) for o in offset)
    m = _TRANSLATE.get(key)
    if m is None:
        if len(_TRANSLATE) > 4096:
            _TRANSLATE.clear()
        m = _TRANSLATE[key] = _translate(key)
        m.setflags(write=False)
    return m


def _translate(offset):
    n = len(offset)
    m = np.zeros((n, n + 1), dtype=np.float64)
    m[np.arange(n), np.arange(n)] = 1.0
    m[:, n] = np.asarray(offset, dtype=np.float64)
    return m


def _compose(first, then):
    """`first` maps space0 -> space1, `then` maps space1 -> space2; returns space0 -> space2.
    Accumulation order per entry follows NDAT:48-95 (offset seeded first, index1 ascending) so that the doubles are
    bit-identical to the reference's."""
    len1, c0 = first.shape
    len0 = c0 - 1
    len2 = then.shape[0]
    if then.shape[1] != len1 + 1:
        raise ValueError("affine transform size mismatch")
    out = np.empty((len2, len0 + 1), dtype=np.float64)
    lin = np.zeros((len2, len0), dtype=np.float64)
    off = np.array(then[:, len1], dtype=np.float64)
    for i1 in range(len1):                       # index1 ascending for every entry, all rows at once
        c = then[:, i1]
        if not c.any():
            continue                             # adding +0.0 products changes nothing (entries are finite)
        lin += c[:, None] * first[i1, :len0]
        off += c * first[i1, len0]
    out[:, :len0] = lin
    out[:, len0] = off
    return out


def _compose_cached(first, then):
    key = (first.shape, first.tobytes(), then.shape, then.tobytes())
    m = _COMPOSE.get(key)
    if m is None:
        if len(_COMPOSE) > 16384:
            _COMPOSE.clear()
        m = _COMPOSE[key] = _compose(first, then)
        m.setflags(write=False)
    return m


def pre_concatenate(matrix01, matrix12):
    """NDAT:29-35."""
    return _compose_cached(matrix01, matrix12)


def concatenate(matrix12, matrix01):
    """NDAT:38-46."""
    return _compose_cached(matrix01, matrix12)


def transform(matrix, source):
    """NDAT:97-127."""
    n = len(source)
    out = np.array(matrix[:, n], dtype=np.float64)
    for x in range(n):
        out = out + matrix[:, x] * float(source[x])
    return out
