"""N-dimensional affine transforms used by every Tensor view (host side, pure numpy).

Mirror of Compute.scala's NDimensionalAffineTransform (NDAT =
Compute.scala/NDimensionalAffineTransform/src/main/scala/com/thoughtworks/compute/NDimensionalAffineTransform.scala).
A matrix maps view coordinates to checkpoint coordinates: rows = rank(checkpoint), cols = rank(view) + 1, the last
column is the offset (Tensors.scala:1289-1293).
"""
import numpy as np


def translate(offset):
    """NDAT:13-26 -- identity plus an offset column."""
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
    out = np.zeros((len2, len0 + 1), dtype=np.float64)
    for i2 in range(len2):
        lin = np.zeros(len0, dtype=np.float64)
        off = then[i2, len1]
        for i1 in range(len1):
            c = then[i2, i1]
            lin = lin + c * first[i1, :len0]
            off = off + c * first[i1, len0]
        out[i2, :len0] = lin
        out[i2, len0] = off
    return out


def pre_concatenate(matrix01, matrix12):
    """NDAT:29-35."""
    return _compose(matrix01, matrix12)


def concatenate(matrix12, matrix01):
    """NDAT:38-46."""
    return _compose(matrix01, matrix12)


def transform(matrix, source):
    """NDAT:97-127."""
    n = len(source)
    out = np.array(matrix[:, n], dtype=np.float64)
    for x in range(n):
        out = out + matrix[:, x] * float(source[x])
    return out
