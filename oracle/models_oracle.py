"""CPU ORACLE (test infrastructure, NOT product code): one SGD step of the BASELINE.json model configs, restated as
plain numpy math from the reference's autodiff rules -- independent of the Tensor algebra, of the Do / Tape machinery
and of every kernel in the product.

Rules restated (file:line into /root/reference):
  * `train` = forward, then backward with the scalar delta 1.0f  => loss = SUM of the outputs      (TL:55, :70-76)
  * delta of `a * b` is delta * other, of `a + b` is delta, reduced back to the operand's shape by
    summing the broadcast dimensions                                                            (TL:110-162, :291-340)
  * `translate` backward translates the delta by -offset, out-of-range reads are the padding 0    (TL:225-238)
  * weight update: data := data - optimizer.delta(delta); here delta' = learning_rate * delta     (TW:50-74, Weights:46-51)
  * convolution index convention out[b,y,x,f] = bias[f] + sum x[b, y-(ky-1), x-(kx-1), c] w[ky,kx,c,f]
    (true convolution, zero 'same' padding)                                                      (TensorsSpec:144-210)
  * `scale` to half size reads in[2y, 2x] (truncating nearest neighbour)                          (Tensors:950-965)
Layer values are rounded to fp32 at every layer boundary (what the device buffers hold); sums inside a layer are
accumulated in fp64, so a correct fp32 kernel agrees to ~1e-6 relative and the tests' 1e-5 / 1e-3 bars are meaningful.
Only tests/, smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

F32 = np.float32


def f32(a):
    return np.asarray(a, dtype=np.float64).astype(F32)


def act_forward(z, activation):
    if activation == "tanh":
        return f32(np.tanh(z.astype(np.float64)))
    if activation == "relu":
        return f32(np.maximum(z, 0))
    return f32(z)


def act_backward(delta, y, activation):
    if activation == "tanh":
        return f32(delta.astype(np.float64) * (1.0 - y.astype(np.float64) ** 2))
    if activation == "relu":
        return f32(delta * (y > 0))
    return f32(delta)


def mlp_step(weights: List[Tuple[np.ndarray, Optional[np.ndarray]]], x: np.ndarray, target: Optional[np.ndarray],
             learning_rate: float, last_activation: str = "none", operand_rounding=None):
    """Feature-major MLP (x: [features, batch]).  Returns (loss, new_weights, activations).
    `operand_rounding(a)` models the tensor-core input rounding of a GEMM mode (None = fp32)."""
    r = operand_rounding or (lambda a: a)
    hs = [f32(x)]
    acts = []
    for i, (w, b) in enumerate(weights):
        act = "tanh" if i + 1 < len(weights) else last_activation
        z = r(w).astype(np.float64) @ r(hs[-1]).astype(np.float64)
        if b is not None:
            z = f32(z).astype(np.float64) + b.astype(np.float64)[:, None]
        hs.append(act_forward(z, act))
        acts.append(act)
    out = hs[-1]
    if target is not None:
        diff = f32(out.astype(np.float64) - target)
        loss_terms = f32(diff.astype(np.float64) ** 2)
        delta = f32(2.0 * diff.astype(np.float64))
    else:
        loss_terms = out
        delta = np.ones_like(out)
    loss = float(loss_terms.astype(np.float64).sum())
    new_weights = []
    for i in reversed(range(len(weights))):
        w, b = weights[i]
        dz = act_backward(delta, hs[i + 1], acts[i])
        dw = f32(r(dz).astype(np.float64) @ r(hs[i]).astype(np.float64).T)
        db = f32(dz.astype(np.float64).sum(axis=1)) if b is not None else None
        delta = f32(r(w).astype(np.float64).T @ r(dz).astype(np.float64))
        new_w = f32(w.astype(np.float64) - f32(dw.astype(np.float64) * F32(learning_rate)))
        new_b = f32(b.astype(np.float64) - f32(db.astype(np.float64) * F32(learning_rate))) if b is not None else None
        new_weights.append((new_w, new_b))
    new_weights.reverse()
    return loss, new_weights, hs


# ---- CNN (C3) --------------------------------------------------------------------------------------------------------

TAPS = [(ky, kx) for ky in range(3) for kx in range(3)]


def shifted(x: np.ndarray, ky: int, kx: int) -> np.ndarray:
    """s[b, y, x, c] = x[b, y-(ky-1), x-(kx-1), c] with zero padding (Tensor.translate with padding 0)."""
    _, h, w, _ = x.shape
    xp = np.pad(x, ((0, 0), (1, 1), (1, 1), (0, 0)))
    oy, ox = ky - 1, kx - 1
    return xp[:, 1 - oy:1 - oy + h, 1 - ox:1 - ox + w, :]


def conv_forward(x, w, b, r):
    out = np.zeros(x.shape[:3] + (w.shape[3],), dtype=np.float64)
    xr, wr = r(x).astype(np.float64), r(w).astype(np.float64)
    for ky, kx in TAPS:
        out += shifted(xr, ky, kx) @ wr[ky, kx]
    return f32(out).astype(np.float64) + b.astype(np.float64)


def conv_backward(x, w, dz, r, need_dx=True):
    xr, wr, dzr = r(x).astype(np.float64), r(w).astype(np.float64), r(dz).astype(np.float64)
    dw = np.zeros(w.shape, dtype=np.float64)
    dx = np.zeros(x.shape, dtype=np.float64) if need_dx else None
    _, h, wd, _ = x.shape
    for ky, kx in TAPS:
        s = shifted(xr, ky, kx)
        dw[ky, kx] = np.tensordot(s, dzr, axes=([0, 1, 2], [0, 1, 2]))
        if need_dx:
            # x[b, y-oy, x-ox] contributed to out[b, y, x]  =>  dx[i] += (dz . w^T)[i + o]
            g = dzr @ wr[ky, kx].T
            gp = np.pad(g, ((0, 0), (1, 1), (1, 1), (0, 0)))
            oy, ox = ky - 1, kx - 1
            dx += gp[:, 1 + oy:1 + oy + h, 1 + ox:1 + ox + wd, :]
    db = dz.astype(np.float64).sum(axis=(0, 1, 2))
    return (f32(dx) if need_dx else None), f32(dw), f32(db)


def cnn_step(convs, fc1, fc2, x, target, learning_rate, operand_rounding=None):
    """C3 model: convs = [(w[3,3,ci,co], b[co])...] with ReLU and 2x down-sampling between them; fc = (W[in,out], b[out]).
    Returns (loss, new_convs, new_fc1, new_fc2)."""
    r = operand_rounding or (lambda a: a)
    lr = F32(learning_rate)
    h = f32(x)
    saved = []
    for i, (w, b) in enumerate(convs):
        y = act_forward(conv_forward(h, w, b, r), "relu")
        saved.append((h, y))
        h = y
        if i + 1 < len(convs):
            h = np.ascontiguousarray(h[:, ::2, ::2, :])
    batch = x.shape[0]
    flat = h.reshape(batch, -1)
    z1 = f32(r(flat).astype(np.float64) @ r(fc1[0]).astype(np.float64)).astype(np.float64) + fc1[1].astype(np.float64)
    h1 = act_forward(z1, "relu")
    z2 = f32(r(h1).astype(np.float64) @ r(fc2[0]).astype(np.float64)).astype(np.float64) + fc2[1].astype(np.float64)
    out = f32(z2)
    diff = f32(out.astype(np.float64) - target)
    loss = float((f32(diff.astype(np.float64) ** 2)).astype(np.float64).sum())
    delta = f32(2.0 * diff.astype(np.float64))

    def sgd(p, g):
        return f32(p.astype(np.float64) - f32(g.astype(np.float64) * lr))

    dw2 = f32(r(h1).astype(np.float64).T @ r(delta).astype(np.float64))
    db2 = f32(delta.astype(np.float64).sum(axis=0))
    d_h1 = f32(r(delta).astype(np.float64) @ r(fc2[0]).astype(np.float64).T)
    dz1 = act_backward(d_h1, h1, "relu")
    dw1 = f32(r(flat).astype(np.float64).T @ r(dz1).astype(np.float64))
    db1 = f32(dz1.astype(np.float64).sum(axis=0))
    d_flat = f32(r(dz1).astype(np.float64) @ r(fc1[0]).astype(np.float64).T)
    new_fc2 = (sgd(fc2[0], dw2), sgd(fc2[1], db2))
    new_fc1 = (sgd(fc1[0], dw1), sgd(fc1[1], db1))
    d = d_flat.reshape(h.shape)
    new_convs = [None] * len(convs)
    for i in reversed(range(len(convs))):
        xin, y = saved[i]
        if i + 1 < len(convs):
            up = np.zeros(y.shape, dtype=F32)
            up[:, ::2, ::2, :] = d
            d = up
        dz = act_backward(d, y, "relu")
        dx, dw, db = conv_backward(xin, convs[i][0], dz, r, need_dx=i > 0)
        new_convs[i] = (sgd(convs[i][0], dw), sgd(convs[i][1], db))
        d = dx
    return loss, new_convs, new_fc1, new_fc2


# ---- the remaining differentiable operations of SURVEY 8(f).2 (mirror of FloatLayers.scala:133-274 per element) -------------
def left_broadcast(a: np.ndarray, shape: Sequence[int]) -> np.ndarray:
    """Tensor.broadcast (Tensors:893-945): dimensions are aligned on the LEFT, missing trailing ones are added."""
    a = np.asarray(a)
    return np.broadcast_to(a.reshape(a.shape + (1,) * (len(shape) - a.ndim)), tuple(shape))


def sum_as(d: np.ndarray, shape: Sequence[int]) -> np.ndarray:
    """TL:124-162: sum a delta back over the dimensions that were broadcast (left-aligned), fp64 accumulation."""
    d = np.asarray(d, dtype=np.float64)
    shape = tuple(shape)
    if d.ndim > len(shape):
        d = d.sum(axis=tuple(range(len(shape), d.ndim)))
    axes = tuple(i for i, s in enumerate(shape) if s == 1 and d.shape[i] != 1)
    if axes:
        d = d.sum(axis=axes, keepdims=True)
    return f32(d)


OP_RULES = {
    # name: (forward(a, b) -> y, backward(delta, a, b, y) -> (delta_a, delta_b)); every product / quotient is one fp32 rounding
    "divide": (lambda a, b: f32(a / b),
               lambda d, a, b, y: (f32(d / b), f32(f32(f32(-d * a)) / f32(b * b)))),            # FloatLayers:133-148
    "min": (lambda a, b: np.minimum(a, b),
            lambda d, a, b, y: (f32(d * (a < b)), f32(d * ~(a < b)))),                          # FloatLayers:150-165
    "max": (lambda a, b: np.maximum(a, b),
            lambda d, a, b, y: (f32(d * (a > b)), f32(d * ~(a > b)))),                          # FloatLayers:167-182
    "abs": (lambda a, b=None: np.abs(a),
            lambda d, a, b, y: (f32(np.where(a < 0, -d, d)), None)),                            # FloatLayers:239-256
    "sqrt": (lambda a, b=None: f32(np.sqrt(a.astype(np.float64))),
             lambda d, a, b, y: (f32(f32(d * F32(0.5)) / y), None)),                            # FloatLayers:258-272
}


def ops_step(a, b, c, learning_rate):
    """One SGD step of a composite that uses every operation of 8(f).2 once (a, b: [n, m]; c: [n], broadcast on the left):
         q = a / b;  r = sqrt(abs(q) + 0.5);  m = max(a, broadcast(c));  k = min(b, m);  p = permute(r * k, [1, 0])
         loss = sum(p * p)
    Returns (loss, new_a, new_b, new_c, p)."""
    a, b, c = f32(a), f32(b), f32(c)
    lr = F32(learning_rate)
    q = OP_RULES["divide"][0](a, b)
    ab = OP_RULES["abs"][0](q)
    s = f32(ab + F32(0.5))
    r = OP_RULES["sqrt"][0](s)
    cb = np.ascontiguousarray(left_broadcast(c, a.shape))
    m = OP_RULES["max"][0](a, cb)
    k = OP_RULES["min"][0](b, m)
    rk = f32(r * k)
    p = np.ascontiguousarray(rk.T)
    pp = f32(p * p)
    loss = float(pp.astype(np.float64).sum())
    # backward: the seed is the scalar 1 (TL:55), `sum` hands it to every element
    d_pp = np.ones_like(pp)
    d_p = f32(f32(d_pp * p) + f32(d_pp * p))                 # `*` sends delta * other to each operand (TL:291-340); both are p
    d_rk = np.ascontiguousarray(d_p.T)                       # inverse permutation
    d_r, d_k = f32(d_rk * k), f32(d_rk * r)
    d_b1, d_m = OP_RULES["min"][1](d_k, b, m, k)
    d_a1, d_cb = OP_RULES["max"][1](d_m, a, cb, m)
    d_c = sum_as(d_cb, c.shape + (1,) * (a.ndim - c.ndim)).reshape(c.shape)
    d_s, _ = OP_RULES["sqrt"][1](d_r, s, None, r)
    d_q, _ = OP_RULES["abs"][1](d_s, q, None, ab)
    d_a2, d_b2 = OP_RULES["divide"][1](d_q, a, b, q)
    d_a, d_b = f32(d_a1 + d_a2), f32(d_b1 + d_b2)
    return loss, f32(a - f32(d_a * lr)), f32(b - f32(d_b * lr)), f32(c - f32(d_c * lr)), p
