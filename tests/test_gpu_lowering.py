"""Host-unrolled chains, written exactly the way the reference's users write them, through the lowering passes:
  * views of a not-yet-materialised elementwise expression are pushed through it (trees.see_through) instead of
    materialising the `[i, j, k]` product of TensorsSpec:337-344;
  * `((t0 + t1) + t2) + ...` runs fold back into one device loop (plan.cpp compress_loops, OP_LOOP in ew_gather_kernel);
  * a launch that is one dense contraction goes to the tcgen05 GEMM in tensor-core modes (Plan::contraction).
fp32 mode must reproduce the oracle's evaluation of the same unrolled tree BIT FOR BIT (same fold order, one rounding per
node); tensor-core modes must stay within the stated tolerance.
"""
import numpy as np
import pytest

STATED_REL = {"tf32": 1e-3, "bf16": 4e-3}


def matmul_broadcast(M, a, b):
    """TensorsSpec:337-344 / benchmarks.scala:188-191: broadcast both, multiply, split the middle dim, left fold."""
    i, j = a.shape
    k = b.shape[1]
    product = a.broadcast([i, j, k]) * b.reshape([1, j, k]).broadcast([i, j, k])
    parts = product.split(1)
    acc = parts[0]
    for p in parts[1:]:
        acc = acc + p
    return acc


def matmul_unrolled(M, a, b):
    """TensorsSpec:371-383 / benchmarks.scala:176-186: one chain per output column, joined."""
    i, j = a.shape
    k = b.shape[1]
    a_cols = a.split(1)
    columns = []
    for c in range(k):
        b_col = b.split(1)[c].split(0)
        acc = a_cols[0] * b_col[0].broadcast([i])
        for t in range(1, j):
            acc = acc + a_cols[t] * b_col[t].broadcast([i])
        columns.append(acc)
    return M.join(columns)


def convolution(M, x, w, bias):
    """TensorsSpec:144-210: out[b,y,x,f] = bias[f] + sum_{ky,kx,c} x[b, y-(ky-1), x-(kx-1), c] * w[ky,kx,c,f], joined over f."""
    bsz, h, wd, cin = x.shape
    kh, kw, _, cout = w.shape
    channels = x.split(3)
    outs = []
    for f in range(cout):
        acc = bias.split(0)[f].broadcast([bsz, h, wd])
        wf = w.split(3)[f]
        for ky in range(kh):
            for kx in range(kw):
                for c in range(cin):
                    tap = wf.split(0)[ky].split(0)[kx].split(0)[c].broadcast([bsz, h, wd])
                    acc = acc + channels[c].translate([0, ky - kh // 2, kx - kw // 2]) * tap
        outs.append(acc)
    return M.join(outs)


def batch_sum(M, delta, other):
    """TensorLayers.sumAs (TL:124-162) of `delta * other` down to the leading dim: the bias / weight gradient chains."""
    t = delta * other
    parts = t.split(1)
    acc = parts[0]
    for p in parts[1:]:
        acc = acc + p
    return acc


def data(seed, *shape):
    return np.random.RandomState(seed).standard_normal(shape).astype(np.float32)


# ---- CPU: the lowering decisions (dls_plan_describe needs no device) --------------------------------------------------

@pytest.fixture(scope="module")
def T():
    import deeplearning.scala_b200.tensors as T
    return T


def describe(T, t):
    return T.default_backend().describe_plan([t.closure()], t.shape)


def test_unrolled_matmuls_lower_to_one_contraction(T):
    a, b = T.random([64, 100], 1), T.random([100, 48], 2)
    d1 = describe(T, matmul_broadcast(T, a, b))
    assert "loops=[macx99:lin*lin,]" in d1 and "contraction=64x48x100NN" in d1 and "code=4 " in d1
    cols = []
    i, j, k = 64, 100, 48
    a_cols = a.split(1)
    for c in range(k):
        b_col = b.split(1)[c].split(0)
        acc = a_cols[0] * b_col[0].broadcast([i])
        for t in range(1, j):
            acc = acc + a_cols[t] * b_col[t].broadcast([i])
        cols.append(acc)
    d2 = T.default_backend().describe_plan([c.closure() for c in cols], [i])
    assert "loops=[48 x macx99:lin*lin" in d2 and "contraction=64x48x100NN" in d2


def test_convolution_chain_folds_into_checked_loops(T):
    x, w, bias = T.random([2, 8, 8, 4], 4), T.random([3, 3, 4, 5], 5), T.random([5], 6)
    bsz, h, wd = 2, 8, 8
    outs = []
    for f in range(5):
        acc = bias.split(0)[f].broadcast([bsz, h, wd])
        wf = w.split(3)[f]
        for ky in range(3):
            for kx in range(3):
                for c in range(4):
                    acc = acc + x.split(3)[c].translate([0, ky - 1, kx - 1]) * wf.split(0)[ky].split(0)[kx].split(0)[c].broadcast([bsz, h, wd])
        outs.append(acc)
    d = T.default_backend().describe_plan([o.closure() for o in outs], [bsz, h, wd])
    assert d.count("macx36:mono*lin") == 5 and "code=16 " in d and "contraction" not in d


def test_short_chains_and_vector_programs_are_left_alone(T):
    a, b, c = (T.random([64, 64], s) for s in (1, 2, 3))
    acc = a * b
    for _ in range(10):
        acc = acc + a * c
    assert "kind=vec" in describe(T, acc)                       # contiguous: stays on the 128-bit kernel
    s = a.split(1)
    assert "loops" not in describe(T, s[0] + s[1] + s[2] + s[3])  # below kLoopMinTerms


# ---- GPU parity -----------------------------------------------------------------------------------------------------------

@pytest.fixture(scope="module")
def backends():
    from oracle import tensors_oracle as O
    import deeplearning.scala_b200.tensors as B
    B.default_backend().ctx
    return O, B


def bits_equal(want, got):
    want, got = np.asarray(want, dtype=np.float32), np.asarray(got, dtype=np.float32)
    return want.shape == got.shape and bool((want.view(np.uint32) == got.view(np.uint32)).all())


@pytest.mark.gpu
@pytest.mark.parametrize("build", [matmul_broadcast, matmul_unrolled], ids=["broadcast", "unrolled"])
def test_unrolled_matmul_fp32_is_bit_exact(backends, build):
    O, B = backends
    a, b = data(1, 48, 100), data(2, 100, 40)
    want = build(O, O.tensor(a), O.tensor(b)).flat_array()
    launches0 = B.default_backend().stats()["kernel_launches"]
    got = build(B, B.tensor(a), B.tensor(b)).flat_array()
    assert B.default_backend().stats()["kernel_launches"] - launches0 == 1     # no materialised [i, j, k] product
    assert bits_equal(want, got)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["tf32", "bf16"])
@pytest.mark.parametrize("build", [matmul_broadcast, matmul_unrolled], ids=["broadcast", "unrolled"])
def test_unrolled_matmul_runs_on_the_tensor_cores(backends, build, mode):
    O, B = backends
    a, b = data(3, 96, 128), data(4, 128, 72)
    be = B.Backend(gemm_mode=mode)
    try:
        got = build(be, be.tensor(a), be.tensor(b)).flat_array().reshape(96, 72)
    finally:
        be.close()
    exact = O.matmul(O.tensor(a), O.tensor(b), operands=mode).flat_array().reshape(96, 72)
    scale = np.abs(a).astype(np.float64) @ np.abs(b).astype(np.float64) + 1.0
    assert (np.abs(got - exact) / scale).max() <= 2e-6           # it really was the GEMM on rounded operands
    chain = a.astype(np.float64) @ b.astype(np.float64)
    assert np.linalg.norm(got - chain) / np.linalg.norm(chain) <= STATED_REL[mode]


@pytest.mark.gpu
def test_convolution_chain_is_bit_exact(backends):
    O, B = backends
    x, w, bias = data(5, 2, 9, 7, 4), data(6, 3, 3, 4, 5), data(7, 5)
    want = convolution(O, O.tensor(x), O.tensor(w), O.tensor(bias)).flat_array()
    got = convolution(B, B.tensor(x), B.tensor(w), B.tensor(bias)).flat_array()
    assert bits_equal(want, got)


@pytest.mark.gpu
def test_batch_sum_chain_is_bit_exact(backends):
    O, B = backends
    d, o = data(8, 33, 300), data(9, 33, 300)
    want = batch_sum(O, O.tensor(d), O.tensor(o)).flat_array()
    got = batch_sum(B, B.tensor(d), B.tensor(o)).flat_array()
    assert bits_equal(want, got)
