"""Parity of the tcgen05 / TMEM contraction kernel (csrc/gemm_tc.cu, through `dls_matmul_ex`) with the CPU oracle.

Two bars per case:
  1. against the oracle evaluated on operands rounded the way the staging kernels round them (bf16 RNE / tf32 round-to-nearest-away):
     only the fp32 accumulation order differs, so the bound is tight -- 2e-6 * sum_k |a_ik| |b_kj| per element.
     This is the check that the TMA boxes, swizzles, UMMA descriptors and the epilogue address every element correctly.
  2. against the plain fp32 chain of the reference (TensorsSpec:337-344): relative Frobenius error below the tolerance
     stated in include/dls_b200.h (TF32 1e-3, BF16 4e-3).
The fp32 mode (CUDA cores) must match the chain bit for bit.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATED_REL = {"tf32": 1e-3, "bf16": 4e-3}


@pytest.fixture(scope="module")
def backends():
    from oracle import tensors_oracle as O
    import deeplearning.scala_b200.tensors as B
    B.default_backend().ctx
    return O, B


def operands(m, n, k, ta, tb, seed):
    rng = np.random.RandomState(seed)
    a = rng.standard_normal((k, m) if ta else (m, k)).astype(np.float32)
    b = rng.standard_normal((n, k) if tb else (k, n)).astype(np.float32)
    return a, b


def run_case(backends, mode, m, n, k, ta, tb, seed=0, **epi):
    O, B = backends
    a, b = operands(m, n, k, ta, tb, seed)
    bias = None
    okw, bkw = {}, {}
    if epi.get("bias_mode", "none") != "none":
        bias = np.random.RandomState(seed + 1).standard_normal(m if epi["bias_mode"] == "row" else n).astype(np.float32)
        okw["bias"], bkw["bias"] = O.tensor(bias), B.tensor(bias)
    got = B.matmul(B.tensor(a), B.tensor(b), transpose_a=ta, transpose_b=tb, mode=mode, **bkw, **epi).flat_array().reshape(m, n)
    exact = O.matmul(O.tensor(a), O.tensor(b), transpose_a=ta, transpose_b=tb, operands=mode, **okw, **epi).flat_array().reshape(m, n)
    chain = O.matmul(O.tensor(a), O.tensor(b), transpose_a=ta, transpose_b=tb, **okw, **epi).flat_array().reshape(m, n)
    A = np.abs(a.T if ta else a).astype(np.float64)
    Bm = np.abs(b.T if tb else b).astype(np.float64)
    scale = A @ Bm + 1.0
    return got, exact, chain, scale


SHAPES = [
    # m, n, k, ta, tb
    (128, 256, 64, False, True),       # exactly one tile, one k block, both K-major
    (128, 256, 256, False, True),      # several k blocks: ring wrap-around
    (256, 512, 128, False, False),     # B MN-major ([K,N] row-major): the forward layer layout
    (256, 512, 128, True, True),       # A MN-major
    (256, 512, 128, True, False),      # both MN-major: the weight-gradient layout
    (300, 520, 200, False, False),     # ragged M, N, K (TMA zero fill / clipped stores)
    (10, 128, 500, False, False),      # C1 layer 2: out[10,128] = W2[10,500] h[500,128]
    (500, 128, 784, False, False),     # C1 layer 1
    (500, 784, 128, False, True),      # C1 dW1 = delta[500,128] x[784,128]^T
    (64, 10, 96, False, False),        # N % 4 != 0: direct-store epilogue
    (1, 1, 1, False, False),
    (129, 257, 65, True, False),
    (1500, 1100, 333, False, True),    # more tiles than one wave of accumulators per CTA
    (27, 32, 20000, True, False),      # split-K: conv1 weight gradient shape (one tile, very long K)
    (288, 64, 8192, True, False),      # split-K with three row tiles
    (40, 100, 5000, False, True),      # split-K, K-major operands, ragged
]


@pytest.mark.parametrize("mode", ["bf16", "tf32"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "%dx%dx%d%s%s" % (s[0], s[1], s[2], "T" if s[3] else "N", "T" if s[4] else "N"))
def test_tensor_core_contraction(backends, mode, shape):
    m, n, k, ta, tb = shape
    got, exact, chain, scale = run_case(backends, mode, m, n, k, ta, tb, seed=m + n + k)
    err = np.abs(got.astype(np.float64) - exact.astype(np.float64)) / scale
    assert err.max() <= 2e-6, "kernel differs from the operand-rounded oracle: %g at %r" % (err.max(), np.unravel_index(err.argmax(), err.shape))
    rel = np.linalg.norm(got.astype(np.float64) - chain) / max(np.linalg.norm(chain), 1e-30)
    assert rel <= STATED_REL[mode], "relative Frobenius error %g above the stated %g" % (rel, STATED_REL[mode])


@pytest.mark.parametrize("mode", ["bf16", "tf32"])
@pytest.mark.parametrize("epi", [dict(bias_mode="row", activation="tanh"), dict(bias_mode="col", activation="relu"),
                                 dict(bias_mode="row", activation="none")], ids=["row+tanh", "col+relu", "row"])
def test_fused_epilogue(backends, mode, epi):
    m, n, k = 200, 264, 160
    got, exact, chain, scale = run_case(backends, mode, m, n, k, False, False, seed=5, **epi)
    # tanh / relu are 1-Lipschitz: the pre-activation bound carries over, plus 2 ulp of tanhf
    err = np.abs(got.astype(np.float64) - exact.astype(np.float64)) / scale
    assert err.max() <= 4e-6, err.max()


@pytest.mark.parametrize("shape", [(96, 80, 200, False, False), (33, 65, 17, True, True), (500, 128, 784, False, False)],
                         ids=["NN", "TT", "c1"])
def test_fp32_contraction_matches_the_chain_bit_for_bit(backends, shape):
    m, n, k, ta, tb = shape
    got, exact, chain, scale = run_case(backends, "fp32", m, n, k, ta, tb, seed=3)
    assert (got.view(np.uint32) == chain.view(np.uint32)).all()


def test_empty_contraction(backends):
    O, B = backends
    a = np.zeros((4, 0), dtype=np.float32)
    b = np.zeros((0, 8), dtype=np.float32)
    for mode in ("fp32", "bf16", "tf32"):
        got = B.matmul(B.tensor(a), B.tensor(b), mode=mode).flat_array()
        assert got.shape == (32,) and (got == 0).all()


@pytest.mark.parametrize("shape", [(1, 70000, 32), (3, 5000, 10), (1, 300, 4096), (64, 1000, 1), (2, 37, 5), (1, 262144, 8)],
                         ids=lambda s: "%dx%dx%d" % s)
def test_sum_axis(backends, shape):
    """`split(axis).reduce(_ + _)` (TL:124-162) as one reduction: column kernel, two-stage tall kernel, row kernel."""
    O, B = backends
    outer, k, inner = shape
    x = np.random.RandomState(k).standard_normal((outer, k, inner)).astype(np.float32)
    got = B.sum_axis(B.tensor(x), 1).flat_array().reshape(outer, inner)
    want = x.astype(np.float64).sum(axis=1)
    scale = np.abs(x).astype(np.float64).sum(axis=1) + 1e-30
    assert (np.abs(got - want) / scale).max() <= 1e-5
    if k <= 5000:
        fold = O.sum_axis(O.tensor(x), 1).flat_array().reshape(outer, inner)
        assert (np.abs(got - fold) / scale).max() <= 1e-5


GATED = [
    # m, n, k, ta, tb
    (256, 512, 128, False, False),
    (300, 520, 200, True, False),      # ragged, A MN-major (the dX layout of a feature-major layer)
    (500, 128, 10, True, False),       # C1: delta of the hidden layer, K = 10
    (64, 250, 4096, False, True),      # few tiles, long K: split-K (the gate runs in the reduce pass)
    (33, 7, 20, False, False),         # N % 4 != 0: direct-store epilogue, scalar gate loads
]


@pytest.mark.parametrize("mode", ["fp32", "tf32", "bf16"])
@pytest.mark.parametrize("gate_mode", ["tanh", "relu"])
@pytest.mark.parametrize("shape", GATED, ids=["%dx%dx%d-%s%s" % (s[0], s[1], s[2], "T" if s[3] else "N", "T" if s[4] else "N") for s in GATED])
def test_gated_product_equals_the_product_then_the_elementwise_tree(backends, mode, gate_mode, shape):
    """dls_matmul_gated: (a b) * act'(g) in the contraction's epilogue must equal, bit for bit, the plain product followed by
    the elementwise tree a layer's backward builds (`d * (1 - y * y)`, `d * min(y * FLT_MAX, 1)`)."""
    O, B = backends
    m, n, k, ta, tb = shape
    a, b = operands(m, n, k, ta, tb, seed=m + n + k)
    rng = np.random.RandomState(7)
    g = rng.standard_normal((m, n)).astype(np.float32)
    if gate_mode == "relu":
        g = np.maximum(g, 0)
        g[0, :3] = [1e-40, 3e-39, 0.0][:min(3, n)]          # subnormal activations: the mask is fractional there, as in the tree
    else:
        g = np.tanh(g)
    at, bt, gt = B.tensor(a), B.tensor(b), B.tensor(g).do_cache()
    got = B.matmul(at, bt, transpose_a=ta, transpose_b=tb, mode=mode, gate=gt, gate_mode=gate_mode).flat_array()
    plain = B.matmul(at, bt, transpose_a=ta, transpose_b=tb, mode=mode).do_cache()
    if gate_mode == "tanh":
        tree = plain * (1.0 - gt * gt)
    else:
        tree = plain * B.minimum(gt * 3.4028234663852886e38, B.scalar(1.0))
    want = tree.flat_array()
    assert (got.view(np.uint32) == want.view(np.uint32)).all()


@pytest.mark.parametrize("shape", [(256, 512, 128), (300, 520, 200), (64, 248, 4096), (128, 10, 500)],
                         ids=["256x512x128", "300x520x200", "splitk-64x248x4096", "128x10x500"])
def test_emitted_operand_copy_is_the_staged_copy(backends, shape):
    """DLS_GEMM_EMIT_OPERAND_COPY: the bf16 copy written by the epilogue is what the staging pass would have produced -- a
    contraction consuming the result gives the same bits with one launch fewer, in both operand orientations."""
    O, B = backends
    m, n, k = shape
    be = B.default_backend()
    a, b = operands(m, n, k, False, False, seed=3)
    rng = np.random.RandomState(4)
    w_right = rng.standard_normal((n, 96)).astype(np.float32)       # c [m, n] . w [n, 96]: c is K-major
    w_left = rng.standard_normal((m, 72)).astype(np.float32)        # c^T [n, m] . w [m, 72]: c is MN-major
    outs, launches = [], []
    for emit in (False, True):
        c = B.matmul(B.tensor(a), B.tensor(b), mode="bf16", activation="tanh", emit_operand_copy=emit).do_cache()
        wr, wl = B.tensor(w_right).do_cache(), B.tensor(w_left).do_cache()
        wr.flat_array(), wl.flat_array()
        n0 = be.stats()["kernel_launches"]
        r1 = B.matmul(c, wr, mode="bf16").flat_array()
        n1 = be.stats()["kernel_launches"]
        r2 = B.matmul(c, wl, transpose_a=True, mode="bf16").flat_array()
        outs.append((c.flat_array(), r1, r2))
        launches.append(n1 - n0)
    for x, y in zip(outs[0], outs[1]):
        assert (x.view(np.uint32) == y.view(np.uint32)).all()
    assert launches[1] == launches[0] - 1, launches


PAIR = [
    # m, n, k, ta, tb: enough 256 x 256 tiles (>= 148) for the CTA-pair kernel, ragged so that the last pair-row has a peer CTA
    # whose rows lie entirely outside the matrix, the last tile column is partial, and K has a partial block
    (2100, 4700, 200, False, True),
    (2100, 4700, 200, True, False),
    (4600, 2200, 136, False, False),
    (3968, 3840, 320, True, True),       # M = 31 x 128: the peer half of the last pair tile is empty
]


@pytest.mark.parametrize("mode", ["bf16", "tf32"])
@pytest.mark.parametrize("shape", PAIR, ids=["%dx%dx%d-%s%s" % (s[0], s[1], s[2], "T" if s[3] else "N", "T" if s[4] else "N") for s in PAIR])
def test_cta_pair_kernel_on_ragged_shapes(backends, mode, shape):
    """gemm2_tc_kernel (cta_group::2): sampled rows of every 128-row slab against the operand-rounded float64 product, with the
    row bias + tanh epilogue; and the single-CTA kernel must agree with it element for element up to accumulation order."""
    O, B = backends
    m, n, k, ta, tb = shape
    a, b = operands(m, n, k, ta, tb, seed=m + n)
    rng = np.random.RandomState(5)
    bias = rng.standard_normal(m).astype(np.float32)
    got = B.matmul(B.tensor(a), B.tensor(b), transpose_a=ta, transpose_b=tb, mode=mode, bias=B.tensor(bias), bias_mode="row",
                   activation="tanh").flat_array().reshape(m, n)
    rows = np.array(sorted({min(m - 1, t * 128 + rng.randint(0, 128)) for t in range((m + 127) // 128)} | {0, m - 1}))
    ar = O.round_operand(a, mode).astype(np.float64)
    br = O.round_operand(b, mode).astype(np.float64)
    A = (ar.T if ta else ar)[rows]
    Bm = br.T if tb else br
    pre = A @ Bm + bias[rows, None]
    scale = np.abs(A) @ np.abs(Bm) + np.abs(bias[rows, None]) + 1.0
    err = np.abs(np.arctanh(np.clip(got[rows].astype(np.float64), -1 + 1e-7, 1 - 1e-7)) - pre)
    ok = np.abs(pre) < 3.0                                  # away from the saturated region, where arctanh amplifies one ulp
    assert (err[ok] / scale[ok]).max() <= 4e-6
    assert np.isfinite(got).all() and np.abs(got).max() <= 1.0


def test_cta_pair_kernel_gate_and_operand_copy(backends):
    """The pair kernel's epilogue carries the same gate and bf16 operand copy as the single-CTA kernel: bit-identical to
    product-then-tree, and a consumer of the result gives the same bits with and without the emitted copy."""
    O, B = backends
    m, n, k = 2304, 4352, 192
    a, b = operands(m, n, k, False, False, seed=9)
    g = np.tanh(np.random.RandomState(8).standard_normal((m, n)).astype(np.float32))
    at, bt, gt = B.tensor(a).do_cache(), B.tensor(b).do_cache(), B.tensor(g).do_cache()
    got = B.matmul(at, bt, mode="bf16", gate=gt, gate_mode="tanh", emit_operand_copy=True).do_cache()
    plain = B.matmul(at, bt, mode="bf16").do_cache()
    want = (plain * (1.0 - gt * gt)).flat_array()
    assert (got.flat_array().view(np.uint32) == want.view(np.uint32)).all()
    w = B.tensor(np.random.RandomState(10).standard_normal((n, 64)).astype(np.float32)).do_cache()
    r1 = B.matmul(got, w, mode="bf16").flat_array()
    r2 = B.matmul(B.tensor(want.reshape(m, n)), w, mode="bf16").flat_array()
    assert (r1.view(np.uint32) == r2.view(np.uint32)).all()


@pytest.mark.parametrize("mode", ["tf32", "bf16"])
@pytest.mark.parametrize("shape", [(256, 512, 128, False, True), (300, 520, 200, False, False), (64, 248, 4096, False, True),
                                   (2304, 4352, 192, True, False), (33, 7, 20, False, False)],
                         ids=["256x512x128", "ragged", "splitk", "pair", "odd-n"])
def test_update_epilogue_equals_gradient_then_update(backends, mode, shape):
    """dls_matmul_sgd: w -= lr * (a b) applied by the contraction's epilogue, in place, equals -- bit for bit -- the
    materialised product followed by `data - delta * learningRate` (TW:63-72), on every kernel path (single CTA, CTA pair,
    split-K, direct-store); in bf16 mode the weight's new operand copy is the staged copy of the new value."""
    O, B = backends
    m, n, k, ta, tb = shape
    be = B.default_backend()
    a, b = operands(m, n, k, ta, tb, seed=m * 3 + n)
    w0 = np.random.RandomState(12).standard_normal((m, n)).astype(np.float32)
    lr = 0.0625 + 1e-3
    at, bt = B.tensor(a).do_cache(), B.tensor(b).do_cache()
    product = B.matmul(at, bt, transpose_a=ta, transpose_b=tb, mode=mode)
    spec = dict(product._ungated)
    want = (B.tensor(w0) - product * lr).flat_array()
    w = B.tensor(w0).do_cache()
    be.matmul_sgd(spec, w, lr)
    got = w.flat_array()
    assert (got.view(np.uint32) == want.view(np.uint32)).all()
    if mode == "bf16":
        probe = B.tensor(np.random.RandomState(13).standard_normal((n, 40)).astype(np.float32)).do_cache()
        n0 = be.stats()["kernel_launches"]
        r1 = B.matmul(w, probe, mode="bf16").flat_array()
        used = be.stats()["kernel_launches"] - n0
        fresh = B.tensor(want.reshape(m, n)).do_cache()
        fresh.flat_array()
        n0 = be.stats()["kernel_launches"]
        r2 = B.matmul(fresh, probe, mode="bf16").flat_array()
        used_fresh = be.stats()["kernel_launches"] - n0
        assert (r1.view(np.uint32) == r2.view(np.uint32)).all()
        assert used == used_fresh          # r1 staged the probe but not the weight; r2 staged the weight but not the probe
