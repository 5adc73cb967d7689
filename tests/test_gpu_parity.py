"""Parity of the CUDA path (through the C ABI) with the CPU oracle on seeded inputs.

Bars (BASELINE.json north_star): index / reshape / permute / view work bit-exact; fp32 `+ - * /` bit-exact (one IEEE
rounding per node on both sides); transcendental chains within 2 ulp; reductions within 1e-5 relative.
Each case is written once against the shared Tensor API and evaluated by both backends.
"""
import numpy as np
import pytest

from golden_cases import ulp_diff

pytestmark = pytest.mark.gpu

ULP_TRANSCENDENTAL = 2          # tolerance stated by north_star for fp32 elementwise ops
REL_REDUCTION = 1e-5            # tolerance stated by north_star for reductions


@pytest.fixture(scope="module")
def backends():
    from oracle import tensors_oracle as O
    import deeplearning.scala_b200.tensors as B
    B.default_backend().ctx
    return O, B


def both(backends, build):
    O, B = backends
    want = build(O)
    got = build(B)
    assert list(want.shape) == list(got.shape)
    return np.asarray(want.flat_array(), dtype=np.float32), got.flat_array(), got


def assert_bits(want, got):
    assert want.shape == got.shape
    same = (want.view(np.uint32) == got.view(np.uint32)) | (np.isnan(want) & np.isnan(got))
    assert same.all(), (np.flatnonzero(~same)[:8], want[~same][:8], got[~same][:8])


def assert_ulp(want, got, max_ulp):
    assert want.shape == got.shape
    bad = ulp_diff(want, got) > max_ulp
    bad &= ~(np.isnan(want) & np.isnan(got))
    assert not bad.any(), (np.flatnonzero(bad)[:8], want[bad][:8], got[bad][:8])


def rnd(seed, *shape, scale=1.0):
    return (np.random.RandomState(seed).standard_normal(shape) * scale).astype(np.float32)


# ---- N1: fused elementwise, vector path ----------------------------------------------------------------------------


@pytest.mark.parametrize("shape", [(64, 256), (3, 8), (128, 1024), (7, 12)])
def test_c2_row_broadcast(backends, shape):
    a, b, c = rnd(1, *shape), rnd(2, *shape), rnd(3, shape[0])

    def build(M):
        return M.tanh(M.tensor(a) * M.tensor(b) + M.tensor(c))     # c: one value per ROW (left-aligned broadcast)
    want, got, t = both(backends, build)
    assert_ulp(want, got, ULP_TRANSCENDENTAL)
    if shape[1] % 4 == 0:
        assert "kind=vec" in t.backend.describe_plan([t.closure()], t.shape)


@pytest.mark.parametrize("shape", [(64, 256), (5, 20)])
def test_c2_column_broadcast(backends, shape):
    a, b, c = rnd(4, *shape), rnd(5, *shape), rnd(6, shape[1])

    def build(M):
        return M.tanh(M.tensor(a) * M.tensor(b) + M.tensor(c).reshape([1, shape[1]]))
    want, got, _ = both(backends, build)
    assert_ulp(want, got, ULP_TRANSCENDENTAL)


@pytest.mark.parametrize("n", [1, 2, 3, 5, 255, 1023, 4097])
def test_ragged_tail_lengths(backends, n):
    a, b = rnd(7, n), rnd(8, n)

    def build(M):
        # fill, not scalar: `[1] + []` raises in the reference too (autoBroadcastShape indexes shape2(0), Tensors:208-222)
        return (M.tensor(a) * M.tensor(b) - M.tensor(a)) / (M.tensor(b) * M.tensor(b) + M.fill(1.0, [n]))
    want, got, _ = both(backends, build)
    assert_bits(want, got)


def test_arithmetic_is_bit_exact_and_not_contracted(backends):
    a, b, c = rnd(9, 4096), rnd(10, 4096), rnd(11, 4096)

    def build(M):
        ta, tb, tc = M.tensor(a), M.tensor(b), M.tensor(c)
        return (ta * tb + tc) - (ta / tb) * tc + (-ta)
    want, got, _ = both(backends, build)
    assert_bits(want, got)          # an FMA contraction of a*b+c would differ in the last bit for some lanes


def test_all_operators_with_special_values(backends):
    special = np.array([0.0, -0.0, 1.0, -1.0, 0.5, 3.0, -3.5, 1e-30, 1e30, np.inf, -np.inf, np.nan, 7.25, -2.0, 88.0, 1e-3],
                       dtype=np.float32)
    a = np.repeat(special, len(special))
    b = np.tile(special, len(special))

    def exact(M):
        ta, tb = M.tensor(a), M.tensor(b)
        parts = [ta + tb, ta - tb, ta * tb, ta / tb, ta % tb, M.minimum(ta, tb), M.maximum(ta, tb), -ta, M.abs_(ta), M.sqrt(ta)]
        return M.join(parts)
    want, got, _ = both(backends, exact)
    assert_bits(want, got)

    def transcendental(M):
        ta = M.tensor(a)
        return M.join([M.tanh(ta), M.exp(ta), M.log(ta)])
    want, got, _ = both(backends, transcendental)
    assert_ulp(want, got, ULP_TRANSCENDENTAL)


def test_shared_subtrees_and_two_compound_operands(backends):
    a, b, c = rnd(12, 32, 64), rnd(13, 32, 64), rnd(14, 32, 64)

    def build(M):
        ta, tb, tc = M.tensor(a), M.tensor(b), M.tensor(c)
        t2 = ta + tb                                  # shared (ExportContext CSE, Trees.scala:220)
        left = t2 * t2 + tc
        right = (ta - tc) * (tb + tc)
        return (left - right) * (t2 / (right * right + M.scalar(1.0)))
    want, got, t = both(backends, build)
    assert_bits(want, got)
    assert "slots=0" not in t.backend.describe_plan([t.closure()], t.shape)


def test_left_fold_chain_of_300_terms(backends):
    xs = [rnd(100 + i, 16, 16) for i in range(20)]

    def build(M):
        ts = [M.tensor(x) for x in xs]
        acc = ts[0] * ts[1]
        for i in range(300):
            acc = acc + ts[i % 20] * ts[(i * 7 + 3) % 20]
        return acc
    want, got, _ = both(backends, build)
    assert_bits(want, got)


def test_fill_and_scalar_leaves(backends):
    a = rnd(15, 8, 16)

    def build(M):
        return (M.tensor(a) + M.fill(2.5, [8, 16])) * M.scalar(3.0) + M.tensor(np.float32(4.0))
    want, got, _ = both(backends, build)
    assert_bits(want, got)


# ---- N2: affine gather ---------------------------------------------------------------------------------------------


@pytest.mark.parametrize("perm", [[1, 0, 2], [2, 1, 0], [1, 2, 0], [0, 2, 1]])
def test_permute_bit_exact(backends, perm):
    x = rnd(20, 5, 6, 8)
    want, got, _ = both(backends, lambda M: M.tensor(x).permute(perm))
    assert_bits(want, got)
    assert_bits(np.ascontiguousarray(x.transpose(perm)).reshape(-1), got)


TILED = [
    # shape, permutation: sizes that reach the tiled path of the gather kernel (the dim the source is contiguous along has
    # >= 32 elements, the output's inner dim >= 16), ragged in both tile directions
    ((70, 45), [1, 0]),
    ((33, 129), [1, 0]),
    ((64, 64), [1, 0]),
    ((3, 40, 50), [0, 2, 1]),
    ((5, 17, 19, 37), [0, 3, 1, 2]),      # NHWC -> NCHW
    ((5, 37, 17, 19), [0, 2, 3, 1]),      # NCHW -> NHWC
    ((36, 7, 48), [2, 1, 0]),
    ((2, 3, 34, 5, 40), [4, 1, 0, 3, 2]), # rank 5: the general-rank instantiation
]


@pytest.mark.parametrize("shape,perm", TILED, ids=["%s-%s" % ("x".join(map(str, s)), "".join(map(str, p))) for s, p in TILED])
def test_tiled_permute_bit_exact(backends, shape, perm):
    """Tensors:1008-1030 `permute` / `transpose` where the source's unit-stride dim is not the output's inner dim: the
    shared-memory tiled path must place every element exactly where the one-lane-per-element path does."""
    O, B = backends
    x = rnd(sum(shape), *shape)
    want, got, _ = both(backends, lambda M: M.tensor(x).permute(perm))
    assert_bits(want, got)
    assert_bits(np.ascontiguousarray(x.transpose(perm)).reshape(-1), got)
    plan = B.default_backend().describe_plan([B.tensor(x).permute(perm).closure()], [shape[p] for p in perm])
    assert "gather" in plan


def test_tiled_path_under_elementwise_ops_and_padding(backends):
    """The tiled path evaluates whole programs: a transposed view with a translate (bounds checks, padding) under
    arithmetic with a second transposed view, a saved sub-expression and a broadcast scalar."""
    a, b = rnd(31, 70, 45), rnd(32, 70, 45)

    def build(M):
        ta = M.tensor(a, padding=-3.5).translate([3, -2]).transpose()
        tb = M.tensor(b).transpose()
        u = ta * tb + M.scalar(0.25)
        return u * u - tb * u                          # `u` is a saved sub-expression (three uses)
    want, got, _ = both(backends, build)
    assert_bits(want, got)
    want, got, _ = both(backends, lambda M: M.tanh(M.tensor(a, padding=-3.5).translate([3, -2]).transpose() * M.tensor(b).transpose()))
    assert_ulp(want, got, ULP_TRANSCENDENTAL)
    want, got, _ = both(backends, lambda M: M.tensor(a, padding=-3.5).translate([3, -2]).transpose() + M.tensor(b).transpose())
    assert_bits(want, got)


def test_mixed_orientation_keeps_the_row_path(backends):
    """One leaf contiguous along the output's inner dim next to a transposed one: not a pure transposition; whatever path
    runs, the bits are the oracle's."""
    a, b = rnd(33, 48, 40), rnd(34, 40, 48)
    want, got, _ = both(backends, lambda M: M.tensor(a).transpose() + M.tensor(b))
    assert_bits(want, got)


@pytest.mark.parametrize("offset", [[0, 0], [1, -2], [-3, 4], [7, 0], [0, -9], [20, 20]])
def test_translate_with_padding_bit_exact(backends, offset):
    x = rnd(21, 9, 11)
    want, got, _ = both(backends, lambda M: M.tensor(x, padding=-7.5).translate(offset))
    assert_bits(want, got)


def test_translate_roundtrip_checkpoint_rule(backends):
    x = rnd(22, 6, 7)

    def view_of_view(M):
        return M.tensor(x, padding=9.0).translate([2, -1]).translate([-2, 1])
    want, got, _ = both(backends, view_of_view)
    assert_bits(want, got)
    assert_bits(x.reshape(-1), got)                   # no border loss: bounds are checked against the checkpoint only

    def through_inline(M):
        return (M.tensor(x, padding=9.0) + M.scalar(0.0)).translate([2, -1]).non_inline().translate([-2, 1])
    want, got, _ = both(backends, through_inline)
    assert_bits(want, got)
    assert (got == 9.0).any()


@pytest.mark.parametrize("new_shape", [[3, 5], [12, 20], [4, 7], [6, 10], [5, 3]])
def test_scale_truncating_resample_bit_exact(backends, new_shape):
    x = rnd(23, 6, 10)
    want, got, _ = both(backends, lambda M: M.tensor(x).scale(new_shape))
    assert_bits(want, got)


@pytest.mark.parametrize("case", ["up2_outer", "up2_inner", "up8", "up4_after_translate", "up2_negative_offsets", "translate_after_scale", "mixed"])
def test_dyadic_scale_rows_bit_exact(backends, case):
    """Coefficients that are multiples of 1/8 run on the 32-bit gather path in integer arithmetic; the result must be
    the `(int)(double)` truncation of KB:386, including the (-1, 0) -> 0 quirk for negative numerators."""
    x = rnd(29, 4, 6, 8)

    def build(M):
        t = M.tensor(x, padding=-7.0)
        if case == "up2_outer":
            return t.scale([8, 12, 8])
        if case == "up2_inner":
            return t.scale([4, 6, 16])
        if case == "up8":
            return t.scale([32, 6, 8])
        if case == "up4_after_translate":
            return t.translate([1, -2, 0]).scale([16, 6, 8])
        if case == "up2_negative_offsets":
            return t.translate([3, 1, 2], [6, 8, 12]).scale([12, 16, 12])
        if case == "translate_after_scale":                  # index = (int)(g * 0.5 - 0.5): g = 0 reads element 0, not padding
            return t.scale([8, 12, 8]).translate([1, 3, 0])
        return (t.scale([8, 12, 8]) * M.tensor(rnd(30, 8, 12, 8))).scale([16, 12, 8])
    want, got, _ = both(backends, build)
    assert_bits(want, got)


def test_split_join_permute_compositions(backends):
    x = rnd(24, 4, 6, 8)

    def build(M):
        t = M.tensor(x)
        slices = t.split(1)                                   # 6 x [4, 8]
        mixed = [s.transpose() * M.scalar(float(i)) for i, s in enumerate(slices)]     # [8, 4]
        return M.join(mixed, 0)                               # [6, 8, 4]
    want, got, _ = both(backends, build)
    assert_bits(want, got)


def test_broadcast_views(backends):
    x = rnd(25, 4, 1, 3)

    def build(M):
        return M.tensor(x).broadcast([4, 5, 3, 2]) + M.tensor(rnd(26, 4)).broadcast([4, 5, 3, 2])
    want, got, _ = both(backends, build)
    assert_bits(want, got)


def test_identity_leaf_smaller_than_launch_reads_padding(backends):
    # KB:440-442: only the upper bound is checked for an identity leaf
    small = rnd(27, 3, 4)

    def build(M):
        big = M.fill(1.0, [5, 6])
        t = M.tensor(small, padding=-1.0)
        return M.join([big, t])            # evaluated at the head's shape [5, 6]
    want, got, _ = both(backends, build)
    assert_bits(want, got)


def test_matmul_by_broadcast_and_unrolled_sum(backends):
    a, b = rnd(28, 12, 20), rnd(29, 20, 8)

    def build(M):
        m1, m2 = M.tensor(a), M.tensor(b)
        product = m1.broadcast([12, 20, 8]) * m2.reshape([1, 20, 8]).broadcast([12, 20, 8])
        parts = product.split(1)
        acc = parts[0]
        for p in parts[1:]:
            acc = acc + p
        return acc
    want, got, _ = both(backends, build)
    assert_bits(want, got)
    np.testing.assert_allclose(got.reshape(12, 8), a @ b, rtol=1e-4, atol=1e-5)


def test_size_one_dim_against_scalar_raises_like_the_reference(backends):
    # Tensors:208-222: shape1(i) == 1 sends the lookup to shape2(i), which does not exist for a rank-0 operand
    for M in backends:
        with pytest.raises(IndexError):
            M.tensor([2.0]) + M.scalar(1.0)


def test_empty_tensor(backends):
    want, got, _ = both(backends, lambda M: M.tensor([]) + M.scalar(1.0))
    assert want.size == 0 and got.size == 0


# ---- N3: reductions ------------------------------------------------------------------------------------------------


@pytest.mark.parametrize("n", [1, 17, 4096, 100003])
def test_sum_of_cached_buffer(backends, n):
    O, B = backends
    x = rnd(30, n)
    want = float(O.tensor(x).sum().flat_array()[0])
    got = float(B.tensor(x).sum().flat_array()[0])
    assert abs(got - want) <= REL_REDUCTION * max(1.0, float(np.abs(x).sum()))
    assert abs(got - float(x.astype(np.float64).sum())) <= REL_REDUCTION * max(1.0, float(np.abs(x).sum()))


@pytest.mark.parametrize("fuse", [True, False])
def test_sum_of_inline_tensor_fused_and_materialised(backends, fuse):
    O, B = backends
    a, b, c = rnd(31, 96, 128), rnd(32, 96, 128), rnd(33, 96)
    want = O.tanh(O.tensor(a) * O.tensor(b) + O.tensor(c)).sum().flat_array()[0]
    be = B.Backend(fuse_reductions=fuse)
    y = be.tanh(be.tensor(a) * be.tensor(b) + be.tensor(c))
    before = be.stats()["kernel_launches"]
    got = y.sum().flat_array()[0]
    launches = be.stats()["kernel_launches"] - before
    assert launches == (1 if fuse else 2)
    ref64 = np.tanh((a * b + c[:, None]).astype(np.float64)).sum()
    assert abs(float(got) - float(want)) <= REL_REDUCTION * np.abs(np.tanh(a * b + c[:, None])).sum()
    assert abs(float(got) - ref64) <= REL_REDUCTION * np.abs(np.tanh(a * b + c[:, None])).sum()
    be.close()


def test_sum_golden_exact(backends):
    _, B = backends
    assert B.fill(15625.0, [8, 8]).sum().to_string() == "1000000.0"
    assert B.fill(1.0, [1024, 1024]).sum().to_string() == "1048576.0"


# ---- N5: RNG -------------------------------------------------------------------------------------------------------


@pytest.mark.parametrize("n,seed", [(1, 0), (1000003, 12345), (4096, -7), (65536, 2 ** 31 - 1)])
def test_random_uniform_bit_exact(backends, n, seed):
    O, B = backends
    assert_bits(np.asarray(O.random([n], seed).flat_array()), B.random([n], seed).flat_array())


@pytest.mark.parametrize("n,seed", [(1, 1), (7, 54321), (100001, 99)])
def test_random_normal_close(backends, n, seed):
    O, B = backends
    want = np.asarray(O.random_normal([n], seed).flat_array())
    got = B.random_normal([n], seed).flat_array()
    with np.errstate(invalid="ignore"):
        ok = (ulp_diff(want, got) <= 4) | (np.abs(want - got) <= 2e-6) | (np.isinf(want) & np.isinf(got))
    ok |= np.isnan(want) & np.isnan(got)          # u1 == 0 is not guarded by the reference (Tensors:409-425)
    assert ok.all(), (want[~ok][:5], got[~ok][:5])


def test_lcg_hash_backend(backends):
    O, B = backends
    be = B.Backend(rng_hash="lcg")
    assert_bits(np.asarray(O.random([1000], 42, hash_fn=O.lcg_hash).flat_array()), be.random([1000], 42).flat_array())
    be.close()


# ---- errors ----------------------------------------------------------------------------------------------------------


def test_shape_errors_are_value_errors(backends):
    _, B = backends
    with pytest.raises(ValueError):
        B.tensor([[1.0, 2.0], [3.0]])
    with pytest.raises(ValueError):
        B.tensor([[1.0, 2.0, 3.0]] * 2) + B.tensor([1.0, 2.0, 3.0])
    with pytest.raises(ValueError):
        B.tensor([1.0, 2.0]).reshape([3])
    with pytest.raises(ValueError):
        B.tensor([1.0, 2.0]).translate([1, 2])
