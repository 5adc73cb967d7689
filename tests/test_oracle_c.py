"""oracle/c2_oracle.c (the timed CPU baseline of bench.py, `cpu_baseline.kind = "port"`) against oracle/tensors_oracle.py,
the numpy restatement that is itself pinned on the reference's golden vectors (tests/test_oracle_golden.py).

  c2_map_rows / c2_map_cols  == Tanh(Plus(Times(a, b), c)) with c broadcast per row (left-aligned, Tensors.scala:208-222)
                                or per column (c.reshape([1, cols])); one rounding per node.  `*` and `+` are bit-exact;
                                tanh differs between libm's tanhf and the oracle's float64-rounded tanh by <= 2 ulp --
                                the tolerance north_star states for fp32 elementwise work.
  c2_sum16                   == the sequential 16-lane `reduce` program of a CPU device (Tensors.scala:313-351), bit-exact.
CPU only: the library is built by `make -C oracle` (also done by __graft_entry__.build()).
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FP = ctypes.POINTER(ctypes.c_float)


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, stdout=subprocess.DEVNULL)
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "libc2oracle.so"))
    for name in ("c2_map_rows", "c2_map_cols"):
        getattr(lib, name).argtypes = [FP, FP, FP, FP, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
    lib.c2_sum16.argtypes = [FP, ctypes.c_size_t]
    lib.c2_sum16.restype = ctypes.c_float
    lib.c2_wang_uniform.argtypes = [FP, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_float, ctypes.c_float, ctypes.c_int]
    lib.c2_fused_sum.argtypes = [FP, FP, FP, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
    lib.c2_fused_sum.restype = ctypes.c_double
    return lib


def ptr(a):
    return a.ctypes.data_as(FP)


def ulp_distance(a, b):
    return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))


@pytest.mark.parametrize("rows,cols,threads", [(2, 3, 1), (7, 33, 3), (64, 257, 16), (300, 100, 5)])
def test_map_kernels_match_the_numpy_oracle(lib, rows, cols, threads):
    from oracle import tensors_oracle as O
    rng = np.random.RandomState(rows * 1000 + cols)
    a = (rng.standard_normal((rows, cols)) * 2).astype(np.float32)
    b = (rng.standard_normal((rows, cols)) * 2).astype(np.float32)
    for per_column in (False, True):
        c = rng.standard_normal(cols if per_column else rows).astype(np.float32)
        y = np.empty_like(a)
        (lib.c2_map_cols if per_column else lib.c2_map_rows)(ptr(a), ptr(b), ptr(c), ptr(y), rows, cols, threads)
        ct = O.tensor(c).reshape([1, cols]) if per_column else O.tensor(c)
        pre = np.asarray((O.tensor(a) * O.tensor(b) + ct).flat_array(), dtype=np.float32).reshape(rows, cols)
        want = np.asarray(O.tanh(O.tensor(a) * O.tensor(b) + ct).flat_array(), dtype=np.float32).reshape(rows, cols)
        # the fp32 product and sum are the same IEEE operations on both sides: check them bit for bit through atanh-free
        # recomputation of the C pipeline's intermediate
        assert (pre.view(np.uint32) == (a * b + (c[None, :] if per_column else c[:, None])).astype(np.float32).view(np.uint32)).all()
        assert ulp_distance(y, want).max() <= 2


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 31, 32, 1000, 4099, 65536 + 5])
def test_sum16_is_the_cpu_reduce_program_bit_for_bit(lib, n):
    from oracle import tensors_oracle as O
    x = np.random.RandomState(n).standard_normal(n).astype(np.float32)
    got = np.float32(lib.c2_sum16(ptr(x), n))
    want = np.float32(O.sum_cpu_order(x))
    assert got.view(np.uint32) == want.view(np.uint32)


@pytest.mark.parametrize("seed,n", [(1, 1), (2, 1000), (17, 70001)])
def test_wang_uniform_inputs_equal_the_reference_rng(lib, seed, n):
    """bench.py's CPU arm generates its inputs with the same Wang-hash uniform generator (Tensors.scala:98-117, :398-407)
    and the same `* 4 - 2` map as the GPU arm, so both arms run the same config on the same values."""
    from oracle import tensors_oracle as O
    out = np.empty(n, dtype=np.float32)
    lib.c2_wang_uniform(ptr(out), n, seed, 4.0, -2.0, 3)
    u = np.asarray(O.random([n], seed=seed).flat_array(), dtype=np.float32)
    want = (u * np.float32(4.0) + np.float32(-2.0)).astype(np.float32)
    assert (out.view(np.uint32) == want.view(np.uint32)).all()


def test_fused_all_core_mode_agrees_with_the_literal_mode(lib):
    """BASELINE.md section 5's best-effort CPU mode (map and sum fused, every core, fp64 partials) against the
    reference-literal materialise-then-sum16: same inputs, sums within the reductions bar (1e-5 relative)."""
    rows, cols = 96, 1031
    rng = np.random.RandomState(3)
    a = rng.standard_normal((rows, cols)).astype(np.float32)
    b = rng.standard_normal((rows, cols)).astype(np.float32)
    c = rng.standard_normal(rows).astype(np.float32)
    y = np.empty_like(a)
    lib.c2_map_rows(ptr(a), ptr(b), ptr(c), ptr(y), rows, cols, 4)
    literal = float(lib.c2_sum16(ptr(y), rows * cols))
    fused = float(lib.c2_fused_sum(ptr(a), ptr(b), ptr(c), rows, cols, 7))
    exact = float(y.astype(np.float64).sum())
    scale = float(np.abs(y).astype(np.float64).sum())
    assert abs(fused - exact) <= 1e-9 * scale
    assert abs(literal - exact) <= 1e-5 * scale
