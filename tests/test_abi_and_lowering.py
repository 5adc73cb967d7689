"""CPU-side checks (no GPU needed): the C-ABI library loads and exports every symbol include/dls_b200.h declares, the
product path fails loudly without a device, the wire format realises the reference's structural tree equality, and the
host-side lowering picks the intended kernel family.  No compute entry point is called here."""
import ctypes
import os
import re
import struct

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def native():
    import __graft_entry__ as G
    if not os.path.exists(os.path.join(G.CSRC, "libdlsb200.so")):
        G.build()
    import deeplearning.scala_b200._native as N
    return N


@pytest.fixture(scope="module")
def T(native):
    import deeplearning.scala_b200.tensors as T
    return T


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol(native):
    header = open(os.path.join(ROOT, "include", "dls_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(dls_[a-z0-9_]+)\s*\(", header)) - {"dls_event_cb"}
    assert len(declared) >= 35
    lib = ctypes.CDLL(native.LIB_PATH)
    missing = [name for name in sorted(declared) if not hasattr(lib, name)]
    assert not missing, missing
    assert set(native.EXPORTS) == declared, set(native.EXPORTS) ^ declared
    assert lib.dls_abi_version() == 1


def test_no_cpu_fallback(native, T):
    if _has_gpu():
        pytest.skip("a GPU is present")
    be = T.Backend()
    with pytest.raises(native.NativeError) as err:
        be.fill(1.0, [4]).flat_array()
    assert err.value.status == native.DLS_ERR_NO_DEVICE
    assert "no CPU fallback" in str(err.value)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "deeplearning", "scala_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("oracle definition", "").replace("pinned by the oracle", "") \
                    .replace("by the oracle)", ""), os.path.join(dirpath, f)


# ---- wire format == structural equality of Trees.scala:68-178 ---------------------------------------------------------


def _blob(T, t):
    from deeplearning.scala_b200.trees import serialize
    return serialize([t.closure()], t.shape)[0]


def test_alpha_equivalent_trees_share_a_blob(T):
    a1, b1 = T.tensor([[1.0, 2.0], [3.0, 4.0]]), T.tensor([[5.0, 6.0], [7.0, 8.0]])
    a2, b2 = T.tensor([[9.0, 9.0], [9.0, 9.0]]), T.tensor([[0.0, 1.0], [2.0, 3.0]])
    assert _blob(T, T.tanh(a1 * b1 + a1)) == _blob(T, T.tanh(a2 * b2 + a2))       # same sharing pattern
    assert _blob(T, T.tanh(a1 * b1 + a1)) != _blob(T, T.tanh(a1 * b1 + b1))       # different sharing pattern
    assert _blob(T, a1 + T.scalar(1.0)) != _blob(T, a1 + T.scalar(2.0))           # different literal
    assert _blob(T, T.tensor([1.0, 2.0]) + T.scalar(1.0)) != _blob(T, T.tensor([1.0, 2.0, 3.0]) + T.scalar(1.0))  # shape
    assert _blob(T, T.tensor([1.0], padding=1.0).translate([1])) != _blob(T, T.tensor([1.0], padding=2.0).translate([1]))
    assert _blob(T, a1.translate([1, 0])) != _blob(T, a1.translate([0, 1]))       # different matrix


def test_wire_header_layout(T):
    t = T.fill(42.0, [2, 3, 5])
    blob = _blob(T, t)
    words = struct.unpack("<%dI" % (len(blob) // 4), blob)
    assert words[0] == 0x54534C44 and words[1] == 1
    assert words[2:6] == (0, 1, 1, 3) and words[6:9] == (2, 3, 5)
    assert words[9] == 1 and struct.unpack("<f", struct.pack("<I", words[10]))[0] == 42.0


# ---- lowering decisions ------------------------------------------------------------------------------------------------


def _plan(T, t, roots=None):
    return t.backend.describe_plan(roots or [t.closure()], t.shape)


def test_contiguous_and_broadcast_leaves_lower_to_the_vector_kernel(T):
    a, b = T.random([1024, 4096], 1), T.random([1024, 4096], 2)
    row, col = T.random([1024], 3), T.random([4096], 4).reshape([1, 4096])
    d = _plan(T, T.tanh(a * b + row))
    assert "kind=vec" in d and "outer=1" in d and "stream:p0" in d and "bcast:p2" in d and "mac=1" in d
    d = _plan(T, T.tanh(a * b + col))
    assert "kind=vec" in d and "vec:p2" in d
    d = _plan(T, a * b + a)
    assert "kind=vec" in d and "outer=0" in d and "dims=[4194304]" in d       # fully collapsed to one flat dim
    d = _plan(T, T.fill(1.0, [7, 9]) + T.scalar(2.0))
    assert "kind=vec" in d and "dims=[63]" in d                               # ragged tail handled by the flat kernel


def test_views_lower_to_the_gather_kernel_with_the_cheapest_leaf_mode(T):
    x = T.random([8, 12, 16], 1)
    assert "strided:p0" in _plan(T, x.permute([2, 0, 1]))                      # proven in-bounds -> no checks
    assert "mono:p0" in _plan(T, x.translate([1, 0, -2]))                      # needs bounds checks + padding
    assert "generic:p0" in _plan(T, x.scale([5, 12, 16]))                      # 8/5 = 1.6 -> double arithmetic
    assert "kind=vec" in _plan(T, x.scale([4, 12, 16]))                        # 8/4 = 2 -> integer, in-bounds, row stride 384
    assert "strided:p0" in _plan(T, x.scale([8, 12, 8]))                       # 16/8 = 2 on the inner dim: stride-2 walk
    assert "kind=vec" in _plan(T, x.split(0)[3] + T.scalar(1.0))               # a slice of the leading dim is contiguous
    s = x.split(2)[5]
    assert "kind=gather" in _plan(T, s + T.scalar(1.0))                        # stride-16 walk


def test_dyadic_scale_rows_take_the_32_bit_gather_path(T):
    x = T.random([8, 12, 16], 1)
    assert "kind=gather4 " in _plan(T, x.scale([16, 24, 16]))                  # 0.5 on two outer dims, 128-bit inner loads
    assert "kind=gather1 " in _plan(T, x.scale([8, 12, 32]))                   # 0.5 along the inner dim: one element per lane
    assert "kind=gather4 " in _plan(T, x.scale([64, 12, 16]))                  # 0.125
    assert "kind=gather " in _plan(T, x.scale([24, 12, 16]))                   # 1/3 -> "0.333": stays on double arithmetic
    assert "kind=gather4 " in _plan(T, x.translate([1, 0, 0]).scale([16, 12, 16]))   # (g * 0.5 - 1), negative numerators


def test_shared_subtrees_get_slots_and_macs_are_fused(T):
    a, b, c = (T.random([64, 64], s) for s in (1, 2, 3))
    t2 = a + b
    d = _plan(T, (t2 * t2 + c) - (a - c) * (b + c))
    assert "slots=0" not in d
    acc = a * b
    for _ in range(10):
        acc = acc + a * c
    d = _plan(T, acc)
    assert "mac=10" in d and "slots=0" in d


def test_deep_left_fold_does_not_recurse(T):
    a = T.random([16], 1)
    acc = a
    for _ in range(20000):
        acc = acc + a
    d = _plan(T, acc)                         # ... and the 20000 identical terms fold back into one device loop
    assert "code=3 " in d and "loops=[addx20000:lin,]" in d


def test_malformed_blobs_are_rejected(native):
    text = ctypes.create_string_buffer(256)
    for blob in (b"", b"\0" * 24, struct.pack("<6I", 0x54534C44, 2, 0, 1, 1, 0), struct.pack("<6I", 0x54534C44, 1, 0, 1, 1, 9)):
        rc = native.lib.dls_plan_describe(None, blob, len(blob), text, len(text))
        assert rc in (native.DLS_ERR_INVALID, native.DLS_ERR_UNSUPPORTED)
        assert native.last_error()


# ---- host algebra (no device work) ---------------------------------------------------------------------------------------


def test_left_aligned_auto_broadcast(T):
    assert T.auto_broadcast_shape([2, 3], [2]) == [2, 3]
    assert T.auto_broadcast_shape([2, 1, 4], [2, 5]) == [2, 5, 4]
    with pytest.raises(ValueError):
        T.auto_broadcast_shape([2, 3], [3])


def test_checkpoint_rule_composes_against_the_original(T):
    x = T.tensor(np.zeros((4, 5), dtype=np.float32))
    v = x.translate([1, 0]).translate([-1, 0])
    assert v.checkpoint is x and np.array_equal(v.matrix, np.array([[1.0, 0, 0], [0, 1.0, 0]]))
    w = (x + T.scalar(0.0)).translate([1, 0])
    assert w.checkpoint is not x and isinstance(w.checkpoint, T.InlineTensor)
    assert isinstance(T.fill(1.0, [2]).broadcast([2, 3]), T.FillTensor)


def test_ndat_matches_oracle():
    from deeplearning.scala_b200 import ndat
    from oracle import tensors_oracle as O
    rng = np.random.RandomState(0)
    for _ in range(20):
        m01 = rng.randint(-5, 6, size=(3, 4)).astype(np.float64)
        m12 = rng.randint(-5, 6, size=(2, 4)).astype(np.float64)
        assert np.array_equal(ndat.pre_concatenate(m01, m12), O.ndat_pre_concatenate(m01, m12))
        assert np.array_equal(ndat.concatenate(m12, m01), O.ndat_concatenate(m12, m01))


def test_float_formatting_matches_oracle():
    from deeplearning.scala_b200._format import float_to_string
    from oracle.tensors_oracle import java_float_to_string
    for v in [0.0, -0.0, 1.0, 1e7, 9999999.0, 1.4561316, 1e-3, 9.99e-4, 3.4e38, 1e-45, float("nan"), float("inf"), 100053.0]:
        assert float_to_string(v) == java_float_to_string(v)
