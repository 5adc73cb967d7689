"""Pins the CPU oracle against the reference's own golden vectors."""
import numpy as np
import pytest

from oracle import plugins_oracle as PO
from oracle import tensors_oracle as O

import golden_cases as GC


@pytest.mark.parametrize("case", GC.TENSOR_CASES, ids=lambda f: f.__name__)
def test_tensor_golden(case):
    case(O)


@pytest.mark.parametrize("case", GC.PLUGIN_CASES, ids=lambda f: f.__name__)
def test_plugin_golden(case):
    case(O, PO.Plugins)


def test_g17_affine_concatenate_matches_awt_semantics():
    # NDATSpec:23-56: preConcatenate(m01, m12) == AffineTransform(m12).concatenate(m01) i.e. apply m01 first.
    rng = np.random.RandomState(0)
    for _ in range(20):
        m01 = rng.randint(-5, 6, size=(2, 3)).astype(np.float64)
        m12 = rng.randint(-5, 6, size=(2, 3)).astype(np.float64)
        m02 = O.ndat_pre_concatenate(m01, m12)
        assert (O.ndat_concatenate(m12, m01) == m02).all()
        for _ in range(5):
            v = rng.randint(-9, 10, size=2).astype(np.float64)
            want = O.ndat_transform(m12, O.ndat_transform(m01, v))
            assert (O.ndat_transform(m02, v) == want).all()


def test_sum_orders_agree_and_match_float64():
    rng = np.random.RandomState(1)
    x = rng.standard_normal(5000).astype(np.float32)
    ref = float(x.astype(np.float64).sum())
    for fn in (O.sum_cpu_order, lambda a: O.sum_gpu_order(a, 64, 8)):
        assert abs(float(fn(x)) - ref) <= 1e-5 * max(1.0, np.abs(x).sum())


def test_translate_roundtrip_keeps_borders_checkpoint_rule():
    # bounds are only checked against the ORIGINAL checkpoint.
    t = O.tensor([[1.0, 2.0], [3.0, 4.0]], padding=9.0)
    assert t.translate([1, 0]).translate([-1, 0]).to_string() == "[[1.0,2.0],[3.0,4.0]]"
    inline = (t + O.scalar(0.0))                        # InlineTensor => becomes a checkpoint, borders lost
    assert inline.translate([1, 0]).non_inline().translate([-1, 0]).to_string() == "[[1.0,2.0],[9.0,9.0]]"


def test_scale_truncates_and_formats_coefficients():
    t = O.tensor([0.0, 1.0, 2.0, 3.0, 4.0, 5.0])
    assert t.scale([3]).to_string() == "[0.0,2.0,4.0]"
    assert t.scale([12]).to_string() == "[0.0,0.0,1.0,1.0,2.0,2.0,3.0,3.0,4.0,4.0,5.0,5.0]"
    # 6/4 = 1.5 -> (int)(g*1.5) = 0,1,3,4
    assert t.scale([4]).to_string() == "[0.0,1.0,3.0,4.0]"
    # 3/7 = 0.428571.. is PRINTED as 0.429 by DecimalFormat (KB:14-32): (int)(g*0.429)
    t3 = O.tensor([10.0, 11.0, 12.0])
    assert t3.scale([7]).to_string() == "[10.0,10.0,10.0,11.0,11.0,12.0,12.0]"


def test_left_aligned_broadcast():
    a = O.tensor([[1.0, 2.0, 3.0], [4.0, 5.0, 6.0]])
    b = O.tensor([10.0, 20.0])
    assert (a + b).to_string() == "[[11.0,12.0,13.0],[24.0,25.0,26.0]]"
    with pytest.raises(ValueError):
        a + O.tensor([1.0, 2.0, 3.0])
