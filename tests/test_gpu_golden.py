"""The reference's own specs (TensorsSpec, cpuSpec, TensorLayersSpec, CumulativeTensorLayersSpec; SURVEY.md section 8c,
G1-G16) run against the CUDA path through the C ABI.  Same transcription that pins the oracle in test_oracle_golden.py."""
import pytest

import golden_cases as GC

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def B():
    import deeplearning.scala_b200.tensors as T
    T.default_backend().ctx          # fails loudly without an sm_100 device / built library
    return T


@pytest.fixture(scope="module")
def P():
    from deeplearning.scala_b200.plugins import Plugins
    return Plugins


@pytest.mark.parametrize("case", GC.TENSOR_CASES, ids=lambda f: f.__name__)
def test_tensor_golden(case, B):
    case(B)


@pytest.mark.parametrize("case", GC.PLUGIN_CASES, ids=lambda f: f.__name__)
def test_plugin_golden(case, B, P):
    case(B, P)


def test_native_library_is_the_one_in_tree(B):
    import deeplearning.scala_b200._native as N
    assert N.LIB_PATH.endswith("deeplearning/scala_b200/csrc/libdlsb200.so")
    before = B.default_backend().stats()["kernel_launches"]
    assert B.fill(1.0, [4]).flat_array().tolist() == [1.0] * 4
    assert B.default_backend().stats()["kernel_launches"] == before + 1
