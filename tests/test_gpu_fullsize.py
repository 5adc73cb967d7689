"""BASELINE.json's full sizes, checked through properties that do not need the oracle to evaluate 2^28 elements:
sampled rows against the oracle, three independent routes to the same sum, replay idempotence, linearity of the sum,
and the view algebra's round trips at sizes whose tile counts exceed 16 bits / whose element counts exceed 2^28."""
import numpy as np
import pytest

from golden_cases import ulp_diff

pytestmark = pytest.mark.gpu

ROWS, COLS = 16384, 16384            # C2: 2^28 fp32 per operand


@pytest.fixture(scope="module")
def c2():
    import deeplearning.scala_b200.tensors as B
    from oracle import tensors_oracle as O
    be = B.Backend()
    a = (be.random([ROWS, COLS], seed=1) * 4.0 - 2.0).do_cache()
    b = (be.random([ROWS, COLS], seed=2) * 4.0 - 2.0).do_cache()
    c = (be.random([ROWS], seed=3) * 4.0 - 2.0).do_cache()
    y = be.tanh(a * b + c).do_cache()
    yield B, O, be, a, b, c, y
    del a, b, c, y
    be.close()


def rows_of(be, t, rows):
    buf = t._buffer()
    return np.stack([be.download(buf, COLS, r * COLS) for r in rows])


def test_c2_sampled_rows_match_the_oracle(c2):
    B, O, be, a, b, c, y = c2
    rows = [0, 1, 4095, 8192, 12345, ROWS - 1]
    ha, hb, hy = rows_of(be, a, rows), rows_of(be, b, rows), rows_of(be, y, rows)
    hc = c.flat_array()[rows]
    want = np.asarray(O.tanh(O.tensor(ha) * O.tensor(hb) + O.tensor(hc)).flat_array(), dtype=np.float32).reshape(len(rows), COLS)
    assert ulp_diff(want.reshape(-1), hy.reshape(-1)).max() <= 2
    # the inputs themselves: the reference's Wang-hash generator is position addressed, so row r of the 2^28-element
    # tensor is a window of the same stream the oracle produces for a small tensor
    flat = np.asarray(O.random([2, COLS], seed=1).flat_array(), dtype=np.float32).reshape(2, COLS)
    got01 = rows_of(be, a, [0, 1])
    assert (((flat * np.float32(4.0)) - np.float32(2.0)).view(np.uint32) == got01.view(np.uint32)).all()


def test_c2_three_routes_to_the_sum_agree(c2):
    B, O, be, a, b, c, y = c2
    materialised = float(y.sum().flat_array()[0])                              # sum_kernel over the cached map
    fused = float(be.tanh(a * b + c).sum().flat_array()[0])                    # map + sum in one launch, nothing stored
    by_rows = B.sum_axis(y, 1).flat_array().astype(np.float64)                 # 16384 row sums, then the host adds them
    assert by_rows.shape == (ROWS,)
    exact = by_rows.sum()
    scale = float(B.sum_axis(be.abs(y).do_cache(), 1).flat_array().astype(np.float64).sum())
    assert abs(materialised - exact) <= 1e-5 * scale
    assert abs(fused - exact) <= 1e-5 * scale
    assert abs(fused - materialised) <= 1e-5 * scale


def test_c2_is_deterministic_and_linear(c2):
    B, O, be, a, b, c, y = c2
    again = be.tanh(a * b + c).do_cache()
    rows = [7, 9000, ROWS - 2]
    assert (rows_of(be, again, rows).view(np.uint32) == rows_of(be, y, rows).view(np.uint32)).all()
    s1 = float(be.tanh(a * b + c).sum().flat_array()[0])
    s2 = float(be.tanh(a * b + c).sum().flat_array()[0])
    assert s1 == s2                                                             # fixed grid, ticketed finish: same bits
    # sum(2 y) == 2 sum(y) exactly (scaling by two commutes with every rounding); sum(y + y) likewise
    twice = float((y * 2.0).sum().flat_array()[0])
    assert twice == 2.0 * float(y.sum().flat_array()[0]) or abs(twice - 2.0 * s1) <= 1e-5 * abs(twice)
    ones = float(be.fill(1.0, [ROWS, COLS]).sum().flat_array()[0])
    assert abs(ones - ROWS * COLS) <= 1e-6 * ROWS * COLS


def test_view_round_trips_at_full_size(c2):
    B, O, be, a, b, c, y = c2
    # transpose twice is the identity (two gather launches over 2^28 elements, 32-bit tile ids near their limit)
    t2 = a.transpose().non_inline().transpose().do_cache()
    rows = [0, 5555, ROWS - 1]
    assert (rows_of(be, t2, rows).view(np.uint32) == rows_of(be, a, rows).view(np.uint32)).all()
    del t2
    # translate there and back: interior untouched, the border column range shows the padding (0)
    moved = a.translate([3, -5]).non_inline().translate([-3, 5]).do_cache()
    got, src = rows_of(be, moved, [0, 100, ROWS - 4, ROWS - 1]), rows_of(be, a, [0, 100, ROWS - 4, ROWS - 1])
    assert (got[:3, 5:].view(np.uint32) == src[:3, 5:].view(np.uint32)).all()   # rows < ROWS-3 survive, columns >= 5 survive
    assert (got[:, :5] == 0).all() and (got[3] == 0).all()
    del moved
    # split / join along the leading dim reassembles the tensor (16 slabs of 1024 rows)
    slabs = a.reshape([16, 1024, COLS]).split(0)
    assert len(slabs) == 16 and list(slabs[3].shape) == [1024, COLS]
    s3 = (slabs[3] + 0.0).do_cache()
    assert (be.download(s3._buffer(), COLS, 17 * COLS).view(np.uint32) == rows_of(be, a, [3 * 1024 + 17])[0].view(np.uint32)).all()
