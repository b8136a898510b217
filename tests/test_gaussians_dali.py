"""N2: the DALI `draw_gaussians` operator semantics as a second renderer mode.

CPU part: oracle/dh_oracle.c:dho_draw_gaussians (restatement of
/root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc:31-75,88-157) reproduces, BIT FOR BIT, the
outputs of the reference's own operator source compiled against oracle/dali_stub (tests/golden/ref_dali.npz, generated
by tests/golden/make_golden_dali.py), plus the analytic known answers the reference's own test checks
(bounding_box_to_heatmap_converter_test.py:100-136: peak == k at the centre, zero outside the radius window).
GPU part: dh_draw_gaussians through the C ABI and the Python API against the goldens and the oracle:
touched masks identical, values <= 1e-6 abs and <= 2 ulp (device expf vs the C library's)."""
import os

import numpy as np
import pytest

import _cabi

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_dali.npz")
CASES = ("typical", "integer_r", "tiny_r", "two_d", "neg_k", "near_int")


def _case(g, name):
    return {k: g[f"{name}_{k}"] for k in ("heat_in", "centers", "radii", "slice_ids", "active", "sizes", "k", "factor", "heat_out")}


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_reference_operator_bit_for_bit(oracle, name):
    c = _case(np.load(GOLD), name)
    heat = c["heat_in"].copy()
    oracle.draw_gaussians(heat, c["centers"], c["radii"], c["sizes"], c["k"], float(c["factor"]), c["slice_ids"], c["active"])
    assert np.array_equal(heat.view(np.uint32), c["heat_out"].view(np.uint32))


def test_oracle_known_answers(oracle):
    heat = np.zeros((1, 2, 21, 21), np.float32)
    centers = np.array([[[10, 10], [3, 4]]], np.int32)
    radii = np.array([[3.0, 2.5]], np.float32)
    oracle.draw_gaussians(heat, centers, radii, np.array([2]), [0.8, 1.0], 1.0 / 3.0, np.array([[0, 1]], np.int32),
                          np.array([[True, True]]))
    assert heat[0, 0, 10, 10] == np.float32(0.8) and heat[0, 1, 4, 3] == np.float32(1.0)
    ys, xs = np.nonzero(heat[0, 0])
    assert ys.min() == 7 and ys.max() == 13 and xs.min() == 7 and xs.max() == 13  # [floor(c-r), ceil(c+r)]
    ys, xs = np.nonzero(heat[0, 1])
    assert ys.min() == 1 and ys.max() == 7 and xs.min() == 0 and xs.max() == 6  # r=2.5 -> +-3
    # sigma = r*factor = 1 -> exp(-d2/2)
    assert abs(heat[0, 0, 10, 11] - 0.8 * np.exp(-0.5)) < 1e-7 and abs(heat[0, 0, 11, 11] - 0.8 * np.exp(-1.0)) < 1e-7
    # inactive objects and padded slots draw nothing
    heat2 = np.zeros((1, 2, 21, 21), np.float32)
    oracle.draw_gaussians(heat2, centers, radii, np.array([1]), [0.8, 1.0], 1.0 / 3.0, np.array([[0, 1]], np.int32),
                          np.array([[False, True]]))
    assert not heat2.any()
