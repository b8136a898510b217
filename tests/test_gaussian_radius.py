"""N4: CenterNet / CenterPoint gaussian_radius(det_size, min_overlap). PARITY UNPINNED: the reference
has no such function and neither CenterNet, CenterPoint nor mmdet is installed; the oracle restates the
published algorithm and is checked here against scalar Python-float evaluations of the same published formulas; the
GPU kernel must then equal the oracle bit for bit (float64 arithmetic, IEEE sqrt, no FMA contraction)."""
import math

import numpy as np
import pytest

import _cabi


def _scalar(h, w, mo, variant):
    """The published function, written the way upstream writes it (Python floats)."""
    a1 = 1
    b1 = (h + w)
    c1 = w * h * (1 - mo) / (1 + mo)
    sq1 = math.sqrt(b1 ** 2 - 4 * a1 * c1)
    a2 = 4
    b2 = 2 * (h + w)
    c2 = (1 - mo) * w * h
    sq2 = math.sqrt(b2 ** 2 - 4 * a2 * c2)
    a3 = 4 * mo
    b3 = -2 * mo * (h + w)
    c3 = (mo - 1) * w * h
    sq3 = math.sqrt(b3 ** 2 - 4 * a3 * c3)
    if variant == "cornernet":
        return min((b1 + sq1) / 2, (b2 + sq2) / 2, (b3 + sq3) / 2)
    return min((b1 - sq1) / (2 * a1), (b2 - sq2) / (2 * a2), (b3 + sq3) / (2 * a3))


@pytest.mark.parametrize("variant", ["cornernet", "mmdet"])
def test_oracle_against_scalar_python(oracle, variant):
    rng = np.random.default_rng(3)
    hw = (rng.random((500, 2)) * 120 + 0.5).astype(np.float32)
    for mo in (0.1, 0.5, 0.7):
        ri, rf = oracle.gaussian_radius_numpy(hw, mo, variant, min_radius=2)
        want = np.array([_scalar(float(h), float(w), mo, variant) for h, w in hw])
        assert np.allclose(rf, want, rtol=1e-6, atol=1e-6)  # (1-mo)*w*h vs (1-mo)*(w*h): last-bit differences only
        exp_i = np.maximum(np.trunc(want).astype(np.int64), 2)
        assert (ri != exp_i).mean() < 0.01
    # a known value: 10 x 10 box, min_overlap 0.5 -> CornerNet form: r3 = (-10 + sqrt(100 + 400)) / 2 ... min is r3
    ri, rf = oracle.gaussian_radius_numpy(np.array([[10, 10]], np.float32), 0.5, "cornernet")
    assert abs(float(rf[0]) - _scalar(10.0, 10.0, 0.5, "cornernet")) < 1e-6 and ri[0] == int(_scalar(10.0, 10.0, 0.5, "cornernet"))


def test_argument_validation_needs_no_device(cabi):
    import ctypes

    one = ctypes.c_void_p(16)
    assert cabi.dh_gaussian_radius(None, one, 0.5, 7, 0, 4, one, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert cabi.dh_gaussian_radius(None, one, 0.5, 0, 0, -1, one, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert cabi.dh_gaussian_radius(None, None, 0.5, 0, 0, 0, None, None) == _cabi.DH_OK


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["cornernet", "mmdet"])
def test_gpu_equals_oracle_bit_for_bit_and_feeds_the_renderer(cuda_dev, oracle, variant):
    import torch
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap import b200

    rng = np.random.default_rng(9)
    B, T = 6, 40
    hw = (rng.random((B, T, 2)) * 60 + 0.25).astype(np.float32)
    hw[0, 0] = [0, 0]
    hw[0, 1] = [1e-3, 500]
    for mo in (0.1, 0.5):
        ri, rf = b200.gaussian_radius(torch.from_numpy(hw).to(cuda_dev), mo, variant, min_radius=1, return_float=True)
        wi, wf = oracle.gaussian_radius_numpy(hw, mo, variant, min_radius=1)
        assert np.array_equal(ri.cpu().numpy().reshape(-1), wi)
        assert np.array_equal(rf.cpu().numpy().reshape(-1).view(np.uint32), wf.view(np.uint32))
    # into the drop-in renderer
    centers = np.stack([rng.integers(0, 128, (B, T)), rng.integers(0, 128, (B, T))], 2).astype(np.int32)
    sizes = rng.integers(0, T + 1, B).astype(np.int64)
    heat = torch.zeros((B, 128, 128), device=cuda_dev)
    rb = lambda t: b200.RaggedBatch(t, sample_sizes=torch.from_numpy(sizes).to(cuda_dev))
    dh.draw_heatmap_batched(heat, rb(torch.from_numpy(centers).to(cuda_dev)), rb(ri), 6.0, 1.0)
    want = np.zeros((B, 128, 128), np.float32)
    oracle.draw_batched(want, centers, wi.reshape(B, T), sizes, 6.0, 1.0)
    assert np.array_equal(heat.cpu().numpy() > 0, want > 0) and np.abs(heat.cpu().numpy() - want).max() <= 1e-6
