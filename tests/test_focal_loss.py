"""N3: fused Gaussian-focal-loss consumer. PARITY UNPINNED — the reference documents the formula
(docs/intro.rst:7-26) but has no implementation. Checked against a float64 numpy evaluation on the ORACLE's target
heatmap: positives counted exactly, loss sum within 2e-6 relative, gradient within 1e-5 relative / 1e-6 absolute
(fp32 1-ulp log / 2-ulp division on the device vs float64), run-to-run bit-identical (fixed-order reduction), and against
torch autograd of the unfused formula on the heatmap our renderer draws."""
import numpy as np
import pytest

import _cabi
import workloads


def test_oracle_formula_known_answers(oracle):
    pred = np.array([[0.5, 0.25]], np.float32)
    tgt = np.array([[1.0, 0.0]], np.float32)
    loss, npos, grad = oracle.gaussian_focal_loss_numpy(pred, tgt, 2.0, 4.0, 0.0)
    want = -(0.5 ** 2) * np.log(0.5) - (0.25 ** 2) * np.log(0.75)
    assert abs(loss - want) < 1e-12 and npos == 1
    assert abs(grad[0, 0] - (2 * 0.5 * np.log(0.5) - 0.25 / 0.5)) < 1e-12
    assert abs(grad[0, 1] - (0.0625 / 0.75 - 2 * 0.25 * np.log(0.75))) < 1e-12


_LOG_POS_COEFFS = (0xbeffffeb, 0x3eaaaa7d, 0xbe800a40, 0x3e4cea14, 0xbe281974, 0x3e0f512f, 0xbe1d9426, 0x3e0fa12d)


def _log_pos_numpy(x):
    """`log_pos` of csrc/dh_gaussians.cu restated operation by operation (float32 FMAs emulated through float64: the
    product of two float32 is exact there)."""
    c = np.array(_LOG_POS_COEFFS, np.uint32).view(np.float32)
    fma = lambda a, b, d: (a.astype(np.float64) * b.astype(np.float64) + d.astype(np.float64)).astype(np.float32)
    ix = x.astype(np.float32).view(np.int32)
    e = (ix - np.int32(0x3F2AAAAB)) & np.int32(-8388608)
    m = (ix - e).view(np.float32)
    i = e.astype(np.float32) * np.float32(2.0 ** -23)
    f = (m - np.float32(1.0)).astype(np.float32)
    s = (f * f).astype(np.float32)
    q = np.full_like(f, c[7])
    for k in range(6, -1, -1):
        q = fma(q, f, np.full_like(f, c[k]))
    return fma(i, np.full_like(f, np.float32(0.693147182)), fma(q, s, f))


def test_log_pos_is_a_one_ulp_log():
    """the loss kernel's logarithm: coefficients as in the CUDA source, <= 1.1 ulp from float64 log on every float32
    in [0.5, 1.0001] (where 1 - pred + eps lives) and on samples of every binade; exactly 0 at 1"""
    import os

    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "accv-lab_b200", "csrc",
                            "dh_gaussians.cu")).read()
    for h in _LOG_POS_COEFFS:
        assert f"0x{h:08x}" in src, f"coefficient 0x{h:08x} not in the CUDA source: test and kernel diverged"
    rng = np.random.default_rng(0)
    xs = [np.arange(0x3F000000, 0x3F800400, dtype=np.uint32).view(np.float32),
          rng.integers(0x00800000, 0x7F7FFFFF, 2_000_000, dtype=np.int64).astype(np.uint32).view(np.float32)]
    for x in xs:
        got = _log_pos_numpy(x).astype(np.float64)
        want = np.log(x.astype(np.float64))
        ulp = np.abs(np.spacing(want.astype(np.float32)).astype(np.float64))
        ulp[ulp == 0] = 1.0
        assert float((np.abs(got - want) / ulp).max()) <= 1.1
    assert _log_pos_numpy(np.array([1.0], np.float32))[0] == 0.0


def test_argument_validation_needs_no_device(cabi):
    import ctypes

    one = ctypes.c_void_p(16)
    f = cabi.dh_gaussian_focal_loss
    assert f(None, one, one, one, one, 0, None, 1, 1, 1, 8, 8, 6.0, 0.0, 2.0, 4.0, 0.0, 0, None, one, one, 1 << 20) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert f(None, one, one, one, one, 0, None, 1, 1, 1, 8, 8, 6.0, 1.0, 2.0, 4.0, 0.0, 0, None, None, one, 1 << 20) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert cabi.dh_focal_loss_workspace_bytes(16, 10, 180, 180) >= 16 * 160 + 16


@pytest.mark.gpu
@pytest.mark.parametrize("name,nb,with_labels", [("cfg1", None, True), ("cfg2", None, True), ("cfg3", 12, True), ("cfg4", 2, False)])
def test_fused_loss_vs_oracle(cuda_dev, oracle, name, nb, with_labels):
    import torch
    from accvlab.draw_heatmap import b200

    w = workloads.make(name, B=nb)
    cfg = w.cfg
    ci, ri = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    shape = (w.B, cfg.C, cfg.H, cfg.W) if with_labels else (w.B, cfg.H, cfg.W)
    target = np.zeros(shape, np.float32)
    oracle.draw_batched(target, ci, ri, w.sizes, 6.0, 1.0, w.labels if with_labels else None)
    rng = np.random.default_rng(17)
    pred = rng.uniform(0.01, 0.99, shape).astype(np.float32)
    want_loss, want_pos, want_grad = oracle.gaussian_focal_loss_numpy(pred, target)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    sizes = dev(w.sizes)
    rb = lambda a: b200.RaggedBatch(dev(a), sample_sizes=sizes)
    p = dev(pred).requires_grad_(True)
    l0 = b200.launch_count()
    loss, npos = b200.gaussian_focal_loss(p, rb(ci), rb(ri), 6.0, 1.0, labels=rb(w.labels) if with_labels else None)
    assert b200.launch_count() - l0 == 2
    assert int(npos) == want_pos
    assert abs(float(loss.detach()) - want_loss) <= 2e-6 * abs(want_loss) + 1e-6
    (loss / max(int(npos), 1)).backward()
    g = p.grad.cpu().numpy().astype(np.float64) * max(want_pos, 1)
    assert np.all(np.abs(g - want_grad) <= 1e-5 * np.abs(want_grad) + 1e-6)
    # special-function-unit variant: same positives, loss within 1e-5 relative
    loss_f, npos_f = b200.gaussian_focal_loss(dev(pred), rb(ci), rb(ri), 6.0, 1.0, labels=rb(w.labels) if with_labels else None,
                                              fast_math=True)
    assert int(npos_f) == want_pos and abs(float(loss_f) - want_loss) <= 1e-5 * abs(want_loss) + 1e-6
    # run-to-run identical (fixed-order reduction)
    loss2, _ = b200.gaussian_focal_loss(dev(pred), rb(ci), rb(ri), 6.0, 1.0, labels=rb(w.labels) if with_labels else None)
    assert float(loss2.detach()) == float(loss.detach())


@pytest.mark.gpu
def test_fused_loss_equals_unfused_torch_on_our_heatmap(cuda_dev):
    """The same loss computed the unfused way: render the target with the product's FRESH renderer, then the formula
    with torch ops (fp32) and autograd — loss within 1e-5 relative, gradient within 1e-4 relative / 1e-6 absolute."""
    import torch
    from accvlab.draw_heatmap import b200

    w = workloads.make("cfg3", B=8)
    cfg = w.cfg
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    sizes, lab = dev(w.sizes), dev(w.labels)
    ci, ri = b200.radius_center(dev(w.centers_f), dev(w.bboxes_f), cfg.osf)
    rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
    target = b200.render_heatmap_batched(rb(ci), rb(ri), cfg.H, cfg.W, num_classes=cfg.C, labels=rb(lab))
    torch.manual_seed(3)
    pred = torch.rand(target.shape, device=cuda_dev) * 0.98 + 0.01
    p1 = pred.clone().requires_grad_(True)
    eps = 1e-12
    pos = target.eq(1)
    ref = torch.where(pos, -(p1 + eps).log() * (1 - p1) ** 2, -(1 - p1 + eps).log() * p1 ** 2 * (1 - target) ** 4).sum()
    ref.backward()
    p2 = pred.clone().requires_grad_(True)
    loss, npos = b200.gaussian_focal_loss(p2, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
    loss.backward()
    assert int(npos) == int(pos.sum())
    assert abs(float(loss.detach()) - float(ref.detach())) <= 1e-5 * abs(float(ref.detach()))
    assert torch.all((p2.grad - p1.grad).abs() <= 1e-4 * p1.grad.abs() + 1e-6)


@pytest.mark.gpu
def test_scale_by_device_scalar(cabi, cuda_dev):
    """backward helper: dst = src * (*scale) with the scalar in device memory; aligned / misaligned / tails / in place"""
    import torch

    for n, off in ((1, 0), (3, 0), (4, 0), (4099, 0), (4099, 1), (2048 * 256 + 5, 0), (100000, 3)):
        buf = torch.randn(n + 8, device=cuda_dev)
        src = buf[off:off + n]
        s = torch.tensor(-0.37, device=cuda_dev)
        dst = torch.full((n + 8,), 9.0, device=cuda_dev)
        out = dst[off:off + n]
        _cabi.check(cabi, cabi.dh_scale_by_device_scalar(_cabi.stream_ptr(), _cabi.ptr(out), _cabi.ptr(src), _cabi.ptr(s), n))
        assert torch.equal(out, src * s)
        assert bool((dst[:off] == 9.0).all()) and bool((dst[off + n:] == 9.0).all())
        want = src * s
        _cabi.check(cabi, cabi.dh_scale_by_device_scalar(_cabi.stream_ptr(), _cabi.ptr(src), _cabi.ptr(src), _cabi.ptr(s), n))
        assert torch.equal(src, want)
    assert cabi.dh_scale_by_device_scalar(None, None, None, None, 0) == _cabi.DH_OK
    assert cabi.dh_scale_by_device_scalar(None, None, None, None, 5) == _cabi.DH_ERR_INVALID_ARGUMENT
