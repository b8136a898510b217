"""Edge cases of the widened rows (N1-N4) on the GPU: empty batches, zero targets, all-empty lists, int32 sizes,
missing optional inputs — the cases the reference's own tests exercise for the main path (empty / ragged / padded)."""
import numpy as np
import pytest

import _cabi

pytestmark = pytest.mark.gpu


def _dev(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def test_draw_gaussians_empty_and_optional_inputs(cabi, cuda_dev, oracle):
    import torch
    from accvlab.draw_heatmap import b200

    H, W = 24, 32
    # no targets at all: FRESH writes zeros, in place leaves the input alone, out of place copies it
    z = torch.zeros((3, 0, 2), dtype=torch.int32, device=cuda_dev)
    r = torch.zeros((3, 0), dtype=torch.float32, device=cuda_dev)
    k = torch.ones(2, device=cuda_dev)
    out = b200.draw_gaussians(None, None, None, z, r, k, 0.3, shape=(3, 2, H, W))
    assert out.shape == (3, 2, H, W) and not out.any()
    bg = torch.rand((3, 2, H, W), device=cuda_dev)
    assert torch.equal(b200.draw_gaussians(bg, None, None, z, r, k, 0.3), bg)
    keep = bg.clone()
    b200.draw_gaussians(bg, None, None, z, r, k, 0.3, out=bg)
    assert torch.equal(bg, keep)
    # empty batch
    e = b200.draw_gaussians(None, None, None, torch.zeros((0, 4, 2), dtype=torch.int32, device=cuda_dev),
                            torch.zeros((0, 4), device=cuda_dev), k, 0.3, shape=(0, 2, H, W))
    assert e.shape == (0, 2, H, W)
    # int32 sizes, no slice ids (single plane), no active flags, [B,H,W] heatmap, all sizes zero for one sample
    rng = np.random.default_rng(0)
    B, T = 4, 9
    centers = np.stack([rng.integers(0, W, (B, T)), rng.integers(0, H, (B, T))], 2).astype(np.int32)
    radii = (1 + rng.random((B, T)) * 4).astype(np.float32)
    sizes = np.array([T, 0, 3, 7], np.int32)
    rb = lambda a: b200.RaggedBatch(_dev(a, cuda_dev), sample_sizes=_dev(sizes, cuda_dev))
    got = b200.draw_gaussians(None, None, None, rb(centers), rb(radii), [0.9], 0.4, shape=(B, H, W))
    want = np.zeros((B, H, W), np.float32)
    oracle.draw_gaussians(want, centers, radii, sizes, [0.9], 0.4)
    assert np.array_equal(got.cpu().numpy() > 0, want > 0) and np.abs(got.cpu().numpy() - want).max() <= 1e-6
    assert not got[1].any()
    # every object inactive
    inactive = np.zeros((B, T), bool)
    got = b200.draw_gaussians(None, rb(inactive), None, rb(centers), rb(radii), [0.9], 0.4, shape=(B, H, W))
    assert not got.any()


def test_pack_edges(cuda_dev):
    import torch
    from accvlab.draw_heatmap import b200

    empties = [torch.zeros((0, 4), device=cuda_dev) for _ in range(3)]
    rb = b200.combine_data(empties)
    assert tuple(rb.tensor.shape) == (3, 0, 4) and rb.sample_sizes.tolist() == [0, 0, 0]
    rb = b200.combine_data(empties, max_sample_size=5)
    assert tuple(rb.tensor.shape) == (3, 5, 4) and not rb.tensor.any()
    one = b200.combine_data([torch.arange(6, device=cuda_dev, dtype=torch.int16).reshape(3, 2)])
    assert one.tensor.cpu().tolist() == [[[0, 1], [2, 3], [4, 5]]] and one.sample_sizes.tolist() == [3]
    with pytest.raises(RuntimeError):
        b200.combine_data([torch.zeros((2, 2), device=cuda_dev)], max_sample_size=1)
    with pytest.raises(RuntimeError):
        b200.combine_data([torch.zeros((2, 2))])  # CPU tensors: there is no CPU path
    c, r = b200.pack_radius_center([torch.zeros((0, 2), device=cuda_dev)], [torch.zeros((0, 4), device=cuda_dev)], 4.0,
                                   max_sample_size=3)
    assert tuple(c.tensor.shape) == (1, 3, 2) and not c.tensor.any() and not r.tensor.any()


def test_converter_and_loss_and_radius_edges(cabi, cuda_dev, oracle):
    import torch
    from accvlab.draw_heatmap import b200

    # converter: T == 0, B == 0
    out = b200.boxes_to_gaussian_params(torch.zeros((2, 0, 4), device=cuda_dev), torch.ones((2, 2), device=cuda_dev), (8, 8))
    assert tuple(out["radii"].tensor.shape) == (2, 0) and tuple(out["center"].tensor.shape) == (2, 0, 2)
    out = b200.boxes_to_gaussian_params(torch.zeros((0, 3, 4), device=cuda_dev), torch.ones((0, 2), device=cuda_dev), (8, 8))
    assert tuple(out["is_active"].tensor.shape) == (0, 3)
    # loss: no targets -> every pixel is a negative with target 0: sum(-p^2 log(1 - p + eps)); empty batch -> 0
    pred = torch.rand((2, 3, 16, 20), device=cuda_dev) * 0.9 + 0.05
    zc = torch.zeros((2, 0, 2), dtype=torch.int32, device=cuda_dev)
    zr = torch.zeros((2, 0), dtype=torch.int32, device=cuda_dev)
    zl = torch.zeros((2, 0), dtype=torch.int32, device=cuda_dev)
    sizes = torch.zeros(2, dtype=torch.int64, device=cuda_dev)
    rbz = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
    loss, npos = b200.gaussian_focal_loss(pred, rbz(zc), rbz(zr), labels=rbz(zl))
    want, wpos, _ = oracle.gaussian_focal_loss_numpy(pred.cpu().numpy(), np.zeros(pred.shape, np.float32))
    assert int(npos) == 0 == wpos and abs(float(loss) - want) <= 2e-6 * want
    loss, npos = b200.gaussian_focal_loss(torch.zeros((0, 16, 20), device=cuda_dev), torch.zeros((0, 4, 2), dtype=torch.int32, device=cuda_dev),
                                          torch.zeros((0, 4), dtype=torch.int32, device=cuda_dev))
    assert float(loss) == 0.0 and int(npos) == 0
    # general exponents (powf variant) and a width that is not a multiple of 4 (scalar variant)
    rng = np.random.default_rng(5)
    B, T, H, W = 3, 6, 17, 23
    centers = np.stack([rng.integers(0, W, (B, T)), rng.integers(0, H, (B, T))], 2).astype(np.int32)
    radii = rng.integers(1, 5, (B, T)).astype(np.int32)
    sz = np.array([6, 2, 0], np.int64)
    target = np.zeros((B, H, W), np.float32)
    oracle.draw_batched(target, centers, radii, sz, 6.0, 1.0)
    p = (rng.random((B, H, W)) * 0.9 + 0.05).astype(np.float32)
    rb = lambda a: b200.RaggedBatch(_dev(a, cuda_dev), sample_sizes=_dev(sz, cuda_dev))
    for alpha, gamma in ((2.0, 4.0), (1.5, 3.0)):
        loss, npos = b200.gaussian_focal_loss(_dev(p, cuda_dev), rb(centers), rb(radii), alpha=alpha, gamma=gamma)
        want, wpos, _ = oracle.gaussian_focal_loss_numpy(p, target, alpha, gamma)
        assert int(npos) == wpos and abs(float(loss) - want) <= 1e-5 * abs(want)
    # gaussian_radius: empty input
    assert b200.gaussian_radius(torch.zeros((0, 2), device=cuda_dev)).numel() == 0


def test_classwise_with_zero_targets_still_covers_every_plane(cuda_dev):
    """Regression: an empty labels tensor has a NULL data pointer, which the C ABI reads as "one plane per sample" —
    the binding must keep the class-wise layout (FRESH: all B*C planes zeroed; loss: all B*C planes summed)."""
    import torch
    from accvlab.draw_heatmap import b200

    B, C, H, W = 3, 5, 12, 16
    sizes = torch.zeros(B, dtype=torch.int64, device=cuda_dev)
    rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
    zc = torch.zeros((B, 0, 2), dtype=torch.int32, device=cuda_dev)
    zr = torch.zeros((B, 0), dtype=torch.int32, device=cuda_dev)
    zl = torch.zeros((B, 0), dtype=torch.int32, device=cuda_dev)
    out = torch.full((B, C, H, W), 7.0, device=cuda_dev)
    b200.render_heatmap_batched(rb(zc), rb(zr), labels=rb(zl), out=out)
    assert not out.any()
    out = torch.full((B, C, H, W), 7.0, device=cuda_dev)
    b200.render_heatmap_from_boxes(torch.zeros((B, 0, 2), device=cuda_dev), torch.zeros((B, 0, 4), device=cuda_dev), 4.0, sizes,
                                   labels=zl, out=out)
    assert not out.any()
