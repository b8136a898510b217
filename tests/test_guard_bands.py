"""Out-of-bounds check of every writing kernel with our own instruments (compute-sanitizer is not available on the
GPU pool): each output lives in the middle of a larger allocation whose borders are filled with a sentinel; after the
call the borders must be untouched and the payload must equal the same call into a plain tensor. Shapes walk every
kernel family of the renderer (main kernel, wide-row kernel, generic pixel-owner kernel, misaligned base, 16-bit
planes), objects hang over every map edge, and the last plane ends exactly at the guard."""
import numpy as np
import pytest

import _cabi

pytestmark = pytest.mark.gpu

GUARD = 4096  # elements on either side
SENTINEL = 12345.0


def _up(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


class Guarded:
    """view of `shape` inside a sentinel-filled allocation, `shift` elements off the 16-byte grid"""

    def __init__(self, shape, dtype, dev, shift=0):
        import torch

        n = int(np.prod(shape))
        self.buf = torch.full((GUARD + shift + n + GUARD,), SENTINEL, dtype=dtype, device=dev)
        self.lo, self.hi = GUARD + shift, GUARD + shift + n
        self.view = self.buf[self.lo:self.hi].view(shape)

    def intact(self):
        b = self.buf
        return bool((b[: self.lo] == SENTINEL).all()) and bool((b[self.hi:] == SENTINEL).all())


def _objects(rng, B, T, C, H, W, rmax):
    centers = np.stack([rng.integers(-rmax, W + rmax, (B, T)), rng.integers(-rmax, H + rmax, (B, T))], 2).astype(np.int32)
    radii = rng.integers(0, rmax + 1, (B, T)).astype(np.int32)
    # corners and edges explicitly
    centers[:, 0] = (0, 0)
    centers[:, 1] = (W - 1, H - 1)
    centers[:, 2] = (W - 1, 0)
    centers[:, 3] = (0, H - 1)
    radii[:, :4] = rmax
    labels = rng.integers(0, C, (B, T)).astype(np.int32)
    labels[:, 1] = C - 1  # the last plane of every sample gets the bottom-right corner object
    sizes = np.full(B, T, np.int64)
    sizes[0] = T // 2
    return centers, radii, labels, sizes


SHAPES = [(128, 128), (23, 40), (30, 180), (7, 1024), (5, 1028), (9, 2052), (11, 6), (13, 17), (2, 4), (232, 400)]


@pytest.mark.parametrize("H,W", SHAPES)
@pytest.mark.parametrize("shift", [0, 1])
def test_renderer_fp32_stays_inside(cabi, cuda_dev, H, W, shift):
    import torch

    rng = np.random.default_rng(H * 7919 + W + shift)
    B, T, C = 3, 24, 2
    centers, radii, labels, sizes = _objects(rng, B, T, C, H, W, max(1, min(H, W, 40) // 2))
    c, r, l, s = (_up(x, cuda_dev) for x in (centers, radii, labels, sizes))
    for mode in (_cabi.DH_MODE_FRESH, _cabi.DH_MODE_ACCUMULATE):
        g = Guarded((B, C, H, W), torch.float32, cuda_dev, shift)
        plain = torch.empty((B, C, H, W), device=cuda_dev)
        if mode == _cabi.DH_MODE_ACCUMULATE:
            g.view.zero_()
            plain.zero_()
        _cabi.draw_batched(cabi, g.view, c, r, s, l, mode=mode)
        _cabi.draw_batched(cabi, plain, c, r, s, l, mode=mode)
        torch.cuda.synchronize()
        assert g.intact(), f"{H}x{W} shift {shift} mode {mode}: wrote outside the heatmap"
        assert torch.equal(g.view, plain)
    # flat entry (counting sort + render) into the same shapes
    idx = _up((np.arange(B)[:, None] * C + labels).astype(np.int32).reshape(-1), cuda_dev)
    g = Guarded((B * C, H, W), torch.float32, cuda_dev, shift)
    g.view.zero_()
    _cabi.draw_flat(cabi, g.view, c.reshape(-1, 2), r.reshape(-1), idx)
    torch.cuda.synchronize()
    assert g.intact()


@pytest.mark.parametrize("H,W", [(128, 128), (23, 40), (30, 180), (7, 1024), (2, 4), (232, 400), (200, 12)])
def test_renderer_16bit_stays_inside(cabi, cuda_dev, H, W):
    import torch

    rng = np.random.default_rng(H * 31 + W)
    B, T, C = 3, 24, 2
    centers, radii, labels, sizes = _objects(rng, B, T, C, H, W, max(1, min(H, W, 40) // 2))
    c, r, l, s = (_up(x, cuda_dev) for x in (centers, radii, labels, sizes))
    for dt in (torch.float16, torch.bfloat16):
        g = Guarded((B, C, H, W), dt, cuda_dev)
        plain = torch.empty((B, C, H, W), dtype=dt, device=cuda_dev)
        _cabi.render16(cabi, g.view, s, l, centers=c, radii=r)
        _cabi.render16(cabi, plain, s, l, centers=c, radii=r)
        torch.cuda.synchronize()
        assert g.intact(), f"{H}x{W} {dt}: wrote outside the heatmap"
        assert torch.equal(g.view.view(torch.int16), plain.view(torch.int16))


@pytest.mark.parametrize("H,W", [(24, 32), (23, 40), (11, 6), (64, 256)])
def test_draw_gaussians_and_loss_stay_inside(cuda_dev, H, W):
    import torch
    from accvlab.draw_heatmap import b200

    rng = np.random.default_rng(H + W)
    B, T, C = 3, 16, 2
    centers, radii, labels, sizes = _objects(rng, B, T, C, H, W, max(1, min(H, W) // 3))
    sz = _up(sizes, cuda_dev)
    rb = lambda a: b200.RaggedBatch(_up(a, cuda_dev), sample_sizes=sz)
    k = torch.tensor([0.9, 1.0], device=cuda_dev)
    rf = (radii.astype(np.float32) + np.float32(0.4)).astype(np.float32)
    # N2: fresh (fast path forced by conftest where the shape allows), in place, out of place
    g = Guarded((B, C, H, W), torch.float32, cuda_dev)
    b200.draw_gaussians(None, None, rb(labels), rb(centers), rb(rf), k, 0.4, out=g.view)
    torch.cuda.synchronize()
    assert g.intact()
    fresh = g.view.clone()
    g2 = Guarded((B, C, H, W), torch.float32, cuda_dev)
    g2.view.zero_()
    b200.draw_gaussians(g2.view, None, rb(labels), rb(centers), rb(rf), k, 0.4, out=g2.view)
    torch.cuda.synchronize()
    assert g2.intact() and torch.equal(g2.view, fresh)
    g3 = Guarded((B, C, H, W), torch.float32, cuda_dev)
    b200.draw_gaussians(torch.zeros((B, C, H, W), device=cuda_dev), None, rb(labels), rb(centers), rb(rf), k, 0.4, out=g3.view)
    torch.cuda.synchronize()
    assert g3.intact() and torch.equal(g3.view, fresh)
    # N3 through the C ABI: the gradient, the two result words and the per-unit partials are everything it writes
    lib = _cabi.load()
    pred = torch.rand((B, C, H, W), device=cuda_dev) * 0.9 + 0.05
    gg = Guarded((B, C, H, W), torch.float32, cuda_dev)
    wbytes = int(lib.dh_focal_loss_workspace_bytes(B, C, H, W))
    ws = Guarded((wbytes // 4,), torch.float32, cuda_dev)
    res = Guarded((4,), torch.float32, cuda_dev)  # double[2]
    ci, ri, li = _up(centers, cuda_dev), _up(radii, cuda_dev), _up(labels, cuda_dev)
    st = lib.dh_gaussian_focal_loss(_cabi.stream_ptr(), _cabi.ptr(pred), _cabi.ptr(ci), _cabi.ptr(ri), _cabi.ptr(sz),
                                    _cabi.DH_SIZES_I64, _cabi.ptr(li), B, T, C, H, W, 6.0, 1.0, 2.0, 4.0, 1e-12, 0,
                                    _cabi.ptr(gg.view), _cabi.ptr(res.view), _cabi.ptr(ws.view), wbytes)
    _cabi.check(lib, st, "dh_gaussian_focal_loss")
    torch.cuda.synchronize()
    assert gg.intact() and ws.intact() and res.intact()
    assert bool(torch.isfinite(gg.view).all())


def test_pack_stays_inside(cuda_dev):
    import torch
    from accvlab.draw_heatmap import b200

    rng = np.random.default_rng(3)
    lens = [0, 5, 17, 1, 9]
    data = [torch.from_numpy(rng.random((n, 4), np.float32)).to(cuda_dev) for n in lens]
    rbt = b200.combine_data(data)
    T = max(lens)
    assert rbt.tensor.shape == (len(lens), T, 4)
    for i, n in enumerate(lens):
        assert torch.equal(rbt.tensor[i, :n], data[i])
        assert not rbt.tensor[i, n:].any()  # padding is zero-filled, nothing spills into the next sample
