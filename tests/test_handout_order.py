"""GPU: the heaviest-first hand-out of groups_kernel (every CTA sorts the samples of the last waves by object count and
maps unit index -> sample through that permutation, csrc/dh_groups_kernel.cuh build_perm) must be invisible in the
result: every plane rendered exactly once, whatever the batch size, the sizes dtype, ties in the object counts, and
whether the batch exceeds the 2048 samples the permutation covers. (A first version of the sort broke ties differently
in different CTAs: planes were rendered twice or never. These are the cases that would catch it again.)"""
import numpy as np
import pytest
import torch

import _cabi

pytestmark = pytest.mark.gpu


def _case(seed, B, T, C, H, W, sizes_kind):
    rng = np.random.default_rng(seed)
    if sizes_kind == "ties":        # many equal object counts: the tie order must be the same in every CTA
        sizes = rng.choice([0, 1, T // 2, T], size=B)
    elif sizes_kind == "all_equal":
        sizes = np.full(B, T // 2)
    else:
        sizes = rng.integers(0, T + 1, size=B)
    centers = np.stack([rng.integers(-3, W + 3, (B, T)), rng.integers(-3, H + 3, (B, T))], 2).astype(np.int32)
    radii = rng.integers(0, max(2, min(H, W) // 3), (B, T)).astype(np.int32)
    labels = rng.integers(0, C, (B, T)).astype(np.int32)
    return centers, radii, labels, sizes.astype(np.int64)


@pytest.mark.parametrize("B,T,C,H,W,kind,sdtype", [
    (900, 12, 3, 24, 32, "random", torch.int64),     # several waves, permutation over all samples
    (900, 12, 3, 24, 32, "ties", torch.int32),
    (700, 8, 4, 16, 64, "all_equal", torch.int64),
    (2600, 6, 2, 16, 16, "ties", torch.int64),       # more samples than the permutation covers (2048)
    (5000, 5, 1, 8, 32, "random", torch.int32),      # one plane per sample (labels given, C = 1)
])
def test_every_plane_rendered_exactly_once(cabi, cuda_dev, oracle, B, T, C, H, W, kind, sdtype):
    centers, radii, labels, sizes = _case(B * 7 + T, B, T, C, H, W, kind)
    dev = cuda_dev
    d = lambda a: torch.from_numpy(a).to(dev)
    ci, ri, lab, sz = d(centers), d(radii), d(labels), d(sizes).to(sdtype)
    want = np.zeros((B, C, H, W), np.float32)
    oracle.draw_batched(want, centers, radii, sizes, 6.0, 0.9, labels)
    want_t = torch.from_numpy(want).to(dev)
    for mode in (_cabi.DH_MODE_FRESH, _cabi.DH_MODE_ACCUMULATE):
        init = 7.0 if mode == _cabi.DH_MODE_FRESH else 0.0  # FRESH must overwrite every pixel of every plane
        heat = torch.full((B, C, H, W), init, device=dev)
        _cabi.draw_batched(cabi, heat, ci, ri, sz, lab, 6.0, 0.9, mode=mode)
        torch.cuda.synchronize()
        assert torch.equal(heat > 0, want_t > 0), f"mode {mode}: touched mask"
        assert float((heat - want_t).abs().max()) <= 1e-6
    # no labels: [B,H,W]
    want1 = np.zeros((B, H, W), np.float32)
    oracle.draw_batched(want1, centers, radii, sizes, 6.0, 0.9, None)
    heat1 = torch.full((B, H, W), 3.0, device=dev)
    _cabi.draw_batched(cabi, heat1, ci, ri, sz, None, 6.0, 0.9, mode=_cabi.DH_MODE_FRESH)
    torch.cuda.synchronize()
    w1 = torch.from_numpy(want1).to(dev)
    assert torch.equal(heat1 > 0, w1 > 0) and float((heat1 - w1).abs().max()) <= 1e-6
    # the fused digest sees every touched pixel once, too
    cf = ci.float() + 0.5
    bf = torch.cat([cf - ri.float()[..., None], cf + ri.float()[..., None]], -1)
    oci, ori = oracle.radius_center(cf.cpu().numpy(), bf.cpu().numpy(), 1.0)  # K1: radii < 1 -> 1, centres truncate toward 0
    want2 = np.zeros((B, C, H, W), np.float32)
    oracle.draw_batched(want2, oci, ori, sizes, 6.0, 0.9, labels)
    heat2 = torch.empty((B, C, H, W), device=dev)
    dg = _cabi.draw_from_boxes_digest(cabi, heat2, cf, bf, 1.0, sz, lab, 6.0, 0.9)
    torch.cuda.synchronize()
    assert torch.equal(dg, _cabi.checksum(cabi, heat2))
    assert int(dg[2]) == int((want2 > 0).sum())
