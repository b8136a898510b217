"""GPU: the CUDA path, called THROUGH THE C ABI (tests/_cabi.py), against the CPU oracle and the committed golden
vectors of the reference. Bars: radii / centres / touched masks bit-exact; fp32 values <= 1e-6 abs (and the ulp
distance is reported). Edge cases follow SURVEY.md section 4 ("gaps the new repo must fill")."""
import numpy as np
import pytest
import torch

import _cabi
import workloads

pytestmark = pytest.mark.gpu

ABS_TOL = 1e-6  # north_star: "within 1e-6 absolute / 1 ulp on fp32 heatmap values"


def _dev(a, dev, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(dev)


def _ulp(a, b):
    a = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    return int(np.abs(a - b).max()) if a.size else 0


def _assert_parity(got, want, what="", max_ulp=2):
    """max_ulp applies against the fp32 "spec" oracle (same arithmetic up to the exp; libdevice expf is within 2 ulp
    of correctly rounded). Against the reference's float64 CPU output only the absolute bar applies: rounding of the
    fp32 exp argument (|arg| <= 9 at factor 6) alone moves small values by a few ulp."""
    got = got.cpu().numpy() if hasattr(got, "cpu") else got
    assert np.array_equal(got > 0, want > 0), f"{what}: touched-pixel mask differs"
    err = float(np.abs(got.astype(np.float64) - want.astype(np.float64)).max()) if got.size else 0.0
    assert err <= ABS_TOL, f"{what}: max abs err {err}"
    if max_ulp is not None:
        assert _ulp(got, want) <= max_ulp, f"{what}: ulp distance {_ulp(got, want)}"


GOLD = ["ref_batched_20x50", "ref_batched_45x63", "ref_classwise_20x50", "ref_classwise_64x64"]


@pytest.mark.parametrize("name", GOLD)
@pytest.mark.parametrize("sizes_dtype", [torch.int32, torch.int64])
def test_golden_batched(cabi, cuda_dev, golden, oracle, name, sizes_dtype):
    g = golden(name)
    labels = g.get("labels")
    k, factor = float(g["k"]), float(g["factor"])
    c, r = _dev(g["centers_int"], cuda_dev), _dev(g["radii"], cuda_dev)
    s = _dev(g["sizes"], cuda_dev, sizes_dtype)
    lab = None if labels is None else _dev(labels, cuda_dev)
    want = np.zeros_like(g["heat"])
    oracle.draw_batched(want, g["centers_int"], g["radii"], g["sizes"], factor, k, labels)
    # reference semantics: in-place max into zeros
    heat = torch.zeros(g["heat"].shape, device=cuda_dev)
    _cabi.draw_batched(cabi, heat, c, r, s, lab, factor, k, _cabi.DH_MODE_ACCUMULATE)
    _assert_parity(heat, want, name + " accumulate vs oracle")
    _assert_parity(heat, g["heat"], name + " accumulate vs reference CPU golden", max_ulp=None)
    # fresh mode overwrites garbage
    heat2 = torch.full(g["heat"].shape, 123.0, device=cuda_dev)
    _cabi.draw_batched(cabi, heat2, c, r, s, lab, factor, k, _cabi.DH_MODE_FRESH)
    assert torch.equal(heat, heat2)
    # K1 fused (float boxes in), with and without the integer side outputs
    cf, bf = _dev(g["centers_f"], cuda_dev), _dev(g["bboxes_f"], cuda_dev)
    heat3 = torch.full(g["heat"].shape, -5.0, device=cuda_dev)
    _cabi.draw_from_boxes(cabi, heat3, cf, bf, float(g["osf"]), s, lab, factor, k)
    assert torch.equal(heat, heat3)
    co = torch.empty_like(c)
    ro = torch.empty_like(r)
    heat4 = torch.empty_like(heat)
    _cabi.draw_from_boxes(cabi, heat4, cf, bf, float(g["osf"]), s, lab, factor, k, centers_out=co, radii_out=ro)
    assert torch.equal(heat, heat4) and torch.equal(co, c) and torch.equal(ro, r)


def test_k1_bit_exact(cabi, cuda_dev, golden):
    g = golden("ref_k1")
    i = 0
    while f"osf_{i}" in g:
        co, ro = _cabi.radius_center(cabi, _dev(g[f"centers_f_{i}"], cuda_dev), _dev(g[f"bboxes_f_{i}"], cuda_dev),
                                     float(g[f"osf_{i}"]))
        assert np.array_equal(co.cpu().numpy(), g[f"centers_int_{i}"])
        assert np.array_equal(ro.cpu().numpy(), g[f"radii_{i}"])
        i += 1


def test_golden_edges_and_flat_entry(cabi, cuda_dev, golden, oracle):
    g = golden("ref_edges")
    objs = g["objs"]
    n = len(objs)
    H, W = g["heat_all"].shape
    k, factor = float(g["k"]), float(g["factor"])
    c, r = _dev(objs[:, :2], cuda_dev), _dev(objs[:, 2], cuda_dev)
    idx = torch.arange(n, dtype=torch.int32, device=cuda_dev)
    for mode in (_cabi.DH_MODE_ACCUMULATE, _cabi.DH_MODE_FRESH):
        heat = torch.zeros((n, H, W), device=cuda_dev)
        _cabi.draw_flat(cabi, heat, c, r, idx, factor, k, mode)
        _assert_parity(heat, g["heat_each"], "edges, one plane per object", max_ulp=None)
        heat1 = torch.zeros((1, H, W), device=cuda_dev)
        _cabi.draw_flat(cabi, heat1, c, r, torch.zeros_like(idx), factor, k, mode)
        _assert_parity(heat1[0], g["heat_all"], "edges, all objects in one plane", max_ulp=None)
    heat_b = torch.zeros((1, H, W), device=cuda_dev)
    _cabi.draw_batched(cabi, heat_b, c[None].contiguous(), r[None].contiguous(),
                       torch.tensor([n], device=cuda_dev, dtype=torch.int32), None, factor, k)
    assert torch.equal(heat_b[0], heat1[0])


def test_golden_reference_examples(cabi, cuda_dev, golden):
    g = golden("ref_examples")
    H, W, C = int(g["H"]), int(g["W"]), int(g["C"])
    B = g["centers"].shape[0]
    k, factor = float(g["k"]), float(g["factor"])
    c, r, s = _dev(g["centers"], cuda_dev), _dev(g["radii"], cuda_dev), _dev(g["sizes"], cuda_dev)
    heat = torch.zeros((B, H, W), device=cuda_dev)
    _cabi.draw_batched(cabi, heat, c, r, s, None, factor, k)
    _assert_parity(heat, g["heat"], "examples", max_ulp=None)
    heat_c = torch.zeros((B, C, H, W), device=cuda_dev)
    _cabi.draw_batched(cabi, heat_c, c, r, s, _dev(g["labels"], cuda_dev), factor, k)
    planes = heat_c.reshape(B * C, H, W).cpu().numpy()
    nz = np.nonzero(planes.reshape(B * C, -1).max(axis=1) > 0)[0]
    assert np.array_equal(nz, g["heat_cls_planes"])
    _assert_parity(planes[nz], g["heat_cls_values"], "examples classwise", max_ulp=None)


@pytest.mark.parametrize("name,B", [("cfg1", None), ("cfg2", None), ("cfg3", None), ("cfg4", 4), ("cfg5", 24)])
def test_baseline_configs_vs_oracle(cabi, cuda_dev, oracle, name, B):
    """BASELINE.json configs (cfg4 / cfg5 on a batch slice the oracle finishes in seconds)."""
    w = workloads.make(name, B=B)
    cfg = w.cfg
    ci, ri = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    want = np.zeros((w.B, cfg.C, cfg.H, cfg.W), np.float32)
    oracle.draw_batched(want, ci, ri, w.sizes, 6.0, 1.0, w.labels)
    cf, bf = _dev(w.centers_f, cuda_dev), _dev(w.bboxes_f, cuda_dev)
    s, lab = _dev(w.sizes, cuda_dev), _dev(w.labels, cuda_dev)
    co, ro = _cabi.radius_center(cabi, cf, bf, float(cfg.osf))
    assert np.array_equal(co.cpu().numpy(), ci) and np.array_equal(ro.cpu().numpy(), ri)
    heat = torch.empty(want.shape, device=cuda_dev)
    _cabi.draw_from_boxes(cabi, heat, cf, bf, float(cfg.osf), s, lab)
    _assert_parity(heat, want, name + " fresh fused")
    heat2 = torch.zeros(want.shape, device=cuda_dev)
    _cabi.draw_batched(cabi, heat2, co, ro, s, lab)
    assert torch.equal(heat, heat2)
    mask = oracle.touched_mask_batched(want.shape, ci, ri, w.sizes, w.labels)
    assert np.array_equal(mask, heat.cpu().numpy() > 0)
    # digest: device kernel == numpy restatement
    d = _cabi.checksum(cabi, heat).cpu().numpy()
    assert np.array_equal(d, oracle.digest_numpy(want)) or _ulp(heat.cpu().numpy(), want) > 0
    assert np.array_equal(d, oracle.digest_numpy(heat.cpu().numpy()))
    if cfg.C == 1:  # non-classwise path on the same data
        heat3 = torch.empty((w.B, cfg.H, cfg.W), device=cuda_dev)
        _cabi.draw_from_boxes(cabi, heat3, cf, bf, float(cfg.osf), s, None)
        assert torch.equal(heat3, heat[:, 0])
