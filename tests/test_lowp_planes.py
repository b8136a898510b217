"""fp16 / bf16 heatmap planes (`dh_render_heatmap_batched_16`). Not in the reference (float32 only),
so the contract is stated against the float32 path, which IS pinned: the 16-bit result equals the float32 result
rounded once to nearest even — bit for bit — and is therefore within one 16-bit ulp of the oracle's float32 values."""
import numpy as np
import pytest

import _cabi
import workloads

pytestmark = pytest.mark.gpu


def _up(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _dtypes():
    import torch

    return (torch.float16, torch.bfloat16)


def _poison(shape, dtype, dev):
    import torch

    return torch.full(shape, float("nan"), dtype=dtype, device=dev)


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg3", "cfg4"])
def test_equals_float32_result_rounded_once(cabi, cuda_dev, oracle, name):
    import torch

    w = workloads.make(name)
    cfg = w.cfg
    cf, bf = _up(w.centers_f, cuda_dev), _up(w.bboxes_f, cuda_dev)
    sizes, labels = _up(w.sizes, cuda_dev), _up(w.labels, cuda_dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    h32 = torch.empty(shape, dtype=torch.float32, device=cuda_dev)
    _cabi.draw_from_boxes(cabi, h32, cf, bf, float(cfg.osf), sizes, labels)
    ci, ri = _cabi.radius_center(cabi, cf, bf, float(cfg.osf))
    # the oracle's float32 values, for the 1-ulp statement
    oci, ori = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    want = np.zeros(shape, np.float32)
    oracle.draw_batched(want, oci, ori, w.sizes, 6.0, 1.0, w.labels)
    want_t = torch.from_numpy(want).to(cuda_dev)
    for dt in _dtypes():
        a = _poison(shape, dt, cuda_dev)  # every element must be overwritten
        _cabi.render16(cabi, a, sizes, labels, centers_f=cf, bboxes_f=bf, osf=float(cfg.osf))
        b = _poison(shape, dt, cuda_dev)
        _cabi.render16(cabi, b, sizes, labels, centers=ci, radii=ri)
        ref = h32.to(dt)
        assert torch.equal(a.view(torch.int16), ref.view(torch.int16)), f"{name} {dt}: from boxes"
        assert torch.equal(b.view(torch.int16), ref.view(torch.int16)), f"{name} {dt}: from integer centres"
        # vs the oracle: non-negative values order like their bit patterns, so |bits difference| is the ulp distance
        d = (a.view(torch.int16).to(torch.int32) - want_t.to(dt).view(torch.int16).to(torch.int32)).abs().max()
        assert int(d) <= 1, f"{name} {dt}: {int(d)} ulp from the oracle"


def test_non_classwise_sizes_i32_and_negative_k(cabi, cuda_dev):
    import torch

    w = workloads.make("cfg1")
    cfg = w.cfg
    cf, bf = _up(w.centers_f, cuda_dev), _up(w.bboxes_f, cuda_dev)
    sizes32 = _up(w.sizes.astype(np.int32), cuda_dev)
    shape = (w.B, cfg.H, cfg.W)
    for k in (0.8, -1.0):
        h32 = torch.empty(shape, dtype=torch.float32, device=cuda_dev)
        _cabi.draw_from_boxes(cabi, h32, cf, bf, float(cfg.osf), sizes32, None, k=k)
        for dt in _dtypes():
            a = _poison(shape, dt, cuda_dev)
            _cabi.render16(cabi, a, sizes32, None, centers_f=cf, bboxes_f=bf, osf=float(cfg.osf), k=k)
            assert torch.equal(a.view(torch.int16), h32.to(dt).view(torch.int16))
            if k < 0:
                assert not a.any()  # max-blend of negative peaks into the fused zero-init


@pytest.mark.parametrize("H,W", [(23, 40), (30, 180), (64, 1024), (2, 4), (200, 12)])
def test_shapes(cabi, cuda_dev, H, W):
    import torch

    rng = np.random.default_rng(H * 1000 + W)
    B, T, C = 5, 40, 3
    centers = np.stack([rng.integers(-3, W + 3, (B, T)), rng.integers(-3, H + 3, (B, T))], 2).astype(np.int32)
    radii = rng.integers(0, max(2, min(H, W) // 2), (B, T)).astype(np.int32)
    labels = rng.integers(0, C, (B, T)).astype(np.int32)
    sizes = np.array([T, 0, 1, T // 2, T], np.int64)
    c, r, l, s = (_up(x, cuda_dev) for x in (centers, radii, labels, sizes))
    h32 = torch.empty((B, C, H, W), dtype=torch.float32, device=cuda_dev)
    _cabi.draw_batched(cabi, h32, c, r, s, l, mode=_cabi.DH_MODE_FRESH)
    for dt in _dtypes():
        a = _poison((B, C, H, W), dt, cuda_dev)
        _cabi.render16(cabi, a, s, l, centers=c, radii=r)
        assert torch.equal(a.view(torch.int16), h32.to(dt).view(torch.int16))


def test_many_objects_per_plane_list_overflow(cabi, cuda_dev):
    """more objects in one unit than the shared-memory list holds (512): the per-group re-scan path"""
    import torch

    rng = np.random.default_rng(7)
    B, T, H, W = 2, 1500, 64, 256
    centers = np.stack([rng.integers(0, W, (B, T)), rng.integers(0, H, (B, T))], 2).astype(np.int32)
    radii = rng.integers(1, 9, (B, T)).astype(np.int32)
    sizes = np.array([T, T - 7], np.int64)
    c, r, s = (_up(x, cuda_dev) for x in (centers, radii, sizes))
    h32 = torch.empty((B, H, W), dtype=torch.float32, device=cuda_dev)
    _cabi.draw_batched(cabi, h32, c, r, s, None, mode=_cabi.DH_MODE_FRESH)
    for dt in _dtypes():
        a = _poison((B, H, W), dt, cuda_dev)
        _cabi.render16(cabi, a, s, None, centers=c, radii=r)
        assert torch.equal(a.view(torch.int16), h32.to(dt).view(torch.int16))


def test_unsupported_shapes_and_arguments(cabi, cuda_dev):
    import torch

    s = torch.zeros(2, dtype=torch.int64, device=cuda_dev)
    c = torch.zeros((2, 1, 2), dtype=torch.int32, device=cuda_dev)
    r = torch.zeros((2, 1), dtype=torch.int32, device=cuda_dev)
    cf = torch.zeros((2, 1, 2), device=cuda_dev)
    bf = torch.zeros((2, 1, 4), device=cuda_dev)
    for H, W in ((3, 4), (8, 6), (4, 2048)):  # H*W % 8, W % 4, W > 1024
        a = torch.zeros((2, H, W), dtype=torch.float16, device=cuda_dev)
        assert _cabi.render16(cabi, a, s, None, centers=c, radii=r, check_status=False) == _cabi.DH_ERR_UNSUPPORTED_SHAPE
    a = torch.zeros((2, 8, 8), dtype=torch.float16, device=cuda_dev)
    # both sources at once / neither / unknown dtype tag
    assert _cabi.render16(cabi, a, s, None, centers=c, radii=r, centers_f=cf, bboxes_f=bf,
                          check_status=False) == _cabi.DH_ERR_INVALID_ARGUMENT
    st = cabi.dh_render_heatmap_batched_16(_cabi.stream_ptr(), _cabi.ptr(a), 7, _cabi.ptr(c), _cabi.ptr(r), None, None, 1.0,
                                           _cabi.ptr(s), _cabi.DH_SIZES_I64, None, 2, 1, 1, 8, 8, 6.0, 1.0, None)
    assert st == _cabi.DH_ERR_INVALID_ARGUMENT
    st = cabi.dh_render_heatmap_batched_16(_cabi.stream_ptr(), _cabi.ptr(a), 1, None, None, None, None, 1.0,
                                           _cabi.ptr(s), _cabi.DH_SIZES_I64, None, 2, 1, 1, 8, 8, 6.0, 1.0, None)
    assert st == _cabi.DH_ERR_INVALID_ARGUMENT
    # empty batch and zero targets are fine; zero targets writes zeros
    assert _cabi.render16(cabi, a[:0], s[:0], None, centers=c[:0], radii=r[:0]) == _cabi.DH_OK
    a.fill_(3.0)
    _cabi.render16(cabi, a, s, None, centers=c[:, :0], radii=r[:, :0])
    assert not a.any()


def test_python_api(cuda_dev):
    import torch
    from accvlab.draw_heatmap import b200

    w = workloads.make("cfg2")
    cfg = w.cfg
    cf, bf = _up(w.centers_f, cuda_dev), _up(w.bboxes_f, cuda_dev)
    sizes, labels = _up(w.sizes, cuda_dev), _up(w.labels, cuda_dev)
    h32 = b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, cfg.H, cfg.W, cfg.C, labels)
    ci, ri = b200.radius_center(cf, bf, cfg.osf)
    rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
    for dt in _dtypes():
        a = b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, cfg.H, cfg.W, cfg.C, labels, dtype=dt)
        b = b200.render_heatmap_batched(rb(ci), rb(ri), cfg.H, cfg.W, cfg.C, labels=rb(labels), dtype=dt)
        assert a.dtype == dt and b.dtype == dt
        assert torch.equal(a, h32.to(dt)) and torch.equal(b, h32.to(dt))
        # accumulate mode has no 16-bit variant: loud error, not a silent cast
        with pytest.raises(RuntimeError):
            b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, labels=labels, out=a, fresh=False)
    import accvlab.draw_heatmap as dh

    with pytest.raises(RuntimeError):  # the reference surface stays float32 only (draw_heatmap_cuda.cu:117)
        dh.draw_heatmap_batched(a, rb(ci), rb(ri), 6.0, 1.0, labels=rb(labels))


def test_full_size_slice_of_cfg5(cabi, cuda_dev):
    """one 8-GPU shard of BASELINE configs[4] (384 samples x 10 classes x 232x400)"""
    import torch

    w = workloads.make("cfg5", B=384)
    cfg = w.cfg
    cf, bf = _up(w.centers_f, cuda_dev), _up(w.bboxes_f, cuda_dev)
    sizes, labels = _up(w.sizes, cuda_dev), _up(w.labels, cuda_dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    h32 = torch.empty(shape, dtype=torch.float32, device=cuda_dev)
    _cabi.draw_from_boxes(cabi, h32, cf, bf, float(cfg.osf), sizes, labels)
    for dt in _dtypes():
        a = _poison(shape, dt, cuda_dev)
        _cabi.render16(cabi, a, sizes, labels, centers_f=cf, bboxes_f=bf, osf=float(cfg.osf))
        assert torch.equal(a.view(torch.int16), h32.to(dt).view(torch.int16))
        del a
