"""GPU: both render paths of the batched entries, on every BASELINE config, against the oracle and against each other.

  ws=False  the workspace-less entries: ONE launch (groups_kernel: in-kernel slot scan, K1 fused) — the product's main
            path (torch binding, bench.py, host entry);
  ws=True   the *_ws entries of include/dh_b200.h with a caller workspace: TWO launches (bin_kernel = K2 as a pre-pass
            that counting-sorts every sample's objects by class into compact per-plane lists, then groups2_kernel, whose
            unit header is one 8-byte load). Bit-identical by contract; round 2 measured it slower on the BASELINE
            configs (profiles/r02/experiments.md), so nothing in the product selects it by default.
"""
import numpy as np
import pytest
import torch

import _cabi
import workloads

pytestmark = pytest.mark.gpu

CASES = [("cfg1", None), ("cfg2", None), ("cfg3", None), ("cfg4", 4), ("cfg5", 24)]


@pytest.mark.parametrize("name,B", CASES)
def test_single_launch_entries_vs_oracle_and_vs_binned_path(cabi, cuda_dev, oracle, name, B):
    w = workloads.make(name, B=B)
    cfg = w.cfg
    dev = cuda_dev
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes, labels = torch.from_numpy(w.sizes).to(dev), torch.from_numpy(w.labels).to(dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    oci, ori = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    want = np.zeros(shape, np.float32)
    oracle.draw_batched(want, oci, ori, w.sizes, 6.0, 0.8, w.labels)
    ci, ri = torch.from_numpy(oci).to(dev), torch.from_numpy(ori).to(dev)
    results = {}
    for ws in (False, True):
        fresh = _cabi.draw_from_boxes(cabi, torch.full(shape, 3.0, device=dev), cf, bf, float(cfg.osf), sizes, labels, 6.0,
                                      0.8, ws=ws)
        acc = _cabi.draw_batched(cabi, torch.zeros(shape, device=dev), ci, ri, sizes, labels, 6.0, 0.8, ws=ws)
        fresh_i = _cabi.draw_batched(cabi, torch.full(shape, 3.0, device=dev), ci, ri, sizes, labels, 6.0, 0.8,
                                     mode=_cabi.DH_MODE_FRESH, ws=ws)
        dg = _cabi.draw_from_boxes_digest(cabi, torch.empty(shape, device=dev), cf, bf, float(cfg.osf), sizes, labels, 6.0,
                                          0.8, ws=ws)
        h16 = torch.empty(shape, device=dev, dtype=torch.float16)
        _cabi.render16(cabi, h16, sizes, labels, centers=ci, radii=ri, factor=6.0, k=0.8, ws=ws)
        torch.cuda.synchronize()
        assert torch.equal(fresh, acc) and torch.equal(fresh, fresh_i)
        assert torch.equal(h16, fresh.to(torch.float16))
        assert torch.equal(dg, _cabi.checksum(cabi, fresh))
        results[ws] = fresh
        got = fresh.cpu().numpy()
        assert np.array_equal(got > 0, want > 0), f"{name} ws={ws}: touched mask"
        assert float(np.abs(got - want).max()) <= 1e-6, f"{name} ws={ws}"
    assert torch.equal(results[False], results[True])
    # one plane per sample (labels == NULL) through both paths
    shape1 = (w.B, cfg.H, cfg.W)
    a = _cabi.draw_batched(cabi, torch.zeros(shape1, device=dev), ci, ri, sizes, None, 6.0, 0.8, ws=False)
    b = _cabi.draw_batched(cabi, torch.zeros(shape1, device=dev), ci, ri, sizes, None, 6.0, 0.8, ws=True)
    want1 = np.zeros(shape1, np.float32)
    oracle.draw_batched(want1, oci, ori, w.sizes, 6.0, 0.8, None)
    assert torch.equal(a, b)
    assert np.array_equal(a.cpu().numpy() > 0, want1 > 0) and float(np.abs(a.cpu().numpy() - want1).max()) <= 1e-6


def test_workspace_fallbacks_never_fail(cabi, cuda_dev):
    """Too small / misaligned workspace, more classes than the pre-pass serves, shapes off the main renderer: the *_ws
    entries silently take the single-launch path and give the same bits."""
    dev = cuda_dev
    w = workloads.make("cfg2", B=4)
    cfg = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes, labels = torch.from_numpy(w.sizes).to(dev), torch.from_numpy(w.labels).to(dev)
    ci, ri = _cabi.radius_center(cabi, cf, bf, float(cfg.osf))
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    want = _cabi.draw_batched(cabi, torch.zeros(shape, device=dev), ci, ri, sizes, labels, ws=False)
    need = int(cabi.dh_batched_workspace_bytes(w.B, w.T, cfg.C))
    buf = torch.empty(need + 64, dtype=torch.uint8, device=dev)

    def call(heat, ptr_off, nbytes, C=cfg.C, lab=labels):
        st = cabi.dh_draw_heatmap_batched_ws(_cabi.stream_ptr(), _cabi.ptr(heat), _cabi.ptr(ci), _cabi.ptr(ri), _cabi.ptr(sizes),
                                             _cabi.DH_SIZES_I64, _cabi.ptr(lab), w.B, w.T, C, cfg.H, cfg.W, 6.0, 1.0,
                                             _cabi.DH_MODE_ACCUMULATE, None, buf.data_ptr() + ptr_off, nbytes)
        _cabi.check(cabi, st)
        torch.cuda.synchronize()
        return heat

    assert torch.equal(call(torch.zeros(shape, device=dev), 0, need), want)        # the real thing
    assert torch.equal(call(torch.zeros(shape, device=dev), 0, need // 2), want)   # too small
    assert torch.equal(call(torch.zeros(shape, device=dev), 4, need), want)        # misaligned
    assert torch.equal(call(torch.zeros(shape, device=dev), 0, 0), want)           # empty
    # 3000 classes (> 2048): labels stay < 10, planes >= 10 stay zero
    big = torch.zeros((w.B, 3000, 8, 12), device=dev)
    need_big = int(cabi.dh_batched_workspace_bytes(w.B, w.T, 3000))
    buf = torch.empty(need_big, dtype=torch.uint8, device=dev)
    st = cabi.dh_draw_heatmap_batched_ws(_cabi.stream_ptr(), _cabi.ptr(big), _cabi.ptr(ci), _cabi.ptr(ri), _cabi.ptr(sizes),
                                         _cabi.DH_SIZES_I64, _cabi.ptr(labels), w.B, w.T, 3000, 8, 12, 6.0, 1.0,
                                         _cabi.DH_MODE_FRESH, None, _cabi.ptr(buf), need_big)
    _cabi.check(cabi, st)
    torch.cuda.synchronize()
    assert float(big[:, 10:].abs().max()) == 0.0 and float(big[:, :10].max()) > 0.0


def test_status_counters_agree_between_the_two_paths(cabi, cuda_dev):
    dev = cuda_dev
    B, T, C, H, W = 5, 40, 6, 48, 64
    g = torch.Generator().manual_seed(3)
    centers = torch.randint(-5, 70, (B, T, 2), generator=g, dtype=torch.int32)
    radii = torch.randint(-2, 9, (B, T), generator=g, dtype=torch.int32)
    labels = torch.randint(-1, C + 1, (B, T), generator=g, dtype=torch.int32)
    radii[0, 3] = 4194304 + 1
    radii[2, 7] = 2**30
    sizes = torch.tensor([T, 0, T - 5, 17, 1], dtype=torch.int32)
    centers, radii, labels, sizes = (t.to(dev) for t in (centers, radii, labels, sizes))
    out = {}
    for ws in (False, True):
        status = torch.zeros(4, dtype=torch.int32, device=dev)
        heat = _cabi.draw_batched(cabi, torch.zeros((B, C, H, W), device=dev), centers, radii, sizes, labels, status=status,
                                  ws=ws)
        torch.cuda.synchronize()
        out[ws] = (heat, status.cpu().numpy().copy())
    assert torch.equal(out[False][0], out[True][0])
    assert np.array_equal(out[False][1], out[True][1]), (out[False][1], out[True][1])
    valid = (torch.arange(T, device=dev)[None, :] < sizes[:, None])
    bad_label = valid & ((labels < 0) | (labels >= C))
    bad_radius = valid & ~bad_label & (radii > 4194304)
    assert int(out[True][1][0]) == int(bad_label.sum()) and int(out[True][1][2]) == int(bad_radius.sum())
