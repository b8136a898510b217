"""dh_draw_heatmap_batched_noalloc: the allocation-free entry (include/dh_b200.h) — no counterpart in the reference.
Its bin-list scratch is a static __device__ array inside libdh_b200.so
(DH_NOALLOC_MAX_SCRATCH_BYTES), shared by every call and keeping the last call's contents, so the checks here are:
two (and more) calls IN A ROW, of different shapes, each bit-identical to the existing entry dh_draw_heatmap_batched on
the same inputs; the size limit is an error, not a fallback. Collected last on purpose (file name), passes alone too.
"""
import numpy as np
import pytest

import _cabi
import workloads

DH_ERR_STATIC_SCRATCH_EXCEEDED = -5


def _noalloc(lib, heat, centers, radii, sizes, labels, factor, k, mode, status=None):
    import torch

    B, T = radii.shape
    if labels is not None:
        C, H, W = heat.shape[1:]
    else:
        C, (H, W) = 1, heat.shape[1:]
    tag = _cabi.DH_SIZES_I64 if sizes.dtype == torch.int64 else _cabi.DH_SIZES_I32
    st = lib.dh_draw_heatmap_batched_noalloc(_cabi.stream_ptr(), _cabi.ptr(heat), _cabi.ptr(centers), _cabi.ptr(radii),
                                             _cabi.ptr(sizes), tag, _cabi.ptr(labels), B, T, C, H, W, factor, k, mode,
                                             _cabi.ptr(status))
    _cabi.check(lib, st, "dh_draw_heatmap_batched_noalloc")
    return heat


# ---- CPU: the limit is checked before any device is touched ---------------------------------------------------------
def test_noalloc_rejects_problems_beyond_the_static_array(cabi):
    cap = int(cabi.dh_noalloc_max_scratch_bytes())
    assert cap == 8 << 20
    # B x T slots x 16 bytes alone exceed the array; pointers are never dereferenced (size check comes first)
    B, T, C = 4096, 200, 10
    assert int(cabi.dh_batched_workspace_bytes(B, T, C)) > cap
    st = cabi.dh_draw_heatmap_batched_noalloc(None, None, None, None, None, _cabi.DH_SIZES_I64, None, B, T, C, 64, 64, 6.0,
                                              1.0, _cabi.DH_MODE_FRESH, None)
    assert st == DH_ERR_STATIC_SCRATCH_EXCEEDED
    msg = cabi.dh_status_string(st).decode()
    assert "static" in msg and "DH_NOALLOC_MAX_SCRATCH_BYTES" in msg, msg
    # sizes whose product would overflow 64 bits are rejected the same way, not wrapped
    st = cabi.dh_draw_heatmap_batched_noalloc(None, None, None, None, None, _cabi.DH_SIZES_I64, None, 1 << 40, 1 << 40, 10, 64,
                                              64, 6.0, 1.0, _cabi.DH_MODE_FRESH, None)
    assert st == DH_ERR_STATIC_SCRATCH_EXCEEDED
    assert cabi.dh_draw_heatmap_batched_noalloc(None, None, None, None, None, _cabi.DH_SIZES_I64, None, -1, 1, 1, 64, 64, 6.0,
                                                1.0, _cabi.DH_MODE_FRESH, None) == _cabi.DH_ERR_INVALID_ARGUMENT


# ---- GPU ------------------------------------------------------------------------------------------------------------
def _inputs(cabi, dev, name, B):
    import torch

    w = workloads.make(name, B=B)
    c = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes, labels = torch.from_numpy(w.sizes).to(dev), torch.from_numpy(w.labels).to(dev)
    ci, ri = _cabi.radius_center(cabi, cf, bf, float(c.osf))
    return w, ci, ri, sizes, labels


@pytest.mark.gpu
def test_noalloc_entry_equals_existing_entry_on_calls_in_a_row(cuda_dev, cabi):
    """Large call, then smaller ones of other shapes (stale records / lists of the previous call stay in the array),
    then the large one again: every result equals dh_draw_heatmap_batched's, bit for bit, in both modes."""
    import torch

    dev = cuda_dev
    seq = [("cfg3", 12), ("cfg1", None), ("cfg2", None), ("cfg4", 2), ("cfg3", 12), ("cfg3", 5)]
    for name, B in seq:
        w, ci, ri, sizes, labels = _inputs(cabi, dev, name, B)
        c = w.cfg
        shape = (w.B, c.C, c.H, c.W)
        assert int(cabi.dh_batched_workspace_bytes(w.B, ci.shape[1], c.C)) <= int(cabi.dh_noalloc_max_scratch_bytes())
        for mode, init in ((_cabi.DH_MODE_FRESH, 3.0), (_cabi.DH_MODE_ACCUMULATE, 0.0), (_cabi.DH_MODE_ACCUMULATE, 0.25)):
            st_a = torch.zeros(4, dtype=torch.int32, device=dev)
            st_b = torch.zeros(4, dtype=torch.int32, device=dev)
            want = _cabi.draw_batched(cabi, torch.full(shape, init, device=dev), ci, ri, sizes, labels, 6.0, 0.8, mode=mode,
                                      status=st_a, ws=False)
            got = _noalloc(cabi, torch.full(shape, init, device=dev), ci, ri, sizes, labels, 6.0, 0.8, mode, status=st_b)
            torch.cuda.synchronize()
            assert torch.equal(got, want), f"{name} B={w.B} mode={mode} init={init}"
            assert torch.equal(st_a, st_b)
            assert float(got.max()) > 0.0
        # one plane per sample (labels == NULL)
        shape1 = (w.B, c.H, c.W)
        want = _cabi.draw_batched(cabi, torch.zeros(shape1, device=dev), ci, ri, sizes, None, 6.0, 1.0, ws=False)
        got = _noalloc(cabi, torch.zeros(shape1, device=dev), ci, ri, sizes, None, 6.0, 1.0, _cabi.DH_MODE_ACCUMULATE)
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"{name} B={w.B} one plane per sample"


@pytest.mark.gpu
def test_noalloc_entry_runs_the_two_launch_path_and_allocates_nothing(cuda_dev, cabi):
    """Same inputs twice in a row: both results equal the existing entry's and the oracle's contract (values within 1e-6
    of each other is implied by bit equality); the call is the two-launch scratch path (launch counter +2) and leaves
    the device's free memory where it was."""
    import torch

    dev = cuda_dev
    w, ci, ri, sizes, labels = _inputs(cabi, dev, "cfg3", 16)
    c = w.cfg
    shape = (w.B, c.C, c.H, c.W)
    want = _cabi.draw_batched(cabi, torch.zeros(shape, device=dev), ci, ri, sizes, labels, 6.0, 1.0,
                              mode=_cabi.DH_MODE_FRESH, ws=False)
    out = torch.empty(shape, device=dev)
    _noalloc(cabi, out, ci, ri, sizes, labels, 6.0, 1.0, _cabi.DH_MODE_FRESH)  # first call: module / symbol set-up
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    free0, _ = torch.cuda.mem_get_info(dev)
    n0 = int(cabi.dh_launch_count())
    out.fill_(7.0)
    _noalloc(cabi, out, ci, ri, sizes, labels, 6.0, 1.0, _cabi.DH_MODE_FRESH)
    n1 = int(cabi.dh_launch_count())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert torch.equal(out, want)
    assert n1 - n0 == 2, "bin_kernel + groups2_kernel expected (the scratch path)"
    assert free1 >= free0, f"device memory was allocated inside the call: free {free0} -> {free1}"
    got = out.cpu().numpy()
    assert np.array_equal(got > 0, want.cpu().numpy() > 0)
