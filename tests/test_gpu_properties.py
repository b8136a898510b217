"""GPU: size-independent properties at full BASELINE sizes, edge cases, the host-buffer entry and the digest."""
import ctypes

import numpy as np
import pytest
import torch

import _cabi
import workloads

pytestmark = pytest.mark.gpu


def _dev(a, dev, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(dev)


def _upload(w, dev):
    return (_dev(w.centers_f, dev), _dev(w.bboxes_f, dev), _dev(w.sizes, dev), _dev(w.labels, dev))


def test_cfg3_full_size_properties(cabi, cuda_dev):
    """fresh == zeros+accumulate; idempotence; accumulate keeps larger existing values; mask == union of windows
    computed independently with torch ops; every value in [0, k]; centre pixels == k exactly."""
    w = workloads.make("cfg3")
    cfg = w.cfg
    cf, bf, s, lab = _upload(w, cuda_dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    fresh = torch.full(shape, 9.0, device=cuda_dev)
    _cabi.draw_from_boxes(cabi, fresh, cf, bf, float(cfg.osf), s, lab, 6.0, 0.8)
    co, ro = _cabi.radius_center(cabi, cf, bf, float(cfg.osf))
    acc = torch.zeros(shape, device=cuda_dev)
    _cabi.draw_batched(cabi, acc, co, ro, s, lab, 6.0, 0.8)
    assert torch.equal(fresh, acc)
    _cabi.draw_batched(cabi, acc, co, ro, s, lab, 6.0, 0.8)  # idempotent
    assert torch.equal(fresh, acc)
    base = torch.full(shape, 0.25, device=cuda_dev)
    _cabi.draw_batched(cabi, base, co, ro, s, lab, 6.0, 0.8)
    assert torch.equal(base, torch.maximum(fresh, torch.tensor(0.25, device=cuda_dev)))
    assert float(fresh.min()) >= 0.0 and float(fresh.max()) == np.float32(0.8)
    # independent mask: scatter window corners into a 2-D difference image per plane, integrate
    T = w.T
    valid = torch.arange(T, device=cuda_dev)[None, :] < s[:, None]
    x, y, r = co[..., 0].long(), co[..., 1].long(), ro.long()
    x0, x1 = (x - r).clamp(0, cfg.W), (x + r + 1).clamp(0, cfg.W)
    y0, y1 = (y - r).clamp(0, cfg.H), (y + r + 1).clamp(0, cfg.H)
    valid &= (x0 < x1) & (y0 < y1)
    plane = (torch.arange(w.B, device=cuda_dev)[:, None] * cfg.C + lab.long())[valid]
    diff = torch.zeros((w.B * cfg.C, cfg.H + 1, cfg.W + 1), device=cuda_dev, dtype=torch.int32)
    one = torch.ones_like(plane, dtype=torch.int32)
    diff.index_put_((plane, y0[valid], x0[valid]), one, accumulate=True)
    diff.index_put_((plane, y1[valid], x1[valid]), one, accumulate=True)
    diff.index_put_((plane, y0[valid], x1[valid]), -one, accumulate=True)
    diff.index_put_((plane, y1[valid], x0[valid]), -one, accumulate=True)
    cover = diff.cumsum(1).cumsum(2)[:, : cfg.H, : cfg.W] > 0
    assert torch.equal(cover.reshape(shape), fresh > 0)
    inside = valid & (x >= 0) & (x < cfg.W) & (y >= 0) & (y < cfg.H)
    pl = (torch.arange(w.B, device=cuda_dev)[:, None] * cfg.C + lab.long())[inside]
    assert bool((fresh.reshape(-1, cfg.H, cfg.W)[pl, y[inside], x[inside]] == np.float32(0.8)).all())


def _window_cover_count(co, ro, s, lab, C, H, W, dev, chunk=256):
    """Number of pixels covered by the union of the clipped (2r+1)^2 windows, computed independently of the renderer
    with torch ops (2-D difference image per plane, integrated), in sample chunks to bound memory."""
    B, T = ro.shape
    total = 0
    for b0 in range(0, B, chunk):
        b1 = min(B, b0 + chunk)
        nb = b1 - b0
        valid = torch.arange(T, device=dev)[None, :] < s[b0:b1, None]
        x, y, r = co[b0:b1, :, 0].long(), co[b0:b1, :, 1].long(), ro[b0:b1].long()
        x0, x1 = (x - r).clamp(0, W), (x + r + 1).clamp(0, W)
        y0, y1 = (y - r).clamp(0, H), (y + r + 1).clamp(0, H)
        valid &= (x0 < x1) & (y0 < y1)
        plane = (torch.arange(nb, device=dev)[:, None] * C + lab[b0:b1].long())[valid]
        diff = torch.zeros((nb * C, H + 1, W + 1), device=dev, dtype=torch.int32)
        one = torch.ones_like(plane, dtype=torch.int32)
        diff.index_put_((plane, y0[valid], x0[valid]), one, accumulate=True)
        diff.index_put_((plane, y1[valid], x1[valid]), one, accumulate=True)
        diff.index_put_((plane, y0[valid], x1[valid]), -one, accumulate=True)
        diff.index_put_((plane, y1[valid], x0[valid]), -one, accumulate=True)
        total += int((diff.cumsum(1, dtype=torch.int32).cumsum(2, dtype=torch.int32)[:, :H, :W] > 0).sum())
        del diff
    return total


@pytest.mark.parametrize("name", ["cfg4", "cfg5"])
def test_full_size_properties_cfg4_cfg5(cabi, cuda_dev, name):
    """BASELINE.json's full sizes (cfg5: 30 720 planes, 11.4 GB) through size-independent properties:
    fresh fused == zeros + accumulate (bitwise); accumulate is idempotent; the digest fused into the render launch ==
    a separate checksum pass == the sum of the digests of 8 independently rendered shards;
    touched-pixel count == an independent union-of-windows count; all values in [0, k]; max == k."""
    w = workloads.make(name)
    cfg = w.cfg
    cf, bf, s, lab = _upload(w, cuda_dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    sample_elems = cfg.C * cfg.H * cfg.W
    fresh = torch.full(shape, 9.0, device=cuda_dev)
    d_fused = _cabi.draw_from_boxes_digest(cabi, fresh, cf, bf, float(cfg.osf), s, lab, 6.0, 1.0)
    d_pass = _cabi.checksum(cabi, fresh)
    assert torch.equal(d_fused, d_pass)
    co, ro = _cabi.radius_center(cabi, cf, bf, float(cfg.osf))
    acc = torch.zeros(shape, device=cuda_dev)
    _cabi.draw_batched(cabi, acc, co, ro, s, lab, 6.0, 1.0)
    assert torch.equal(fresh, acc)
    _cabi.draw_batched(cabi, acc, co, ro, s, lab, 6.0, 1.0)  # idempotent
    assert torch.equal(fresh, acc)
    del acc
    d = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    for g in range(8):
        lo, hi = workloads.shard_range(w.B, g, 8)
        part = torch.empty((hi - lo, cfg.C, cfg.H, cfg.W), device=cuda_dev)
        _cabi.draw_from_boxes(cabi, part, cf[lo:hi].contiguous(), bf[lo:hi].contiguous(), float(cfg.osf),
                              s[lo:hi].contiguous(), lab[lo:hi].contiguous())
        _cabi.checksum(cabi, part, lo * sample_elems, d)
        del part
    assert torch.equal(d, d_pass)
    assert int(d_pass[2]) == _window_cover_count(co, ro, s, lab, cfg.C, cfg.H, cfg.W, cuda_dev)
    assert float(fresh.min()) >= 0.0 and float(fresh.max()) == 1.0


@pytest.mark.parametrize("name", ["cfg4", "cfg5"])
def test_full_size_values_vs_oracle_cfg4_cfg5(cabi, cuda_dev, oracle, name):
    """BASELINE.json's FULL sizes against the ORACLE itself (not against the renderer's other mode): cfg4 at B=32 and
    cfg5 at B=3072 (30 720 planes, 11.4 GB). The oracle's threaded "refcpu" family (bit-identical to the reference's
    Python CPU path on every golden, tests/test_oracle_golden.py) renders the batch in 384-sample chunks; every chunk
    is compared with the GPU result — touched mask exact, values <= 1e-6 — and its digest (dh_checksum of the ORACLE's
    pixels, with the chunk's index base) is accumulated: the mask-hash and count words must equal those of the
    digest fused into the full-size render launch."""
    import os

    w = workloads.make(name)
    cfg = w.cfg
    cf, bf, s, lab = _upload(w, cuda_dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    sample_elems = cfg.C * cfg.H * cfg.W
    got = torch.full(shape, 9.0, device=cuda_dev)
    d_gpu = _cabi.draw_from_boxes_digest(cabi, got, cf, bf, float(cfg.osf), s, lab, 6.0, 1.0)
    oci, ori = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    co, ro = _cabi.radius_center(cabi, cf, bf, float(cfg.osf))
    assert np.array_equal(co.cpu().numpy(), oci) and np.array_equal(ro.cpu().numpy(), ori)  # K1 bit-exact, full size
    threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    chunk = 384
    d_or = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    max_err = 0.0
    host = np.empty((min(chunk, w.B), cfg.C, cfg.H, cfg.W), np.float32)
    for b0 in range(0, w.B, chunk):
        b1 = min(w.B, b0 + chunk)
        want = host[: b1 - b0]
        oracle.refcpu_draw_batched(want, oci[b0:b1], ori[b0:b1], w.sizes[b0:b1], 6.0, 1.0, w.labels[b0:b1],
                                   zero_first=True, nthreads=threads)
        want_d = torch.from_numpy(want).to(cuda_dev)
        part = got[b0:b1]
        assert torch.equal(part > 0, want_d > 0), f"touched mask differs in samples [{b0}, {b1})"
        max_err = max(max_err, float((part - want_d).abs().max()))
        _cabi.checksum(cabi, want_d, b0 * sample_elems, d_or)
        del want_d
    assert max_err <= 1e-6, max_err
    assert int(d_or[1]) == int(d_gpu[1]) and int(d_or[2]) == int(d_gpu[2])  # mask hash, touched count
    # value-bits word: sum of fp32 bit patterns; a 2-ulp budget per touched pixel bounds the difference
    assert abs(int(d_or[0]) - int(d_gpu[0])) <= 2 * int(d_gpu[2])


def test_sharded_equals_unsharded_digest(cabi, cuda_dev):
    """G contiguous batch slices rendered independently give the unsharded result (digest additive)."""
    w = workloads.make("cfg5", B=96)
    cfg = w.cfg
    cf, bf, s, lab = _upload(w, cuda_dev)
    whole = torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=cuda_dev)
    _cabi.draw_from_boxes(cabi, whole, cf, bf, float(cfg.osf), s, lab)
    d_whole = _cabi.checksum(cabi, whole)
    sample_elems = cfg.C * cfg.H * cfg.W
    for G in (2, 8):
        d = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
        for g in range(G):
            lo, hi = workloads.shard_range(w.B, g, G)
            part = torch.empty((hi - lo, cfg.C, cfg.H, cfg.W), device=cuda_dev)
            _cabi.draw_from_boxes(cabi, part, cf[lo:hi].contiguous(), bf[lo:hi].contiguous(), float(cfg.osf),
                                  s[lo:hi].contiguous(), lab[lo:hi].contiguous())
            assert torch.equal(part, whole[lo:hi])
            _cabi.checksum(cabi, part, lo * sample_elems, d)
        assert torch.equal(d, d_whole)


def test_plane_offsets_are_64_bit(cabi, cuda_dev):
    """F5: B*C*H*W >= 2^31 elements in one call (the reference's int index overflows there)."""
    H = W = 1024
    B = 2049  # 2049 * 2^20 > 2^31 elements, 8.6 GB
    c = torch.zeros((B, 1, 2), dtype=torch.int32, device=cuda_dev)
    c[:, 0, 0] = 1000
    c[:, 0, 1] = 1020
    r = torch.full((B, 1), 2, dtype=torch.int32, device=cuda_dev)
    s = torch.ones(B, dtype=torch.int32, device=cuda_dev)
    heat = torch.empty((B, H, W), device=cuda_dev)
    _cabi.draw_batched(cabi, heat, c, r, s, None, 6.0, 1.0, _cabi.DH_MODE_FRESH)
    assert float(heat[B - 1, 1020, 1000]) == 1.0 and float(heat[B - 1, 1020, 1003]) == 0.0
    assert int((heat[B - 1] > 0).sum()) == 5 * 5 and int((heat[0] > 0).sum()) == 25
    assert int(_cabi.checksum(cabi, heat)[2]) == 25 * B
    del heat


def test_edge_cases(cabi, cuda_dev, oracle):
    dev = cuda_dev
    H, W = 37, 53  # W % 4 != 0 -> scalar store path
    objs = np.array([[5, 5, 2], [0, 0, 4], [52, 36, 3], [-2, 10, 3], [60, 10, 3], [20, -9, 3], [20, 20, 100],
                     [10, 30, 0], [30, 10, -1], [53, 36, 1], [25, 18, 7]], np.int32)
    n = len(objs)
    c, r = _dev(objs[None, :, :2], dev), _dev(objs[None, :, 2], dev)
    for k in (1.0, 0.8, 0.0, -0.5):
        for factor in (6.0, 1.0, 12.0, 40.0):
            for init in (0.0, 0.3):
                want = np.full((1, H, W), init, np.float32)
                oracle.draw_batched(want, objs[None, :, :2], objs[None, :, 2], np.array([n]), factor, k)
                heat = torch.full((1, H, W), init, device=dev)
                _cabi.draw_batched(cabi, heat, c, r, torch.tensor([n], device=dev), None, factor, k)
                got = heat.cpu().numpy()
                assert np.abs(got - want).max() <= 1e-6, (k, factor, init)
                # masks must agree except where the value is an fp32 denormal (exp argument < -87.3, reachable only
                # with factor >~ 13): there a 1-ulp difference between libdevice expf and the oracle's correctly
                # rounded exp can flip v > 0. Against the reference GPU build (same expf) masks are always equal.
                flip = (got > 0) != (want > 0)
                assert not (flip & (np.maximum(got, want) >= 1e-37)).any(), (k, factor, init)
    # sizes: 0, > T (clamped), negative (clamped to 0); T == 0; B == 0
    heat = torch.full((3, H, W), 5.0, device=dev)
    c3, r3 = c.expand(3, n, 2).contiguous(), r.expand(3, n).contiguous()
    _cabi.draw_batched(cabi, heat, c3, r3, torch.tensor([0, n + 50, -3], device=dev), None, 6.0, 1.0, _cabi.DH_MODE_FRESH)
    assert float(heat[0].abs().max()) == 0.0 and float(heat[2].abs().max()) == 0.0 and float(heat[1].max()) == 1.0
    heat = torch.full((2, H, W), 5.0, device=dev)
    e2 = torch.empty((2, 0, 2), dtype=torch.int32, device=dev)
    _cabi.draw_batched(cabi, heat, e2, e2[..., 0].contiguous(), torch.zeros(2, dtype=torch.int64, device=dev), None, 6.0, 1.0,
                       _cabi.DH_MODE_FRESH)
    assert float(heat.abs().max()) == 0.0
    # labels out of range never write out of bounds: skipped and counted
    C = 3
    lab = torch.tensor([[0, 1, 2, 3, -1, 7, 0, 1, 2, 0, 1]], dtype=torch.int32, device=dev)
    guard = torch.full((1 + 2, C, H, W), 2.0, device=dev)
    status = torch.zeros(4, dtype=torch.int32, device=dev)
    _cabi.draw_batched(cabi, guard[1:2], c, r, torch.tensor([n], device=dev), lab, 6.0, 1.0, _cabi.DH_MODE_FRESH, status)
    assert float(guard[0].min()) == 2.0 and float(guard[2].min()) == 2.0
    assert status.tolist()[0] == 3
    want = np.zeros((1, C, H, W), np.float32)
    oracle.draw_batched(want, objs[None, :, :2], objs[None, :, 2], np.array([n]), 6.0, 1.0, lab.cpu().numpy())
    assert np.abs(guard[1:2].cpu().numpy() - want).max() <= 1e-6
    # flat entry: bad plane indices skipped and counted; big N goes through the counting-sort path
    rng = np.random.default_rng(5)
    N, M = 20000, 300
    fc = np.stack([rng.integers(-5, W + 5, N), rng.integers(-5, H + 5, N)], 1).astype(np.int32)
    fr = rng.integers(0, 6, N).astype(np.int32)
    fi = rng.integers(-2, M + 2, N).astype(np.int32)
    want = np.zeros((M, H, W), np.float32)
    oracle.draw_flat(want, fc, fr, fi, 6.0, 1.0)
    status.zero_()
    heat = torch.zeros((M, H, W), device=dev)
    _cabi.draw_flat(cabi, heat, _dev(fc, dev), _dev(fr, dev), _dev(fi, dev), 6.0, 1.0, _cabi.DH_MODE_ACCUMULATE, status)
    got = heat.cpu().numpy()
    assert np.abs(got - want).max() <= 1e-6 and np.array_equal(got > 0, want > 0)
    assert status.tolist()[1] == int(((fi < 0) | (fi >= M)).sum())


def test_more_objects_than_the_shared_list(cabi, cuda_dev, oracle):
    """> 1024 overlapping objects in one plane: the renderer falls back to chunked passes over the source."""
    rng = np.random.default_rng(9)
    T, H, W = 3000, 64, 96
    c = np.stack([rng.integers(0, W, T), rng.integers(0, H, T)], 1).astype(np.int32)[None]
    r = rng.integers(20, 70, T).astype(np.int32)[None]
    want = np.zeros((1, H, W), np.float32)
    oracle.draw_batched(want, c, r, np.array([T]), 6.0, 1.0)
    heat = torch.empty((1, H, W), device=cuda_dev)
    _cabi.draw_batched(cabi, heat, _dev(c, cuda_dev), _dev(r, cuda_dev), torch.tensor([T], device=cuda_dev), None, 6.0, 1.0,
                       _cabi.DH_MODE_FRESH)
    assert np.abs(heat.cpu().numpy() - want).max() <= 1e-6


def test_host_buffer_entry(cabi, cuda_dev, oracle):
    w = workloads.make("cfg2")
    cfg = w.cfg
    ci, ri = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    want = np.zeros((w.B, cfg.C, cfg.H, cfg.W), np.float32)
    oracle.draw_batched(want, ci, ri, w.sizes, 6.0, 1.0, w.labels)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    cf, bf, s, lab = pin(w.centers_f), pin(w.bboxes_f), pin(w.sizes), pin(w.labels)
    args = (w.B, w.T, cfg.C, cfg.H, cfg.W)
    st = _cabi.stream_ptr()
    for staging, chunk in ((1, 3), (0, 5), (1, 0)):
        nbytes = cabi.dh_host_scratch_bytes(*args, 1, chunk, staging)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=cuda_dev)
        out = torch.full(want.shape, 3.0).pin_memory()
        ck = torch.zeros(3, dtype=torch.int64).pin_memory()
        hd = None if staging else torch.empty(want.shape, device=cuda_dev)
        rc = cabi.dh_draw_heatmap_batched_from_boxes_host(
            st, _cabi.ptr(out), _cabi.ptr(hd), _cabi.ptr(ck), _cabi.ptr(cf), _cabi.ptr(bf), float(cfg.osf), _cabi.ptr(s),
            _cabi.ptr(lab), *args, 6.0, 1.0, _cabi.ptr(scratch), nbytes, chunk)
        _cabi.check(cabi, rc, "host entry")
        torch.cuda.synchronize()
        assert np.abs(out.numpy() - want).max() <= 1e-6 and np.array_equal(out.numpy() > 0, want > 0)
        assert np.array_equal(ck.numpy(), oracle.digest_numpy(out.numpy()))
        if hd is not None:
            assert np.array_equal(hd.cpu().numpy(), out.numpy())
    # device-resident result + digest only
    nbytes = cabi.dh_host_scratch_bytes(*args, 1, 0, 0)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=cuda_dev)
    hd = torch.empty(want.shape, device=cuda_dev)
    ck2 = torch.zeros(3, dtype=torch.int64).pin_memory()
    rc = cabi.dh_draw_heatmap_batched_from_boxes_host(
        st, None, _cabi.ptr(hd), _cabi.ptr(ck2), _cabi.ptr(cf), _cabi.ptr(bf), float(cfg.osf), _cabi.ptr(s), _cabi.ptr(lab),
        *args, 6.0, 1.0, _cabi.ptr(scratch), nbytes, 0)
    _cabi.check(cabi, rc, "host entry (device result)")
    torch.cuda.synchronize()
    assert np.array_equal(ck2.numpy(), ck.numpy())
    assert cabi.dh_draw_heatmap_batched_from_boxes_host(
        st, None, _cabi.ptr(hd), None, _cabi.ptr(cf), _cabi.ptr(bf), float(cfg.osf), _cabi.ptr(s), _cabi.ptr(lab), *args,
        6.0, 1.0, _cabi.ptr(scratch), 16, 0) == _cabi.DH_ERR_WORKSPACE_TOO_SMALL


def test_host_entry_pinned_gather_equals_pageable_copies(cabi, cuda_dev, oracle):
    """Small inputs: pinned arrays are read in place by one gather launch, pageable ones (or a mix) go through
    cudaMemcpyAsync; odd offsets exercise the 4-byte path of the gather. Same heatmap, same digest, no labels too."""
    w = workloads.make("cfg2")
    cfg = w.cfg
    args = (w.B, w.T, cfg.C, cfg.H, cfg.W)
    nbytes = cabi.dh_host_scratch_bytes(*args, 1, 0, 0)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=cuda_dev)
    st = _cabi.stream_ptr()

    def shifted_pin(a):  # pinned, but 4 bytes off the 16-byte grid
        flat = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
        t = torch.empty(flat.size + 4, dtype=torch.uint8).pin_memory()
        t[4:] = torch.from_numpy(flat.copy())
        return t[4:]

    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    plain = lambda a: torch.from_numpy(np.ascontiguousarray(a).copy())
    results = []
    for mk, launches in ((pin, 2), (plain, 1), (shifted_pin, 2), (None, 1)):
        if mk is None:  # a mix: one array pageable -> the whole call takes the copy path
            cf, bf, s, lab = pin(w.centers_f), plain(w.bboxes_f), pin(w.sizes), pin(w.labels)
        else:
            cf, bf, s, lab = mk(w.centers_f), mk(w.bboxes_f), mk(w.sizes), mk(w.labels)
        for with_labels in (True, False):
            shape = (w.B, cfg.C, cfg.H, cfg.W) if with_labels else (w.B, cfg.H, cfg.W)
            hd = torch.full(shape, 7.0, device=cuda_dev)
            ck = torch.zeros(3, dtype=torch.int64).pin_memory()
            l0 = cabi.dh_launch_count()
            rc = cabi.dh_draw_heatmap_batched_from_boxes_host(
                st, None, _cabi.ptr(hd), _cabi.ptr(ck), _cabi.ptr(cf), _cabi.ptr(bf), float(cfg.osf), _cabi.ptr(s),
                _cabi.ptr(lab) if with_labels else None, *args, 6.0, 1.0, _cabi.ptr(scratch), nbytes, 0)
            _cabi.check(cabi, rc, "host entry")
            torch.cuda.synchronize()
            assert cabi.dh_launch_count() - l0 == launches
            got = hd.cpu().numpy()
            assert np.array_equal(ck.numpy(), oracle.digest_numpy(got))
            results.append((with_labels, got))
    for wl, got in results[2:]:
        ref = results[0][1] if wl else results[1][1]
        assert np.array_equal(got, ref)
    ci, ri = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    want = np.zeros(results[0][1].shape, np.float32)
    oracle.draw_batched(want, ci, ri, w.sizes, 6.0, 1.0, w.labels)
    assert np.abs(results[0][1] - want).max() <= 1e-6 and np.array_equal(results[0][1] > 0, want > 0)


def test_host_entry_pipelines_large_inputs(cabi, cuda_dev, oracle):
    """Large inputs (>= 1.5 MiB; here 3.5 MiB): the host entry copies them in geometrically growing chunks on its copy stream while earlier chunks render on two
    alternating streams. Result and digest must equal the single-launch path (chunk_samples < 0) and the oracle."""
    cfg = workloads.Config("pipe", 2048, 2, 32, 64, 4, 0, 64, 64, 96, 777)
    w = workloads.make(cfg)
    assert w.B * w.T * 28 >= (2 << 20)
    ci, ri = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    want = np.zeros((w.B, cfg.C, cfg.H, cfg.W), np.float32)
    oracle.draw_batched(want, ci, ri, w.sizes, 6.0, 1.0, w.labels)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    cf, bf, s, lab = pin(w.centers_f), pin(w.bboxes_f), pin(w.sizes), pin(w.labels)
    args = (w.B, w.T, cfg.C, cfg.H, cfg.W)
    st = _cabi.stream_ptr()
    nbytes = cabi.dh_host_scratch_bytes(*args, 1, 0, 0)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=cuda_dev)
    ref_ck = oracle.digest_numpy(want)
    results = []
    for chunk in (-1, 0, 100, 1, 0):  # single launch | default chunks | 21 -> 16 chunks | B chunks -> 16 | again (reuse)
        hd = torch.full(want.shape, 7.0, device=cuda_dev)
        ck = torch.zeros(3, dtype=torch.int64).pin_memory()
        l0 = cabi.dh_launch_count()
        rc = cabi.dh_draw_heatmap_batched_from_boxes_host(
            st, None, _cabi.ptr(hd), _cabi.ptr(ck), _cabi.ptr(cf), _cabi.ptr(bf), float(cfg.osf), _cabi.ptr(s),
            _cabi.ptr(lab), *args, 6.0, 1.0, _cabi.ptr(scratch), nbytes, chunk)
        _cabi.check(cabi, rc, "host entry (pipelined)")
        torch.cuda.synchronize()
        nl = cabi.dh_launch_count() - l0
        assert (nl == 2) if chunk < 0 else (2 <= nl <= 16)  # single shot = zero-copy gather of the pinned inputs + render
        results.append(hd.cpu().numpy())
        assert np.array_equal(ck.numpy(), oracle.digest_numpy(results[-1]))
    for r in results[1:]:
        assert np.array_equal(r, results[0])
    assert np.abs(results[0] - want).max() <= 1e-6 and np.array_equal(results[0] > 0, want > 0)
    del ref_ck


def test_digest_index_bases_beyond_32_bits(cabi, cuda_dev, oracle):
    """dh_checksum and the digest fused into the renderer agree with the numpy restatement when pixel indices need
    more than 32 bits (index_base of a late shard of a very large job)."""
    rng = np.random.default_rng(11)
    v = (rng.random(50001, dtype=np.float32) * (rng.random(50001) < 0.3)).astype(np.float32)
    for base in (0, (1 << 32) - 7, (1 << 33) + 5, (1 << 40) + 123):
        d = _cabi.checksum(cabi, _dev(v, cuda_dev), base).cpu().numpy()
        assert np.array_equal(d, oracle.digest_numpy(v, base))
    w = workloads.make("cfg1")
    cfg = w.cfg
    heat = torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=cuda_dev)
    for base in (0, (1 << 32) - 1000, (1 << 35) + 64):
        dg = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
        rc = cabi.dh_draw_heatmap_batched_from_boxes_digest(
            _cabi.stream_ptr(), _cabi.ptr(heat), _cabi.ptr(_dev(w.centers_f, cuda_dev)), _cabi.ptr(_dev(w.bboxes_f, cuda_dev)),
            float(cfg.osf), _cabi.ptr(_dev(w.sizes, cuda_dev)), _cabi.DH_SIZES_I64, _cabi.ptr(_dev(w.labels, cuda_dev)),
            w.B, w.T, cfg.C, cfg.H, cfg.W, 6.0, 1.0, base, _cabi.ptr(dg), None)
        _cabi.check(cabi, rc, "fused digest")
        torch.cuda.synchronize()
        assert np.array_equal(dg.cpu().numpy(), oracle.digest_numpy(heat.cpu().numpy(), base))


def test_expf_is_monotone_and_close_to_correctly_rounded(cabi, cuda_dev):
    """Exhaustive over every fp32 in [-104, -0]: the renderer's expf is monotone (so one expf per pixel is bit-identical
    to max over per-object expf) and within 2 ulp of the correctly rounded value (the oracle's exp)."""
    lo, hi = 0x80000000, 0xC2D00000  # -0.0 ... -104.0 (bit patterns increase as the value decreases)
    step = 1 << 26
    prev_last = None
    worst = 0
    for start in range(lo, hi, step):
        n = min(step, hi - start) + 1
        out = torch.empty(n, device=cuda_dev)
        _cabi.check(cabi, cabi.dh_expf_probe(_cabi.stream_ptr(), ctypes.c_uint32(start), n, _cabi.ptr(out)), "expf_probe")
        assert bool((out[1:] <= out[:-1]).all()), "expf not monotone"
        if prev_last is not None:
            assert float(out[0]) <= prev_last
        prev_last = float(out[-1])
        x = (torch.arange(n, device=cuda_dev, dtype=torch.int64) + start).to(torch.int32)  # wraps to the same bits
        ref = torch.exp(x.view(torch.float32).double()).float()
        normal = ref >= 1.1754944e-38
        d = (out.view(torch.int32).long() - ref.view(torch.int32).long()).abs()[normal]
        worst = max(worst, int(d.max()))
        del out, x, ref, d
    assert worst <= 2, worst


@pytest.mark.parametrize("H,W,T", [(40, 1024, 300), (9, 4096, 50), (24, 640, 1500), (5, 8192, 40), (12, 2048, 1500)])
def test_wide_maps_use_the_block_cooperative_and_generic_kernels(cabi, cuda_dev, oracle, H, W, T):
    """W <= 1024 -> groups_kernel (T = 1500 overflows its shared list: warps scan the source); W in (1024, 4096] ->
    rows_kernel (block-cooperative bands; T = 1500: multi-pass with a persistent tile); W > 4096 -> generic kernel."""
    rng = np.random.default_rng(H * W + T)
    B, C = 3, 2
    c = np.stack([rng.integers(-20, W + 20, (B, T)), rng.integers(-5, H + 5, (B, T))], 2).astype(np.int32)
    r = rng.integers(0, 40, (B, T)).astype(np.int32)
    lab = rng.integers(0, C, (B, T)).astype(np.int32)
    sizes = np.array([T, T // 2, 0], np.int64)
    for k, init in ((1.0, 0.0), (0.7, 0.2), (-1.0, 0.0)):
        want = np.full((B, C, H, W), init, np.float32)
        oracle.draw_batched(want, c, r, sizes, 6.0, k, lab)
        heat = torch.full((B, C, H, W), init, device=cuda_dev)
        _cabi.draw_batched(cabi, heat, _dev(c, cuda_dev), _dev(r, cuda_dev), _dev(sizes, cuda_dev), _dev(lab, cuda_dev), 6.0, k)
        got = heat.cpu().numpy()
        assert np.abs(got - want).max() <= 1e-6 and np.array_equal(got > 0, want > 0), (k, init)
        if init == 0.0:
            fresh = torch.full((B, C, H, W), 7.0, device=cuda_dev)
            _cabi.draw_batched(cabi, fresh, _dev(c, cuda_dev), _dev(r, cuda_dev), _dev(sizes, cuda_dev), _dev(lab, cuda_dev),
                               6.0, k, _cabi.DH_MODE_FRESH)
            assert torch.equal(fresh, heat)


@pytest.mark.parametrize("H,W", [(16, 4), (300, 8), (33, 128), (7, 512), (64, 256)])
def test_small_and_odd_shapes_on_the_warp_kernel(cabi, cuda_dev, oracle, H, W):
    """Row groups of several rows (W < 512), partial last bands/groups, H smaller than a band."""
    rng = np.random.default_rng(H * 1000 + W)
    B, T = 5, 40
    c = np.stack([rng.integers(-3, W + 3, (B, T)), rng.integers(-3, H + 3, (B, T))], 2).astype(np.int32)
    r = rng.integers(0, 12, (B, T)).astype(np.int32)
    sizes = rng.integers(0, T + 1, B).astype(np.int64)
    want = np.zeros((B, H, W), np.float32)
    oracle.draw_batched(want, c, r, sizes, 6.0, 1.0)
    for mode in (_cabi.DH_MODE_FRESH, _cabi.DH_MODE_ACCUMULATE):
        heat = torch.zeros((B, H, W), device=cuda_dev) if mode == _cabi.DH_MODE_ACCUMULATE else torch.full((B, H, W), 3.0, device=cuda_dev)
        _cabi.draw_batched(cabi, heat, _dev(c, cuda_dev), _dev(r, cuda_dev), _dev(sizes, cuda_dev), None, 6.0, 1.0, mode)
        got = heat.cpu().numpy()
        assert np.abs(got - want).max() <= 1e-6 and np.array_equal(got > 0, want > 0), mode


@pytest.mark.parametrize("H,W", [(232, 400), (64, 256), (50, 1024), (31, 36)])
def test_sparse_register_path_all_signs_and_modes(cabi, cuda_dev, oracle, H, W):
    """<= 2 objects per row group: the register-only path of groups_kernel (no staged read-modify-write), for
    k > 0, k < 0 and k == 0, FRESH and ACCUMULATE into non-zero planes, clipped at every border."""
    rng = np.random.default_rng(H * 7 + W)
    B, T, C = 8, 3, 2
    c = np.stack([rng.integers(-4, W + 4, (B, T)), rng.integers(-4, H + 4, (B, T))], 2).astype(np.int32)
    r = rng.integers(0, 25, (B, T)).astype(np.int32)
    lab = rng.integers(0, C, (B, T)).astype(np.int32)
    sizes = rng.integers(0, T + 1, B).astype(np.int32)
    args = (_dev(c, cuda_dev), _dev(r, cuda_dev), _dev(sizes, cuda_dev), _dev(lab, cuda_dev))
    for k in (1.0, 0.8, -0.5, 0.0):
        for init in (0.0, 0.3, -0.2):
            want = np.full((B, C, H, W), init, np.float32)
            oracle.draw_batched(want, c, r, sizes.astype(np.int64), 6.0, k, lab)
            heat = torch.full((B, C, H, W), init, device=cuda_dev)
            _cabi.draw_batched(cabi, heat, *args, 6.0, k)
            got = heat.cpu().numpy()
            assert np.abs(got - want).max() <= 1e-6 and np.array_equal(got > 0, want > 0), (k, init)
        fresh = torch.full((B, C, H, W), 5.0, device=cuda_dev)
        _cabi.draw_batched(cabi, fresh, *args, 6.0, k, _cabi.DH_MODE_FRESH)
        zero = torch.zeros((B, C, H, W), device=cuda_dev)
        _cabi.draw_batched(cabi, zero, *args, 6.0, k)
        assert torch.equal(fresh, zero), k


@pytest.mark.parametrize("name,nb", [("cfg1", None), ("cfg2", None), ("cfg3", 24), ("cfg4", 3)])
def test_fused_digest_equals_checksum_pass(cabi, cuda_dev, name, nb):
    """dh_draw_heatmap_batched_from_boxes_digest: the digest accumulated inside the render launch equals dh_checksum
    of the rendered heatmap bit for bit (sparse and dense paths), and is additive over batch shards."""
    w = workloads.make(name, B=nb)
    cfg = w.cfg
    cf, bf, s, lab = _upload(w, cuda_dev)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    plain = torch.empty(shape, device=cuda_dev)
    _cabi.draw_from_boxes(cabi, plain, cf, bf, float(cfg.osf), s, lab, 6.0, 0.9)
    want = _cabi.checksum(cabi, plain)
    fused = torch.full(shape, 2.0, device=cuda_dev)
    got = _cabi.draw_from_boxes_digest(cabi, fused, cf, bf, float(cfg.osf), s, lab, 6.0, 0.9)
    assert torch.equal(fused, plain)
    assert torch.equal(got, want), (got.tolist(), want.tolist())
    assert int(want[2]) == int((plain > 0).sum())
    sample_elems = cfg.C * cfg.H * cfg.W
    d = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    for g in range(2):
        lo, hi = workloads.shard_range(w.B, g, 2)
        part = torch.empty((hi - lo,) + shape[1:], device=cuda_dev)
        _cabi.draw_from_boxes_digest(cabi, part, cf[lo:hi].contiguous(), bf[lo:hi].contiguous(), float(cfg.osf),
                                     s[lo:hi].contiguous(), lab[lo:hi].contiguous(), 6.0, 0.9, lo * sample_elems, d)
    assert torch.equal(d, want)
