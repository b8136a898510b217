"""Round-2 additions at the drop-in boundary:
  * slot arrays that are only 4-byte aligned (a contiguous torch view at an odd storage offset) — the reference reads
    them with scalar int loads (cuh:57-58,86-87) and accepts them; so must we;
  * dh_status.skipped_radius / skipped_label count objects of EVERY class;
  * keyword-only `assume_zero_init` on the reference-named entry points (Missing #5);
  * libdh_ref_shim.so: the reference's C++-linkage launchers (draw_heatmap_cuda.cu:29-31,43-46) by mangled name
    (Missing #3).
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import _cabi
import workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "accv-lab_b200", "accvlab", "draw_heatmap")
SHIM = os.path.join(PKG, "libdh_ref_shim.so")
SYM_FLAT = "_Z17draw_heatmap_cudaP11CUstream_stPfPKiS3_S3_iiffii"
SYM_BATCHED = "_Z25draw_heatmap_batched_cudaP11CUstream_stPfPKiS3_S3_iiiffiiS3_"


# ---- CPU ---------------------------------------------------------------------------------------------------------------
def test_shim_exports_the_reference_launchers_by_mangled_name():
    """What g++ emits for the declarations in DH/benchmark_cpp/bench_*.cpp:24-27 is what the shim must define."""
    assert os.path.exists(SHIM), "libdh_ref_shim.so missing: run accv-lab_b200/build.py"
    out = subprocess.run(["nm", "-D", "--defined-only", SHIM], capture_output=True, text=True, check=True).stdout
    assert SYM_FLAT in out and SYM_BATCHED in out, out
    # the names are derived, not guessed: compile the reference-style declarations and read the undefined symbols
    src = ("#include <cuda_runtime.h>\n"
           "void draw_heatmap_cuda(cudaStream_t, float*, const int*, const int*, const int*, int, int, float, float, int, int);\n"
           "void draw_heatmap_batched_cuda(cudaStream_t, float*, const int*, const int*, const int*, int, int, int, float, float,"
           " int, const int = 0, const int* = nullptr);\n"
           "void f() { draw_heatmap_cuda(0,0,0,0,0,0,0,0,0,0,0); draw_heatmap_batched_cuda(0,0,0,0,0,0,0,0,0,0,0); }\n")
    import tempfile

    with tempfile.TemporaryDirectory() as td:
        cpp, obj = os.path.join(td, "d.cpp"), os.path.join(td, "d.o")
        open(cpp, "w").write(src)
        subprocess.run(["g++", "-c", cpp, "-o", obj, "-I/usr/local/cuda/include"], check=True)
        und = subprocess.run(["nm", "-u", obj], capture_output=True, text=True, check=True).stdout
    assert SYM_FLAT in und and SYM_BATCHED in und, und


def test_wrapper_signature_keeps_reference_positionals_and_adds_keyword_only_flag():
    import inspect

    import accvlab.draw_heatmap as dh

    sig = inspect.signature(dh.draw_heatmap_batched)
    names = list(sig.parameters)
    assert names[:6] == ["heatmap", "centers", "radii", "diameter_to_sigma_factor", "k_scale", "labels"]
    p = sig.parameters["assume_zero_init"]
    assert p.kind is inspect.Parameter.KEYWORD_ONLY and p.default is False
    from accvlab.draw_heatmap import draw_heatmap_ext as ext

    for fn in (ext.draw_heatmap, ext.draw_heatmap_batched_impl, ext.draw_heatmap_batched_classwise_impl):
        assert "*, assume_zero_init: bool = False" in fn.__doc__, fn.__doc__


# ---- GPU ---------------------------------------------------------------------------------------------------------------
def _misaligned_copy(t, torch):
    """Same values in a contiguous tensor whose data pointer is 4 (mod 8 and mod 16)."""
    flat = torch.empty(t.numel() + 4, dtype=t.dtype, device=t.device)  # 4-byte elements
    off = ((4 - flat.data_ptr() % 16) % 16) // 4  # element offset that puts the view at 4 (mod 16)
    v = flat[off:off + t.numel()].view(t.shape)
    v.copy_(t)
    assert v.is_contiguous() and v.data_ptr() % 8 == 4
    return v


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", ["cfg1", "cfg3"])
def test_misaligned_slot_arrays_are_accepted(cuda_dev, cabi, cfg):
    import torch

    w = workloads.make(cfg, B=6 if cfg == "cfg3" else None)
    c = w.cfg
    dev = cuda_dev
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes, labels = torch.from_numpy(w.sizes).to(dev), torch.from_numpy(w.labels).to(dev)
    ci, ri = _cabi.radius_center(cabi, cf, bf, float(c.osf))
    shape = (w.B, c.C, c.H, c.W)
    want = _cabi.draw_batched(cabi, torch.zeros(shape, device=dev), ci, ri, sizes, labels)
    # integer centres at 4 (mod 8)
    ci_m = _misaligned_copy(ci, torch)
    for mode in (_cabi.DH_MODE_ACCUMULATE, _cabi.DH_MODE_FRESH):
        got = _cabi.draw_batched(cabi, torch.zeros(shape, device=dev), ci_m, ri, sizes, labels, mode=mode)
        assert torch.equal(got, want)
    # float centres / boxes at 4 (mod 8) / 4 (mod 16), K1 fused
    cf_m, bf_m = _misaligned_copy(cf, torch), _misaligned_copy(bf, torch)
    got = _cabi.draw_from_boxes(cabi, torch.empty(shape, device=dev), cf_m, bf_m, float(c.osf), sizes, labels)
    assert torch.equal(got, want)
    got16 = torch.empty(shape, device=dev, dtype=torch.bfloat16)
    _cabi.render16(cabi, got16, sizes, labels, centers_f=cf_m, bboxes_f=bf_m, osf=float(c.osf))
    assert torch.equal(got16, want.to(torch.bfloat16))
    # the Python drop-in too
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap.b200 import RaggedBatch

    heat = torch.zeros(shape, device=dev)
    rb = lambda t: RaggedBatch(t, sample_sizes=sizes)
    dh.draw_heatmap_batched(heat, rb(ci_m), rb(ri), 6.0, 1.0, rb(labels))
    torch.cuda.synchronize()
    assert torch.equal(heat, want)


@pytest.mark.gpu
@pytest.mark.parametrize("source", ["int", "boxes"])
def test_status_counts_skips_of_every_class(cuda_dev, cabi, source):
    """Oversize radii on labels 0, 3 and 7, bad labels -1 and C: all five are counted, once each, nothing is drawn for them."""
    import torch

    dev = cuda_dev
    B, T, C, H, W = 3, 12, 8, 64, 96
    big = 4194304 + 5  # > DH_MAX_RADIUS
    centers = torch.full((B, T, 2), 20, dtype=torch.int32)
    radii = torch.full((B, T), 2, dtype=torch.int32)
    labels = (torch.arange(T, dtype=torch.int32) % C).repeat(B, 1)
    sizes = torch.tensor([T, T - 2, 0], dtype=torch.int64)
    radii[0, 0], labels[0, 0] = big, 0
    radii[0, 3], labels[0, 3] = big, 3
    radii[1, 7], labels[1, 7] = big, 7
    labels[0, 5] = -1
    labels[1, 2] = C
    radii[1, T - 1] = big  # padded slot (t >= size): never counted, never drawn
    radii[2, :] = big      # empty sample
    centers, radii, labels, sizes = (t.to(dev) for t in (centers, radii, labels, sizes))
    status = torch.zeros(4, dtype=torch.int32, device=dev)
    heat = torch.zeros((B, C, H, W), device=dev)
    if source == "int":
        _cabi.draw_batched(cabi, heat, centers, radii, sizes, labels, status=status)
    else:  # float boxes whose K1 radius is `radii` at osf 1: centre c, box c -/+ r
        cf = centers.float() + 0.5
        r = radii.float()[..., None]
        bf = torch.cat([cf - r, cf + r], dim=-1)
        st = cabi.dh_draw_heatmap_batched_from_boxes(_cabi.stream_ptr(), _cabi.ptr(heat), _cabi.ptr(cf), _cabi.ptr(bf), 1.0,
                                                     _cabi.ptr(sizes), _cabi.DH_SIZES_I64, _cabi.ptr(labels), B, T, C, H, W,
                                                     6.0, 1.0, _cabi.DH_MODE_ACCUMULATE, None, None, _cabi.ptr(status))
        _cabi.check(cabi, st)
    torch.cuda.synchronize()
    s = status.cpu().numpy().view(np.uint32)
    assert int(s[0]) == 2, f"skipped_label {s}"
    assert int(s[2]) == 3, f"skipped_radius {s}"
    # what was drawn: exactly the valid small objects
    keep = (torch.arange(T, device=dev)[None, :] < sizes[:, None]) & (radii <= 4194304) & (labels >= 0) & (labels < C)
    want = torch.zeros_like(heat)
    r2 = torch.where(keep, radii, torch.full_like(radii, -1))
    _cabi.draw_batched(cabi, want, centers, r2, sizes, torch.clamp(labels, 0, C - 1))
    assert torch.equal(heat, want)


@pytest.mark.gpu
def test_assume_zero_init_on_the_reference_named_entries(cuda_dev):
    """Garbage in, fresh heatmap out; without the flag the same call max-blends into the garbage (reference semantics)."""
    import torch

    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap import draw_heatmap_ext as ext
    from accvlab.draw_heatmap.b200 import RaggedBatch, radius_center

    dev = cuda_dev
    w = workloads.make("cfg2", B=5)
    c = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes, labels = torch.from_numpy(w.sizes).to(dev), torch.from_numpy(w.labels).to(dev)
    ci, ri = radius_center(cf, bf, c.osf)
    rb = lambda t: RaggedBatch(t, sample_sizes=sizes)
    shape = (w.B, c.C, c.H, c.W)
    want = torch.zeros(shape, device=dev)
    dh.draw_heatmap_batched(want, rb(ci), rb(ri), 6.0, 0.8, rb(labels))
    garbage = torch.full(shape, 7.0, device=dev)
    heat = garbage.clone()
    assert dh.draw_heatmap_batched(heat, rb(ci), rb(ri), 6.0, 0.8, rb(labels), assume_zero_init=True) is None
    assert torch.equal(heat, want)
    heat = garbage.clone()
    dh.draw_heatmap_batched(heat, rb(ci), rb(ri), 6.0, 0.8, rb(labels))
    assert torch.equal(heat, garbage)  # 7.0 >= every Gaussian value: accumulate keeps the old contents (cuh:30)
    # no labels, and the impl functions directly
    want1 = torch.zeros((w.B, c.H, c.W), device=dev)
    ext.draw_heatmap_batched_impl(want1, ci, ri, sizes, 6.0, 0.8)
    heat1 = torch.full_like(want1, -3.0)
    ext.draw_heatmap_batched_impl(heat1, ci, ri, sizes, 6.0, 0.8, assume_zero_init=True)
    assert torch.equal(heat1, want1)
    with pytest.raises(TypeError):
        ext.draw_heatmap_batched_impl(heat1, ci, ri, sizes, 6.0, 0.8, True)  # keyword-only: the positional list is the reference's
    # flat entry
    valid = torch.arange(w.T, device=dev)[None, :] < sizes[:, None]
    idx = torch.arange(w.B, device=dev, dtype=torch.int32)[:, None].expand(w.B, w.T)[valid].contiguous()
    heatf = torch.full_like(want1, 5.0)
    ext.draw_heatmap(heatf, ci[valid].contiguous(), ri[valid].contiguous(), idx, 6.0, 0.8, assume_zero_init=True)
    assert torch.equal(heatf, want1)
    # 16-bit planes stay out of the reference-named entries
    with pytest.raises(RuntimeError):
        ext.draw_heatmap_batched_impl(heat1.to(torch.bfloat16), ci, ri, sizes, 6.0, 0.8, assume_zero_init=True)


@pytest.mark.gpu
def test_shim_launchers_called_like_the_reference_cpp_benchmarks(cuda_dev, cabi):
    """ctypes call of the two mangled C++ symbols with the argument lists of bench_batched_classwise.cpp:123-126 and
    bench_flattened.cpp; compared with the C ABI on the same inputs."""
    import torch

    dev = cuda_dev
    shim = ctypes.CDLL(SHIM)
    p, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    fb = getattr(shim, SYM_BATCHED)
    fb.restype, fb.argtypes = None, [p, p, p, p, p, i, i, i, f, f, i, i, p]
    ff = getattr(shim, SYM_FLAT)
    ff.restype, ff.argtypes = None, [p, p, p, p, p, i, i, f, f, i, i]
    w = workloads.make("cfg2", B=6)
    c = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes32, labels = torch.from_numpy(w.sizes.astype(np.int32)).to(dev), torch.from_numpy(w.labels).to(dev)
    ci, ri = _cabi.radius_center(cabi, cf, bf, float(c.osf))
    sp = _cabi.stream_ptr()
    # class-wise
    want = _cabi.draw_batched(cabi, torch.zeros((w.B, c.C, c.H, c.W), device=dev), ci, ri, sizes32, labels)
    got = torch.zeros_like(want)
    fb(sp, _cabi.ptr(got), _cabi.ptr(ci), _cabi.ptr(ri), _cabi.ptr(sizes32), c.H, c.W, w.T, 6.0, 1.0, w.B, c.C, _cabi.ptr(labels))
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    # one plane per sample: max_num_classes = 0, labels = nullptr (the declaration's defaults)
    want1 = _cabi.draw_batched(cabi, torch.zeros((w.B, c.H, c.W), device=dev), ci, ri, sizes32)
    got1 = torch.zeros_like(want1)
    fb(sp, _cabi.ptr(got1), _cabi.ptr(ci), _cabi.ptr(ri), _cabi.ptr(sizes32), c.H, c.W, w.T, 6.0, 1.0, w.B, 0, None)
    torch.cuda.synchronize()
    assert torch.equal(got1, want1)
    # flat, small (direct scan) and large enough to need the counting-sort workspace (taken from the stream's pool)
    for reps in (1, 40):
        valid = torch.arange(w.T, device=dev)[None, :] < torch.from_numpy(w.sizes).to(dev)[:, None]
        idx = torch.arange(w.B, device=dev, dtype=torch.int32)[:, None].expand(w.B, w.T)[valid]
        cc, rr = ci[valid], ri[valid]
        M = w.B * reps
        idx = torch.cat([idx + w.B * k for k in range(reps)]).contiguous()
        cc, rr = cc.repeat(reps, 1).contiguous(), rr.repeat(reps).contiguous()
        gotf = torch.zeros((M, c.H, c.W), device=dev)
        ff(sp, _cabi.ptr(gotf), _cabi.ptr(cc), _cabi.ptr(rr), _cabi.ptr(idx), c.H, c.W, 6.0, 1.0, int(rr.numel()), M)
        torch.cuda.synchronize()
        assert torch.equal(gotf, want1.repeat(reps, 1, 1))
