"""ctypes/numpy front end of the CPU oracle (oracle/dh_oracle.c).  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (accv-lab_b200/) must never import this module.

Functions mirror the reference (paths relative to /root/reference/packages/draw_heatmap):
  radius_center           tests/_test_helpers.py:20-28
  draw_batched / draw_flat   accvlab/draw_heatmap/include/draw_heatmap_cuda_kernel.cuh:36-109 (fp32 "spec" family)
  refcpu_draw_batched     tests/_gaussian_drawing.py:215-240 + tests/_test_helpers.py:41-74 (float64 "refcpu" family)
  *_numpy                 independent numpy restatements used to cross-check the C code on small cases
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libdh_oracle.so")
        if not os.path.exists(path):
            from . import build_oracle  # type: ignore

            build_oracle.build_c_oracle()
        L = ctypes.CDLL(path)
        c_f, c_i32, c_i64, c_p = ctypes.c_float, ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p
        L.dho_radius_center.argtypes = [c_p, c_p, c_f, c_i64, c_p, c_p]
        L.dho_radius_center.restype = None
        L.dho_draw_batched.argtypes = [c_p, c_p, c_p, c_p, c_p] + [c_i64] * 5 + [c_f, c_f]
        L.dho_draw_batched.restype = None
        L.dho_draw_flat.argtypes = [c_p, c_p, c_p, c_p] + [c_i64] * 4 + [c_f, c_f]
        L.dho_draw_flat.restype = None
        L.dho_refcpu_draw_batched.argtypes = [c_p, c_p, c_p, c_p, c_p] + [c_i64] * 5 + [ctypes.c_double, c_f,
                                                                                       ctypes.c_int, ctypes.c_int]
        L.dho_refcpu_draw_batched.restype = None
        L.dho_draw_gaussians.argtypes = [c_p] * 7 + [c_i64] * 6 + [c_f]
        L.dho_draw_gaussians.restype = None
        _LIB = L
    return _LIB


def _c(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def radius_center(centers, bboxes, out_size_factor):
    """centers f32 [...,2], bboxes f32 [...,4] -> (centers_int i32 [...,2], radii i32 [...])."""
    centers = _c(centers, np.float32)
    bboxes = _c(bboxes, np.float32)
    n = centers.size // 2
    co = np.empty(centers.shape, np.int32)
    ro = np.empty(centers.shape[:-1], np.int32)
    lib().dho_radius_center(_ptr(centers), _ptr(bboxes), float(out_size_factor), n, _ptr(co), _ptr(ro))
    return co, ro


def draw_batched(heat, centers, radii, sizes, factor=6.0, k_scale=1.0, labels=None):
    """In-place max-blend into `heat` (f32 [B,H,W] or [B,C,H,W] C-contiguous numpy array). Returns heat."""
    assert heat.dtype == np.float32 and heat.flags.c_contiguous
    centers = _c(centers, np.int32)
    radii = _c(radii, np.int32)
    sizes = _c(sizes, np.int64)
    B, T = radii.shape
    if labels is not None:
        labels = _c(labels, np.int32)
        assert heat.ndim == 4
        C, H, W = heat.shape[1:]
    else:
        assert heat.ndim == 3
        C, (H, W) = 1, heat.shape[1:]
    assert heat.shape[0] == B
    lib().dho_draw_batched(_ptr(heat), _ptr(centers), _ptr(radii), _ptr(sizes), _ptr(labels), B, T, C, H, W,
                           float(factor), float(k_scale))
    return heat


def draw_flat(heat, centers, radii, heatmap_idxes, factor=6.0, k_scale=1.0):
    assert heat.dtype == np.float32 and heat.flags.c_contiguous and heat.ndim == 3
    centers = _c(centers, np.int32)
    radii = _c(radii, np.int32)
    idx = _c(heatmap_idxes, np.int32)
    M, H, W = heat.shape
    lib().dho_draw_flat(_ptr(heat), _ptr(centers), _ptr(radii), _ptr(idx), radii.shape[0], M, H, W, float(factor),
                        float(k_scale))
    return heat


def refcpu_draw_batched(heat, centers, radii, sizes, factor=6.0, k_scale=1.0, labels=None, zero_first=False,
                        nthreads=1):
    """The reference's float64 CPU path over a padded [B,T] layout (C port), `nthreads` host threads over samples."""
    assert heat.dtype == np.float32 and heat.flags.c_contiguous
    centers = _c(centers, np.int32)
    radii = _c(radii, np.int32)
    sizes = _c(sizes, np.int64)
    B, T = radii.shape
    if labels is not None:
        labels = _c(labels, np.int32)
        C, H, W = heat.shape[1:]
    else:
        C, (H, W) = 1, heat.shape[1:]
    lib().dho_refcpu_draw_batched(_ptr(heat), _ptr(centers), _ptr(radii), _ptr(sizes), _ptr(labels), B, T, C, H, W,
                                  float(factor), float(k_scale), int(bool(zero_first)), int(nthreads))
    return heat


# ---- independent numpy restatements (small cases only) ------------------------------------------------
def spec_draw_one_numpy(plane, x, y, r, factor, k_scale):
    """cuh:36-49,61-68 for one object, vectorised over the window, fp32 throughout except the exp (f64->f32)."""
    H, W = plane.shape
    x, y, r = int(x), int(y), int(r)
    if r < 0:
        return
    left, right = min(x, r), min(W - x, r + 1)
    top, bottom = min(y, r), min(H - y, r + 1)
    if top + bottom <= 0 or left + right <= 0:
        return
    f32 = np.float32
    sigma = f32(2 * r + 1) / f32(factor)
    var = f32(2.0) * sigma * sigma
    var_inv = f32(1.0) / var
    i = np.arange(-top, bottom, dtype=np.int64)
    j = np.arange(-left, right, dtype=np.int64)
    ii = (i * i).astype(f32)[:, None]
    jj = (j * j).astype(f32)[None, :]
    arg = (-(ii + jj)) * var_inv
    assert arg.dtype == np.float32
    g = np.exp(arg.astype(np.float64)).astype(f32) * f32(k_scale)
    win = plane[y - top : y + bottom, x - left : x + right]
    keep = win >= g
    win[...] = np.where(keep, win, g)


def draw_batched_numpy(heat, centers, radii, sizes, factor=6.0, k_scale=1.0, labels=None):
    B, T = np.asarray(radii).shape
    for s in range(B):
        n = int(min(max(int(sizes[s]), 0), T))
        for t in range(n):
            if labels is not None:
                lab = int(labels[s][t])
                if lab < 0 or lab >= heat.shape[1]:
                    continue
                plane = heat[s, lab]
            else:
                plane = heat[s]
            spec_draw_one_numpy(plane, centers[s][t][0], centers[s][t][1], radii[s][t], factor, k_scale)
    return heat


def radius_center_numpy(centers, bboxes, osf):
    centers = np.asarray(centers, np.float32)
    bboxes = np.asarray(bboxes, np.float32)
    tl = centers - bboxes[..., :2]
    br = bboxes[..., 2:] - centers
    m = np.minimum(np.minimum(tl[..., 0], tl[..., 1]), np.minimum(br[..., 0], br[..., 1]))
    r = np.ceil(m / np.float32(osf)).astype(np.int32)
    r[r < 1] = 1
    c = np.trunc(centers / np.float32(osf)).astype(np.int32)
    return c, r


def touched_mask_batched(shape, centers, radii, sizes, labels=None):
    """Union of the clipped (2r+1)^2 windows: the pixels the reference kernel touches (cuh:40-47,65-68)."""
    mask = np.zeros(shape, dtype=bool)
    H, W = shape[-2:]
    B, T = np.asarray(radii).shape
    for s in range(B):
        n = int(min(max(int(sizes[s]), 0), T))
        for t in range(n):
            x, y, r = int(centers[s][t][0]), int(centers[s][t][1]), int(radii[s][t])
            if r < 0:
                continue
            x0, x1 = max(x - r, 0), min(x + r + 1, W)
            y0, y1 = max(y - r, 0), min(y + r + 1, H)
            if x0 >= x1 or y0 >= y1:
                continue
            if labels is not None:
                lab = int(labels[s][t])
                if lab < 0 or lab >= shape[1]:
                    continue
                mask[s, lab, y0:y1, x0:x1] = True
            else:
                mask[s, y0:y1, x0:x1] = True
    return mask


# ---- digest (include/dh_b200.h: dh_checksum), numpy restatement --------------------------------------
def _mix_index(i):
    """Pixel index (uint64) -> 64-bit contribution: a 32-bit multiply/xorshift mixer (murmur3 finaliser) over the low
    word folded with the high word, widened with one more multiply. Cheap on the GPU (32-bit IMADs only)."""
    i = np.asarray(i, dtype=np.uint64)
    lo = (i & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    hi = (i >> np.uint64(32)).astype(np.uint32)
    h = lo ^ (hi * np.uint32(0x9E3779B1))
    h = h * np.uint32(0x85EBCA6B)
    h = h ^ (h >> np.uint32(13))
    h = h * np.uint32(0xC2B2AE35)
    h = h ^ (h >> np.uint32(16))
    low = h * np.uint32(0x27D4EB2F) + lo
    return (h.astype(np.uint64) << np.uint64(32)) | low.astype(np.uint64)


def digest_numpy(values, index_base=0):
    """(sum of fp32 bit patterns, sum of mix_index(index) over v>0, count of v>0), all wrapping uint64 -> int64 view."""
    v = np.ascontiguousarray(values, dtype=np.float32).reshape(-1)
    with np.errstate(over="ignore"):
        s_bits = v.view(np.uint32).astype(np.uint64).sum(dtype=np.uint64)
        idx = np.nonzero(v > 0)[0].astype(np.uint64) + np.uint64(index_base)
        s_hash = _mix_index(idx).sum(dtype=np.uint64) if idx.size else np.uint64(0)
    return np.array([s_bits, s_hash, np.uint64(idx.size)], dtype=np.uint64).view(np.int64)


# ---- N1: list -> RaggedBatch packing, numpy restatement ---------------------------------------------------
def combine_data_numpy(samples, max_sample_size=None):
    """combine_data, flattening mode (/root/reference/packages/batching_helpers/accvlab/batching_helpers/
    batched_processing_py.py:398-426): zeros [B, max n_i, *data], sample i's n_i = min(shape[0], size) rows copied to
    the front; dtype and trailing dims from the first non-empty sample; int64 sample sizes."""
    samples = [np.asarray(s) for s in samples]
    sizes = np.array([min(s.shape[0], s.size) if s.ndim else 0 for s in samples], dtype=np.int64)
    ne = [s for s in samples if s.size > 0]
    sample = ne[0] if ne else samples[0]
    T = int(sizes.max()) if max_sample_size is None else int(max_sample_size)
    out = np.zeros((len(samples), T) + tuple(sample.shape[1:]), dtype=sample.dtype)
    for i, s in enumerate(samples):
        if sizes[i] > 0:
            out[i, : sizes[i]] = s
    return out, sizes


def pack_radius_center_numpy(centers_list, bboxes_list, osf, max_sample_size=None):
    """combine_data + get_centers_and_radii (DH/tests/_test_helpers.py:20-28) on the valid slots; padded slots = 0."""
    c, sizes = combine_data_numpy([np.asarray(a, np.float32).reshape(-1, 2) for a in centers_list], max_sample_size)
    b, _ = combine_data_numpy([np.asarray(a, np.float32).reshape(-1, 4) for a in bboxes_list], max_sample_size)
    ci, ri = radius_center(c, b, osf)
    pad = np.arange(c.shape[1])[None, :] >= sizes[:, None]
    ci[pad] = 0
    ri[pad] = 0
    return ci, ri, sizes


# ---- N2: DALI draw_gaussians semantics (dh_oracle.c: dho_draw_gaussians) -----------------------------------
def draw_gaussians(heat, centers, radii, sizes, k_for_classes, radius_to_sigma_factor, slice_ids=None, active=None):
    """In place on heat f32 [B,C,H,W] (or [B,H,W] -> C=1). centers i32 [B,T,2], radii f32 [B,T], slice_ids i32 [B,T],
    active bool [B,T], sizes [B]. Restates DrawGaussians.cc:31-75,88-157 (see dh_oracle.c)."""
    assert heat.dtype == np.float32 and heat.flags.c_contiguous
    shape4 = heat.shape if heat.ndim == 4 else (heat.shape[0], 1) + heat.shape[1:]
    B, C, H, W = shape4
    centers = _c(centers, np.int32)
    radii = _c(radii, np.float32)
    T = radii.shape[1] if radii.ndim == 2 else 0
    sizes = _c(sizes, np.int64)
    k = _c(k_for_classes, np.float32).reshape(-1)
    sl = None if slice_ids is None else _c(slice_ids, np.int32)
    ac = None if active is None else _c(np.asarray(active).astype(np.uint8), np.uint8)
    lib().dho_draw_gaussians(_ptr(heat), _ptr(centers), _ptr(radii), _ptr(sl), _ptr(ac), _ptr(sizes), _ptr(k), k.size,
                             B, T, C, H, W, float(radius_to_sigma_factor))
    return heat


# ---- N2 front half: BoundingBoxToHeatmapConverter up to draw_gaussians, numpy restatement ----------------------
def boxes_to_gaussian_params_numpy(bboxes, centers, classes, valid, sizes, scale, heatmap_hw, radius_scaling_factor,
                                   min_radius, max_radius, min_fraction_area, num_classes=None, per_class_min_size=None,
                                   min_object_size=None):
    """RaggedBatch layout [B,T]. Restates, with one fp32 rounding per reference operation
    (paths under /root/reference/packages/dali_pipeline_framework/accvlab/dali_pipeline_framework):
      centre from box           operators_impl/numba_operators/numba_operators.py:811-815   (x1+x2)/2, (y1+y2)/2
      scale + clip              operators_impl/python_operator_functions/python_operator_functions.py:140-171
                                (trafo = diag(sx, sy): x' = fl(sx*x); clip to [0, W-1] / [0, H-1]; |.| sizes; area fraction)
      floor / offset            processing_steps/bounding_box_to_heatmap_converter.py:352-355
      is_active                 python_operator_functions.py:213-251, & is_valid (converter :369-371)
      radius                    numba_operators.py:865-904 (corner order resolved, min distance to the border, max(0,.),
                                float64 product with the scaling factor stored to fp32), clamp converter :376
    centers / classes / valid may be None (centre from the box / no class check / all valid).
    Returns dict(bboxes_clipped, center_int, center_offset, radii, is_active, hw_clipped, fraction); padded slots are 0."""
    f32 = np.float32
    bb = np.asarray(bboxes, f32)
    B, T = bb.shape[:2]
    sizes = np.asarray(sizes, np.int64)
    scale = np.asarray(scale, f32).reshape(B, 2)
    Hh, Wh = int(heatmap_hw[0]), int(heatmap_hw[1])
    if centers is None:
        cen = np.stack([(bb[..., 0] + bb[..., 2]) / f32(2.0), (bb[..., 1] + bb[..., 3]) / f32(2.0)], -1).astype(f32)
    else:
        cen = np.asarray(centers, f32)
    sx, sy = scale[:, None, 0], scale[:, None, 1]
    xs = [(sx * bb[..., 0]).astype(f32), (sx * bb[..., 2]).astype(f32)]
    ys = [(sy * bb[..., 1]).astype(f32), (sy * bb[..., 3]).astype(f32)]
    clipx = lambda v: np.minimum(np.maximum(v, f32(0)), f32(Wh - 1)).astype(f32)
    clipy = lambda v: np.minimum(np.maximum(v, f32(0)), f32(Hh - 1)).astype(f32)
    x1, x2, y1, y2 = clipx(xs[0]), clipx(xs[1]), clipy(ys[0]), clipy(ys[1])
    h_c, w_c = np.abs(y2 - y1).astype(f32), np.abs(x2 - x1).astype(f32)
    h_o, w_o = np.abs(ys[1] - ys[0]).astype(f32), np.abs(xs[1] - xs[0]).astype(f32)
    with np.errstate(invalid="ignore", divide="ignore"):
        frac = ((h_c * w_c).astype(f32) / (h_o * w_o).astype(f32)).astype(f32)
    cx, cy = clipx((sx * cen[..., 0]).astype(f32)), clipy((sy * cen[..., 1]).astype(f32))
    cix, ciy = np.floor(cx).astype(np.int32), np.floor(cy).astype(np.int32)
    offx, offy = (cx - cix.astype(f32)).astype(f32), (cy - ciy.astype(f32)).astype(f32)
    # is_active
    act = np.ones((B, T), bool)
    if classes is not None:
        cls = np.asarray(classes, np.int64)
        act &= cls < int(num_classes)
        if per_class_min_size is not None and len(per_class_min_size):
            tab = np.asarray(per_class_min_size, np.float64)
            cc = np.where(cls < int(num_classes), cls, 0)
            cc = np.clip(cc, 0, len(tab) - 1)  # negative ids index from the end in the reference; not exercised
            act &= (h_c.astype(np.float64) >= tab[cc, 0]) & (w_c.astype(np.float64) >= tab[cc, 1])
    elif min_object_size is not None and len(min_object_size):
        act &= (h_c.astype(np.float64) >= float(min_object_size[0])) & (w_c.astype(np.float64) >= float(min_object_size[1]))
    act &= frac.astype(np.float64) >= float(min_fraction_area)
    if valid is not None:
        act &= np.asarray(valid, bool)
    # radius on the CLIPPED box around the CLIPPED centre
    l, r = np.minimum(x1, x2), np.maximum(x1, x2)
    t, b = np.minimum(y1, y2), np.maximum(y1, y2)
    dist = np.minimum(np.minimum((cx - l).astype(f32), (cy - t).astype(f32)), np.minimum((r - cx).astype(f32), (b - cy).astype(f32)))
    dist = np.maximum(dist.astype(np.float64), 0.0)
    rad = (dist * float(radius_scaling_factor)).astype(f32)
    rad = np.minimum(np.maximum(f32(min_radius), rad), f32(max_radius)).astype(f32)
    pad = np.arange(T)[None, :] >= sizes[:, None]
    res = dict(bboxes_clipped=np.stack([x1, y1, x2, y2], -1).astype(f32), center_int=np.stack([cix, ciy], -1).astype(np.int32),
               center_offset=np.stack([offx, offy], -1).astype(f32), radii=rad, is_active=act,
               hw_clipped=np.stack([h_c, w_c], -1).astype(f32), fraction=frac)
    for k, v in res.items():
        v[pad] = 0
    return res


# ---- N4: CenterNet / CenterPoint gaussian_radius — PARITY UNPINNED (no implementation in /root/reference) -------
def gaussian_radius_numpy(det_hw, min_overlap=0.5, variant="cornernet", min_radius=0):
    """Published algorithm, float64 like upstream (Python floats). variant "cornernet": CornerNet `gaussian_radius`
    (kept verbatim by CenterNet src/lib/utils/image.py and CenterPoint det3d/core/utils/center_utils.py):
    r = (b + sqrt(b^2 - 4ac)) / 2; variant "mmdet": mmdet/models/utils/gaussian_target.py: (b -/+ sqrt(.)) / (2a).
    Returns (int radii = max(min_radius, int(r)), float r). Neither package is installed here and the reference has no
    such function, so this restatement is checked against hand-computed values only."""
    d = np.asarray(det_hw, np.float32).astype(np.float64).reshape(-1, 2)
    h, w = d[:, 0], d[:, 1]
    mo = float(min_overlap)
    with np.errstate(invalid="ignore"):
        if variant == "cornernet":
            c1 = w * h * (1 - mo) / (1 + mo)
            r1 = ((h + w) + np.sqrt((h + w) * (h + w) - 4 * c1)) / 2
            b2 = 2 * (h + w)
            c2 = (1 - mo) * (w * h)
            r2 = (b2 + np.sqrt(b2 * b2 - 16 * c2)) / 2
            a3, b3, c3 = 4 * mo, -2 * mo * (h + w), (mo - 1) * (w * h)
            r3 = (b3 + np.sqrt(b3 * b3 - 4 * a3 * c3)) / 2
        else:
            c1 = w * h * (1 - mo) / (1 + mo)
            r1 = ((h + w) - np.sqrt((h + w) * (h + w) - 4 * 1.0 * c1)) / (2 * 1.0)
            b2 = 2 * (h + w)
            c2 = (1 - mo) * (w * h)
            r2 = (b2 - np.sqrt(b2 * b2 - 4 * 4.0 * c2)) / (2 * 4.0)
            a3, b3, c3 = 4 * mo, -2 * mo * (h + w), (mo - 1) * (w * h)
            r3 = (b3 + np.sqrt(b3 * b3 - 4 * a3 * c3)) / (2 * a3)
        r = np.minimum(np.minimum(r1, r2), r3)
    ri = np.where(np.isnan(r), -2**31, np.clip(np.trunc(np.nan_to_num(r)), -2**31, 2**31 - 1)).astype(np.int64)
    return np.maximum(ri, int(min_radius)).astype(np.int32), r.astype(np.float32)


# ---- N3: Gaussian focal loss on the oracle's target — PARITY UNPINNED (formula only in the reference) -----------
def gaussian_focal_loss_numpy(pred, target, alpha=2.0, gamma=4.0, eps=1e-12):
    """/root/reference/packages/draw_heatmap/docs/intro.rst:7-26 (== mmdet gaussian_focal_loss, reduction 'sum'),
    evaluated in float64 on fp32 inputs: returns (loss sum, number of target == 1 pixels, d loss / d pred as float64)."""
    p = np.asarray(pred, np.float32).astype(np.float64)
    t = np.asarray(target, np.float32).astype(np.float64)
    pos = t == 1.0
    with np.errstate(divide="ignore", invalid="ignore"):
        lp, ln = np.log(p + eps), np.log(1 - p + eps)
        w = (1 - t) ** gamma
        loss = np.where(pos, -lp * (1 - p) ** alpha, -ln * p ** alpha * w)
        grad = np.where(pos, alpha * (1 - p) ** (alpha - 1) * lp - (1 - p) ** alpha / (p + eps),
                        w * (p ** alpha / (1 - p + eps) - alpha * p ** (alpha - 1) * ln))
    return float(loss.sum()), int(pos.sum()), grad
