"""ctypes binding of the C ABI (include/dh_b200.h) used by the parity tests, bench.py and smoke().

This is the binding a non-torch caller would write (INTEGRATION.md shows the same stub); the tests call the
product THROUGH it so that parity is established at the drop-in boundary, not at a Python convenience layer.
"""
from __future__ import annotations

import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "accv-lab_b200", "accvlab", "draw_heatmap", "libdh_b200.so")
HEADER = os.path.join(ROOT, "include", "dh_b200.h")

DH_OK = 0
DH_ERR_INVALID_ARGUMENT = -1
DH_ERR_WORKSPACE_TOO_SMALL = -2
DH_ERR_UNSUPPORTED_SHAPE = -3
DH_MODE_ACCUMULATE = 0
DH_MODE_FRESH = 1
DH_SIZES_I32 = 0
DH_SIZES_I64 = 1

_p, _i, _i64, _f, _sz, _u64 = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t,
                               ctypes.c_uint64)

SIGNATURES = {
    "dh_version": (_i, []),
    "dh_build_info": (ctypes.c_char_p, []),
    "dh_status_string": (ctypes.c_char_p, [_i]),
    "dh_launch_count": (_u64, []),
    "dh_radius_center": (_i, [_p, _p, _p, _f, _i64, _p, _p]),
    "dh_draw_heatmap_batched": (_i, [_p, _p, _p, _p, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f, _i, _p]),
    "dh_draw_heatmap_batched_from_boxes": (_i, [_p, _p, _p, _p, _f, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f,
                                                _i, _p, _p, _p]),
    "dh_draw_heatmap_batched_from_boxes_digest": (_i, [_p, _p, _p, _p, _f, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f,
                                                       _f, _i64, _p, _p]),
    "dh_render_heatmap_batched_16": (_i, [_p, _p, _i, _p, _p, _p, _p, _f, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f,
                                          _p]),
    "dh_batched_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "dh_draw_heatmap_batched_ws": (_i, [_p, _p, _p, _p, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f, _i, _p, _p, _sz]),
    "dh_draw_heatmap_batched_from_boxes_ws": (_i, [_p, _p, _p, _p, _f, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f,
                                                   _i, _p, _p, _p, _p, _sz]),
    "dh_draw_heatmap_batched_from_boxes_digest_ws": (_i, [_p, _p, _p, _p, _f, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f,
                                                          _f, _i64, _p, _p, _p, _sz]),
    "dh_render_heatmap_batched_16_ws": (_i, [_p, _p, _i, _p, _p, _p, _p, _f, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f,
                                             _f, _p, _p, _sz]),
    "dh_noalloc_max_scratch_bytes": (_sz, []),
    "dh_draw_heatmap_batched_noalloc": (_i, [_p, _p, _p, _p, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f, _i, _p]),
    "dh_draw_heatmap_workspace_bytes": (_sz, [_i64, _i64]),
    "dh_draw_heatmap": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _f, _f, _i, _p, _sz, _p]),
    "dh_checksum": (_i, [_p, _p, _i64, _i64, _p]),
    "dh_host_scratch_bytes": (_sz, [_i64, _i64, _i64, _i64, _i64, _i, _i64, _i]),
    "dh_draw_heatmap_batched_from_boxes_host": (_i, [_p, _p, _p, _p, _p, _p, _f, _p, _p, _i64, _i64, _i64, _i64, _i64,
                                                     _f, _f, _p, _sz, _i64]),
    "dh_pack_ragged": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i]),
    "dh_pack_radius_center": (_i, [_p, _p, _p, _p, _f, _i64, _i64, _p, _p]),
    "dh_draw_gaussians": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _i64, _f, _p, _p,
                               _sz]),
    "dh_draw_gaussians_workspace_bytes": (_sz, [_i64, _i64]),
    "dh_boxes_to_gaussian_params": (_i, [_p, _p, _p, _p, _p, _p, _i, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64,
                                         ctypes.c_double, _f, _f, ctypes.c_double, _p, _p, _p, _p, _p, _p, _p]),
    "dh_gaussian_radius": (_i, [_p, _p, ctypes.c_double, _i, ctypes.c_int32, _i64, _p, _p]),
    "dh_focal_loss_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64]),
    "dh_gaussian_focal_loss": (_i, [_p, _p, _p, _p, _p, _i, _p, _i64, _i64, _i64, _i64, _i64, _f, _f, _f, _f, _f, _i, _p,
                                    _p, _p, _sz]),
    "dh_scale_by_device_scalar": (_i, [_p, _p, _p, _p, _i64]),
    "dh_fill_probe": (_i, [_p, _p, _sz]),
    "dh_expf_probe": (_i, [_p, ctypes.c_uint32, _i64, _p]),
}


def declared_functions():
    """Names of every function declared in include/dh_b200.h (prototype = identifier followed by '(' at top level)."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", "", src, flags=re.M)
    names = re.findall(r"\b(dh_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def load() -> ctypes.CDLL:
    path = os.environ.get("DH_B200_LIB") or LIB_PATH  # override = A/B builds of the same ABI (tuning runs only)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run `python accv-lab_b200/build.py` (there is no CPU fallback)")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def ptr(t):
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return ctypes.c_void_p(t.data_ptr())
    return ctypes.c_void_p(t.ctypes.data)


def check(lib, st, what=""):
    if st != DH_OK:
        raise RuntimeError(f"{what} failed: {lib.dh_status_string(st).decode()} (status {st})")


def stream_ptr():
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---- convenience wrappers over torch CUDA tensors (still only raw pointers cross the ABI) -------------
# `ws`: False = the workspace-less entry (single-launch renderer: the product's main path — what the torch binding, bench.py
#               and the host entry call);
#       True  = the *_ws entry with a bin-list workspace (K2 pre-pass + groups2 renderer, two launches);
#       None  = USE_WS (env DH_TEST_WS, default 0). tests/test_two_render_paths.py covers both explicitly.
USE_WS = os.environ.get("DH_TEST_WS", "0") != "0"


def _workspace(lib, B, T, C, device, ws):
    """(workspace tensor or None, bytes) for the batched entries."""
    import torch

    if ws is None:
        ws = USE_WS
    if not ws:
        return None, 0
    n = int(lib.dh_batched_workspace_bytes(B, T, C))
    return torch.empty(max(n, 16), dtype=torch.uint8, device=device), n


def draw_batched(lib, heat, centers, radii, sizes, labels=None, factor=6.0, k=1.0, mode=DH_MODE_ACCUMULATE, status=None,
                 ws=None):
    import torch

    B, T = radii.shape
    if labels is not None:
        C, H, W = heat.shape[1:]
    else:
        C, (H, W) = 1, heat.shape[1:]
    tag = DH_SIZES_I64 if sizes.dtype == torch.int64 else DH_SIZES_I32
    wt, wn = _workspace(lib, B, T, C, heat.device, ws)
    if wt is None:
        st = lib.dh_draw_heatmap_batched(stream_ptr(), ptr(heat), ptr(centers), ptr(radii), ptr(sizes), tag, ptr(labels), B,
                                         T, C, H, W, factor, k, mode, ptr(status))
    else:
        st = lib.dh_draw_heatmap_batched_ws(stream_ptr(), ptr(heat), ptr(centers), ptr(radii), ptr(sizes), tag, ptr(labels),
                                            B, T, C, H, W, factor, k, mode, ptr(status), ptr(wt), wn)
    check(lib, st, "dh_draw_heatmap_batched")
    return heat


def draw_from_boxes(lib, heat, centers_f, bboxes_f, osf, sizes, labels=None, factor=6.0, k=1.0, mode=DH_MODE_FRESH,
                    centers_out=None, radii_out=None, ws=None, status=None):
    import torch

    B, T = centers_f.shape[:2]
    if labels is not None:
        C, H, W = heat.shape[1:]
    else:
        C, (H, W) = 1, heat.shape[1:]
    tag = DH_SIZES_I64 if sizes.dtype == torch.int64 else DH_SIZES_I32
    wt, wn = _workspace(lib, B, T, C, heat.device, ws)
    if wt is None:
        st = lib.dh_draw_heatmap_batched_from_boxes(stream_ptr(), ptr(heat), ptr(centers_f), ptr(bboxes_f), osf, ptr(sizes),
                                                    tag, ptr(labels), B, T, C, H, W, factor, k, mode, ptr(centers_out),
                                                    ptr(radii_out), ptr(status))
    else:
        st = lib.dh_draw_heatmap_batched_from_boxes_ws(stream_ptr(), ptr(heat), ptr(centers_f), ptr(bboxes_f), osf,
                                                       ptr(sizes), tag, ptr(labels), B, T, C, H, W, factor, k, mode,
                                                       ptr(centers_out), ptr(radii_out), ptr(status), ptr(wt), wn)
    check(lib, st, "dh_draw_heatmap_batched_from_boxes")
    return heat


DH_OUT_F16, DH_OUT_BF16 = 1, 2


def render16(lib, heat, sizes, labels=None, centers=None, radii=None, centers_f=None, bboxes_f=None, osf=1.0, factor=6.0,
             k=1.0, status=None, check_status=True, ws=None):
    """dh_render_heatmap_batched_16 into a float16 / bfloat16 CUDA tensor (FRESH)."""
    import torch

    src = radii if radii is not None else centers_f
    B, T = src.shape[:2]
    if labels is not None:
        C, H, W = heat.shape[1:]
    else:
        C, (H, W) = 1, heat.shape[1:]
    tag = DH_SIZES_I64 if sizes.dtype == torch.int64 else DH_SIZES_I32
    out = DH_OUT_F16 if heat.dtype == torch.float16 else DH_OUT_BF16
    wt, wn = _workspace(lib, B, T, C, heat.device, ws)
    if wt is None:
        st = lib.dh_render_heatmap_batched_16(stream_ptr(), ptr(heat), out, ptr(centers), ptr(radii), ptr(centers_f),
                                              ptr(bboxes_f), osf, ptr(sizes), tag, ptr(labels), B, T, C, H, W, factor, k,
                                              ptr(status))
    else:
        st = lib.dh_render_heatmap_batched_16_ws(stream_ptr(), ptr(heat), out, ptr(centers), ptr(radii), ptr(centers_f),
                                                 ptr(bboxes_f), osf, ptr(sizes), tag, ptr(labels), B, T, C, H, W, factor, k,
                                                 ptr(status), ptr(wt), wn)
    if check_status:
        check(lib, st, "dh_render_heatmap_batched_16")
    return st


def draw_from_boxes_digest(lib, heat, centers_f, bboxes_f, osf, sizes, labels=None, factor=6.0, k=1.0, index_base=0,
                           digest=None, ws=None):
    """FRESH render with the digest fused into the launch; returns the int64[3] device digest (wrapping uint64)."""
    import torch

    B, T = centers_f.shape[:2]
    if labels is not None:
        C, H, W = heat.shape[1:]
    else:
        C, (H, W) = 1, heat.shape[1:]
    if digest is None:
        digest = torch.zeros(3, dtype=torch.int64, device=heat.device)
    tag = DH_SIZES_I64 if sizes.dtype == torch.int64 else DH_SIZES_I32
    wt, wn = _workspace(lib, B, T, C, heat.device, ws)
    if wt is None:
        st = lib.dh_draw_heatmap_batched_from_boxes_digest(stream_ptr(), ptr(heat), ptr(centers_f), ptr(bboxes_f), osf,
                                                           ptr(sizes), tag, ptr(labels), B, T, C, H, W, factor, k,
                                                           index_base, ptr(digest), None)
    else:
        st = lib.dh_draw_heatmap_batched_from_boxes_digest_ws(stream_ptr(), ptr(heat), ptr(centers_f), ptr(bboxes_f), osf,
                                                              ptr(sizes), tag, ptr(labels), B, T, C, H, W, factor, k,
                                                              index_base, ptr(digest), None, ptr(wt), wn)
    check(lib, st, "dh_draw_heatmap_batched_from_boxes_digest")
    return digest


def draw_flat(lib, heat, centers, radii, idx, factor=6.0, k=1.0, mode=DH_MODE_ACCUMULATE, status=None):
    import torch

    N = radii.shape[0]
    M, H, W = heat.shape
    nbytes = lib.dh_draw_heatmap_workspace_bytes(N, M)
    ws = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=heat.device)
    st = lib.dh_draw_heatmap(stream_ptr(), ptr(heat), ptr(centers), ptr(radii), ptr(idx), N, M, H, W, factor, k, mode,
                             ptr(ws), nbytes, ptr(status))
    check(lib, st, "dh_draw_heatmap")
    return heat


def radius_center(lib, centers_f, bboxes_f, osf):
    import torch

    n = centers_f.numel() // 2
    co = torch.empty(centers_f.shape, dtype=torch.int32, device=centers_f.device)
    ro = torch.empty(centers_f.shape[:-1], dtype=torch.int32, device=centers_f.device)
    check(lib, lib.dh_radius_center(stream_ptr(), ptr(centers_f), ptr(bboxes_f), osf, n, ptr(co), ptr(ro)),
          "dh_radius_center")
    return co, ro


def checksum(lib, values, index_base=0, digest=None):
    import torch

    if digest is None:
        digest = torch.zeros(3, dtype=torch.int64, device=values.device)
    check(lib, lib.dh_checksum(stream_ptr(), ptr(values), values.numel(), index_base, ptr(digest)), "dh_checksum")
    return digest
