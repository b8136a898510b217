"""Generate tests/golden/ref_converter.npz by RUNNING THE REFERENCE'S OWN code for the front half of its
BoundingBoxToHeatmapConverter (the step before draw_gaussians), in the build container.

    python tests/golden/make_golden_converter.py     # needs /root/reference; not needed on the GPU box

Functions executed from where they lie under /root/reference/packages/dali_pipeline_framework/accvlab/dali_pipeline_framework:
  operators_impl/python_operator_functions/python_operator_functions.py:103-171  apply_clipping_and_get_with_clipping_info
  operators_impl/python_operator_functions/python_operator_functions.py:175-253  get_is_active
  operators_impl/numba_operators/numba_operators.py:811-815   get_center  (inner run_fn of get_center_from_bboxes)
  operators_impl/numba_operators/numba_operators.py:865-904   get_radii   (inner run_fn of get_radii_from_bboxes)
The two inner functions are what DALI's numba_function operator compiles; they are captured through a stub of
`nvidia.dali.plugin.numba.fn.experimental.numba_function` and compiled with numba.njit here (numba typing matters: the
radius is a float64 product stored to float32). `nvidia.dali` and `cupy` are not installed: empty stand-in modules
satisfy the imports (cupy.get_array_module -> numpy). The glue between these functions follows
processing_steps/bounding_box_to_heatmap_converter.py:321-399 (floor -> int32 centre, offset, min/max radius clamp,
`& is_valid`), evaluated with numpy float32 operations as DALI's arithmetic operators do.
"""
from __future__ import annotations

import enum
import importlib.util
import os
import sys
import types

import numba
import numpy as np

REF = "/root/reference/packages/dali_pipeline_framework/accvlab/dali_pipeline_framework/operators_impl"
OUT = os.path.dirname(os.path.abspath(__file__))


def install_stubs():
    def mod(name):
        m = types.ModuleType(name)
        sys.modules[name] = m
        return m

    nv, dali = mod("nvidia"), mod("nvidia.dali")
    nv.dali = dali
    t = mod("nvidia.dali.types")
    t.DALIDataType = enum.Enum("DALIDataType", "FLOAT INT32 INT64 BOOL UINT8 FLOAT64 INT16 UINT16 UINT32 UINT64 INT8 FLOAT16")
    dali.types = t
    dali.fn = mod("nvidia.dali.fn")
    pl = mod("nvidia.dali.pipeline")
    pl.do_not_convert = lambda f: f
    pl.DataNode = object
    dali.pipeline = pl
    mod("nvidia.dali.plugin")
    mod("nvidia.dali.plugin.numba")
    mod("nvidia.dali.plugin.numba.fn")
    fnex = mod("nvidia.dali.plugin.numba.fn.experimental")
    fnex.numba_function = lambda *ins, run_fn=None, **kw: run_fn  # hand the operator body back to the caller
    cupy = mod("cupy")
    cupy.get_array_module = lambda *a: np


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    install_stubs()
    pof = load(os.path.join(REF, "python_operator_functions", "python_operator_functions.py"), "ref_pof")
    nop = load(os.path.join(REF, "numba_operators", "numba_operators.py"), "ref_nop")
    get_center = numba.njit(nop.get_center_from_bboxes("bboxes"))
    rng = np.random.default_rng(20263)
    out = {}

    def case(name, n_list, image_hw, heatmap_hw, sf, min_r, max_r, min_frac, use_classes, per_class, min_size, num_classes,
             own_centers, with_valid):
        get_radii = numba.njit(nop.get_radii_from_bboxes("bboxes", centers="centers", scaling_factor=sf))
        B, T = len(n_list), max(max(n_list), 1)
        Himg, Wimg = image_hw
        scale = np.array([heatmap_hw[1] / Wimg, heatmap_hw[0] / Himg], np.float32)
        trafo = np.array([[scale[0], 0, 0], [0, scale[1], 0]], np.float32)  # fn.transforms.scale: 2 x 3, float32
        arr = dict(bboxes=np.zeros((B, T, 4), np.float32), centers=np.zeros((B, T, 2), np.float32),
                   classes=np.zeros((B, T), np.int32), valid=np.ones((B, T), bool),
                   bboxes_clipped=np.zeros((B, T, 4), np.float32), center_int=np.zeros((B, T, 2), np.int32),
                   center_offset=np.zeros((B, T, 2), np.float32), radii=np.zeros((B, T), np.float32),
                   is_active=np.zeros((B, T), bool), hw_clipped=np.zeros((B, T, 2), np.float32),
                   fraction=np.zeros((B, T), np.float32))
        for s, n in enumerate(n_list):
            if n == 0:
                continue
            c = np.stack([rng.uniform(-0.2, 1.2, n) * Wimg, rng.uniform(-0.2, 1.2, n) * Himg], 1)
            d = rng.uniform(0, 0.25, (n, 4)) * min(Himg, Wimg)
            d[rng.random(n) < 0.1] = 0.0                                   # degenerate (zero-area) boxes
            bb = np.concatenate([c - d[:, :2], c + d[:, 2:]], 1).astype(np.float32)
            swap = rng.random(n) < 0.3                                      # corners in any order / on either diagonal
            bb[swap] = bb[swap][:, [2, 1, 0, 3]]
            swap = rng.random(n) < 0.3
            bb[swap] = bb[swap][:, [0, 3, 2, 1]]
            cls = rng.integers(0, num_classes + 2, n).astype(np.int32)      # some classes are not considered
            valid = rng.random(n) < 0.9
            if own_centers:
                cen = (c + rng.normal(0, 0.05, (n, 2)) * min(Himg, Wimg)).astype(np.float32)
            else:
                cen = np.zeros((n, 2), np.float32)
                get_center(cen, bb)
            rc, cc, hw, fr = pof.apply_clipping_and_get_with_clipping_info(bb, cen, trafo, heatmap_hw)
            ci = np.floor(cc).astype(np.int32)
            off = (cc - ci.astype(np.float32)).astype(np.float32)
            if use_classes:
                act = pof.get_is_active(hw, cls, fr, None, None if per_class is None else np.array(per_class), num_classes, min_frac)
            else:
                act = pof.get_is_active(hw, None, fr, None if min_size is None else np.array(min_size), None, None, min_frac)
            if with_valid:
                act = act & valid
            rad = np.zeros((n,), np.float32)
            get_radii(rad, rc, cc)
            rad = np.minimum(np.maximum(np.float32(min_r), rad), np.float32(max_r)).astype(np.float32)
            for k, v in dict(bboxes=bb, centers=cen, classes=cls, valid=valid, bboxes_clipped=rc, center_int=ci,
                             center_offset=off, radii=rad, is_active=act, hw_clipped=hw, fraction=fr).items():
                arr[k][s, :n] = v
        arr.update(sizes=np.array(n_list, np.int64), scale=np.tile(scale, (B, 1)),
                   params=np.array([heatmap_hw[0], heatmap_hw[1], sf, min_r, max_r, min_frac, num_classes, int(use_classes),
                                    int(own_centers), int(with_valid)], np.float64),
                   per_class=np.array(per_class if per_class is not None else np.zeros((0, 2)), np.float64).reshape(-1, 2),
                   min_size=np.array(min_size if min_size is not None else [], np.float64))
        out.update({f"{name}_{k}": v for k, v in arr.items()})

    case("perclass", [40, 0, 17, 64], (900, 1600), (112, 200), 0.8, 2.0, 20.0, 0.3, True, [[2, 2], [1, 3], [4, 1], [0, 0]], None, 4,
         False, True)
    case("classes_nothresh", [33, 5], (512, 512), (128, 128), 0.5, 1.0, 12.5, 0.0, True, None, None, 3, True, False)
    case("noclass_minsize", [50, 21, 1], (928, 1600), (232, 400), 0.8, 2.0, 1e9, 0.25, False, None, [3, 2], 1, False, True)
    case("noclass_plain", [25], (300, 500), (75, 125), 1.0 / 3.0, 0.0, 7.0, 0.5, False, None, None, 1, True, True)
    np.savez_compressed(os.path.join(OUT, "ref_converter.npz"), **out)
    print("wrote ref_converter.npz:", sorted({k.split('_bboxes')[0] for k in out if k.endswith('_bboxes')}),
          {k: int(v.sum()) for k, v in out.items() if k.endswith("is_active")})


if __name__ == "__main__":
    main()
