"""Generate tests/golden/ref_dali.npz by RUNNING THE REFERENCE'S OWN draw_gaussians operator source
(/root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc, compiled unmodified from where it lies
against oracle/dali_stub into oracle/_ref/libdg_ref.so by oracle/build_oracle.py:build_reference_dali).

    python tests/golden/make_golden_dali.py     # needs /root/reference; not needed on the GPU box

Cases (seeded): typical converter output (integer-ish and fractional radii, per-class k, inactive objects, non-zero
input heatmap), windows clipped at / wholly outside every border, tiny / zero / negative radii, radii a hair above and
below integers at large coordinates (fp32 rounding of centre -/+ radius makes the window asymmetric), 2-D heatmaps.
Stored per case: inputs in the RaggedBatch layout [B,T] + sizes, the input heatmap, the operator's output."""
from __future__ import annotations

import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import build_oracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def run_reference(lib, heat_in, centers, radii, slice_ids, active, sizes, k, factor):
    """RaggedBatch layout -> the operator's per-sample lists (valid slots only) -> output heatmap."""
    B = heat_in.shape[0]
    C = heat_in.shape[1] if heat_in.ndim == 4 else 0
    H, W = heat_in.shape[-2:]
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    cat = lambda a: np.ascontiguousarray(np.concatenate([a[s, : sizes[s]] for s in range(B)], axis=0))
    c, r, sl, ac = cat(centers).astype(np.int32), cat(radii).astype(np.float32), cat(slice_ids).astype(np.int32), \
        cat(active).astype(np.uint8)
    out = np.empty_like(heat_in)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    k = np.ascontiguousarray(k, np.float32)
    rc = lib.dgref_draw_gaussians(p(heat_in), p(out), B, C, H, W, p(ac), p(sl), p(c), p(r), p(offs), p(k), k.size,
                                  ctypes.c_float(factor))
    assert rc == 0, rc
    return out


def main():
    path = build_oracle.build_reference_dali()
    lib = ctypes.CDLL(path)
    rng = np.random.default_rng(20262)
    out = {}

    def case(name, B, C, H, W, T, rad_fn, factor, k, p_active=0.85, bg=True, two_d=False, center_fn=None):
        sizes = rng.integers(0, T + 1, B).astype(np.int64)
        sizes[0] = T
        centers = np.stack([rng.integers(-12, W + 12, (B, T)), rng.integers(-12, H + 12, (B, T))], axis=2).astype(np.int32)
        if center_fn is not None:
            centers = center_fn(centers)
        radii = rad_fn((B, T)).astype(np.float32)
        slice_ids = rng.integers(0, max(C, 1), (B, T)).astype(np.int32)
        active = rng.random((B, T)) < p_active
        shape = (B, H, W) if two_d else (B, C, H, W)
        heat_in = (rng.random(shape, dtype=np.float32) * 0.3 * (rng.random(shape) < 0.2)).astype(np.float32) if bg \
            else np.zeros(shape, np.float32)
        k = np.asarray(k, np.float32)
        got = run_reference(lib, heat_in, centers, radii, slice_ids, active, sizes, k, factor)
        out.update({f"{name}_{n}": v for n, v in dict(heat_in=heat_in, centers=centers, radii=radii, slice_ids=slice_ids,
                                                      active=active, sizes=sizes, k=k, factor=np.float32(factor),
                                                      heat_out=got).items()})

    case("typical", 6, 4, 48, 80, 24, lambda s: 1.0 + rng.random(s) * 9.0, 1.0 / 3.0, [1.0, 0.8, 1.0, 0.5])
    case("integer_r", 4, 3, 40, 56, 16, lambda s: rng.integers(1, 7, s).astype(np.float32), 0.35, [1.0, 1.0, 0.9], bg=False)
    case("tiny_r", 3, 2, 24, 24, 20, lambda s: rng.choice([0.0, 1e-30, 1e-3, 0.25, 0.5, 0.999, 1.0, -0.5, -0.25, -3.0], s), 0.5,
         [1.0, 0.7])
    case("two_d", 3, 0, 33, 47, 12, lambda s: 0.5 + rng.random(s) * 6.0, 0.4, [0.9], two_d=True)
    case("neg_k", 2, 2, 32, 32, 10, lambda s: 1.0 + rng.random(s) * 5.0, 0.3, [-0.5, 0.0])
    eps = np.float32(2.0 ** -20)
    case("near_int", 2, 1, 16, 3000, 64,
         lambda s: (rng.integers(1, 5, s) + rng.choice([-3, -2, -1, 0, 1, 2, 3], s) * eps).astype(np.float32), 0.4, [1.0],
         center_fn=lambda c: np.stack([rng.integers(2040, 2990, c.shape[:2]), c[..., 1]], axis=2).astype(np.int32))
    np.savez_compressed(os.path.join(OUT, "ref_dali.npz"), **out)
    print("wrote ref_dali.npz", sorted({k.split('_heat')[0] for k in out if k.endswith('heat_out')}),
          {k: float(v.max()) for k, v in out.items() if k.endswith("heat_out")})


if __name__ == "__main__":
    main()
