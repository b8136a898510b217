"""Synthetic workloads of BASELINE.json `configs` (recipes: SURVEY.md 8(d)).  Measurement/test infrastructure.

The generator is modelled on the reference's test-data generator
(/root/reference/packages/draw_heatmap/tests/_test_helpers.py:77-108: per sample N boxes, centres uniform over the
image, four centre-to-border distances uniform in [0, max_bbox/2)), but vectorised and driven by a seeded
numpy PCG64 stream so every machine and every rank count sees identical data. Padded slots carry the drawable
decoy box of the reference's padding-robustness fixture (tests/test_draw_heatmap_batched.py:52-58): a renderer
that ignores `sample_sizes` fails parity.

Everything here is numpy on the host; nothing imports the oracle or the CUDA extension.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np


@dataclass(frozen=True)
class Config:
    name: str
    B: int           # samples (batch x cameras, flattened)
    C: int           # classes (planes per sample)
    H: int
    W: int
    osf: float       # out_size_factor (image px per heatmap px)
    n_min: int
    n_max: int
    T: Optional[int]  # padded targets per sample (None: max N_i)
    max_bbox: float  # distances ~ U[0, max_bbox/2)   (ignored if sym_lo/hi set)
    seed: int
    sym_lo: float = 0.0  # cfg 4: one symmetric distance ~ U[sym_lo, sym_hi] on all four sides
    sym_hi: float = 0.0
    describe: str = ""


CONFIGS = {
    "cfg1": Config("cfg1", 4, 10, 128, 128, 8, 40, 60, None, 96, 20261,
                   describe="CenterPoint nuScenes BEV: batch 4, 10 classes, 128x128, ~50 boxes/sample"),
    "cfg2": Config("cfg2", 16, 10, 180, 180, 8, 0, 200, 200, 128, 20262,
                   describe="CenterPoint/BEVFusion BEV: batch 16, 10 classes, 180x180, ragged 0-200 boxes/sample"),
    "cfg3": Config("cfg3", 48, 10, 232, 400, 4, 0, 100, 100, 480, 20263,
                   describe="multi-camera CenterNet: batch 8 x 6 cams, 10 classes, 232x400, ragged 0-100 boxes"),
    "cfg4": Config("cfg4", 32, 1, 512, 512, 1, 1024, 1024, 1024, 0, 20264, sym_lo=16.0, sym_hi=64.0,
                   describe="dense-overlap stress: batch 32, 1 class, 512x512, 1024 boxes/sample, r in 16..64"),
    "cfg5": Config("cfg5", 3072, 10, 232, 400, 4, 0, 100, 100, 480, 20265,
                   describe="scaling sweep: batch 512 x 6 cams, 10 classes, 232x400, sharded over the GPUs"),
}


@dataclass
class Workload:
    cfg: Config
    centers_f: np.ndarray  # f32 [B,T,2] (x,y) image coordinates
    bboxes_f: np.ndarray   # f32 [B,T,4] (x1,y1,x2,y2)
    labels: np.ndarray     # i32 [B,T]
    sizes: np.ndarray      # i64 [B]

    @property
    def B(self):
        return self.centers_f.shape[0]

    @property
    def T(self):
        return self.centers_f.shape[1]

    @property
    def planes(self):
        return self.B * self.cfg.C

    @property
    def out_bytes(self):
        return self.planes * self.cfg.H * self.cfg.W * 4

    def slice(self, lo: int, hi: int) -> "Workload":
        return Workload(self.cfg, np.ascontiguousarray(self.centers_f[lo:hi]), np.ascontiguousarray(self.bboxes_f[lo:hi]),
                        np.ascontiguousarray(self.labels[lo:hi]), np.ascontiguousarray(self.sizes[lo:hi]))


def make(cfg: Config | str, B: Optional[int] = None) -> Workload:
    """Generate the (whole, unsharded) workload of a config; `B` overrides the sample count (same stream prefix)."""
    if isinstance(cfg, str):
        cfg = CONFIGS[cfg]
    B = cfg.B if B is None else B
    rng = np.random.Generator(np.random.PCG64(cfg.seed))
    sizes = rng.integers(cfg.n_min, cfg.n_max + 1, size=B).astype(np.int64)
    T = int(cfg.T if cfg.T is not None else max(int(sizes.max()) if B else 0, 1))
    img_w, img_h = cfg.W * cfg.osf, cfg.H * cfg.osf
    cx = (rng.random((B, T), dtype=np.float32) * np.float32(img_w)).astype(np.float32)
    cy = (rng.random((B, T), dtype=np.float32) * np.float32(img_h)).astype(np.float32)
    if cfg.sym_hi > 0:
        d1 = (np.float32(cfg.sym_lo) + rng.random((B, T), dtype=np.float32) * np.float32(cfg.sym_hi - cfg.sym_lo))
        d = np.repeat(d1[:, :, None], 4, axis=2).astype(np.float32)
    else:
        d = (rng.random((B, T, 4), dtype=np.float32) * np.float32(cfg.max_bbox / 2)).astype(np.float32)
    labels = rng.integers(0, cfg.C, size=(B, T)).astype(np.int32)
    centers = np.stack([cx, cy], axis=2).astype(np.float32)
    bboxes = np.concatenate([centers - d[:, :, 0:2], centers + d[:, :, 2:4]], axis=2).astype(np.float32)
    # padded slots -> drawable decoy (centre (2,2)*osf, bbox (1,1,3,3)*osf, label 0)
    pad = np.arange(T)[None, :] >= sizes[:, None]
    centers[pad] = np.float32(2 * cfg.osf)
    bboxes[pad] = np.array([1, 1, 3, 3], np.float32) * np.float32(cfg.osf)
    labels[pad] = 0
    return Workload(cfg, np.ascontiguousarray(centers), np.ascontiguousarray(bboxes), np.ascontiguousarray(labels), sizes)


def shard_range(batch_size: int, rank: int, world_size: int):
    """Contiguous slice [lo, hi) of the flattened batch x camera dimension owned by `rank`."""
    return batch_size * rank // world_size, batch_size * (rank + 1) // world_size
