"""GPU: the object-heavy renderer (csrc/dh_tiles_kernel.cuh: register tiles; compiled into the groups_kernel instantiation
the host selects for T >= 256 padded slots and k >= 0, taken PER UNIT when at least 48 objects land on the unit) against
the oracle and against the staged renderer on THE SAME objects: padding the slot arrays from T = 255 to T >= 256 changes
the kernel, never the result. Shapes chosen to hit every edge of the tile geometry: H not a multiple of the 8-row
strip, W not a multiple of the 64-column tile, W > 512 (more tiles than warps: two rounds), T > 1024 (several list
passes), empty strips / samples (bulk zero path), objects centred off the map, negative and oversize radii, both
sizes dtypes, both modes, K1 fused (float boxes) with the fused digest, skip counters counted once.
"""
import numpy as np
import pytest
import torch

import _cabi
import workloads

pytestmark = pytest.mark.gpu


def _objects(seed, B, n_max, C, H, W, r_hi, empty_every=0):
    rng = np.random.default_rng(seed)
    sizes = rng.integers(0, n_max + 1, size=B)
    if empty_every:
        sizes[::empty_every] = 0
    sizes[-1] = n_max
    centers = np.stack([rng.integers(-r_hi, W + r_hi, (B, n_max)), rng.integers(-r_hi, H + r_hi, (B, n_max))], 2).astype(np.int32)
    radii = rng.integers(-1, r_hi + 1, (B, n_max)).astype(np.int32)  # -1: draws nothing (reference: empty loops)
    labels = rng.integers(0, C, (B, n_max)).astype(np.int32)
    return centers, radii, labels, sizes.astype(np.int64)


def _pad(a, T, fill):
    """[B, n, ...] -> [B, T, ...]; padded slots hold a drawable decoy that must never be drawn (t >= size)."""
    out = np.full((a.shape[0], T) + a.shape[2:], fill, a.dtype)
    out[:, :a.shape[1]] = a
    return out


CASES = [
    # B, n_max, C, H,  W,    r_hi, T_small, T_big, sizes dtype
    (3, 255, 4, 70, 200, 40, 255, 256, torch.int64),   # H % 8 = 6, W % 64 = 8: partial strip and partial tile
    (2, 200, 1, 19, 1024, 90, 200, 300, torch.int32),  # 16 tiles per strip: two rounds; objects wider than a tile
    (2, 255, 2, 64, 128, 12, 255, 2500, torch.int64),  # small objects, hugely padded T: three list passes
    (5, 120, 3, 33, 64, 70, 120, 512, torch.int64),    # objects larger than the map; one tile per strip
    (4, 250, 2, 8, 512, 5, 250, 256, torch.int32),     # a single strip per plane
    (3, 20, 2, 64, 128, 10, 20, 300, torch.int64),     # few objects in widely padded slots: every unit stays staged
    (4, 140, 1, 96, 256, 30, 140, 400, torch.int32),   # dense and sparse units side by side (sample sizes 0..140)
]


@pytest.mark.parametrize("B,n_max,C,H,W,r_hi,T_small,T_big,sdtype", CASES)
def test_tile_renderer_equals_staged_renderer_and_oracle(cabi, cuda_dev, oracle, B, n_max, C, H, W, r_hi, T_small, T_big,
                                                         sdtype):
    centers, radii, labels, sizes = _objects(1000 + H + W, B, n_max, C, H, W, r_hi, empty_every=3)
    dev = cuda_dev
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    sz = d(sizes).to(sdtype)
    want = np.zeros((B, C, H, W), np.float32)
    oracle.draw_batched(want, centers, radii, sizes, 6.0, 0.8, labels)
    want_t = torch.from_numpy(want).to(dev)
    rng = np.random.default_rng(5)
    prefill = torch.from_numpy(rng.uniform(-0.5, 0.9, (B, C, H, W)).astype(np.float32)).to(dev)
    res = {}
    for T in (T_small, T_big):
        assert (T >= 256) == (T == T_big)
        ci, ri, lab = d(_pad(centers, T, 2)), d(_pad(radii, T, 1)), d(_pad(labels, T, 0))
        fresh = _cabi.draw_batched(cabi, torch.full((B, C, H, W), 3.0, device=dev), ci, ri, sz, lab, 6.0, 0.8,
                                   mode=_cabi.DH_MODE_FRESH)
        acc0 = _cabi.draw_batched(cabi, torch.zeros((B, C, H, W), device=dev), ci, ri, sz, lab, 6.0, 0.8)
        accp = _cabi.draw_batched(cabi, prefill.clone(), ci, ri, sz, lab, 6.0, 0.8)
        one = _cabi.draw_batched(cabi, torch.full((B, H, W), 3.0, device=dev), ci, ri, sz, None, 6.0, 0.8,
                                 mode=_cabi.DH_MODE_FRESH)
        torch.cuda.synchronize()
        assert torch.equal(fresh, acc0), f"T={T}: FRESH vs ACCUMULATE into zeros"
        assert torch.equal(fresh > 0, want_t > 0), f"T={T}: touched mask"
        assert float((fresh - want_t).abs().max()) <= 1e-6, f"T={T}"
        res[T] = (fresh, accp, one)
    for a, b, what in zip(res[T_small], res[T_big], ("fresh", "accumulate into a prefilled map", "one plane per sample")):
        assert torch.equal(a, b), f"{what}: staged renderer (T={T_small}) != tile renderer (T={T_big})"
    # accumulate into a prefilled map (negative values included) == elementwise reference blend of cuh:30
    fresh, accp, one = res[T_big]
    touched = want_t > 0
    # where nothing was drawn the map is untouched; where something was drawn: keep old iff old >= v
    # (pixels drawn with a value that underflowed to 0 are not in `touched`; they follow the same rule: max(old, 0))
    blend = torch.where(prefill >= fresh, prefill, fresh)
    assert torch.equal(accp[touched], blend[touched])
    assert torch.equal(accp[~touched & (prefill >= 0)], prefill[~touched & (prefill >= 0)])
    want1 = np.zeros((B, H, W), np.float32)
    oracle.draw_batched(want1, centers, radii, sizes, 6.0, 0.8, None)
    w1 = torch.from_numpy(want1).to(dev)
    assert torch.equal(one > 0, w1 > 0) and float((one - w1).abs().max()) <= 1e-6


def test_tile_renderer_cfg4_boxes_digest_and_status(cabi, cuda_dev, oracle):
    """BASELINE cfg4 (2 samples of it): K1 fused, fused digest == dh_checksum of the output == oracle values; negative k
    (served by the staged renderer) and k = 0 on the same inputs."""
    dev = cuda_dev
    w = workloads.make("cfg4", B=2)
    c = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    sizes, labels = torch.from_numpy(w.sizes).to(dev), torch.from_numpy(w.labels).to(dev)
    shape = (w.B, c.C, c.H, c.W)
    oci, ori = oracle.radius_center(w.centers_f, w.bboxes_f, c.osf)
    for k in (1.0, 0.0, -0.7):
        want = np.zeros(shape, np.float32)
        oracle.draw_batched(want, oci, ori, w.sizes, 6.0, k, w.labels)
        want_t = torch.from_numpy(want).to(dev)
        heat = torch.full(shape, 5.0, device=dev)
        dg = _cabi.draw_from_boxes_digest(cabi, heat, cf, bf, float(c.osf), sizes, labels, 6.0, k)
        plain = _cabi.draw_from_boxes(cabi, torch.full(shape, 5.0, device=dev), cf, bf, float(c.osf), sizes, labels, 6.0, k)
        acc = _cabi.draw_batched(cabi, torch.zeros(shape, device=dev), torch.from_numpy(oci).to(dev),
                                 torch.from_numpy(ori).to(dev), sizes, labels, 6.0, k)
        torch.cuda.synchronize()
        assert torch.equal(heat, plain) and torch.equal(heat, acc), f"k={k}"
        assert torch.equal(dg, _cabi.checksum(cabi, heat)), f"k={k}"
        assert torch.equal(heat > 0, want_t > 0), f"k={k}"
        assert float((heat - want_t).abs().max()) <= 1e-6, f"k={k}"


@pytest.mark.parametrize("W", [96, 640])
def test_tile_renderer_counts_skips_once(cabi, cuda_dev, W):
    """Oversize radii and bad labels are counted once per object, whatever the number of strips, passes and rounds."""
    dev = cuda_dev
    B, T, C, H = 2, 1300, 3, 40
    big = 4194304 + 9
    centers = torch.full((B, T, 2), 10, dtype=torch.int32)
    radii = torch.full((B, T), 1, dtype=torch.int32)
    labels = (torch.arange(T, dtype=torch.int32) % C).repeat(B, 1)
    sizes = torch.tensor([T, 1100], dtype=torch.int64)
    for t in (0, 5, 1024, 1299):
        radii[0, t] = big
    radii[1, 1030] = big
    radii[1, 1200] = big  # padded slot of sample 1: not counted
    labels[0, 7] = -1
    labels[1, 1050] = C
    labels[1, 1250] = C + 3  # padded
    centers, radii, labels, sizes = (t.to(dev) for t in (centers, radii, labels, sizes))
    for mode in (_cabi.DH_MODE_FRESH, _cabi.DH_MODE_ACCUMULATE):
        status = torch.zeros(4, dtype=torch.int32, device=dev)
        heat = torch.zeros((B, C, H, W), device=dev)
        _cabi.draw_batched(cabi, heat, centers, radii, sizes, labels, mode=mode, status=status)
        torch.cuda.synchronize()
        s = status.cpu().numpy().view(np.uint32)
        assert int(s[0]) == 2 and int(s[2]) == 5, f"mode {mode}: {s}"
        assert float(heat.max()) == 1.0 and int((heat > 0).sum()) == 9 * B * C


def _bits(t):
    return t.contiguous().view(torch.int32)


@pytest.mark.parametrize("T", [1500, 300])  # per unit: more objects than the shared list holds (slot passes) / fewer
@pytest.mark.parametrize("factor,odd_slots", [
    (6.0, ()),                  # every object plain: arithmetic sweep only (odd and even hit counts per tile)
    (6.0, (3, 77)),             # r > 2896 in the first list pass: predicated sweep from the start
    (6.0, (1100, 1490)),        # ... only in a LATER pass: the tile switches sweeps with accumulators in flight
    (1e21, ()),                 # var underflows to 0, 1/var = inf: m = inf off centre, NaN at the centre (cuh:45)
    (1e-21, ()),                # var = inf, 1/var = 0: m = 0 over the whole window
    (3e16, ()),                 # 1/var ~ 1e31 > 1e30 but finite: m overflows for distant pixels only
])
def test_tile_renderer_unusual_objects_bit_equal_staged(cabi, cuda_dev, oracle, factor, odd_slots, T):
    """The tile renderer's two sweeps (packed-fp32 arithmetic masking for plain objects, per-pixel predication for the
    rest) against the staged renderer on the same T (the *_ws entry never selects the tile renderer): same bits, NaNs
    included, FRESH and ACCUMULATE into a prefilled map; against the oracle where the values are ordinary."""
    dev = cuda_dev
    B, C, H, W = 2, 2, 44, 200
    odd_slots = tuple(t * T // 1500 for t in odd_slots)
    rng = np.random.default_rng(77)
    centers = np.stack([rng.integers(-20, W + 20, (B, T)), rng.integers(-20, H + 20, (B, T))], 2).astype(np.int32)
    radii = rng.integers(0, 31, (B, T)).astype(np.int32)
    labels = rng.integers(0, C, (B, T)).astype(np.int32)
    for i, t in enumerate(odd_slots):
        radii[:, t] = (2897, 5000)[i % 2]
        centers[:, t] = (W // 2 + 2900 * (i % 2), H // 2)  # 2897 covers the map; 5000 from 2900 px to the right too
        labels[:, t] = i % C
    sizes = np.array([T, T - 1 if odd_slots else T * 4 // 5], np.int64)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    ci, ri, lab, sz = d(centers), d(radii), d(labels), d(sizes)
    prefill = torch.from_numpy(rng.uniform(-0.5, 0.9, (B, C, H, W)).astype(np.float32)).to(dev)
    out = {}
    for ws in (False, True):
        fresh = _cabi.draw_batched(cabi, torch.full((B, C, H, W), 3.0, device=dev), ci, ri, sz, lab, factor, 0.8,
                                   mode=_cabi.DH_MODE_FRESH, ws=ws)
        accp = _cabi.draw_batched(cabi, prefill.clone(), ci, ri, sz, lab, factor, 0.8, ws=ws)
        torch.cuda.synchronize()
        out[ws] = (fresh, accp)
    assert torch.equal(_bits(out[False][0]), _bits(out[True][0])), "fresh: tile renderer != staged renderer"
    assert torch.equal(_bits(out[False][1]), _bits(out[True][1])), "accumulate: tile renderer != staged renderer"
    if factor == 6.0:
        want = np.zeros((B, C, H, W), np.float32)
        oracle.draw_batched(want, centers, radii, sizes, factor, 0.8, labels)
        want_t = torch.from_numpy(want).to(dev)
        assert torch.equal(out[False][0] > 0, want_t > 0)
        assert float((out[False][0] - want_t).abs().max()) <= 1e-6


@pytest.mark.parametrize("H,W", [(37, 200), (8, 64), (50, 1000)])
def test_tile_renderer_writes_nothing_outside_the_planes(cabi, cuda_dev, H, W):
    """Guard bands around a densely hit output (partial last strip, partial last tile): the sentinels before and after
    the planes stay intact in both modes and the planes equal the staged renderer's (no compute-sanitizer on this
    pool: bounds are checked this way)."""
    dev = cuda_dev
    B, C, T, n = 2, 2, 320, 300
    rng = np.random.default_rng(H * W)
    centers = np.stack([rng.integers(-30, W + 30, (B, T)), rng.integers(-30, H + 30, (B, T))], 2).astype(np.int32)
    radii = rng.integers(1, 40, (B, T)).astype(np.int32)
    labels = rng.integers(0, C, (B, T)).astype(np.int32)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    ci, ri, lab, sz = d(centers), d(radii), d(labels), d(np.array([n, n - 7], np.int64))
    N, pad = B * C * H * W, 4096
    for mode in (_cabi.DH_MODE_FRESH, _cabi.DH_MODE_ACCUMULATE):
        out = {}
        for ws in (False, True):
            buf = torch.full((N + 2 * pad,), -7.5, device=dev)
            heat = buf[pad:pad + N].view(B, C, H, W)
            if mode == _cabi.DH_MODE_ACCUMULATE:
                heat.zero_()
            _cabi.draw_batched(cabi, heat, ci, ri, sz, lab, 6.0, 1.0, mode=mode, ws=ws)
            torch.cuda.synchronize()
            assert bool((buf[:pad] == -7.5).all()) and bool((buf[pad + N:] == -7.5).all()), f"mode {mode}, ws {ws}"
            out[ws] = heat.clone()
        assert torch.equal(_bits(out[False]), _bits(out[True])), f"mode {mode}"
        assert float(out[False].min()) >= 0.0 and float(out[False].max()) == 1.0


@pytest.mark.parametrize("seed", range(8))
def test_tile_renderer_random_shapes_bit_equal_staged(cabi, cuda_dev, seed):
    """Differential run over random shapes, slot counts, object counts and radius ranges (dense units, sparse units,
    lists beyond the shared capacity, radii beyond the exact-integer bound): the per-unit renderer choice never changes
    a bit against the staged renderer of the *_ws entry, FRESH and ACCUMULATE into a prefilled map."""
    dev = cuda_dev
    rng = np.random.default_rng(9000 + seed)
    for _ in range(5):
        B, C = int(rng.integers(1, 5)), int(rng.integers(1, 4))
        H, W = int(rng.integers(1, 140)), 4 * int(rng.integers(1, 257))
        T = int(rng.integers(256, 1400))
        r_hi = int(rng.choice([3, 12, 40, 120, 3500]))
        sizes = rng.integers(0, T + 1, B).astype(np.int64)
        sizes[rng.integers(0, B)] = T if rng.random() < 0.5 else int(rng.integers(0, 40))
        centers = np.stack([rng.integers(-r_hi, W + r_hi, (B, T)), rng.integers(-r_hi, H + r_hi, (B, T))], 2).astype(np.int32)
        radii = rng.integers(-1, r_hi + 1, (B, T)).astype(np.int32)
        labels = rng.integers(0, C, (B, T)).astype(np.int32)
        factor = float(rng.choice([6.0, 2.5, 11.0]))
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        ci, ri, lab, sz = d(centers), d(radii), d(labels), d(sizes)
        prefill = torch.from_numpy(rng.uniform(-0.5, 0.9, (B, C, H, W)).astype(np.float32)).to(dev)
        res = []
        for ws in (False, True):
            fresh = _cabi.draw_batched(cabi, torch.full((B, C, H, W), 3.0, device=dev), ci, ri, sz, lab, factor, 0.9,
                                       mode=_cabi.DH_MODE_FRESH, ws=ws)
            accp = _cabi.draw_batched(cabi, prefill.clone(), ci, ri, sz, lab, factor, 0.9, ws=ws)
            torch.cuda.synchronize()
            res.append((fresh, accp))
        what = f"B={B} C={C} H={H} W={W} T={T} r_hi={r_hi} sizes={sizes.tolist()} factor={factor}"
        assert torch.equal(_bits(res[0][0]), _bits(res[1][0])), "fresh: " + what
        assert torch.equal(_bits(res[0][1]), _bits(res[1][1])), "accumulate: " + what


def test_flat_entry_dense_planes_bit_equal_staged_batched(cabi, cuda_dev, oracle):
    """The flat entry (reference draw_heatmap: concatenated objects + a plane index each) with hundreds of objects per
    plane also reaches the per-unit tile renderer (selected from 16 objects per plane on average): same bits as the
    staged renderer of the batched *_ws entry on the same objects, both modes, and the oracle's values."""
    dev = cuda_dev
    w = workloads.make("cfg4", B=3)
    c = w.cfg
    oci, ori = oracle.radius_center(w.centers_f, w.bboxes_f, c.osf)
    n = [int(s) for s in w.sizes]
    cen = np.concatenate([oci[b, :n[b]] for b in range(w.B)]).astype(np.int32)
    rad = np.concatenate([ori[b, :n[b]] for b in range(w.B)]).astype(np.int32)
    idx = np.concatenate([np.full(n[b], b, np.int32) for b in range(w.B)])
    perm = np.random.default_rng(4).permutation(len(idx))  # the flat entry takes the objects in any order
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    sizes = d(w.sizes)
    prefill = torch.from_numpy(np.random.default_rng(6).uniform(-0.5, 0.9, (w.B, c.H, c.W)).astype(np.float32)).to(dev)
    for mode in (_cabi.DH_MODE_FRESH, _cabi.DH_MODE_ACCUMULATE):
        start = (lambda: torch.full((w.B, c.H, c.W), 2.0, device=dev)) if mode == _cabi.DH_MODE_FRESH else prefill.clone
        flat = _cabi.draw_flat(cabi, start(), d(cen[perm]), d(rad[perm]), d(idx[perm]), 6.0, 1.0, mode=mode)
        staged = _cabi.draw_batched(cabi, start(), d(oci), d(ori), sizes, None, 6.0, 1.0, mode=mode, ws=True)
        torch.cuda.synchronize()
        assert torch.equal(_bits(flat), _bits(staged)), f"mode {mode}"
    want = np.zeros((w.B, c.H, c.W), np.float32)
    oracle.draw_batched(want, oci, ori, w.sizes, 6.0, 1.0, None)
    fresh = _cabi.draw_flat(cabi, torch.empty((w.B, c.H, c.W), device=dev), d(cen), d(rad), d(idx), 6.0, 1.0,
                            mode=_cabi.DH_MODE_FRESH)
    torch.cuda.synchronize()
    want_t = torch.from_numpy(want).to(dev)
    assert torch.equal(fresh > 0, want_t > 0) and float((fresh - want_t).abs().max()) <= 1e-6
