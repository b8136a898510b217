"""CPU: pin the oracle (oracle/dh_oracle.c + numpy restatements) against golden vectors produced by RUNNING the
reference's own Python CPU path (tests/golden/make_golden.py; reference files
packages/draw_heatmap/tests/_test_helpers.py:20-74 and _gaussian_drawing.py:215-240)."""
import numpy as np
import pytest

BATCHED = ["ref_batched_20x50", "ref_batched_45x63"]
CLASSWISE = ["ref_classwise_20x50", "ref_classwise_64x64"]


def _ulp_diff(a, b):
    a = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    return np.abs(a - b)


def test_k1_bit_exact_vs_reference(oracle, golden):
    g = golden("ref_k1")
    i = 0
    while f"osf_{i}" in g:
        osf = float(g[f"osf_{i}"])
        c, r = oracle.radius_center(g[f"centers_f_{i}"], g[f"bboxes_f_{i}"], osf)
        assert np.array_equal(c, g[f"centers_int_{i}"]), f"centres differ at osf={osf}"
        assert np.array_equal(r, g[f"radii_{i}"]), f"radii differ at osf={osf}"
        c2, r2 = oracle.radius_center_numpy(g[f"centers_f_{i}"], g[f"bboxes_f_{i}"], osf)
        assert np.array_equal(c2, c) and np.array_equal(r2, r)
        i += 1
    assert i == 8


@pytest.mark.parametrize("name", BATCHED + CLASSWISE)
def test_k1_on_padded_batches(oracle, golden, name):
    g = golden(name)
    c, r = oracle.radius_center(g["centers_f"], g["bboxes_f"], float(g["osf"]))
    assert np.array_equal(c, g["centers_int"]) and np.array_equal(r, g["radii"])


@pytest.mark.parametrize("name", BATCHED + CLASSWISE)
def test_refcpu_family_equals_reference_cpu(oracle, golden, name):
    """The C port of the reference's float64 CPU path reproduces the reference's own output bit for bit."""
    g = golden(name)
    labels = g.get("labels")
    heat = np.zeros_like(g["heat"])
    oracle.refcpu_draw_batched(heat, g["centers_int"], g["radii"], g["sizes"], float(g["factor"]), float(g["k"]), labels)
    assert _ulp_diff(heat, g["heat"]).max() == 0
    # threads over samples and the fused zeroing give the same result
    heat2 = np.full_like(g["heat"], 7.0)
    oracle.refcpu_draw_batched(heat2, g["centers_int"], g["radii"], g["sizes"], float(g["factor"]), float(g["k"]), labels,
                               zero_first=True, nthreads=3)
    assert np.array_equal(heat2, heat)


@pytest.mark.parametrize("name", BATCHED + CLASSWISE)
def test_spec_family_vs_reference_cpu(oracle, golden, name):
    """fp32 kernel arithmetic (cuh:36-49) vs the reference CPU output: same touched mask, <= 1e-6 abs."""
    g = golden(name)
    labels = g.get("labels")
    heat = np.zeros_like(g["heat"])
    oracle.draw_batched(heat, g["centers_int"], g["radii"], g["sizes"], float(g["factor"]), float(g["k"]), labels)
    assert np.array_equal(heat > 0, g["heat"] > 0)
    assert np.abs(heat.astype(np.float64) - g["heat"]).max() <= 1e-6
    # mask == union of clipped windows
    mask = oracle.touched_mask_batched(heat.shape, g["centers_int"], g["radii"], g["sizes"], labels)
    assert np.array_equal(mask, heat > 0)
    # C vs independent numpy restatement: bit-exact
    heat_np = np.zeros_like(heat)
    oracle.draw_batched_numpy(heat_np, g["centers_int"], g["radii"], g["sizes"], float(g["factor"]), float(g["k"]), labels)
    assert np.array_equal(heat_np, heat)


def test_edges_vs_reference_cpu(oracle, golden):
    g = golden("ref_edges")
    objs = g["objs"]
    H, W = g["heat_all"].shape
    k, factor = float(g["k"]), float(g["factor"])
    for i, (x, y, r) in enumerate(objs):
        heat = np.zeros((1, H, W), np.float32)
        oracle.draw_batched(heat, np.array([[[x, y]]], np.int32), np.array([[r]], np.int32), np.array([1]), factor, k)
        assert np.array_equal(heat[0] > 0, g["heat_each"][i] > 0), f"mask differs for object {(x, y, r)}"
        assert np.abs(heat[0] - g["heat_each"][i]).max() <= 1e-6
        heat64 = np.zeros((1, H, W), np.float32)
        oracle.refcpu_draw_batched(heat64, np.array([[[x, y]]], np.int32), np.array([[r]], np.int32), np.array([1]), factor, k)
        assert np.array_equal(heat64[0], g["heat_each"][i])
    heat = np.zeros((1, H, W), np.float32)
    n = len(objs)
    oracle.draw_batched(heat, objs[None, :, :2], objs[None, :, 2], np.array([n]), factor, k)
    assert np.abs(heat[0] - g["heat_all"]).max() <= 1e-6
    # flat entry with every object in its own plane
    heat_f = np.zeros((n, H, W), np.float32)
    oracle.draw_flat(heat_f, objs[:, :2], objs[:, 2], np.arange(n, dtype=np.int32), factor, k)
    assert np.abs(heat_f - g["heat_each"]).max() <= 1e-6


def test_reference_example_inputs(oracle, golden):
    """examples/input_data.py:35-59 of the reference, rendered by its CPU path."""
    g = golden("ref_examples")
    H, W, C = int(g["H"]), int(g["W"]), int(g["C"])
    B = g["centers"].shape[0]
    k, factor = float(g["k"]), float(g["factor"])
    heat = np.zeros((B, H, W), np.float32)
    oracle.draw_batched(heat, g["centers"], g["radii"], g["sizes"], factor, k)
    assert np.array_equal(heat > 0, g["heat"] > 0)
    assert np.abs(heat - g["heat"]).max() <= 1e-6
    heat_c = np.zeros((B, C, H, W), np.float32)
    oracle.draw_batched(heat_c, g["centers"], g["radii"], g["sizes"], factor, k, g["labels"])
    planes = heat_c.reshape(B * C, H, W)
    nz = np.nonzero(planes.reshape(B * C, -1).max(axis=1) > 0)[0]
    assert np.array_equal(nz, g["heat_cls_planes"])
    assert np.abs(planes[nz] - g["heat_cls_values"]).max() <= 1e-6


def test_known_answers(oracle):
    """Analytic pins: centre == k exactly; r=1, factor=6: edge e^-2, corner e^-4; square support."""
    heat = np.zeros((1, 9, 9), np.float32)
    oracle.draw_batched(heat, np.array([[[4, 4]]], np.int32), np.array([[1]], np.int32), np.array([1]), 6.0, 0.8)
    assert heat[0, 4, 4] == np.float32(0.8)
    assert abs(heat[0, 4, 5] - 0.8 * np.exp(-2.0)) < 1e-7 and abs(heat[0, 5, 5] - 0.8 * np.exp(-4.0)) < 1e-7
    assert (heat > 0).sum() == 9
    # existing contents survive where larger (in-place max, F2)
    heat = np.full((1, 9, 9), 0.5, np.float32)
    oracle.draw_batched(heat, np.array([[[4, 4]]], np.int32), np.array([[1]], np.int32), np.array([1]), 6.0, 0.8)
    assert heat[0, 4, 4] == np.float32(0.8) and heat[0, 5, 5] == np.float32(0.5) and heat[0, 0, 0] == np.float32(0.5)


def test_digest_is_additive_over_shards(oracle):
    rng = np.random.default_rng(1)
    v = rng.random(10_000, dtype=np.float32)
    v[rng.random(10_000) < 0.5] = 0
    whole = oracle.digest_numpy(v)
    with np.errstate(over="ignore"):
        parts = oracle.digest_numpy(v[:3333]).view(np.uint64) + oracle.digest_numpy(v[3333:], 3333).view(np.uint64)
    assert np.array_equal(parts.view(np.int64), whole)
    v2 = v.copy()
    i = int(np.nonzero(v2 > 0)[0][0])
    v2[i], v2[i + 1] = v2[i + 1], v2[i]  # same multiset of values, different mask positions -> hash must differ
    if (v2 > 0)[i] != (v > 0)[i]:
        assert oracle.digest_numpy(v2)[1] != whole[1]
