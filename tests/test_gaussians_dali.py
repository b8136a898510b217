"""N2: the DALI `draw_gaussians` operator semantics as a second renderer mode.

CPU part: oracle/dh_oracle.c:dho_draw_gaussians (restatement of
/root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc:31-75,88-157) reproduces, BIT FOR BIT, the
outputs of the reference's own operator source compiled against oracle/dali_stub (tests/golden/ref_dali.npz, generated
by tests/golden/make_golden_dali.py), plus the analytic known answers the reference's own test checks
(bounding_box_to_heatmap_converter_test.py:100-136: peak == k at the centre, zero outside the radius window).
GPU part: dh_draw_gaussians through the C ABI and the Python API against the goldens and the oracle:
touched masks identical, values <= 1e-6 abs and <= 2 ulp (device expf vs the C library's)."""
import os

import numpy as np
import pytest

import _cabi

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_dali.npz")
CASES = ("typical", "integer_r", "tiny_r", "two_d", "neg_k", "near_int")


def _case(g, name):
    return {k: g[f"{name}_{k}"] for k in ("heat_in", "centers", "radii", "slice_ids", "active", "sizes", "k", "factor", "heat_out")}


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_reference_operator_bit_for_bit(oracle, name):
    c = _case(np.load(GOLD), name)
    heat = c["heat_in"].copy()
    oracle.draw_gaussians(heat, c["centers"], c["radii"], c["sizes"], c["k"], float(c["factor"]), c["slice_ids"], c["active"])
    assert np.array_equal(heat.view(np.uint32), c["heat_out"].view(np.uint32))


def test_oracle_known_answers(oracle):
    heat = np.zeros((1, 2, 21, 21), np.float32)
    centers = np.array([[[10, 10], [3, 4]]], np.int32)
    radii = np.array([[3.0, 2.5]], np.float32)
    oracle.draw_gaussians(heat, centers, radii, np.array([2]), [0.8, 1.0], 1.0 / 3.0, np.array([[0, 1]], np.int32),
                          np.array([[True, True]]))
    assert heat[0, 0, 10, 10] == np.float32(0.8) and heat[0, 1, 4, 3] == np.float32(1.0)
    ys, xs = np.nonzero(heat[0, 0])
    assert ys.min() == 7 and ys.max() == 13 and xs.min() == 7 and xs.max() == 13  # [floor(c-r), ceil(c+r)]
    ys, xs = np.nonzero(heat[0, 1])
    assert ys.min() == 1 and ys.max() == 7 and xs.min() == 0 and xs.max() == 6  # r=2.5 -> +-3
    # sigma = r*factor = 1 -> exp(-d2/2)
    assert abs(heat[0, 0, 10, 11] - 0.8 * np.exp(-0.5)) < 1e-7 and abs(heat[0, 0, 11, 11] - 0.8 * np.exp(-1.0)) < 1e-7
    # inactive objects and padded slots draw nothing
    heat2 = np.zeros((1, 2, 21, 21), np.float32)
    oracle.draw_gaussians(heat2, centers, radii, np.array([1]), [0.8, 1.0], 1.0 / 3.0, np.array([[0, 1]], np.int32),
                          np.array([[False, True]]))
    assert not heat2.any()


# ---------------------------------------------------------------------------------------------- GPU
def _ulp(a, b):
    ai, bi = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    return int(np.abs(ai - bi).max()) if a.size else 0


def _check(got, want, name):
    assert np.array_equal(got > 0, want > 0), f"{name}: touched mask differs"
    assert np.array_equal(np.isnan(got), np.isnan(want))
    err = float(np.nanmax(np.abs(got - want))) if got.size else 0.0
    assert err <= 1e-6, f"{name}: max abs err {err}"
    # a denormal can land one step across a power of two: compare ulps on normal values only
    big = np.abs(want) > 1e-30
    assert _ulp(got[big], want[big]) <= 2, f"{name}: more than 2 ulp"


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_matches_reference_operator_goldens(cabi, cuda_dev, name):
    import torch
    from accvlab.draw_heatmap import b200

    c = _case(np.load(GOLD), name)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    rb = lambda a: b200.RaggedBatch(dev(a), sample_sizes=dev(c["sizes"]))
    hin = dev(c["heat_in"])
    l0 = b200.launch_count()
    out = b200.draw_gaussians(hin, rb(c["active"]), rb(c["slice_ids"]), rb(c["centers"]), rb(c["radii"]), dev(c["k"]),
                              float(c["factor"]))
    assert b200.launch_count() - l0 == 1
    assert out.data_ptr() != hin.data_ptr() and np.array_equal(hin.cpu().numpy(), c["heat_in"])  # out of place, input intact
    _check(out.cpu().numpy(), c["heat_out"], name)
    # in place gives the same bits
    h2 = hin.clone()
    b200.draw_gaussians(h2, rb(c["active"]), rb(c["slice_ids"]), rb(c["centers"]), rb(c["radii"]), dev(c["k"]),
                        float(c["factor"]), out=h2)
    assert torch.equal(h2, out)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_fresh_fast_path_on_the_golden_objects(cuda_dev, oracle, name):
    """FRESH rendering of every golden case's objects through the Python API (classify -> main renderer -> exact fix-up
    of asymmetric / negative-radius / degenerate-sigma objects) == the oracle on zeros (itself bit-identical to the
    reference operator) and == the single exact kernel bit for bit."""
    import torch
    from accvlab.draw_heatmap import b200
    from accvlab.draw_heatmap import draw_heatmap_ext as ext

    c = _case(np.load(GOLD), name)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    sizes = dev(c["sizes"])
    rb = lambda a: b200.RaggedBatch(dev(a), sample_sizes=sizes)
    shape = c["heat_in"].shape
    want = np.zeros(shape, np.float32)
    oracle.draw_gaussians(want, c["centers"], c["radii"], c["sizes"], c["k"], float(c["factor"]), c["slice_ids"], c["active"])
    l0 = b200.launch_count()
    fast = b200.draw_gaussians(None, rb(c["active"]), rb(c["slice_ids"]), rb(c["centers"]), rb(c["radii"]), dev(c["k"]),
                               float(c["factor"]), shape=shape, fast=True)
    W = shape[-1]
    assert b200.launch_count() - l0 == (3 if (W % 4 == 0 and W <= 1024) else 1)
    _check(fast.cpu().numpy(), want, name + " fresh fast")
    zeros = torch.zeros(shape, device=cuda_dev)
    exact = b200.draw_gaussians(zeros, rb(c["active"]), rb(c["slice_ids"]), rb(c["centers"]), rb(c["radii"]), dev(c["k"]),
                                float(c["factor"]))
    # max(0, .) over a zero input == the fresh result, except that a negative k leaves -0.0 vs 0.0 apart: compare values
    assert np.array_equal(fast.cpu().numpy(), exact.cpu().numpy())


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,T,C", [(33, 47, 40, 3), (8, 4100, 30, 1), (232, 400, 100, 10), (64, 64, 700, 2)])
def test_gpu_vs_oracle_shapes(cabi, cuda_dev, oracle, H, W, T, C):
    """Scalar path (W % 4 != 0), rows wider than one pass, a BASELINE-sized map, and T beyond one list chunk (700 > 512);
    fresh (no input) and accumulate into a non-zero input."""
    import torch

    rng = np.random.default_rng(H * 1000 + W)
    B = 5
    sizes = rng.integers(0, T + 1, B).astype(np.int64)
    sizes[1] = T
    centers = np.stack([rng.integers(-8, W + 8, (B, T)), rng.integers(-8, H + 8, (B, T))], 2).astype(np.int32)
    radii = (0.3 + rng.random((B, T)) * 12.0).astype(np.float32)
    radii[rng.random((B, T)) < 0.2] = np.float32(3.0)
    slice_ids = rng.integers(0, C, (B, T)).astype(np.int32)
    active = rng.random((B, T)) < 0.8
    k = (0.5 + rng.random(C) * 0.5).astype(np.float32)
    factor = 0.3
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    d = dict(c=dev(centers), r=dev(radii), s=dev(slice_ids), a=dev(active), n=dev(sizes), k=dev(k))

    wsb = cabi.dh_draw_gaussians_workspace_bytes(B, T)
    ws = torch.empty(wsb, dtype=torch.uint8, device=cuda_dev)

    def run(hin, out, fast=False):
        rc = cabi.dh_draw_gaussians(_cabi.stream_ptr(), _cabi.ptr(hin), _cabi.ptr(out), _cabi.ptr(d["c"]), _cabi.ptr(d["r"]),
                                    _cabi.ptr(d["s"]), _cabi.ptr(d["a"]), _cabi.ptr(d["n"]), _cabi.DH_SIZES_I64,
                                    _cabi.ptr(d["k"]), C, B, T, C, H, W, factor, None, _cabi.ptr(ws) if fast else None,
                                    wsb if fast else 0)
        _cabi.check(cabi, rc, "dh_draw_gaussians")
        torch.cuda.synchronize()

    want = np.zeros((B, C, H, W), np.float32)
    oracle.draw_gaussians(want, centers, radii, sizes, k, factor, slice_ids, active)
    out = torch.full((B, C, H, W), 9.0, device=cuda_dev)
    run(None, out)  # fresh, single exact kernel
    _check(out.cpu().numpy(), want, "fresh")
    out_f = torch.full((B, C, H, W), 9.0, device=cuda_dev)
    l0 = cabi.dh_launch_count()
    run(None, out_f, fast=True)  # fresh, classify -> main renderer -> exact fix-up (when the map qualifies)
    assert cabi.dh_launch_count() - l0 == (3 if (W % 4 == 0 and W <= 1024) else 1)
    _check(out_f.cpu().numpy(), want, "fresh fast path")
    assert torch.equal(out_f, out), "fast path and exact kernel must agree bit for bit"
    bg = (rng.random((B, C, H, W), dtype=np.float32) * 0.6 * (rng.random((B, C, H, W)) < 0.3)).astype(np.float32)
    want2 = bg.copy()
    oracle.draw_gaussians(want2, centers, radii, sizes, k, factor, slice_ids, active)
    out2 = torch.empty((B, C, H, W), device=cuda_dev)
    run(dev(bg), out2)
    _check(out2.cpu().numpy(), want2, "accumulate")


@pytest.mark.gpu
def test_gpu_out_of_range_classes_and_radii_are_skipped_and_counted(cabi, cuda_dev, oracle):
    import torch

    B, T, C, H, W = 1, 4, 2, 16, 16
    centers = np.array([[[4, 4], [8, 8], [12, 12], [5, 9]]], np.int32)
    radii = np.array([[2.0, np.nan, 3.0, 1e9]], np.float32)
    slice_ids = np.array([[0, 1, 7, 1]], np.int32)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    status = torch.zeros(4, dtype=torch.int32, device=cuda_dev)
    out = torch.empty((B, C, H, W), device=cuda_dev)
    keep = [dev(centers), dev(radii), dev(slice_ids), dev(np.array([T], np.int64)), dev(np.array([1.0, 1.0], np.float32))]
    rc = cabi.dh_draw_gaussians(_cabi.stream_ptr(), None, _cabi.ptr(out), _cabi.ptr(keep[0]), _cabi.ptr(keep[1]),
                                _cabi.ptr(keep[2]), None, _cabi.ptr(keep[3]), _cabi.DH_SIZES_I64, _cabi.ptr(keep[4]), 2,
                                B, T, C, H, W, 0.3, _cabi.ptr(status), None, 0)
    _cabi.check(cabi, rc, "dh_draw_gaussians")
    torch.cuda.synchronize()
    want = np.zeros((B, C, H, W), np.float32)
    oracle.draw_gaussians(want, centers, radii, np.array([T]), [1.0, 1.0], 0.3, slice_ids, None)
    _check(out.cpu().numpy(), want, "skips")
    st = status.cpu().numpy()
    assert st[0] == 1 and st[2] == 2  # one bad class; NaN and 1e9 radii


# ------------------------------------------------------------------------- converter front half (boxes -> params)
CONV_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_converter.npz")
CONV_CASES = ("perclass", "classes_nothresh", "noclass_minsize", "noclass_plain")
CONV_OUT = ("bboxes_clipped", "center_int", "center_offset", "radii", "is_active", "hw_clipped", "fraction")


def _conv_case(name):
    g = np.load(CONV_GOLD)
    c = {k[len(name) + 1:]: g[k] for k in g.files if k.startswith(name + "_")}
    P = c["params"]
    c["kw"] = dict(heatmap_hw=(int(P[0]), int(P[1])), radius_scaling_factor=float(P[2]), min_radius=float(P[3]),
                   max_radius=float(P[4]), min_fraction_area=float(P[5]), num_classes=int(P[6]),
                   per_class_min_size=c["per_class"] if len(c["per_class"]) else None,
                   min_object_size=c["min_size"] if len(c["min_size"]) else None)
    c["use_classes"], c["own_centers"], c["with_valid"] = bool(P[7]), bool(P[8]), bool(P[9])
    return c


def _same(a, b, name):
    if a.dtype.kind == "f":
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) or np.array_equal(a, b, equal_nan=True), name
    else:
        assert np.array_equal(a, b), name


@pytest.mark.parametrize("name", CONV_CASES)
def test_oracle_converter_front_half_reproduces_reference_bit_for_bit(oracle, name):
    c = _conv_case(name)
    got = oracle.boxes_to_gaussian_params_numpy(c["bboxes"], c["centers"] if c["own_centers"] else None,
                                                c["classes"] if c["use_classes"] else None,
                                                c["valid"] if c["with_valid"] else None, c["sizes"], c["scale"], **c["kw"])
    for k in CONV_OUT:
        _same(got[k], c[k], f"{name}/{k}")


@pytest.mark.gpu
@pytest.mark.parametrize("name", CONV_CASES)
def test_gpu_converter_front_half_matches_reference_bit_for_bit(cuda_dev, name):
    """boxes -> (clipped boxes, int centre, offset, radius, is_active, hw, fraction) on the GPU == the reference's own
    Python / numba functions (tests/golden/ref_converter.npz), every output bit-exact; then straight into draw_gaussians."""
    import torch
    from accvlab.draw_heatmap import b200

    c = _conv_case(name)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    sizes = dev(c["sizes"])
    rb = lambda a: b200.RaggedBatch(dev(a), sample_sizes=sizes)
    kw = c["kw"]
    l0 = b200.launch_count()
    got = b200.boxes_to_gaussian_params(
        rb(c["bboxes"]), dev(c["scale"]), kw["heatmap_hw"], centers=rb(c["centers"]) if c["own_centers"] else None,
        classes=rb(c["classes"]) if c["use_classes"] else None, is_valid=rb(c["valid"]) if c["with_valid"] else None,
        radius_scaling_factor=kw["radius_scaling_factor"], min_radius=kw["min_radius"], max_radius=kw["max_radius"],
        min_fraction_area=kw["min_fraction_area"], num_classes=kw["num_classes"],
        per_class_min_object_sizes=kw["per_class_min_size"], min_object_size=kw["min_object_size"])
    assert b200.launch_count() - l0 == 1
    names = dict(bboxes_clipped="bboxes_clipped", center="center_int", center_offset="center_offset", radii="radii",
                 is_active="is_active", hw_clipped="hw_clipped", fraction_area="fraction")
    for k, gk in names.items():
        _same(got[k].tensor.cpu().numpy(), c[gk], f"{name}/{k}")
    # and on into the renderer: same heatmap as the oracle on the reference's parameters
    C = kw["num_classes"] if c["use_classes"] else 1
    H, W = kw["heatmap_hw"]
    kcls = np.linspace(0.7, 1.0, C).astype(np.float32)
    slice_ids = rb(np.clip(c["classes"], 0, C - 1).astype(np.int32)) if c["use_classes"] else None
    heat = b200.draw_gaussians(None, got["is_active"], slice_ids, got["center"], got["radii"], dev(kcls), 1.0 / 3.0,
                               shape=(len(c["sizes"]), C, H, W))
    from oracle import oracle as orc
    want = np.zeros(heat.shape, np.float32)
    orc.draw_gaussians(want, c["center_int"], c["radii"], c["sizes"], kcls, 1.0 / 3.0,
                       np.clip(c["classes"], 0, C - 1).astype(np.int32) if c["use_classes"] else None, c["is_active"])
    _check(heat.cpu().numpy(), want, name + " render")


@pytest.mark.gpu
def test_gpu_fast_path_is_automatic_from_256_mib(cuda_dev):
    """The product's own rule, no forcing: below 256 MiB of output b200.draw_gaussians is ONE launch (the exact kernel),
    from 256 MiB on THREE (classify -> main renderer -> exact fix-up), and both give the same bits as the exact kernel."""
    import torch
    from accvlab.draw_heatmap import b200

    C, H, W, T = 10, 232, 400, 100
    for B, launches in ((8, 1), (80, 3)):  # 29.7 MB / 297 MB of output
        g = torch.Generator().manual_seed(B)
        sizes = torch.randint(0, T + 1, (B,), generator=g, dtype=torch.int64)
        centers = torch.stack([torch.randint(-5, W + 5, (B, T), generator=g), torch.randint(-5, H + 5, (B, T), generator=g)],
                              2).int()
        radii = (torch.rand(B, T, generator=g) * 30 + 0.4).float()
        radii[torch.rand(B, T, generator=g) < 0.3] = 7.0  # integer radii: the symmetric-window class of the fast path
        slice_ids = torch.randint(0, C, (B, T), generator=g).int()
        k = (0.5 + 0.5 * torch.rand(C, generator=g)).float().to(cuda_dev)
        d = [t.to(cuda_dev) for t in (centers, radii, slice_ids, sizes)]
        rb = lambda t: b200.RaggedBatch(t, sample_sizes=d[3])
        l0 = b200.launch_count()
        auto = b200.draw_gaussians(None, None, rb(d[2]), rb(d[0]), rb(d[1]), k, 0.35, shape=(B, C, H, W))
        assert b200.launch_count() - l0 == launches
        exact = b200.draw_gaussians(None, None, rb(d[2]), rb(d[0]), rb(d[1]), k, 0.35, shape=(B, C, H, W), fast=False)
        torch.cuda.synchronize()
        assert torch.equal(auto, exact)
        del auto, exact
