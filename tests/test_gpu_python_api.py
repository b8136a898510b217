"""GPU: the Python drop-in surface (accvlab.draw_heatmap), written like the reference's own tests
(packages/draw_heatmap/tests/test_draw_heatmap.py:34-85, test_draw_heatmap_batched.py:36-97,
test_draw_heatmap_batched_classwise.py:59-130: seed 7, IMG 320x800, osf 16, k 0.8, factor 6, B=48, C=20, decoy
padding), plus a direct comparison with the UNMODIFIED reference CUDA extension when oracle/_ref/ holds its build
(compiled in the build container from /root/reference by oracle/build_oracle.py; the .so travels to the GPU box)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
IMG_SHAPE, OSF, K, FACTOR, B, T, C, MAX_BBOX = (320, 800, 3), 16, 0.8, 6, 48, 50, 20, 120
H, W = IMG_SHAPE[0] // OSF, IMG_SHAPE[1] // OSF


def _load_ref(name):
    path = os.path.join(ROOT, "oracle", "_ref", name + ".so")
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (needs /root/reference at build time)")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _generate(dev, seed=7, with_labels=True):
    """generate_gt_bboxes(_with_labels) of the reference (_test_helpers.py:77-108) with a CPU generator."""
    gen = torch.Generator().manual_seed(seed)
    centers = torch.ones((B, T, 2)) * 2 * OSF  # decoy padding (test_draw_heatmap_batched.py:52-58)
    bboxes = torch.zeros((B, T, 4))
    bboxes[:, :, 0:2] = OSF
    bboxes[:, :, 2:4] = 3 * OSF
    labels = torch.zeros((B, T), dtype=torch.int32)
    sizes = torch.zeros(B, dtype=torch.int64)
    for i in range(B):
        n = int(torch.randint(0, T + 1, (1,), generator=gen))
        c = torch.stack([torch.rand(n, generator=gen) * IMG_SHAPE[1], torch.rand(n, generator=gen) * IMG_SHAPE[0]], 1)
        d = torch.rand(n, 4, generator=gen) * MAX_BBOX / 2
        centers[i, :n] = c
        bboxes[i, :n] = torch.cat([c - d[:, :2], c + d[:, 2:]], 1)
        labels[i, :n] = torch.randint(0, C, (n,), generator=gen).int()
        sizes[i] = n
    return centers.to(dev), bboxes.to(dev), labels.to(dev), sizes.to(dev)


def _oracle_heat(oracle, ci, ri, sizes, labels=None):
    shape = (B, H, W) if labels is None else (B, C, H, W)
    want = np.zeros(shape, np.float32)
    oracle.draw_batched(want, ci.cpu().numpy(), ri.cpu().numpy(), sizes.cpu().numpy(), FACTOR, K,
                        None if labels is None else labels.cpu().numpy())
    return want


def test_draw_heatmap_batched_like_reference_test(cuda_dev, oracle):
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap.b200 import RaggedBatch, radius_center

    cf, bf, labels, sizes = _generate(cuda_dev)
    ci, ri = radius_center(cf, bf, OSF)
    heat = torch.zeros((B, H, W), dtype=torch.float32, device=cuda_dev)
    ret = dh.draw_heatmap_batched(heat, RaggedBatch(ci, sample_sizes=sizes), RaggedBatch(ri, sample_sizes=sizes),
                                  diameter_to_sigma_factor=FACTOR, k_scale=K)
    assert ret is None
    want = _oracle_heat(oracle, ci, ri, sizes)
    mse = float(((heat.cpu() - torch.from_numpy(want)) ** 2).mean())
    assert mse < 1e-3  # the reference's own bar (test_draw_heatmap_batched.py:95-97)
    assert np.abs(heat.cpu().numpy() - want).max() <= 1e-6 and np.array_equal(heat.cpu().numpy() > 0, want > 0)
    # int32 sample_sizes (examples/draw_heatmap_batched.py:45) give the same result
    heat2 = torch.zeros_like(heat)
    s32 = sizes.to(torch.int32)
    dh.draw_heatmap_batched(heat2, RaggedBatch(ci, sample_sizes=s32), RaggedBatch(ri, sample_sizes=s32), FACTOR, K)
    assert torch.equal(heat, heat2)


def test_draw_heatmap_batched_classwise_like_reference_test(cuda_dev, oracle):
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap.b200 import RaggedBatch, radius_center, render_heatmap_batched, render_heatmap_from_boxes

    cf, bf, labels, sizes = _generate(cuda_dev)
    ci, ri = radius_center(cf, bf, OSF)
    rb = lambda t: RaggedBatch(t, sample_sizes=sizes)
    heat = torch.zeros((B, C, H, W), dtype=torch.float32, device=cuda_dev)
    dh.draw_heatmap_batched(heat, rb(ci), rb(ri), diameter_to_sigma_factor=FACTOR, k_scale=K, labels=rb(labels))
    want = _oracle_heat(oracle, ci, ri, sizes, labels)
    assert float(((heat.cpu() - torch.from_numpy(want)) ** 2).mean()) < 1e-3
    assert np.abs(heat.cpu().numpy() - want).max() <= 1e-6 and np.array_equal(heat.cpu().numpy() > 0, want > 0)
    # additions: fresh render (allocating) and K1-fused render give identical bits
    assert torch.equal(render_heatmap_batched(rb(ci), rb(ri), H, W, C, rb(labels), FACTOR, K), heat)
    assert torch.equal(render_heatmap_from_boxes(cf, bf, OSF, sizes, H, W, C, labels, FACTOR, K), heat)


def test_draw_heatmap_flat_like_reference_test(cuda_dev, oracle):
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap.b200 import radius_center

    cf, bf, labels, sizes = _generate(cuda_dev)
    ci, ri = radius_center(cf, bf, OSF)
    valid = torch.arange(T, device=cuda_dev)[None, :] < sizes[:, None]
    idx = torch.arange(B, device=cuda_dev, dtype=torch.int32)[:, None].expand(B, T)[valid].contiguous()
    heat = torch.zeros((B, H, W), dtype=torch.float32, device=cuda_dev)
    dh.draw_heatmap(heat, ci[valid].contiguous(), ri[valid].contiguous(), idx, FACTOR, K)
    want = _oracle_heat(oracle, ci, ri, sizes)
    assert np.abs(heat.cpu().numpy() - want).max() <= 1e-6 and np.array_equal(heat.cpu().numpy() > 0, want > 0)
    # classwise through the flat API: idx = sample*C + label (docs/intro.rst:88-93)
    idx_c = (torch.arange(B, device=cuda_dev, dtype=torch.int32)[:, None] * C + labels)[valid].contiguous()
    heat_c = torch.zeros((B * C, H, W), dtype=torch.float32, device=cuda_dev)
    dh.draw_heatmap(heat_c, ci[valid].contiguous(), ri[valid].contiguous(), idx_c, diameter_to_sigma_factor=FACTOR, k_scale=K)
    want_c = _oracle_heat(oracle, ci, ri, sizes, labels)
    assert np.abs(heat_c.cpu().numpy().reshape(want_c.shape) - want_c).max() <= 1e-6


def test_error_conventions_on_device(cuda_dev):
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap.b200 import RaggedBatch

    heat = torch.zeros((2, 8, 8), device=cuda_dev)
    c = torch.zeros((2, 3, 2), dtype=torch.int32, device=cuda_dev)
    r = torch.zeros((2, 3), dtype=torch.int32, device=cuda_dev)
    n = torch.tensor([1, 2], device=cuda_dev)
    rb = lambda t: RaggedBatch(t, sample_sizes=n)
    with pytest.raises(RuntimeError, match="contiguous"):
        dh.draw_heatmap_batched(heat.transpose(1, 2), rb(c), rb(r))
    with pytest.raises(RuntimeError, match="heatmap must be of shape"):
        dh.draw_heatmap_batched(heat[:, None].contiguous(), rb(c), rb(r))
    with pytest.raises(RuntimeError, match="batch_size"):
        dh.draw_heatmap_batched(torch.zeros((3, 8, 8), device=cuda_dev), rb(c), rb(r))
    with pytest.raises(RuntimeError, match="dtype"):
        dh.draw_heatmap_batched(heat, rb(c.long()), rb(r))
    with pytest.raises(RuntimeError, match="dtype"):
        dh.draw_heatmap_batched(heat.double(), rb(c), rb(r))
    with pytest.raises(RuntimeError, match="max_num_classes"):
        dh.draw_heatmap_batched(heat, rb(c), rb(r), labels=rb(r))


@pytest.mark.parametrize("build,abs_tol", [("dh_ref_ieee", 1e-6), ("dh_ref_fast", 1e-6)])
def test_against_unmodified_reference_cuda_extension(cuda_dev, build, abs_tol):
    """Same inputs into the reference's own kernel (draw_heatmap_cuda_kernel.cuh:76-109) and into ours.
    IEEE build (USE_FAST_MATH=0): same arithmetic -> expect bit-identical values. Default fast-math build (F3):
    touched masks identical, values within 1e-6 abs."""
    import workloads
    from accvlab.draw_heatmap import draw_heatmap_ext as ours

    ref = _load_ref(build)
    for name, nb in (("cfg1", None), ("cfg2", None), ("cfg3", 12), ("cfg4", 2)):
        w = workloads.make(name, B=nb)
        cfg = w.cfg
        cf, bf = torch.from_numpy(w.centers_f).to(cuda_dev), torch.from_numpy(w.bboxes_f).to(cuda_dev)
        lab, s32 = torch.from_numpy(w.labels).to(cuda_dev), torch.from_numpy(w.sizes).to(cuda_dev).int()
        ci, ri = ours.radius_center(cf, bf, float(cfg.osf))
        a = torch.zeros((w.B, cfg.C, cfg.H, cfg.W), device=cuda_dev)
        b = torch.zeros_like(a)
        ref.draw_heatmap_batched_classwise_impl(a, ci, ri, s32, lab, 6.0, 1.0)
        ours.draw_heatmap_batched_classwise_impl(b, ci, ri, s32, lab, 6.0, 1.0)
        torch.cuda.synchronize()
        assert torch.equal(a > 0, b > 0), f"{name}: touched mask differs from the reference {build} build"
        err = float((a - b).abs().max())
        assert err <= abs_tol, f"{name}: {err}"
        if build == "dh_ref_ieee":
            ulp = int((a.view(torch.int32).long() - b.view(torch.int32).long()).abs().max())
            assert ulp <= 1, f"{name}: {ulp} ulp vs the IEEE reference build"
        # non-classwise + flat entries
        a2 = torch.zeros((w.B, cfg.H, cfg.W), device=cuda_dev)
        b2 = torch.zeros_like(a2)
        ref.draw_heatmap_batched_impl(a2, ci, ri, s32, 6.0, 1.0)
        ours.draw_heatmap_batched_impl(b2, ci, ri, s32, 6.0, 1.0)
        assert torch.equal(a2 > 0, b2 > 0) and float((a2 - b2).abs().max()) <= abs_tol
        valid = torch.arange(w.T, device=cuda_dev)[None, :] < s32[:, None]
        idx = torch.arange(w.B, device=cuda_dev, dtype=torch.int32)[:, None].expand(w.B, w.T)[valid].contiguous()
        a3 = torch.zeros_like(a2)
        b3 = torch.zeros_like(a2)
        ref.draw_heatmap(a3, ci[valid].contiguous(), ri[valid].contiguous(), idx, 6.0, 1.0)
        ours.draw_heatmap(b3, ci[valid].contiguous(), ri[valid].contiguous(), idx, 6.0, 1.0)
        assert torch.equal(b3, b2) and float((a3 - b3).abs().max()) <= abs_tol


def test_one_process_two_devices(cuda_dev):
    """One process driving several GPUs: kernel attributes (dynamic shared memory opt-in) are per device, streams and
    allocations follow the tensors' device. Skipped on single-GPU boxes."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import workloads
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap import b200

    w = workloads.make("cfg2")
    cfg = w.cfg
    outs = []
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        heat = b200.render_heatmap_from_boxes(up(w.centers_f), up(w.bboxes_f), cfg.osf, up(w.sizes), cfg.H, cfg.W, cfg.C, up(w.labels))
        assert heat.device == dev
        acc = torch.zeros_like(heat)
        ci, ri = b200.radius_center(up(w.centers_f), up(w.bboxes_f), cfg.osf)
        rb = lambda t: b200.RaggedBatch(t, sample_sizes=up(w.sizes))
        dh.draw_heatmap_batched(acc, rb(ci), rb(ri), 6.0, 1.0, labels=rb(up(w.labels)))
        torch.cuda.synchronize(dev)
        assert torch.equal(acc, heat)
        outs.append(heat.cpu())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
