"""Measurement of the two "next" rows on one B200. Measurement script, not part of the product.

N1  per-sample box lists -> RaggedBatch (+K1): ours (pack_radius_center + combine_data: 2 launches + 2 small H2D) vs the
    reference's way (torch.zeros + one slice-assign per sample and array, DH/examples/draw_heatmap_batched.py:42-50,
    then the 8-op torch chain of get_centers_and_radii, DH/tests/_test_helpers.py:20-28), same GPU, host wall clock.
N2  DALI draw_gaussians semantics: dh_draw_gaussians (fresh / out-of-place accumulate) vs the reference operator's CPU
    arithmetic (oracle port, bit-identical to DrawGaussians.cc; 1 thread) on a bounded sample."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import workloads
from accvlab.draw_heatmap import b200
from oracle import oracle as orc

dev = torch.device("cuda", 0)
ROWS = os.environ.get("DH_ROWS", "N1,N2,N3").split(",")  # e.g. DH_ROWS=N3 for one row

def wall(fn, n):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) * 1e3 / n

def ev(fn, n):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

# ------------------------------------------------------------------ N1
for name, B in ((("cfg2", None), ("cfg3", None), ("cfg5", 3072)) if "N1" in ROWS else ()):
    w = workloads.make(name, B=B); cfg = w.cfg
    cf, bf, lab = (torch.from_numpy(a).to(dev) for a in (w.centers_f, w.bboxes_f, w.labels))
    cl = [cf[i, : w.sizes[i]] for i in range(w.B)]
    bl = [bf[i, : w.sizes[i]] for i in range(w.B)]
    ll = [lab[i, : w.sizes[i]] for i in range(w.B)]
    T = w.T
    def ref_way():
        c = torch.zeros((w.B, T, 2), dtype=torch.float32, device=dev)
        b = torch.zeros((w.B, T, 4), dtype=torch.float32, device=dev)
        l = torch.zeros((w.B, T), dtype=torch.int32, device=dev)
        n = torch.zeros((w.B,), dtype=torch.int32, device=dev)
        for i in range(w.B):
            k = len(cl[i]); c[i, :k] = cl[i]; b[i, :k] = bl[i]; l[i, :k] = ll[i]; n[i] = k
        tl = c - b[..., :2]; br = b[..., 2:] - c                      # get_centers_and_radii as torch ops
        r = torch.ceil(torch.cat([tl, br], -1).min(-1).values / cfg.osf).to(torch.int32); r[r < 1] = 1
        return (c / cfg.osf).to(torch.int32), r, l, n
    def ours():
        crb, rrb = b200.pack_radius_center(cl, bl, cfg.osf, max_sample_size=T)
        lrb = b200.combine_data(ll, other_with_same_sample_sizes=crb, max_sample_size=T)
        return crb, rrb, lrb
    a, b_ = ref_way(), ours()
    valid = torch.arange(T, device=dev)[None] < b_[0].sample_sizes[:, None]
    assert torch.equal(a[0][valid], b_[0].tensor[valid]) and torch.equal(a[1][valid], b_[1].tensor[valid])
    reps = 3 if w.B > 1000 else 30
    r = {"row": "N1", "config": name, "samples": w.B, "boxes": int(w.sizes.sum()),
         "reference_style_loop_ms": wall(ref_way, reps), "ours_ms": wall(ours, reps * 5)}
    r["speedup"] = r["reference_style_loop_ms"] / r["ours_ms"]
    print(json.dumps(r), flush=True)

# ------------------------------------------------------------------ N2
for name, B in ((("cfg2", None), ("cfg3", None), ("cfg5", 384)) if "N2" in ROWS else ()):
    w = workloads.make(name, B=B); cfg = w.cfg
    ci, ri = orc.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    radii_f = (ri.astype(np.float32) + np.float32(0.37)).astype(np.float32)   # float radii, fractional
    k = np.linspace(0.6, 1.0, cfg.C).astype(np.float32)
    factor = 1.0 / 3.0
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    sizes = d(w.sizes)
    rb = lambda a: b200.RaggedBatch(d(a), sample_sizes=sizes)
    C_, R_, S_, K_ = rb(ci), rb(radii_f), rb(w.labels), d(k)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    out_bytes = int(np.prod(shape)) * 4
    nbuf = max(2, min(32, -(-(512 << 20) // out_bytes)))
    outs = [torch.empty(shape, device=dev) for _ in range(nbuf)]
    bg = torch.zeros(shape, device=dev)
    it = [0]
    def fresh():
        it[0] += 1; b200.draw_gaussians(None, None, S_, C_, R_, K_, factor, out=outs[it[0] % nbuf])
    def accum():
        it[0] += 1; b200.draw_gaussians(bg, None, S_, C_, R_, K_, factor, out=outs[it[0] % nbuf])
    reps = 20 if out_bytes > (1 << 30) else 200
    t_f, t_a = ev(fresh, reps), ev(accum, reps)
    # CPU: the reference operator's arithmetic (oracle port == DrawGaussians.cc bit for bit), one thread, bounded sample
    nb = min(w.B, 48)
    want = np.zeros((nb,) + shape[1:], np.float32)
    t0 = time.perf_counter()
    orc.draw_gaussians(want, ci[:nb], radii_f[:nb], w.sizes[:nb], k, factor, w.labels[:nb], None)
    t_cpu = (time.perf_counter() - t0) * 1e3 * (w.B / nb)
    got = outs[it[0] % nbuf]; fresh(); torch.cuda.synchronize()
    got = outs[it[0] % nbuf][:nb].cpu().numpy()
    assert np.array_equal(got > 0, want > 0) and float(np.abs(got - want).max()) <= 1e-6
    r = {"row": "N2", "config": name, "samples": w.B, "planes": w.planes, "out_MB": out_bytes / 1e6,
         "gpu_fresh_ms": t_f, "gpu_fresh_GBps": out_bytes / t_f / 1e6, "gpu_accumulate_out_of_place_ms": t_a,
         "gpu_accumulate_GBps_rw": 2 * out_bytes / t_a / 1e6, "cpu_reference_arithmetic_1thread_ms_scaled": t_cpu,
         "cpu_sample": f"first {nb} samples, scaled to {w.B}", "speedup_fresh_vs_cpu_1thread": t_cpu / t_f}
    print(json.dumps(r), flush=True)

# ------------------------------------------------------------------ N3
for name, B in ((("cfg3", None), ("cfg5", 384)) if "N3" in ROWS else ()):
    w = workloads.make(name, B=B); cfg = w.cfg
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    sizes, lab = d(w.sizes), d(w.labels)
    ci, ri = b200.radius_center(d(w.centers_f), d(w.bboxes_f), cfg.osf)
    rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    pred = torch.rand(shape, device=dev) * 0.98 + 0.01
    px_bytes = int(np.prod(shape)) * 4
    def fused_fwd():
        b200.gaussian_focal_loss(pred, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
    def fused_fwd_fast():
        b200.gaussian_focal_loss(pred, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab), fast_math=True)
    pg = pred.clone().requires_grad_(True)
    def fused_fwd_bwd():
        pg.grad = None
        loss, npos = b200.gaussian_focal_loss(pg, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
        loss.backward()
    tgt_buf = torch.empty(shape, device=dev)
    def unfused_fwd_bwd():
        pg.grad = None
        t = b200.render_heatmap_batched(rb(ci), rb(ri), labels=rb(lab), out=tgt_buf)          # our fast renderer
        pos = t.eq(1)
        l = torch.where(pos, -(pg + 1e-12).log() * (1 - pg) ** 2, -(1 - pg + 1e-12).log() * pg ** 2 * (1 - t) ** 4).sum()
        l.backward()
    reps = 10
    r = {"row": "N3", "config": name, "samples": w.B, "planes": w.planes, "pred_MB": px_bytes / 1e6,
         "fused_forward_ms": ev(fused_fwd, reps), "fused_forward_fast_math_ms": ev(fused_fwd_fast, reps), "fused_forward_backward_ms": ev(fused_fwd_bwd, reps),
         "unfused_render_plus_torch_loss_fwd_bwd_ms": ev(unfused_fwd_bwd, 3)}
    r["fused_forward_GBps_read"] = px_bytes / r["fused_forward_ms"] / 1e6
    r["speedup_fwd_bwd"] = r["unfused_render_plus_torch_loss_fwd_bwd_ms"] / r["fused_forward_backward_ms"]
    print(json.dumps(r), flush=True)
