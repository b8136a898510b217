"""Ours vs the UNMODIFIED reference CUDA extension (oracle/_ref/dh_ref_fast.so, its default --use_fast_math build) on
the BASELINE configs, same device, same inputs. Measurement script (SURVEY.md 8(d) "Reference GPU timed beside it").
Reference cost of a fresh heatmap = torch.zeros_ + draw_heatmap_batched_classwise_impl (its API accumulates into a
caller-zeroed tensor); ours = one FRESH launch with K1 fused, and the drop-in accumulate path (zero_ + ours) as well.
cfg5 runs on a 1536-sample slice: the reference indexes pixels with int32 (overflows at B*C*H*W >= 2^31)."""
import importlib.util, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import torch
import workloads
from accvlab.draw_heatmap import draw_heatmap_ext as ours
from accvlab.draw_heatmap import b200

def load_ref(name):
    path = os.path.join(ROOT, "oracle", "_ref", name + ".so")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    return mod

def ev(fn, n):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

ref = load_ref("dh_ref_fast")
dev = torch.device("cuda", 0)
rows = []
for name, nb, reps in (("cfg1", None, 300), ("cfg2", None, 300), ("cfg3", None, 200), ("cfg4", None, 30), ("cfg5", 1536, 10)):
    w = workloads.make(name, B=nb); cfg = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    lab, s64 = torch.from_numpy(w.labels).to(dev), torch.from_numpy(w.sizes).to(dev)
    s32 = s64.int()
    ci, ri = ours.radius_center(cf, bf, float(cfg.osf))
    out_bytes = w.B * cfg.C * cfg.H * cfg.W * 4
    nbuf = max(1, min(64, -(-(512 << 20) // out_bytes)))
    bufs = [torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=dev) for _ in range(nbuf)]
    it = [0]
    def nxt():
        it[0] += 1; return bufs[it[0] % nbuf]
    def f_ref():
        h = nxt(); h.zero_(); ref.draw_heatmap_batched_classwise_impl(h, ci, ri, s32, lab, 6.0, 1.0)
    def f_ref_kernel_only():
        ref.draw_heatmap_batched_classwise_impl(nxt(), ci, ri, s32, lab, 6.0, 1.0)
    def f_ours_dropin():
        h = nxt(); h.zero_(); ours.draw_heatmap_batched_classwise_impl(h, ci, ri, s32, lab, 6.0, 1.0)
    def f_ours_fresh():
        b200.render_heatmap_from_boxes(cf, bf, float(cfg.osf), s64, labels=lab, out=nxt(), fresh=True)
    r = {"config": name, "samples": w.B, "planes": w.planes, "out_MB": out_bytes / 1e6,
         "ref_zeros_plus_draw_ms": ev(f_ref, reps), "ref_draw_only_ms": ev(f_ref_kernel_only, reps),
         "ours_zero_plus_accumulate_ms": ev(f_ours_dropin, reps), "ours_fresh_fused_ms": ev(f_ours_fresh, reps)}
    r["speedup_fresh_vs_ref"] = r["ref_zeros_plus_draw_ms"] / r["ours_fresh_fused_ms"]
    r["speedup_dropin_vs_ref"] = r["ref_zeros_plus_draw_ms"] / r["ours_zero_plus_accumulate_ms"]
    rows.append(r); print(json.dumps(r), flush=True)

# ---- flat entry (draw_heatmap: concatenated objects + plane index), accumulate semantics like the reference
for name, nb, reps in (("cfg2", None, 200), ("cfg3", None, 100), ("cfg5", 1536, 10)):
    w = workloads.make(name, B=nb); cfg = w.cfg
    cf, bf = torch.from_numpy(w.centers_f).to(dev), torch.from_numpy(w.bboxes_f).to(dev)
    lab, s64 = torch.from_numpy(w.labels).to(dev), torch.from_numpy(w.sizes).to(dev)
    ci, ri = ours.radius_center(cf, bf, float(cfg.osf))
    valid = torch.arange(w.T, device=dev)[None, :] < s64[:, None]
    idx = (torch.arange(w.B, device=dev, dtype=torch.int32)[:, None] * cfg.C + lab)[valid].contiguous()
    cflat, rflat = ci[valid].contiguous(), ri[valid].contiguous()
    M = w.B * cfg.C
    out_bytes = M * cfg.H * cfg.W * 4
    nbuf = max(1, min(64, -(-(512 << 20) // out_bytes)))
    bufs = [torch.empty((M, cfg.H, cfg.W), device=dev) for _ in range(nbuf)]
    it = [0]
    def nxt():
        it[0] += 1; return bufs[it[0] % nbuf]
    def f_ref():
        h = nxt(); h.zero_(); ref.draw_heatmap(h, cflat, rflat, idx, 6.0, 1.0)
    def f_ours():
        h = nxt(); h.zero_(); ours.draw_heatmap(h, cflat, rflat, idx, 6.0, 1.0)
    def f_ours_fresh():
        b200.render_heatmap_flat(cflat, rflat, idx, M, cfg.H, cfg.W, out=nxt())
    r = {"entry": "flat", "config": name, "objects": int(idx.numel()), "planes": M, "ref_zeros_plus_draw_ms": ev(f_ref, reps),
         "ours_zero_plus_accumulate_ms": ev(f_ours, reps), "ours_fresh_ms": ev(f_ours_fresh, reps)}
    r["speedup_dropin_vs_ref"] = r["ref_zeros_plus_draw_ms"] / r["ours_zero_plus_accumulate_ms"]
    r["speedup_fresh_vs_ref"] = r["ref_zeros_plus_draw_ms"] / r["ours_fresh_ms"]
    print(json.dumps(r), flush=True)
