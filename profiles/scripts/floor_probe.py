"""Structural floor of the FRESH renderer: the same launch with NO objects (all sample sizes 0: every unit is one run of
bulk stores from the zero tile), beside torch.zero_ and the real workload. Measurement script.
    python profiles/scripts/floor_probe.py cfg3 [B]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads
name = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else None
lib = _cabi.load(); dev = torch.device("cuda", 0)
w = workloads.make(name, B=B); cfg = w.cfg
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cf, bf, sizes, lab = d(w.centers_f), d(w.bboxes_f), d(w.sizes), d(w.labels)
zero_sizes = torch.zeros_like(sizes)
one = sizes.clamp(max=1)
nbuf = max(1, min(16, -(-(512 << 20) // w.out_bytes)))
heats = [torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=dev) for _ in range(nbuf)]
def ev(fn, n=200):
    for i in range(20): fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n): fn(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
def render(sz):
    return lambda i: _cabi.draw_from_boxes(lib, heats[i % nbuf], cf, bf, float(cfg.osf), sz, lab, ws=False)
print(f"{name} B={w.B} {w.out_bytes/1e6:.1f} MB  zero_ {ev(lambda i: heats[i % nbuf].zero_()):.1f} us | empty render {ev(render(zero_sizes)):.1f} us"
      f" | <=1 object per sample {ev(render(one)):.1f} us | real {ev(render(sizes)):.1f} us")
