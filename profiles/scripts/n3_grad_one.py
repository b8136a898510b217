"""A few forward+gradient launches of the fused focal loss on the cfg5 slice (3840 planes) — target for ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import workloads
from accvlab.draw_heatmap import b200
dev = torch.device("cuda", 0)
w = workloads.make("cfg5", B=384); cfg = w.cfg
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
sizes, lab = d(w.sizes), d(w.labels)
ci, ri = b200.radius_center(d(w.centers_f), d(w.bboxes_f), cfg.osf)
rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
pred = (torch.rand((w.B, cfg.C, cfg.H, cfg.W), device=dev) * 0.98 + 0.01).requires_grad_(True)
ts = []
for i in range(8):
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    pred.grad = None
    e0.record()
    loss, npos = b200.gaussian_focal_loss(pred, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
    e1.record()
    loss.backward()
    e2.record()
    torch.cuda.synchronize()
    ts.append((e0.elapsed_time(e1), e1.elapsed_time(e2)))
print("forward+grad kernel ms, backward ms:", ts[-3:])
