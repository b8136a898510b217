"""A few launches of the fused focal loss on the cfg5 slice (3840 planes) — target for ncu. Measurement script."""
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
pred = torch.rand((w.B, cfg.C, cfg.H, cfg.W), device=dev) * 0.98 + 0.01
for i in range(5):
    loss, npos = b200.gaussian_focal_loss(pred, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
torch.cuda.synchronize()
print("ok", float(loss), int(npos))
