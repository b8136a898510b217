"""A few FRESH launches of dh_draw_gaussians on the cfg5 slice (3840 planes) — target for ncu. Measurement script."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import workloads
from accvlab.draw_heatmap import b200
from oracle import oracle as orc
dev = torch.device("cuda", 0)
w = workloads.make("cfg5", B=384); cfg = w.cfg
ci, ri = orc.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
radii_f = (ri.astype(np.float32) + np.float32(0.37)).astype(np.float32)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
sizes = d(w.sizes)
rb = lambda a: b200.RaggedBatch(d(a), sample_sizes=sizes)
C_, R_, S_, K_ = rb(ci), rb(radii_f), rb(w.labels), d(np.linspace(0.6, 1.0, cfg.C).astype(np.float32))
outs = [torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=dev) for _ in range(2)]
for i in range(6):
    b200.draw_gaussians(None, None, S_, C_, R_, K_, 1.0 / 3.0, out=outs[i % 2])
torch.cuda.synchronize()
print("ok")
