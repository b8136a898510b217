"""A few FRESH bf16 launches of the main renderer on the whole of cfg5 — target for ncu. Measurement script."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import workloads
from accvlab.draw_heatmap import b200
dev = torch.device("cuda", 0)
w = workloads.make("cfg5"); cfg = w.cfg
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cf, bf, sizes, lab = d(w.centers_f), d(w.bboxes_f), d(w.sizes), d(w.labels)
out = torch.empty((w.B, cfg.C, cfg.H, cfg.W), dtype=torch.bfloat16, device=dev)
for i in range(5):
    b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, labels=lab, out=out, fresh=True)
torch.cuda.synchronize()
print("ok")
