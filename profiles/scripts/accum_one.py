"""One drop-in (accumulate) call pattern for ncu: zero_() + dh_draw_heatmap_batched(ACCUMULATE) on the first 1536 samples of
cfg5, a few times. Measurement script."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads
lib = _cabi.load()
dev = torch.device("cuda", 0)
w = workloads.make("cfg5", B=1536); cfg = w.cfg
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cf, bf, sizes, lab = d(w.centers_f), d(w.bboxes_f), d(w.sizes), d(w.labels)
ci, ri = _cabi.radius_center(lib, cf, bf, float(cfg.osf))
heat = torch.zeros((w.B, cfg.C, cfg.H, cfg.W), device=dev)
for i in range(5):
    heat.zero_()
    _cabi.draw_batched(lib, heat, ci, ri, sizes, lab, mode=_cabi.DH_MODE_ACCUMULATE)
torch.cuda.synchronize()
print("touched", int(_cabi.checksum(lib, heat)[2]))
