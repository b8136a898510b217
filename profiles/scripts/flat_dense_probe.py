"""Flat entry (dh_draw_heatmap: concatenated objects + plane index) on BASELINE cfg4 and cfg3: ms per call, FRESH and
ACCUMULATE. Measurement script."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads
lib = _cabi.load()
dev = torch.device("cuda", 0)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
for name in ("cfg4", "cfg3"):
    w = workloads.make(name); c = w.cfg
    ci, ri = _cabi.radius_center(lib, d(w.centers_f), d(w.bboxes_f), float(c.osf))
    ci, ri = ci.cpu().numpy(), ri.cpu().numpy()
    n = [int(s) for s in w.sizes]
    planes = [b * c.C + int(w.labels[b, t]) for b in range(w.B) for t in range(n[b])]
    cen = np.concatenate([ci[b, :n[b]] for b in range(w.B)]); rad = np.concatenate([ri[b, :n[b]] for b in range(w.B)])
    cen_d, rad_d, idx_d = d(cen.astype(np.int32)), d(rad.astype(np.int32)), d(np.array(planes, np.int32))
    heat = torch.zeros((w.B * c.C, c.H, c.W), device=dev)
    res = {}
    for mode, label in ((_cabi.DH_MODE_FRESH, "fresh_ms"), (_cabi.DH_MODE_ACCUMULATE, "accumulate_ms")):
        ts = []
        for i in range(25):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); _cabi.draw_flat(lib, heat, cen_d, rad_d, idx_d, 6.0, 1.0, mode=mode); e1.record()
            torch.cuda.synchronize()
            if i >= 5: ts.append(e0.elapsed_time(e1))
        res[label] = round(float(np.median(ts)), 4)
    print(json.dumps({"workload": name, "objects": len(planes), "planes": w.B * c.C, **res}), flush=True)
