"""Drop-in (accumulate) path: zero_() + dh_draw_heatmap_batched(ACCUMULATE), CUDA events, median of 30. Also the draw
alone into zeros (re-zeroed outside the timed region). Measurement script.
    [DH_B200_LIB=...] python profiles/scripts/accum_bench.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads

lib = _cabi.load()
dev = torch.device("cuda", 0)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
for name, B in (("cfg5", 1536), ("cfg5", 384), ("cfg3", None), ("cfg2", None), ("cfg4", None)):
    w = workloads.make(name, B=B) if B else workloads.make(name)
    cfg = w.cfg
    cf, bf, sizes, lab = d(w.centers_f), d(w.bboxes_f), d(w.sizes), d(w.labels)
    ci, ri = _cabi.radius_center(lib, cf, bf, float(cfg.osf))
    heat = torch.zeros((w.B, cfg.C, cfg.H, cfg.W), device=dev)
    both, draw = [], []
    for i in range(35):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        heat.zero_()
        e[1].record()
        _cabi.draw_batched(lib, heat, ci, ri, sizes, lab, mode=_cabi.DH_MODE_ACCUMULATE)
        e[2].record()
        torch.cuda.synchronize()
        if i >= 5:
            both.append(e[0].elapsed_time(e[2]))
            draw.append(e[1].elapsed_time(e[2]))
    print(json.dumps({"workload": name, "B": w.B, "zero_plus_accumulate_ms": round(float(np.median(both)), 4),
                      "accumulate_only_ms": round(float(np.median(draw)), 4)}), flush=True)
