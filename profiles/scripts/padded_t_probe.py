"""Sparse data in widely padded slot arrays (e.g. CenterPoint's max_objs = 500 with ~50 real objects per sample): the same
objects rendered with T = 100 (as generated) and padded to T = 256 / 500, FRESH from boxes and ACCUMULATE. Shows what the
T-based choice of the object-heavy renderer costs or saves. Measurement script.
    [DH_B200_LIB=...] python profiles/scripts/padded_t_probe.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads

lib = _cabi.load()
dev = torch.device("cuda", 0)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

def pad(a, T):
    out = np.zeros((a.shape[0], T) + a.shape[2:], a.dtype)
    out[:, :a.shape[1]] = a
    return out

for name, B in (("cfg5", 384), ("cfg3", None), ("cfg4", None)):
    w = workloads.make(name, B=B) if B else workloads.make(name)
    cfg = w.cfg
    ref = None
    for T in sorted({w.T, 256, 500, 1024}):
        if T < w.T:
            continue
        cf, bf, lab, sizes = d(pad(w.centers_f, T)), d(pad(w.bboxes_f, T)), d(pad(w.labels, T)), d(w.sizes)
        heat = torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=dev)
        ci, ri = _cabi.radius_center(lib, cf, bf, float(cfg.osf))
        ts, ta = [], []
        for i in range(25):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record()
            _cabi.draw_from_boxes(lib, heat, cf, bf, float(cfg.osf), sizes, lab)
            e[1].record()
            torch.cuda.synchronize()
            if ref is None:
                ref = heat.clone()
            same = bool(torch.equal(ref, heat))
            heat.zero_()
            e[2].record()
            _cabi.draw_batched(lib, heat, ci, ri, sizes, lab, mode=_cabi.DH_MODE_ACCUMULATE)
            e[3].record()
            torch.cuda.synchronize()
            if i >= 5:
                ts.append(e[0].elapsed_time(e[1]))
                ta.append(e[2].elapsed_time(e[3]))
        print(json.dumps({"workload": name, "B": w.B, "T": T, "fresh_ms": round(float(np.median(ts)), 4),
                          "accumulate_ms": round(float(np.median(ta)), 4), "same_bits_as_unpadded": same}), flush=True)
