"""Where the register-tile renderer starts to pay, per unit: n objects of radius r on every plane (232x400 planes, one class,
slot arrays padded to T = 256), FRESH and ACCUMULATE, timed with whatever selection rule the loaded library has.
Run twice on a measurement build: DH_G_DENSE_MIN=1 DH_G_DENSE_COST16=0 (every unit with an object: tiles) and
DH_G_DENSE_COST16=2000000000 (never). Measurement script.
    DH_B200_LIB=... DH_G_DENSE_COST16=... python profiles/scripts/dense_rule_probe.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi

lib = _cabi.load()
dev = torch.device("cuda", 0)
B, H, W, T = 512, 232, 400, 256
rng = np.random.default_rng(3)
heat = torch.empty((B, H, W), device=dev)
for r in (2, 6, 15, 40):
    for n in (4, 12, 32, 96, 256):
        centers = np.stack([rng.integers(0, W, (B, T)), rng.integers(0, H, (B, T))], 2).astype(np.int32)
        radii = np.full((B, T), r, np.int32)
        sizes = np.full((B,), n, np.int64)
        ci, ri, sz = (torch.from_numpy(a).to(dev) for a in (centers, radii, sizes))
        tf, ta = [], []
        for i in range(12):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record()
            _cabi.draw_batched(lib, heat, ci, ri, sz, None, mode=_cabi.DH_MODE_FRESH)
            e[1].record()
            heat.zero_()
            e[2].record()
            _cabi.draw_batched(lib, heat, ci, ri, sz, None, mode=_cabi.DH_MODE_ACCUMULATE)
            e[3].record()
            torch.cuda.synchronize()
            if i >= 2:
                tf.append(e[0].elapsed_time(e[1]))
                ta.append(e[2].elapsed_time(e[3]))
        d = 2 * r + 1
        # per unit (2 per plane, 116 rows): about n * (d + 116) / (H + d) objects land on it
        per_unit = n * min(1.0, (d + 116) / (H + d))
        print(json.dumps({"r": r, "n": n, "objects_per_unit": round(per_unit, 1),
                          "cost_over_area": round(per_unit * d * (d + 64) / (116 * W), 2),
                          "fresh_ms": round(float(np.median(tf)), 4), "accumulate_ms": round(float(np.median(ta)), 4)}), flush=True)
