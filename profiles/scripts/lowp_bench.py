"""fp32 vs fp16 / bf16 heatmap planes (N4): FRESH render from float boxes, CUDA events, median of 30 launches after 5
warm-ups. Output buffers are far larger than L2 (cfg5: 11.4 / 5.7 GB; cfg3 is rotated over 8 buffers). Prints one
JSON line per (workload, dtype). Measurement script."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import workloads
from accvlab.draw_heatmap import b200

dev = torch.device("cuda", 0)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
for name, B, nbuf in (("cfg5", None, 1), ("cfg5", 384, 2), ("cfg3", None, 8), ("cfg4", None, 8)):
    w = workloads.make(name, B=B) if B else workloads.make(name)
    cfg = w.cfg
    cf, bf, sizes, lab = d(w.centers_f), d(w.bboxes_f), d(w.sizes), d(w.labels)
    for dt in (torch.float32, torch.float16, torch.bfloat16):
        outs = [torch.empty((w.B, cfg.C, cfg.H, cfg.W), dtype=dt, device=dev) for _ in range(nbuf)]
        ts = []
        for i in range(35):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, labels=lab, out=outs[i % nbuf], fresh=True)
            e1.record()
            torch.cuda.synchronize()
            if i >= 5:
                ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        nbytes = outs[0].numel() * outs[0].element_size()
        print(json.dumps({"workload": name, "B": w.B, "dtype": str(dt).split(".")[-1], "ms": round(ms, 4),
                          "min_ms": round(min(ts), 4), "out_GB": round(nbytes / 1e9, 3),
                          "write_GBps": round(nbytes / ms / 1e6, 1), "planes_per_s": round(w.planes / ms * 1e3)}))
        del outs
        torch.cuda.empty_cache()
