"""Per-unit timeline of the groups kernel (profiling build with -DDH_TRACE_UNITS, see the nvcc line in
profiles/r01/experiments.md): where a launch loses time to ramp-up and tail. Measurement script.
    DH_B200_LIB=profiles/trace_build/libdh_b200_trace.so python profiles/scripts/unit_trace.py cfg5 384"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads

name = sys.argv[1] if len(sys.argv) > 1 else "cfg5"
B = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2] != "-" else None
MODE = _cabi.DH_MODE_ACCUMULATE if len(sys.argv) > 3 and sys.argv[3] == "accum" else _cabi.DH_MODE_FRESH  # accum: re-draw
lib = _cabi.load()
lib.dh_debug_set_trace.argtypes = [ctypes.c_void_p, ctypes.c_int64]
dev = torch.device("cuda", 0)
w = workloads.make(name, B=B); cfg = w.cfg
d = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (w.centers_f, w.bboxes_f, w.sizes, w.labels)]
heat = [torch.empty((w.B, cfg.C, cfg.H, cfg.W), device=dev) for _ in range(2)]
NU = 1 << 20
trace = torch.zeros(NU * 4, dtype=torch.int64, device=dev)
sp = _cabi.stream_ptr()

WSB = int(lib.dh_batched_workspace_bytes(w.B, w.T, cfg.C))
WS = torch.empty(max(WSB, 16), dtype=torch.uint8, device=dev)

def run(i):
    st = lib.dh_draw_heatmap_batched_from_boxes_ws(sp, _cabi.ptr(heat[i % 2]), _cabi.ptr(d[0]), _cabi.ptr(d[1]), float(cfg.osf),
        _cabi.ptr(d[2]), _cabi.DH_SIZES_I64, _cabi.ptr(d[3]), w.B, w.T, cfg.C, cfg.H, cfg.W, 6.0, 1.0, MODE, None, None, None,
        _cabi.ptr(WS), WSB)  # DH_G_NOBIN=1 in the environment selects the single-launch renderer
    assert st == 0

for h in heat: h.zero_()
for i in range(5): run(i)
if MODE == _cabi.DH_MODE_ACCUMULATE: heat[0].zero_()  # the traced launch draws into zeros, like the drop-in call
torch.cuda.synchronize()
assert lib.dh_debug_set_trace(_cabi.ptr(trace), NU) == 0
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); run(0); b.record(); torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(NU, 4)
n = int((t[:, 0] != 0).sum())
t = t[:n]
t0 = t[:, 0].min()
start, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
dur = end - start
ucnt, ngroups = t[:, 3] & 0xffffffff, t[:, 3] >> 32
print(f"{name} B={w.B}: {n} units, kernel (events) {a.elapsed_time(b)*1e3:.1f} us, first start 0, last start {start.max():.1f}, last end {end.max():.1f} us")
print(f"unit duration us: mean {dur.mean():.2f} p50 {np.percentile(dur,50):.2f} p90 {np.percentile(dur,90):.2f} p99 {np.percentile(dur,99):.2f} max {dur.max():.2f}")
T = end.max()
edges = np.linspace(0, T, 41)
act = [(np.minimum(end, hi) - np.maximum(start, lo)).clip(0).sum() / (hi - lo) for lo, hi in zip(edges[:-1], edges[1:])]
print("active units per 1/40 of the launch:", " ".join(f"{x:.0f}" for x in act))
first = np.sort(start)[: 444]
print(f"first-wave starts: min {first.min():.2f} max {first.max():.2f} us; first-wave durations mean {dur[np.argsort(start)[:444]].mean():.2f}")
last = np.argsort(end)[-10:]
print("last 10 units to end: (start, dur, objects, touched groups)", [(round(float(start[i]), 1), round(float(dur[i]), 1), int(ucnt[i]), int(ngroups[i])) for i in last])
big = np.argsort(dur)[-5:]
print("5 longest units: (start, dur, objects, touched groups)", [(round(float(start[i]), 1), round(float(dur[i]), 1), int(ucnt[i]), int(ngroups[i])) for i in big])
for lo_, hi_ in ((0, 1), (1, 4), (4, 16), (16, 1 << 20)):
    m = (ucnt >= lo_) & (ucnt < hi_)
    if m.any(): print(f"units with {lo_}..{hi_-1} objects: {int(m.sum())}, mean duration {dur[m].mean():.2f} us")
if os.environ.get("DH_TRACE_SAVE"):  # raw (start ns, end ns, SM, objects | touched groups << 32) per unit, for offline analysis
    np.save(os.environ["DH_TRACE_SAVE"], t)
