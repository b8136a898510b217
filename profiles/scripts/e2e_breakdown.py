"""Where the end-to-end step of cfg5 spends its time (measurement script, not part of the product)."""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "accv-lab_b200")]
import numpy as np, torch
import _cabi, workloads

B = int(sys.argv[1]) if len(sys.argv) > 1 else None
lib = _cabi.load()
dev = torch.device("cuda", 0)
w = workloads.make("cfg5", B=B)
cfg = w.cfg
nb, T, C, H, W = w.B, w.T, cfg.C, cfg.H, cfg.W
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
h = [pin(a) for a in (w.centers_f, w.bboxes_f, w.sizes, w.labels)]
d = [t.to(dev) for t in h]
heat = torch.empty((nb, C, H, W), device=dev)
stream = torch.cuda.current_stream()
sp = ctypes.c_void_p(stream.cuda_stream)
ckd = torch.zeros(3, dtype=torch.int64, device=dev)
ck = torch.zeros(3, dtype=torch.int64).pin_memory()
nbytes = lib.dh_host_scratch_bytes(nb, T, C, H, W, 1, 0, 0)
scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)

def plain():
    lib.dh_draw_heatmap_batched_from_boxes(sp, _cabi.ptr(heat), _cabi.ptr(d[0]), _cabi.ptr(d[1]), float(cfg.osf), _cabi.ptr(d[2]),
        _cabi.DH_SIZES_I64, _cabi.ptr(d[3]), nb, T, C, H, W, 6.0, 1.0, _cabi.DH_MODE_FRESH, None, None, None)
def digest():
    lib.dh_draw_heatmap_batched_from_boxes_digest(sp, _cabi.ptr(heat), _cabi.ptr(d[0]), _cabi.ptr(d[1]), float(cfg.osf), _cabi.ptr(d[2]),
        _cabi.DH_SIZES_I64, _cabi.ptr(d[3]), nb, T, C, H, W, 6.0, 1.0, 0, _cabi.ptr(ckd), None)
def h2d():
    for a, b in zip(d, h):
        a.copy_(b, non_blocking=True)
def host(chunk, with_ck=True):
    def f():
        st = lib.dh_draw_heatmap_batched_from_boxes_host(sp, None, _cabi.ptr(heat), _cabi.ptr(ck) if with_ck else None, _cabi.ptr(h[0]), _cabi.ptr(h[1]),
            float(cfg.osf), _cabi.ptr(h[2]), _cabi.ptr(h[3]), nb, T, C, H, W, 6.0, 1.0, _cabi.ptr(scratch), nbytes, chunk)
        assert st == 0, st
    return f

def ev(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(n): fn()
    b.record(stream); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
def wall(fn, n=50):
    for _ in range(5): fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn(); stream.synchronize()
    return (time.perf_counter() - t0) * 1e3 / n

print(f"B={nb} inputs {sum(t.numel()*t.element_size() for t in h)/1e6:.2f} MB")
print(f"plain render  (events) {ev(plain):.4f} ms   wall+sync {wall(plain):.4f}")
print(f"digest render (events) {ev(digest):.4f} ms   wall+sync {wall(digest):.4f}")
print(f"h2d 4 arrays  (events) {ev(h2d):.4f} ms   wall+sync {wall(h2d):.4f}")
for chunk in (-1, 0, nb // 2, nb // 4, nb // 8, nb // 16):
    print(f"host entry chunk_samples={chunk:5d}: digest wall+sync {wall(host(chunk)):.4f} ms | no digest {wall(host(chunk, False)):.4f} ms")
