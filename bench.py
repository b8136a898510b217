#!/usr/bin/env python
"""bench.py — Gaussian heatmaps/sec + achieved HBM GB/s of the draw_heatmap hot path on 1/2/4/8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg5] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A "step" is one pass of the hot path over one batch of synthetic boxes: float boxes [B,T] (RaggedBatch layout) ->
radius/centre (K1) -> binning (K2) -> max-blend (K3) -> single write of the fresh heatmap [B,C,H,W] (K4), ONE launch
of dh_draw_heatmap_batched_from_boxes through the C ABI (include/dh_b200.h). Default workload: BASELINE.json
configs[4] "cfg5" (batch 512x6 cams, 10 classes, 232x400: 30720 planes, 11.4 GB of output), the configuration the
metric's 1/2/4/8-GPU sweep is quoted on; it fits one GPU, and with N ranks each rank renders its contiguous batch
slice (strong scaling, no collective on the timed path; NCCL only for the max-over-ranks time and an untimed digest).

Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "accv-lab_b200"))

import numpy as np  # noqa: E402

import workloads  # noqa: E402

FACTOR, K_SCALE = 6.0, 1.0
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="cfg5", choices=sorted(workloads.CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="override the sample count of the workload")
    ap.add_argument("--cpu-seconds", type=float, default=8.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--two-launch", action="store_true", help="A/B: K2 pre-pass + list-fed renderer (the *_ws entries)")
    ap.add_argument("--no-side-legs", action="store_true", help="skip cfg1-4 / accumulate / reference-GPU side records")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, read+write copy)", d
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)", {}


class ClockSampler(threading.Thread):
    """Polls SM clock + throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v for v in vis.split(",") if v.strip() != ""]
                if index < len(ids) and ids[index].strip().isdigit():
                    phys = int(ids[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # no NVML: clocks are reported as unavailable, the run is not invalidated here
            self.err = repr(e)

    def run(self):
        while self.ok and not self._stop_evt.is_set():
            try:
                self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                bits = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.004)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable or no samples"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------------------
def cpu_reference_leg(w, seconds: float, threads: int, max_samples: int = 384):
    """Time the reference's CPU path (C port `dho_refcpu_draw_batched` of tests/_gaussian_drawing.py:215-240 +
    tests/_test_helpers.py:41-74, including the per-sample zeroing its drivers do with torch.zeros) on a bounded
    sample of the workload: the first `max_samples` samples, repeated until `seconds` have passed (>= 1 pass).
    Radius/centre integers come from the oracle's K1 (not timed on either arm's CPU leg: the reference's CPU drivers
    take integer centres/radii as inputs too)."""
    from oracle import build_oracle, oracle as orc  # the ONLY use of oracle/ in bench.py: as the timed CPU baseline

    build_oracle.build_c_oracle()
    cfg = w.cfg
    nb = min(w.B, max_samples)
    s = w.slice(0, nb)
    ci, ri = orc.radius_center(s.centers_f, s.bboxes_f, cfg.osf)
    heat = np.empty((nb, cfg.C, cfg.H, cfg.W), np.float32)
    heat.fill(1.0)  # touch the pages before timing
    passes, t0 = 0, time.perf_counter()
    while True:
        orc.refcpu_draw_batched(heat, ci, ri, s.sizes, FACTOR, K_SCALE, s.labels, zero_first=True, nthreads=threads)
        passes += 1
        dt = time.perf_counter() - t0
        if dt >= seconds or passes >= 50:
            break
    planes = nb * cfg.C * passes
    return {"value": planes / dt, "unit": "heatmaps/s", "cores": threads, "kind": "port",
            "sample": f"first {nb} of {w.B} samples of {cfg.name} ({nb * cfg.C} planes {cfg.H}x{cfg.W}), "
                      f"{passes} passes in {dt:.2f} s, {threads} pthreads over samples, C port of the reference's "
                      f"float64 CPU drawer incl. its per-sample zero fill",
            "ms_per_pass": 1e3 * dt / passes, "planes_per_pass": nb * cfg.C}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores (oracle port; the
    reference's CPU path is Python/numpy under /root/reference, which does not exist on the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = workloads.make(args.workload, B=args.batch)
    threads = os.cpu_count() or 1
    try:
        threads = len(os.sched_getaffinity(0))
    except Exception:
        pass
    cfg = w.cfg
    nb = min(w.B, 384)
    warm = max(args.warmup, 1)
    # one "step" = one pass over the bounded sample
    leg = cpu_reference_leg(w, seconds=0.0, threads=threads, max_samples=nb)  # warm-up pass (page faults, caches)
    for _ in range(warm - 1):
        cpu_reference_leg(w, seconds=0.0, threads=threads, max_samples=nb)
    steps = max(1, min(args.steps, 20))
    t0 = time.perf_counter()
    from oracle import oracle as orc

    s = w.slice(0, nb)
    ci, ri = orc.radius_center(s.centers_f, s.bboxes_f, cfg.osf)
    heat = np.empty((nb, cfg.C, cfg.H, cfg.W), np.float32)
    heat.fill(1.0)
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.refcpu_draw_batched(heat, ci, ri, s.sizes, FACTOR, K_SCALE, s.labels, zero_first=True, nthreads=threads)
    dt = time.perf_counter() - t0
    value = nb * cfg.C * steps / dt
    line = {
        "impl": "reference", "metric": "gaussian_heatmaps_per_sec", "value": value, "unit": "heatmaps/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": 1e3 * dt / steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64->f32", "data": "synthetic",
        "config": workload_config(w, args.gpus, note=f"each step = first {nb} samples ({nb * cfg.C} planes)"),
        "cpu_baseline": {"value": value, "unit": "heatmaps/s", "cores": threads, "kind": "port",
                         "sample": leg["sample"].split(",")[0] + f", {steps} timed passes"},
        "e2e": {"value": value, "unit": "heatmaps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(w, n_gpus, note=""):
    cfg = w.cfg
    return {"workload": f"{cfg.name}: {cfg.describe}", "samples": w.B, "classes": cfg.C, "height": cfg.H, "width": cfg.W,
            "max_targets": w.T, "objects": int(w.sizes.sum()), "planes": w.planes, "output_bytes": w.out_bytes,
            "diameter_to_sigma_factor": FACTOR, "k_scale": K_SCALE, "mode": "fresh (zero-init fused), K1 fused",
            "parallelism": f"batch sharded over {n_gpus} GPU(s), no collective on the timed path",
            "l2": "per-rank output per step is larger than the 126 MB L2 or rotated over buffers totalling > 512 MB",
            "note": note}


# ------------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch

    import _cabi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        # NCCL writes its version banner (any NCCL_DEBUG level >= VERSION) to STDOUT, where the contract is ONE JSON
        # line: point fd 1 at stderr while the communicator comes up, then restore it.
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    lib = _cabi.load()  # raises if libdh_b200.so is missing: no fallback

    w_all = workloads.make(args.workload, B=args.batch)
    cfg = w_all.cfg
    lo, hi = workloads.shard_range(w_all.B, rank, world)
    w = w_all.slice(lo, hi)
    nb, T, C, H, W = w.B, w.T, cfg.C, cfg.H, cfg.W
    sample_elems = C * H * W

    # ---- device-resident inputs (for `value`) and pinned host inputs (for `e2e`) ----
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    h_cf, h_bf, h_sz, h_lab = pin(w.centers_f), pin(w.bboxes_f), pin(w.sizes), pin(w.labels)
    d_cf, d_bf, d_sz, d_lab = (t.to(dev) for t in (h_cf, h_bf, h_sz, h_lab))
    out_bytes = nb * sample_elems * 4
    nbuf = 1 if out_bytes >= (256 << 20) else int(min(64, -(-(512 << 20) // max(out_bytes, 1))))
    heats = [torch.empty((nb, C, H, W), dtype=torch.float32, device=dev) for _ in range(nbuf)]
    stream = torch.cuda.current_stream()
    sp = ctypes.c_void_p(stream.cuda_stream)

    # --two-launch: the *_ws entry with a caller-held bin-list workspace (K2 pre-pass + list-fed renderer) instead of the
    # single-launch renderer; A/B measurements only (profiles/r02/experiments.md)
    ws_bytes = int(lib.dh_batched_workspace_bytes(nb, T, C)) if args.two_launch else 0
    ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=dev) if args.two_launch else None

    def step(i, sp=sp):
        st = lib.dh_draw_heatmap_batched_from_boxes_ws(sp, _cabi.ptr(heats[i % nbuf]), _cabi.ptr(d_cf), _cabi.ptr(d_bf),
                                                       float(cfg.osf), _cabi.ptr(d_sz), _cabi.DH_SIZES_I64,
                                                       _cabi.ptr(d_lab), nb, T, C, H, W, FACTOR, K_SCALE,
                                                       _cabi.DH_MODE_FRESH, None, None, None, _cabi.ptr(ws), ws_bytes)
        if st != 0:
            raise RuntimeError(f"dh_draw_heatmap_batched_from_boxes: {lib.dh_status_string(st).decode()}")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for i in range(warm):
        step(i)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = lib.dh_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(args.steps):
        step(i)
    e1.record(stream)
    barrier()
    launches = int(lib.dh_launch_count() - launches0)
    ms_total = e0.elapsed_time(e1)
    if sampler and len(sampler.samples) < 8:  # very short timed region: keep the GPU under the same load a bit longer
        t_end = time.perf_counter() + 0.25
        while time.perf_counter() < t_end:
            step(0)
            torch.cuda.synchronize()
    if sampler:
        sampler.stop()
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total_max = float(t.item())
    ms_step = ms_total_max / args.steps
    planes_total = w_all.planes
    value = planes_total / (ms_step * 1e-3)

    # ---- launch-bound configs: the same steps replayed from a CUDA graph (one graph = one pass over
    # the rotating buffers), so the per-launch host cost drops out; reported beside `value`, never instead of it
    graph_ms = None
    if rank == 0 and out_bytes < (256 << 20):
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(stream)
            with torch.cuda.graph(g, stream=side):
                csp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
                for i in range(nbuf):
                    step(i, csp)
            reps = max(3, args.steps // nbuf)
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                g.replay()
            b.record()
            torch.cuda.synchronize()
            graph_ms = a.elapsed_time(b) / (reps * nbuf)
        except Exception as e:  # measurement extra only
            print(f"[bench] CUDA-graph replay skipped: {e!r}", file=sys.stderr)

    # ---- untimed digest of the result: additive over shards, so every N must print the same three numbers ----
    digest = torch.zeros(3, dtype=torch.int64, device=dev)
    _cabi.checksum(lib, heats[(args.steps - 1) % nbuf], lo * sample_elems, digest)
    if dist is not None:
        dist.all_reduce(digest, op=dist.ReduceOp.SUM)  # NCCL over NVLink; wrapping int64 == wrapping uint64 adds
    digest = [int(x) for x in digest.cpu().tolist()]

    # ---- roofline of the (only) kernel of the step, from this rank's own launch durations ----
    peak, peak_src, _ = load_peaks()
    algo_bytes = out_bytes + nb * T * (8 + 16 + 4) + nb * 8  # one fp32 store per pixel + the padded input slots
    # one launch per step (groups_kernel, K1 fused); with --two-launch the step is bin_kernel + groups2_kernel and the
    # step time bounds the renderer's duration from above
    launch_ms = ms_total / max(args.steps, 1)
    achieved = algo_bytes / (launch_ms * 1e-3) / 1e9
    # `traffic` is an OFFLINE figure (one ncu --set full capture per kernel change, profiles/roofline_traffic.json names
    # the capture and the commit it was taken at); everything else in this object is measured in this run
    traffic, traffic_source = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        traffic = tj.get(f"{cfg.name}_n{world}")
        traffic_source = tj.get("source") if traffic is not None else None
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_source, "kernel": ("dh::groups_kernel<false,SRC_BOXES,false,OUT_F32>" if not args.two_launch else
                           "dh::groups2_kernel<false,false,OUT_F32> (+ dh::bin_kernel<SRC_BOXES> in the same step)"),
                "peak_source": peak_src, "launches_per_step": launches / max(args.steps, 1),
                "algorithmic_bytes_per_launch": algo_bytes, "launch_ms": launch_ms,
                "frac_of_nominal_8TBs": achieved / 8000.0, "cuda_graph_replay_ms_per_step": graph_ms,
                "cuda_graph_replay_GBps": (algo_bytes / (graph_ms * 1e-3) / 1e9) if graph_ms else None}
    # ---- write-only ceilings on the same buffer (not part of any reported throughput) ----
    ceil = {}
    if rank == 0:
        buf = heats[0]
        for name, fn in (("fill_probe_st_v4", lambda: lib.dh_fill_probe(sp, _cabi.ptr(buf), buf.numel() * 4)),
                         ("torch_zero_", lambda: buf.zero_())):
            reps = max(3, min(50, int(2e9 // max(out_bytes, 1))))
            for _ in range(3):
                fn()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record(stream)
            for _ in range(reps):
                fn()
            b.record(stream)
            torch.cuda.synchronize()
            ceil[name + "_GBps"] = out_bytes * reps / (a.elapsed_time(b) * 1e-3) / 1e9
        step(0)  # restore a rendered result in buffer 0
        torch.cuda.synchronize()
    roofline["write_only_ceilings"] = ceil
    if ceil.get("torch_zero__GBps"):  # the copy peak is a read+write figure; a write-only kernel is better judged against this
        roofline["frac_of_torch_zero_"] = achieved / ceil["torch_zero__GBps"]
    return dict(torch=torch, lib=lib, _cabi=_cabi, dist=dist, dev=dev, rank=rank, world=world, w=w, w_all=w_all, lo=lo,
                host=(h_cf, h_bf, h_sz, h_lab), heats=heats, value=value, ms_step=ms_step, launches=launches,
                roofline=roofline, digest=digest, sampler=sampler, warm=warm, barrier=barrier)


# ------------------------------------------------------------------------------------------------------------------
# Side legs (rank 0, N=1 only; each bounded to about a second): the other BASELINE configs, the drop-in accumulate path
# and the unmodified reference CUDA extension, so that these numbers are driver-visible and not only builder records.
def _ev_ms(torch, fn, reps, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def _fresh_call(lib, _cabi, sp, heat, d, cfg, nb, T, ws, ws_bytes):
    st = lib.dh_draw_heatmap_batched_from_boxes_ws(sp, _cabi.ptr(heat), _cabi.ptr(d[0]), _cabi.ptr(d[1]), float(cfg.osf),
                                                   _cabi.ptr(d[2]), _cabi.DH_SIZES_I64, _cabi.ptr(d[3]), nb, T, cfg.C, cfg.H,
                                                   cfg.W, FACTOR, K_SCALE, _cabi.DH_MODE_FRESH, None, None, None,
                                                   _cabi.ptr(ws), ws_bytes)
    if st != 0:
        raise RuntimeError(lib.dh_status_string(st).decode())


def side_configs_leg(ctx, peak):
    """cfg1-cfg4: FRESH render with K1 fused, CUDA events over back-to-back steps on buffers rotated
    over > 512 MB, plus the same steps replayed from a CUDA graph (launch-bound configs)."""
    torch, lib, _cabi, dev = (ctx[k] for k in ("torch", "lib", "_cabi", "dev"))
    sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = {}
    for name in ("cfg1", "cfg2", "cfg3", "cfg4"):
        w = workloads.make(name)
        cfg = w.cfg
        d = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (w.centers_f, w.bboxes_f, w.sizes, w.labels)]
        nbuf = int(min(64, -(-(512 << 20) // w.out_bytes)))
        heats = [torch.empty((w.B, cfg.C, cfg.H, cfg.W), dtype=torch.float32, device=dev) for _ in range(nbuf)]
        wsb, ws = 0, None  # single-launch renderer
        it = [0]

        def step(sp_=sp):
            it[0] += 1
            _fresh_call(lib, _cabi, sp_, heats[it[0] % nbuf], d, cfg, w.B, w.T, ws, wsb)

        reps = 400 if name != "cfg4" else 100
        ms = _ev_ms(torch, step, reps, warm=20)
        graph_ms = None
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(g, stream=side):
                csp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
                for _ in range(nbuf):
                    step(csp)
            graph_ms = _ev_ms(torch, g.replay, max(3, reps // nbuf), warm=3) / nbuf
        except Exception as e:
            print(f"[bench] {name}: CUDA-graph replay skipped: {e!r}", file=sys.stderr)
        algo = w.out_bytes + w.B * w.T * 28 + w.B * 8
        r = {"planes": w.planes, "output_bytes": w.out_bytes, "ms_per_step": ms, "heatmaps_per_s": w.planes / (ms * 1e-3),
             "GBps": algo / (ms * 1e-3) / 1e9, "frac": algo / (ms * 1e-3) / 1e9 / peak, "cuda_graph_replay_ms": graph_ms,
             "steps": reps, "rotating_buffers": nbuf}
        if name == "cfg4":  # not an HBM config: 1024 boxes of r 16..64 on every 512x512 plane -> pixel-object pairs
            ci, ri = _cabi.radius_center(lib, d[0], d[1], float(cfg.osf))
            x, y, rr = ci[..., 0].long(), ci[..., 1].long(), ri.long()
            pairs = int((((x + rr).clamp(max=cfg.W - 1) - (x - rr).clamp(min=0) + 1).clamp(min=0) *
                         ((y + rr).clamp(max=cfg.H - 1) - (y - rr).clamp(min=0) + 1).clamp(min=0)).sum())
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            issue_peak = sms * 4 * 1.965e9  # warp instructions / s: 4 schedulers per SM at the 1965 MHz the clocks record shows
            r.update({"bound": "instruction issue (not HBM)", "pixel_object_pairs": pairs,
                      "pairs_per_s": pairs / (ms * 1e-3), "issue_peak_warp_inst_per_s": issue_peak,
                      "pairs_per_issue_slot": pairs / (ms * 1e-3) / issue_peak,
                      "note": "register-tile renderer chosen per unit on the device: a packed-fp32 instruction handles 64 "
                              "(pixel, object) pairs, ~4 issue slots per object-row of 64 pixels, and pairs of dominated "
                              "objects are never evaluated; profiles/r02/tiles holds the ncu instruction count"})
        out[name] = r
        del heats
    return out


def accumulate_leg(ctx, peak):
    """The drop-in path a reference user calls (max-blend into the caller's tensor, DH_MODE_ACCUMULATE) on the first
    1536 samples of cfg5: accumulate alone into an already zeroed heatmap, and zero_() + accumulate (what a fresh heatmap
    costs through the reference-named API). Roofline of this mode: 8 B per TOUCHED pixel (read + write)."""
    torch, lib, _cabi, dev, w_all = (ctx[k] for k in ("torch", "lib", "_cabi", "dev", "w_all"))
    nb = min(w_all.B, 1536)
    w = w_all.slice(0, nb)
    cfg = w.cfg
    d = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (w.centers_f, w.bboxes_f, w.sizes, w.labels)]
    ci, ri = _cabi.radius_center(lib, d[0], d[1], float(cfg.osf))
    heat = ctx["heats"][0][:nb]
    wsb, ws = 0, None  # single-launch renderer
    sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def acc():
        st = lib.dh_draw_heatmap_batched_ws(sp, _cabi.ptr(heat), _cabi.ptr(ci), _cabi.ptr(ri), _cabi.ptr(d[2]),
                                            _cabi.DH_SIZES_I64, _cabi.ptr(d[3]), nb, w.T, cfg.C, cfg.H, cfg.W, FACTOR, K_SCALE,
                                            _cabi.DH_MODE_ACCUMULATE, None, _cabi.ptr(ws), wsb)
        if st != 0:
            raise RuntimeError(lib.dh_status_string(st).decode())

    def zero_acc():
        heat.zero_()
        acc()

    def fresh():
        _fresh_call(lib, _cabi, sp, heat, d, cfg, nb, w.T, ws, wsb)

    heat.zero_()
    acc()
    touched = int(_cabi.checksum(lib, heat)[2])
    reps = 20
    ms_acc, ms_zero_acc, ms_fresh = _ev_ms(torch, acc, reps), _ev_ms(torch, zero_acc, reps), _ev_ms(torch, fresh, reps)
    algo = 8 * touched
    return {"samples": nb, "planes": nb * cfg.C, "touched_pixels": touched, "accumulate_ms": ms_acc,
            "zero_plus_accumulate_ms": ms_zero_acc, "fresh_fused_ms": ms_fresh,
            "roofline": {"bound": "hbm", "algorithmic_bytes": algo, "achieved": algo / (ms_acc * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": algo / (ms_acc * 1e-3) / 1e9 / peak,
                         "note": "8 B per touched pixel; the launch is bound by per-unit latency, not by bytes"},
            "heatmaps_per_s_zero_plus_accumulate": nb * cfg.C / (ms_zero_acc * 1e-3)}


def reference_gpu_leg(ctx):
    """The UNMODIFIED reference CUDA extension (oracle/_ref/dh_ref_fast.so: the reference's own sources compiled for
    sm_100a with its default --use_fast_math, oracle/build_oracle.py) on the same device and inputs — a BASELINE that is
    measured, like the CPU arm; the product never loads it. Reference cost of a fresh heatmap = zero_() + its draw call
    (its API max-blends into a caller-zeroed tensor). First 1536 samples of cfg5: the reference indexes pixels with
    int32 and overflows on the whole."""
    import importlib.util

    torch, lib, _cabi, dev, w_all = (ctx[k] for k in ("torch", "lib", "_cabi", "dev", "w_all"))
    path = os.path.join(ROOT, "oracle", "_ref", "dh_ref_fast.so")
    if not os.path.exists(path):
        return {"unavailable": "oracle/_ref/dh_ref_fast.so not built (needs /root/reference at build time)"}
    spec = importlib.util.spec_from_file_location("dh_ref_fast", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    nb = min(w_all.B, 1536)
    w = w_all.slice(0, nb)
    cfg = w.cfg
    d = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (w.centers_f, w.bboxes_f, w.sizes, w.labels)]
    ci, ri = _cabi.radius_center(lib, d[0], d[1], float(cfg.osf))
    s32 = d[2].int()
    heat = ctx["heats"][0][:nb]

    def f_ref():
        heat.zero_()
        ref.draw_heatmap_batched_classwise_impl(heat, ci, ri, s32, d[3], FACTOR, K_SCALE)

    ms_ref = _ev_ms(torch, f_ref, 5, warm=2)
    ref_digest = [int(x) for x in _cabi.checksum(lib, heat).cpu().tolist()]
    wsb, ws = 0, None  # single-launch renderer
    sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ms_fresh = _ev_ms(torch, lambda: _fresh_call(lib, _cabi, sp, heat, d, cfg, nb, w.T, ws, wsb), 20)
    our_digest = [int(x) for x in _cabi.checksum(lib, heat).cpu().tolist()]
    return {"samples": nb, "planes": nb * cfg.C, "reference_zero_plus_draw_ms": ms_ref,
            "reference_heatmaps_per_s": nb * cfg.C / (ms_ref * 1e-3), "ours_fresh_fused_ms": ms_fresh,
            "ours_heatmaps_per_s": nb * cfg.C / (ms_fresh * 1e-3), "speedup_fresh": ms_ref / ms_fresh,
            "touched_mask_identical": ref_digest[1:] == our_digest[1:],
            "kind": "unmodified reference extension, --use_fast_math (its default), sm_100a"}



def e2e_leg(ctx, steps, warm):
    """End to end through the C-ABI host entry: every step copies the step's inputs (boxes, labels, sizes) from
    PINNED HOST memory to the device, renders the fresh heatmap (K1 fused), digests it on the device and reads the
    24-byte digest back to the host (the step's "result read": the reference API's output is a device tensor consumed
    on the GPU by the loss — draw_heatmap_cuda.cu:91-165 — so the heatmap itself is not shipped to the host here; the
    variant that does ship it is reported separately as `e2e_full_copy`). Host wall clock, max over ranks."""
    torch, lib, _cabi, dist, dev, w = (ctx[k] for k in ("torch", "lib", "_cabi", "dist", "dev", "w"))
    cfg = w.cfg
    h_cf, h_bf, h_sz, h_lab = ctx["host"]
    nb, T, C, H, W = w.B, w.T, cfg.C, cfg.H, cfg.W
    nbytes = lib.dh_host_scratch_bytes(nb, T, C, H, W, 1, 0, 0)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ck = torch.zeros(3, dtype=torch.int64).pin_memory()
    heat = ctx["heats"][0]
    stream = torch.cuda.current_stream()
    sp = ctypes.c_void_p(stream.cuda_stream)

    # the argument list is constant over the steps: marshalled once, as a C caller would hold it (7 data_ptr() calls and
    # the ctypes boxing cost ~20 us of interpreter time per step otherwise — harness time, not the library's)
    args = (sp, None, _cabi.ptr(heat), _cabi.ptr(ck), _cabi.ptr(h_cf), _cabi.ptr(h_bf), ctypes.c_float(float(cfg.osf)),
            _cabi.ptr(h_sz), _cabi.ptr(h_lab), nb, T, C, H, W, ctypes.c_float(FACTOR), ctypes.c_float(K_SCALE),
            _cabi.ptr(scratch), nbytes, 0)
    ck_host = ck.numpy()  # view of the pinned result words
    entry = lib.dh_draw_heatmap_batched_from_boxes_host
    sync = torch.cuda.current_stream().synchronize

    def step():
        st = entry(*args)
        if st != 0:
            raise RuntimeError(f"host entry: {lib.dh_status_string(st).decode()}")
        sync()
        return int(ck_host[2])  # host read of the result

    for _ in range(warm):
        step()
    ctx["barrier"]()
    l0 = lib.dh_launch_count()
    t0 = time.perf_counter()
    for _ in range(steps):
        touched = step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    launches = int(lib.dh_launch_count() - l0)
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    h2d = h_cf.numel() * 4 + h_bf.numel() * 4 + h_sz.numel() * 8 + h_lab.numel() * 4
    return {"value": ctx["w_all"].planes * steps / dt, "unit": "heatmaps/s", "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": 24, "ms_per_step": 1e3 * dt / steps, "gpu_launches_per_step": launches / steps,
            "touched_pixels_this_rank": touched,
            "what": "pinned host boxes -> H2D -> fused render -> device digest -> D2H digest, per step, per rank"}


def e2e_full_copy_leg(ctx, seconds=3.0):
    """Same entry with a HOST output buffer: the whole heatmap crosses PCIe back to pinned host memory (chunked,
    render of chunk i overlapping the copy of chunk i-1). Rank 0, bounded to <= 96 samples. PCIe-bound by design."""
    torch, lib, _cabi, dev, w = (ctx[k] for k in ("torch", "lib", "_cabi", "dev", "w"))
    cfg = w.cfg
    nb = min(w.B, 96)
    s = w.slice(0, nb)
    T, C, H, W = s.T, cfg.C, cfg.H, cfg.W
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    h_cf, h_bf, h_sz, h_lab = pin(s.centers_f), pin(s.bboxes_f), pin(s.sizes), pin(s.labels)
    out = torch.empty((nb, C, H, W), dtype=torch.float32).pin_memory()
    nbytes = lib.dh_host_scratch_bytes(nb, T, C, H, W, 1, 0, 1)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    sp = ctypes.c_void_p(stream.cuda_stream)

    def step():
        st = lib.dh_draw_heatmap_batched_from_boxes_host(
            sp, _cabi.ptr(out), None, None, _cabi.ptr(h_cf), _cabi.ptr(h_bf), float(cfg.osf), _cabi.ptr(h_sz),
            _cabi.ptr(h_lab), nb, T, C, H, W, FACTOR, K_SCALE, _cabi.ptr(scratch), nbytes, 0)
        if st != 0:
            raise RuntimeError(f"host entry: {lib.dh_status_string(st).decode()}")
        stream.synchronize()

    for _ in range(3):
        step()
    n, t0 = 0, time.perf_counter()
    while n < 3 or (time.perf_counter() - t0 < seconds and n < 200):
        step()
        n += 1
    dt = time.perf_counter() - t0
    d2h = out.numel() * 4
    return {"value": nb * C * n / dt, "unit": "heatmaps/s", "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * dt / n,
            "d2h_GBps": d2h * n / dt / 1e9, "samples": nb, "steps": n, "rank": 0,
            "what": "as e2e but the whole heatmap is copied to pinned host memory (PCIe-bound)"}


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    ctx = run_ours(args)
    torch, rank, world = ctx["torch"], ctx["rank"], ctx["world"]
    e2e = None if args.no_e2e else e2e_leg(ctx, args.steps, ctx["warm"])
    e2e_full = None
    cpu = None
    if rank == 0 and world == 1:
        if not args.no_e2e:
            e2e_full = e2e_full_copy_leg(ctx)
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            try:
                threads = len(os.sched_getaffinity(0))
            except Exception:
                pass
            cpu = cpu_reference_leg(ctx["w_all"], args.cpu_seconds, threads)
            one = cpu_reference_leg(ctx["w_all"], min(args.cpu_seconds, 4.0), 1, max_samples=48)
            cpu["single_thread_value"] = one["value"]
    side = {}
    if rank == 0 and world == 1 and not args.no_side_legs and args.workload == "cfg5" and args.batch is None:
        peak = ctx["roofline"]["peak"]
        for key, fn in (("configs", lambda: side_configs_leg(ctx, peak)), ("accumulate", lambda: accumulate_leg(ctx, peak)),
                        ("reference_gpu", lambda: reference_gpu_leg(ctx))):
            try:
                side[key] = fn()
            except Exception as e:  # side records never take the headline down
                side[key] = {"error": repr(e)}
    if ctx["dist"] is not None:
        ctx["dist"].barrier()
    if rank == 0:
        line = {
            "metric": "gaussian_heatmaps_per_sec", "value": ctx["value"], "unit": "heatmaps/s", "n_gpus": world,
            "steps": args.steps, "warmup": ctx["warm"], "ms_per_step": ctx["ms_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(ctx["w_all"], world),
            "roofline": ctx["roofline"], "cpu_baseline": cpu, "e2e": e2e, "e2e_full_copy": e2e_full,
            "gpu_launches": ctx["launches"], "clocks": ctx["sampler"].summary() if ctx["sampler"] else None,
            "digest": ctx["digest"], "hbm_write_GBps_aggregate": ctx["w_all"].out_bytes / (ctx["ms_step"] * 1e-3) / 1e9,
            "build": ctx["lib"].dh_build_info().decode(),
        }
        line.update(side)
        print(json.dumps(line), flush=True)
    if ctx["dist"] is not None:
        ctx["dist"].destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
