"""Threading / stream behaviour of the boundary (SURVEY.md 8(b) "Threading / streams": the reference runs on the current
stream of the heatmap's device, is fully asynchronous, re-entrant, and safe from several Python threads on different
streams, `draw_heatmap_cuda.cu:64-65,94-95,130-131`) — plus CUDA-graph capture, which the bench uses for small configs."""
import threading

import numpy as np
import pytest

import workloads

pytestmark = pytest.mark.gpu


def _up(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _inputs(name, dev, B=None):
    from accvlab.draw_heatmap import b200

    w = workloads.make(name, B=B) if B else workloads.make(name)
    cf, bf, sizes, lab = (_up(a, dev) for a in (w.centers_f, w.bboxes_f, w.sizes, w.labels))
    ci, ri = b200.radius_center(cf, bf, w.cfg.osf)
    return w, cf, bf, sizes, lab, ci, ri


def test_python_threads_on_their_own_streams(cuda_dev):
    """4 threads x 25 iterations, each thread on its own stream with its own workload, all entry points mixed: every
    result equals the one computed alone on the default stream."""
    import torch
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap import b200

    names = ["cfg1", "cfg2", "cfg1", "cfg2"]
    packs = [_inputs(n, cuda_dev) for n in names]
    want = []
    for w, cf, bf, sizes, lab, ci, ri in packs:
        rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
        fresh = b200.render_heatmap_from_boxes(cf, bf, w.cfg.osf, sizes, w.cfg.H, w.cfg.W, w.cfg.C, lab)
        half = b200.render_heatmap_batched(rb(ci), rb(ri), w.cfg.H, w.cfg.W, w.cfg.C, labels=rb(lab), dtype=torch.bfloat16)
        want.append((fresh, half))
    torch.cuda.synchronize()
    errors = []

    def work(i):
        try:
            w, cf, bf, sizes, lab, ci, ri = packs[i]
            rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
            s = torch.cuda.Stream(device=cuda_dev)
            with torch.cuda.stream(s):
                for it in range(25):
                    a = b200.render_heatmap_from_boxes(cf, bf, w.cfg.osf, sizes, w.cfg.H, w.cfg.W, w.cfg.C, lab)
                    acc = torch.zeros_like(a)
                    dh.draw_heatmap_batched(acc, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
                    h = b200.render_heatmap_batched(rb(ci), rb(ri), w.cfg.H, w.cfg.W, w.cfg.C, labels=rb(lab),
                                                    dtype=torch.bfloat16)
                    s.synchronize()
                    if not (torch.equal(a, want[i][0]) and torch.equal(acc, want[i][0]) and torch.equal(h, want[i][1])):
                        errors.append((i, it))
                        return
        except Exception as e:  # noqa: BLE001 - reported through the list
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_calls_are_ordered_on_the_current_stream(cuda_dev):
    """the call is enqueued behind earlier work of the current (non-default) stream and ahead of later work on it"""
    import torch
    from accvlab.draw_heatmap import b200

    w, cf, bf, sizes, lab, ci, ri = _inputs("cfg2", cuda_dev)
    want = b200.render_heatmap_from_boxes(cf, bf, w.cfg.osf, sizes, w.cfg.H, w.cfg.W, w.cfg.C, lab)
    torch.cuda.synchronize()
    s = torch.cuda.Stream(device=cuda_dev)
    out = torch.empty_like(want)
    with torch.cuda.stream(s):
        torch.cuda._sleep(20_000_000)  # ~10 ms of earlier work on s
        out.fill_(5.0)                 # ordered before the render: must be overwritten
        b200.render_heatmap_from_boxes(cf, bf, w.cfg.osf, sizes, labels=lab, out=out)
        total = out.sum()              # ordered after it
    s.synchronize()
    assert torch.equal(out, want) and float(total) == float(want.sum())


def test_cuda_graph_capture_and_replay(cuda_dev):
    """every device-side entry is capturable (no synchronisation, no host round trip inside): capture once, replay on
    new input CONTENTS, compare with the eager result"""
    import torch
    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap import b200

    w, cf, bf, sizes, lab, ci, ri = _inputs("cfg2", cuda_dev)
    cfg = w.cfg
    rb = lambda t: b200.RaggedBatch(t, sample_sizes=sizes)
    shape = (w.B, cfg.C, cfg.H, cfg.W)
    fresh, acc, half = torch.empty(shape, device=cuda_dev), torch.zeros(shape, device=cuda_dev), torch.empty(
        shape, dtype=torch.float16, device=cuda_dev)
    pred = torch.rand(shape, device=cuda_dev) * 0.9 + 0.05
    kcls = torch.linspace(0.6, 1.0, cfg.C, device=cuda_dev)
    rf = ri.to(torch.float32) + 0.25
    s = torch.cuda.Stream(device=cuda_dev)
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):  # warm-up outside the capture (lazy attribute set-up, allocator)
        b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, labels=lab, out=fresh)
        b200.gaussian_focal_loss(pred, rb(ci), rb(ri), labels=rb(lab))
        b200.draw_gaussians(None, None, rb(lab), rb(ci), rb(rf), kcls, 0.3, shape=shape)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, labels=lab, out=fresh)
        acc.zero_()
        dh.draw_heatmap_batched(acc, rb(ci), rb(ri), 6.0, 1.0, labels=rb(lab))
        b200.render_heatmap_batched(rb(ci), rb(ri), labels=rb(lab), out=half)
        loss, npos = b200.gaussian_focal_loss(pred, rb(ci), rb(ri), labels=rb(lab))
        dg = b200.draw_gaussians(None, None, rb(lab), rb(ci), rb(rf), kcls, 0.3, shape=shape)
    # new contents in the captured buffers: shift every box and shuffle the predictions
    cf += 8.0
    bf += 8.0
    ci2, ri2 = b200.radius_center(cf, bf, cfg.osf)
    ci.copy_(ci2)
    ri.copy_(ri2)
    rf.copy_(ri2.to(torch.float32) + 0.25)
    pred.copy_(pred.flip(0))
    g.replay()
    torch.cuda.synchronize()
    want = b200.render_heatmap_from_boxes(cf, bf, cfg.osf, sizes, cfg.H, cfg.W, cfg.C, lab)
    assert torch.equal(fresh, want) and torch.equal(acc, want) and torch.equal(half, want.to(torch.float16))
    loss2, npos2 = b200.gaussian_focal_loss(pred, rb(ci), rb(ri), labels=rb(lab))
    assert float(loss) == float(loss2) and int(npos) == int(npos2)
    dg2 = b200.draw_gaussians(None, None, rb(lab), rb(ci), rb(rf), kcls, 0.3, shape=shape)
    assert torch.equal(dg, dg2)
