"""CPU: workload generator, sharding and the N>1 digest exchange (gloo, world_size 2)."""
import os
import socket
import sys

import numpy as np
import pytest

import workloads


def test_workloads_are_deterministic_and_padded_with_decoys():
    for name in ("cfg1", "cfg2", "cfg3"):
        a, b = workloads.make(name), workloads.make(name)
        assert np.array_equal(a.centers_f, b.centers_f) and np.array_equal(a.bboxes_f, b.bboxes_f)
        assert np.array_equal(a.sizes, b.sizes) and np.array_equal(a.labels, b.labels)
        cfg = a.cfg
        assert a.centers_f.shape == (cfg.B, a.T, 2) and a.bboxes_f.shape == (cfg.B, a.T, 4)
        assert a.sizes.min() >= cfg.n_min and a.sizes.max() <= cfg.n_max and a.sizes.max() <= a.T
        pad = np.arange(a.T)[None, :] >= a.sizes[:, None]
        if pad.any():
            assert np.all(a.centers_f[pad] == 2 * cfg.osf) and np.all(a.labels[pad] == 0)
        assert a.out_bytes == cfg.B * cfg.C * cfg.H * cfg.W * 4
    w = workloads.make("cfg2")
    assert w.T == 200 and w.out_bytes == 20_736_000
    assert workloads.make("cfg3").out_bytes == 178_176_000


def test_cfg4_radii_and_cfg_sizes(oracle):
    w = workloads.make("cfg4", B=2)
    _, r = oracle.radius_center(w.centers_f, w.bboxes_f, w.cfg.osf)
    assert r.min() >= 16 and r.max() <= 64
    w3 = workloads.make("cfg3", B=4)
    _, r3 = oracle.radius_center(w3.centers_f, w3.bboxes_f, w3.cfg.osf)
    assert r3.min() >= 1 and r3.max() <= 60


def test_shard_ranges_partition_the_batch():
    for B in (0, 1, 7, 48, 3072):
        for G in (1, 2, 3, 4, 8):
            r = [workloads.shard_range(B, g, G) for g in range(G)]
            assert r[0][0] == 0 and r[-1][1] == B
            assert all(r[i][1] == r[i + 1][0] for i in range(G - 1))
            assert max(hi - lo for lo, hi in r) - min(hi - lo for lo, hi in r) <= 1
    from accvlab.draw_heatmap.b200 import shard_range

    assert shard_range(48, 1, 2) == workloads.shard_range(48, 1, 2)
    with pytest.raises(ValueError):
        shard_range(48, 2, 2)


def test_ragged_batch_stand_in():
    import torch
    from accvlab.draw_heatmap.b200 import RaggedBatch

    t = torch.arange(24).reshape(2, 4, 3)
    rb = RaggedBatch(t, sample_sizes=torch.tensor([1, 3]))
    assert rb.tensor is t and rb.max_sample_size == 4 and rb.non_uniform_dim == 1
    assert rb.mask.tolist() == [[True, False, False, False], [True, True, True, False]]
    rb2 = RaggedBatch(t, mask=rb.mask)
    assert rb2.sample_sizes.tolist() == [1, 3]
    with pytest.raises(ValueError):
        RaggedBatch(t)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, ret):
    """One rank of the sharded path on CPU: render the rank's batch slice (oracle stands in for the GPU renderer —
    this test covers the HOST logic: slicing, index bases and the digest all_reduce), then all_reduce the digest."""
    import torch
    import torch.distributed as dist

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import oracle as orc

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = workloads.make("cfg1", B=6)
    cfg = w.cfg
    lo, hi = workloads.shard_range(w.B, rank, world)
    s = w.slice(lo, hi)
    c, r = orc.radius_center(s.centers_f, s.bboxes_f, cfg.osf)
    heat = np.zeros((hi - lo, cfg.C, cfg.H, cfg.W), np.float32)
    orc.draw_batched(heat, c, r, s.sizes, 6.0, 1.0, s.labels)
    d = torch.from_numpy(orc.digest_numpy(heat, index_base=lo * cfg.C * cfg.H * cfg.W).copy())
    dist.all_reduce(d, op=dist.ReduceOp.SUM)  # wrapping int64 adds == wrapping uint64 adds
    if rank == 0:
        ret.put(d.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_digest_equals_unsharded_gloo_world2(oracle):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    got = ret.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    w = workloads.make("cfg1", B=6)
    cfg = w.cfg
    c, r = oracle.radius_center(w.centers_f, w.bboxes_f, cfg.osf)
    heat = np.zeros((w.B, cfg.C, cfg.H, cfg.W), np.float32)
    oracle.draw_batched(heat, c, r, w.sizes, 6.0, 1.0, w.labels)
    assert np.array_equal(got, oracle.digest_numpy(heat))
