"""Generate tests/golden/*.npz by RUNNING THE REFERENCE'S OWN PYTHON CODE (CPU tensors) in the build container.

    python tests/golden/make_golden.py          # needs /root/reference; not needed on the GPU box

Nothing of the reference is copied: its modules are imported from where they lie
(/root/reference/packages/draw_heatmap/tests/_test_helpers.py, _gaussian_drawing.py) and its example inputs
(examples/input_data.py) are executed with `device` forced to "cpu". The recorded outputs are:
  centers_int / radii   <- get_centers_and_radii                     (_test_helpers.py:20-28)
  heat                  <- get_heatmap_multiple_samples /
                           get_heatmaps_multiple_samples_with_labels  (_test_helpers.py:41-74 -> _gaussian_drawing.py:215-240)
Data generation follows generate_gt_bboxes (_test_helpers.py:77-94) but with a CPU torch.Generator so the vectors are
reproducible on any machine; padded slots carry the drawable decoy box of test_draw_heatmap_batched.py:52-58.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF_DH = "/root/reference/packages/draw_heatmap"
sys.path.insert(0, os.path.join(REF_DH, "tests"))
import _test_helpers as th  # noqa: E402  (the reference's module)

OUT = os.path.dirname(os.path.abspath(__file__))


def gen_boxes(gen, B, max_targets, img_hw, max_bbox, min_targets=0):
    """generate_gt_bboxes (_test_helpers.py:77-94) with an explicit CPU generator."""
    centers_list, bboxes_list = [], []
    for _ in range(B):
        n = int(torch.randint(min_targets, max_targets + 1, (1,), generator=gen))
        cx = torch.rand(n, generator=gen) * img_hw[1]
        cy = torch.rand(n, generator=gen) * img_hw[0]
        c = torch.stack([cx, cy], dim=1)
        d = torch.rand(n, 4, generator=gen) * max_bbox / 2
        bb = torch.cat([c - d[:, 0:2], c + d[:, 2:4]], dim=1)
        centers_list.append(c)
        bboxes_list.append(bb)
    return centers_list, bboxes_list


def pad(centers_list, bboxes_list, T, osf, labels_list=None):
    """Padded tensors with the drawable decoy in invalid slots (test_draw_heatmap_batched.py:52-64)."""
    B = len(centers_list)
    centers = torch.ones((B, T, 2), dtype=torch.float32) * 2 * osf
    bboxes = torch.zeros((B, T, 4), dtype=torch.float32)
    bboxes[:, :, 0:2] = 1 * osf
    bboxes[:, :, 2:4] = 3 * osf
    labels = torch.zeros((B, T), dtype=torch.int32)
    sizes = torch.zeros(B, dtype=torch.int64)
    for i in range(B):
        n = centers_list[i].shape[0]
        centers[i, :n] = centers_list[i]
        bboxes[i, :n] = bboxes_list[i]
        if labels_list is not None:
            labels[i, :n] = labels_list[i].to(torch.int32)
        sizes[i] = n
    return centers, bboxes, labels, sizes


def case_batched(name, seed, B, T, img_hw, osf, max_bbox, k, factor, C=None):
    gen = torch.Generator().manual_seed(seed)
    cl, bl = gen_boxes(gen, B, T, img_hw, max_bbox)
    ll = None
    if C is not None:
        ll = [torch.randint(0, C, (c.shape[0],), generator=gen) for c in cl]
    centers, bboxes, labels, sizes = pad(cl, bl, T, osf, ll)
    H, W = img_hw[0] // osf, img_hw[1] // osf
    # reference K1 on the padded tensors (as test_draw_heatmap_batched.py:83 does)
    ci, ri = th.get_centers_and_radii(centers, bboxes, osf)
    # reference CPU heatmaps from the per-sample lists
    cli, rli = th.get_centers_and_radii_multiple_samples(cl, bl, osf)
    rli = [t.numpy().tolist() for t in rli]
    if C is None:
        heat = torch.stack(th.get_heatmap_multiple_samples(cli, rli, k_scale=k, diameter_to_sigma_factor=factor,
                                                           heatmap_size=[H, W]))
    else:
        lli = [l.numpy().tolist() for l in ll]
        heat = torch.stack(th.get_heatmaps_multiple_samples_with_labels(
            cli, rli, lli, k_scale=k, diameter_to_sigma_factor=factor, heatmap_size=[H, W], max_num_classes=C))
    out = dict(centers_f=centers.numpy(), bboxes_f=bboxes.numpy(), sizes=sizes.numpy(), osf=np.float32(osf),
               k=np.float32(k), factor=np.float32(factor), centers_int=ci.numpy(), radii=ri.numpy(),
               heat=heat.numpy().astype(np.float32))
    if C is not None:
        out["labels"] = labels.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, heat.shape, "objects", int(sizes.sum()), "max", float(heat.max()))


def case_k1(name, seed):
    """K1 only: non-power-of-two out_size_factors, centres outside their box, large coordinates."""
    gen = torch.Generator().manual_seed(seed)
    out = {}
    for i, osf in enumerate([1, 3, 4, 5, 7.5, 8, 16, 0.7]):
        n = 4000
        c = (torch.rand(n, 2, generator=gen) - 0.1) * 3000.0
        d = (torch.rand(n, 4, generator=gen) - 0.05) * 400.0  # some negative distances -> radius clamps to 1
        bb = torch.cat([c - d[:, 0:2], c + d[:, 2:4]], dim=1)
        ci, ri = th.get_centers_and_radii(c, bb, osf)
        out[f"osf_{i}"] = np.float32(osf)
        out[f"centers_f_{i}"] = c.numpy()
        out[f"bboxes_f_{i}"] = bb.numpy()
        out[f"centers_int_{i}"] = ci.numpy()
        out[f"radii_{i}"] = ri.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "ok")


def case_edges(name):
    """Integer inputs straight into the reference CPU drawer: borders, outside centres, r larger than the map."""
    H, W = 13, 18
    objs = [  # (x, y, r)
        (0, 0, 3), (W - 1, H - 1, 4), (W, 5, 2), (-1, 5, 3), (5, -2, 2), (5, H + 1, 3), (9, 6, 30), (-4, -4, 3),
        (W + 3, H + 3, 3), (3, 3, 1), (3, 3, 0), (17, 0, 1), (0, 12, 2), (40, 40, 5), (-30, 2, 5), (8, 7, 6),
    ]
    centers = torch.tensor([[o[0], o[1]] for o in objs], dtype=torch.int32)
    k, factor = 0.8, 6
    heats = []
    for i, o in enumerate(objs):  # one plane per object so every case is pinned individually
        heats.append(th.get_heatmap_single(centers[i : i + 1], [o[2]], k, factor, [H, W]))
    together = th.get_heatmap_single(centers, [o[2] for o in objs], k, factor, [H, W])
    np.savez_compressed(os.path.join(OUT, name + ".npz"), objs=np.array(objs, np.int32), k=np.float32(k),
                        factor=np.float32(factor), heat_each=torch.stack(heats).numpy(), heat_all=together.numpy())
    print(name, "ok")


def case_example_inputs(name):
    """The reference's fixed example inputs (examples/input_data.py:35-59), executed with device='cpu'."""
    src = open(os.path.join(REF_DH, "examples", "input_data.py")).read().replace('device = "cuda:0"', 'device = "cpu"')
    ns: dict = {}
    exec(compile(src, "input_data.py", "exec"), ns)
    cl, rl, ll = ns["centers_list"], ns["radii_list"], ns["labels_list"]
    H, W = ns["HEATMAP_SIZE"]
    C, T = ns["MAX_NUM_CLASSES"], ns["MAX_NUM_TARGET"]
    k, factor = ns["k_scale"], ns["diameter_to_sigma_factor"]
    B = len(cl)
    centers = np.zeros((B, T, 2), np.int32)
    radii = np.zeros((B, T), np.int32)
    labels = np.zeros((B, T), np.int32)
    sizes = np.zeros(B, np.int64)
    for i in range(B):
        n = len(cl[i])
        centers[i, :n] = cl[i].numpy()
        radii[i, :n] = rl[i].numpy()
        labels[i, :n] = ll[i].numpy()
        sizes[i] = n
    rli = [r.numpy().tolist() for r in rl]
    lli = [l.numpy().tolist() for l in ll]
    heat = torch.stack(th.get_heatmap_multiple_samples(cl, rli, k, factor, [H, W])).numpy()
    heat_cls = torch.stack(th.get_heatmaps_multiple_samples_with_labels(cl, rli, lli, k, factor, [H, W], C)).numpy()
    # keep the file small: classwise planes are mostly empty -> store only the non-empty ones
    nz = np.argwhere(heat_cls.reshape(B * C, -1).max(axis=1) > 0)[:, 0]
    np.savez_compressed(os.path.join(OUT, name + ".npz"), centers=centers, radii=radii, labels=labels, sizes=sizes,
                        k=np.float32(k), factor=np.float32(factor), H=H, W=W, C=C, heat=heat,
                        heat_cls_planes=nz.astype(np.int64), heat_cls_values=heat_cls.reshape(B * C, H, W)[nz])
    print(name, heat.shape, heat_cls.shape, "non-empty class planes", len(nz))


if __name__ == "__main__":
    torch.set_num_threads(1)
    # the reference's own test shape family: IMG 320x800, osf 16 -> 20x50, k=0.8, factor=6
    case_batched("ref_batched_20x50", seed=7, B=12, T=50, img_hw=(320, 800), osf=16, max_bbox=120, k=0.8, factor=6)
    case_batched("ref_classwise_20x50", seed=7, B=6, T=50, img_hw=(320, 800), osf=16, max_bbox=120, k=0.8, factor=6, C=8)
    # larger radii, width not a multiple of 4, non-power-of-two osf
    case_batched("ref_batched_45x63", seed=11, B=5, T=30, img_hw=(225, 315), osf=5, max_bbox=200, k=1.0, factor=6)
    case_batched("ref_classwise_64x64", seed=13, B=4, T=60, img_hw=(512, 512), osf=8, max_bbox=300, k=1.0, factor=4, C=3)
    case_k1("ref_k1", seed=3)
    case_edges("ref_edges")
    case_example_inputs("ref_examples")
