"""Generate tests/golden/ref_pack.npz by RUNNING THE REFERENCE'S OWN combine_data (CPU tensors) in the build container.

    python tests/golden/make_golden_pack.py     # needs /root/reference; not needed on the GPU box

The reference package accvlab.batching_helpers is imported from where it lies
(/root/reference/packages/batching_helpers); its two compiled extension modules (batched_indexing_access_cuda/_cpu,
not built here, and not used by combine_data) are replaced by empty stub modules for the import only.
Recorded: the per-sample inputs (concatenated + offsets) and combine_data's outputs `.tensor` / `.sample_sizes`
(batched_processing_py.py:275-438) for float32 [n,2] / [n,4], int32 [n], uint8 [n,3], a nested list (flattening
mode, :373-397), and lists with empty samples; plus get_centers_and_radii (DH/tests/_test_helpers.py:20-28) applied to
the combined centres/boxes (what pack_radius_center must reproduce on the valid slots).
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

REF_BH = "/root/reference/packages/batching_helpers"
REF_DH = "/root/reference/packages/draw_heatmap"
for name in ("accvlab.batching_helpers.batched_indexing_access_cuda", "accvlab.batching_helpers.batched_indexing_access_cpu"):
    sys.modules[name] = types.ModuleType(name)
sys.path.insert(0, REF_BH)
sys.path.insert(0, os.path.join(REF_DH, "tests"))
import accvlab.batching_helpers as bh  # noqa: E402  (the reference's package)
import _test_helpers as th  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def cat(lst):
    nz = [t for t in lst]
    offs = np.cumsum([0] + [int(t.shape[0]) if t.numel() > 0 else 0 for t in nz]).astype(np.int64)
    ne = [t for t in nz if t.numel() > 0]
    flat = torch.cat(ne, dim=0).numpy() if ne else np.zeros((0,), np.float32)
    return flat, offs


def main():
    gen = torch.Generator().manual_seed(4242)
    out = {}
    sizes = [5, 0, 17, 1, 9, 0, 3, 33]
    cases = {
        "f32x2": [torch.rand(n, 2, generator=gen) * 800 for n in sizes],
        "f32x4": [torch.rand(n, 4, generator=gen) * 800 for n in sizes],
        "i32": [torch.randint(-50, 50, (n,), generator=gen, dtype=torch.int32) for n in sizes],
        "u8x3": [torch.randint(0, 255, (n, 3), generator=gen, dtype=torch.uint8) for n in sizes],
        "f64x2x2": [torch.rand(n, 2, 2, generator=gen, dtype=torch.float64) for n in sizes],
    }
    for k, lst in cases.items():
        rb = bh.combine_data(lst)
        flat, offs = cat(lst)
        out[f"{k}_flat"], out[f"{k}_offs"] = flat, offs
        out[f"{k}_tensor"], out[f"{k}_sizes"] = rb.tensor.numpy(), rb.sample_sizes.numpy()
    # nested (flattening mode): [[a, b], [c], [[d], e]]
    a, b, c, d, e = (torch.rand(n, 2, generator=gen) for n in (4, 2, 7, 0, 5))
    rb = bh.combine_data([[a, b], [c], [[d], e]])
    flat, offs = cat([a, b, c, d, e])
    out["nested_flat"], out["nested_offs"] = flat, offs
    out["nested_tensor"], out["nested_sizes"] = rb.tensor.numpy(), rb.sample_sizes.numpy()
    # K1 on the combined centres / boxes (valid slots are what matters; padded slots are zeros -> r=1, c=0 in the
    # reference helper, the packed kernel writes 0 there)
    osf = 8.0
    centers = [torch.rand(n, 2, generator=gen) * torch.tensor([1440.0, 1440.0]) for n in sizes]
    dist = [torch.rand(n, 4, generator=gen) * 64 for n in sizes]
    bboxes = [torch.cat([c_ - d_[:, :2], c_ + d_[:, 2:]], dim=1) for c_, d_ in zip(centers, dist)]
    crb, brb = bh.combine_data(centers), bh.combine_data(bboxes)
    ci, ri = th.get_centers_and_radii(crb.tensor, brb.tensor, osf)
    fc, oc = cat(centers)
    fb, _ = cat(bboxes)
    out.update(k1_centers_flat=fc, k1_bboxes_flat=fb, k1_offs=oc, k1_osf=np.float32(osf),
               k1_centers_int=ci.numpy(), k1_radii=ri.numpy(), k1_sizes=crb.sample_sizes.numpy())
    np.savez_compressed(os.path.join(OUT, "ref_pack.npz"), **out)
    print("wrote ref_pack.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
