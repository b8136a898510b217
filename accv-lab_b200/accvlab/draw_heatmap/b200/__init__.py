"""Additions of the B200 build that the reference does not have (the reference surface stays untouched).

* :func:`render_heatmap_batched` — FRESH mode: allocates (or overwrites) the heatmap, zero-init fused into the render
  kernel, so the caller-side ``torch.zeros`` of the reference (e.g.
  /root/reference/packages/draw_heatmap/tests/test_draw_heatmap_batched_classwise.py:114-116) disappears.
* :func:`radius_center` — K1, bit-exact to tests/_test_helpers.py:20-28 of the reference.
* :func:`render_heatmap_from_boxes` — K1 fused into the renderer: float boxes in, heatmap out, one launch.
* :class:`RaggedBatch` — minimal stand-in for accvlab.batching_helpers.RaggedBatch (fields ``tensor``,
  ``sample_sizes``; /root/reference/packages/batching_helpers/accvlab/batching_helpers/data_format/ragged_batch.py:109-171).
* :func:`combine_data`, :func:`pack_radius_center` — per-sample tensor lists -> RaggedBatch in ONE launch (the
  reference: one slice-assign kernel per sample, batching_helpers/batched_processing_py.py:275-438,
  DH/examples/draw_heatmap_batched.py:42-50); the second applies K1 on the way.
* :func:`draw_gaussians` — the DALI ``draw_gaussians`` operator semantics (CPU-only in the reference:
  dali_pipeline_framework/ext_impl/DrawGaussians.cc:31-157) as a GPU op on RaggedBatch inputs.
* :func:`shard_range` — contiguous batch slice of a rank (multi-GPU: no collective on the data path).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple, Union

import torch

from accvlab.draw_heatmap import draw_heatmap_ext as _ext

__all__ = [
    "RaggedBatch",
    "radius_center",
    "render_heatmap_batched",
    "render_heatmap_from_boxes",
    "render_heatmap_flat",
    "draw_gaussians",
    "boxes_to_gaussian_params",
    "gaussian_focal_loss",
    "gaussian_radius",
    "combine_data",
    "pack_radius_center",
    "shard_range",
    "launch_count",
    "fill_probe",
]


class RaggedBatch:
    """Right-padded batch: ``tensor [B, max_sample_size, *data]`` + ``sample_sizes [B]``; padding content undefined."""

    def __init__(self, tensor: torch.Tensor, mask: Optional[torch.Tensor] = None,
                 sample_sizes: Optional[torch.Tensor] = None, non_uniform_dim: Optional[int] = None):
        if sample_sizes is None:
            if mask is None:
                raise ValueError("Either `mask` or `sample_sizes` must be given")
            sample_sizes = mask.sum(dim=1)
        if non_uniform_dim not in (None, 1):
            raise NotImplementedError("only one batch dimension (non_uniform_dim=1) is supported by draw_heatmap")
        self._tensor = tensor
        self._sample_sizes = sample_sizes
        self._mask = mask

    @property
    def tensor(self) -> torch.Tensor:
        return self._tensor

    @property
    def sample_sizes(self) -> torch.Tensor:
        return self._sample_sizes

    @property
    def mask(self) -> torch.Tensor:
        if self._mask is None:
            t = torch.arange(self._tensor.shape[1], device=self._tensor.device)
            self._mask = t[None, :] < self._sample_sizes[:, None]
        return self._mask

    @property
    def max_sample_size(self) -> int:
        return self._tensor.shape[1]

    @property
    def non_uniform_dim(self) -> int:
        return 1


def radius_center(centers: torch.Tensor, bboxes: torch.Tensor, out_size_factor: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(int32 centres [...,2], int32 radii [...])`` from float centres ``[...,2]`` and boxes ``[...,4]`` (x1,y1,x2,y2)."""
    return _ext.radius_center(centers, bboxes, float(out_size_factor))


def _tensor_and_sizes(x):
    return (x.tensor, x.sample_sizes) if hasattr(x, "sample_sizes") else (x, None)


def render_heatmap_batched(centers, radii, height: Optional[int] = None, width: Optional[int] = None,
                           num_classes: Optional[int] = None, labels=None, diameter_to_sigma_factor: float = 6.0,
                           k_scale: float = 1.0, out: Optional[torch.Tensor] = None,
                           dtype: torch.dtype = torch.float32, use_workspace: bool = False) -> torch.Tensor:
    """Render a fresh heatmap (every pixel written once; zeros where nothing is drawn). Returns the heatmap.

    ``centers``/``radii``/``labels`` are RaggedBatches (or anything with ``.tensor`` and ``.sample_sizes``).
    Either pass ``out`` (any contents, fully overwritten) or ``height``/``width`` (+ ``num_classes`` with labels).
    ``dtype`` (or the dtype of ``out``) may be ``torch.float16`` / ``torch.bfloat16``: the values are the float32
    result rounded once to nearest even, at half the output bytes (not in the reference, whose heatmaps are float32).
    """
    ct, sizes = _tensor_and_sizes(centers)
    rt, _ = _tensor_and_sizes(radii)
    lt = None if labels is None else _tensor_and_sizes(labels)[0]
    if out is None:
        shape = (ct.shape[0], height, width) if lt is None else (ct.shape[0], num_classes, height, width)
        out = torch.empty(shape, dtype=dtype, device=ct.device)
    # use_workspace=True: the two-launch path of the *_ws C-ABI entries (K2 pre-pass into a bin-list workspace from the
    # caching allocator, then the list-fed renderer); bit-identical results, measured slower on the BASELINE configs
    _ext.render_heatmap_batched(out, ct, rt, sizes, lt, diameter_to_sigma_factor, k_scale, True, use_workspace)
    return out


def render_heatmap_from_boxes(centers: torch.Tensor, bboxes: torch.Tensor, out_size_factor: float,
                              sample_sizes: torch.Tensor, height: Optional[int] = None, width: Optional[int] = None,
                              num_classes: Optional[int] = None, labels: Optional[torch.Tensor] = None,
                              diameter_to_sigma_factor: float = 6.0, k_scale: float = 1.0,
                              out: Optional[torch.Tensor] = None, fresh: bool = True,
                              dtype: torch.dtype = torch.float32, use_workspace: bool = False) -> torch.Tensor:
    """Float centres ``[B,T,2]`` + boxes ``[B,T,4]`` -> heatmap, radius/centre computation fused (one launch).
    float16 / bfloat16 heatmaps (``dtype`` or ``out``) are served in fresh mode only."""
    if out is None:
        shape = (centers.shape[0], height, width) if labels is None else (centers.shape[0], num_classes, height, width)
        out = torch.empty(shape, dtype=dtype, device=centers.device)
        fresh = True
    _ext.render_heatmap_from_boxes(out, centers, bboxes, float(out_size_factor), sample_sizes, labels,
                                   diameter_to_sigma_factor, k_scale, fresh, use_workspace)
    return out


def render_heatmap_flat(centers: torch.Tensor, radii: torch.Tensor, heatmap_idxes: torch.Tensor, num_heatmaps: int,
                        height: int, width: int, diameter_to_sigma_factor: float = 6.0, k_scale: float = 1.0,
                        out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fresh-mode variant of ``draw_heatmap`` (concatenated objects + plane index)."""
    if out is None:
        out = torch.empty((num_heatmaps, height, width), dtype=torch.float32, device=centers.device)
    _ext.render_heatmap_flat(out, centers, radii, heatmap_idxes, diameter_to_sigma_factor, k_scale, True)
    return out


def draw_gaussians(heatmap: Optional[torch.Tensor], active, slice_ids, centers, radii, k_for_classes,
                   radius_to_sigma_factor: float, out: Optional[torch.Tensor] = None,
                   shape: Optional[Tuple[int, ...]] = None, fast: Optional[bool] = None) -> torch.Tensor:
    """GPU version of the reference's DALI operator ``draw_gaussians(heatmap, active, slice_ids, centers, radii,
    k_for_classes=..., radius_to_sigma_factor=...)`` (ext_impl/DrawGaussians.cc:88-157; Python call site
    processing_steps/bounding_box_to_heatmap_converter.py:321-399): returns ``max(heatmap, per-class Gaussians)``.

    Inputs are in the RaggedBatch layout instead of DALI's per-sample lists: ``centers`` int32 ``[B,T,2]``, ``radii``
    float32 ``[B,T]``, ``slice_ids`` int32 ``[B,T]`` (or ``None``: class 0), ``active`` bool ``[B,T]`` (or ``None``),
    each a ``RaggedBatch`` (``sample_sizes`` taken from ``radii``) or a plain tensor meaning "all T slots valid".
    ``heatmap`` float32 ``[B,C,H,W]`` / ``[B,H,W]``; ``None`` renders a fresh heatmap of ``shape`` (zero-init fused).
    ``out`` may be ``heatmap`` itself (in place); default: a new tensor, like the operator.
    ``fast`` (fresh rendering only): True / False force / forbid the three-launch path through the main renderer; None
    takes it from 256 MiB of output on. Results are identical either way."""
    r_t, sizes = _tensor_and_sizes(radii)
    c_t, _ = _tensor_and_sizes(centers)
    a_t = None if active is None else _tensor_and_sizes(active)[0]
    s_t = None if slice_ids is None else _tensor_and_sizes(slice_ids)[0]
    if sizes is None:
        sizes = torch.full((r_t.shape[0],), r_t.shape[1], dtype=torch.int64, device=r_t.device)
    if not isinstance(k_for_classes, torch.Tensor):
        k_for_classes = torch.tensor(list(k_for_classes), dtype=torch.float32, device=r_t.device)
    if out is None:
        if heatmap is None:
            assert shape is not None, "`shape` is needed when no input heatmap is given"
            out = torch.empty(tuple(shape), dtype=torch.float32, device=r_t.device)
        else:
            out = torch.empty_like(heatmap)
    return _ext.draw_gaussians(heatmap, out, a_t, s_t, c_t, r_t, sizes, k_for_classes, float(radius_to_sigma_factor),
                               -1 if fast is None else int(bool(fast)))


def boxes_to_gaussian_params(bboxes, scale_xy: torch.Tensor, heatmap_hw: Tuple[int, int], centers=None, classes=None,
                             is_valid=None, radius_scaling_factor: float = 0.8, min_radius: float = 0.0,
                             max_radius: float = float("inf"), min_fraction_area: float = 0.0, num_classes: int = 1,
                             per_class_min_object_sizes=None, min_object_size=None) -> dict:
    """Everything the reference's ``BoundingBoxToHeatmapConverter`` computes between the annotation boxes and
    ``draw_gaussians`` (processing_steps/bounding_box_to_heatmap_converter.py:321-376; ~10 DALI CPU operators per sample
    there), in one launch. ``bboxes`` float32 ``[B,T,4]`` RaggedBatch (image coordinates, corners in any order),
    ``scale_xy`` float32 ``[B,2]`` = heatmap size / image size per sample. Returns RaggedBatches ``bboxes_clipped``,
    ``center`` (int32), ``center_offset``, ``radii`` (float32), ``is_active`` (bool), ``hw_clipped``, ``fraction_area``:
    feed ``is_active`` / ``classes`` / ``center`` / ``radii`` to :func:`draw_gaussians`."""
    b_t, sizes = _tensor_and_sizes(bboxes)
    if sizes is None:
        sizes = torch.full((b_t.shape[0],), b_t.shape[1], dtype=torch.int64, device=b_t.device)
    tens = lambda x: None if x is None else _tensor_and_sizes(x)[0]
    table = None
    if classes is not None and per_class_min_object_sizes is not None:
        table = torch.as_tensor(per_class_min_object_sizes, dtype=torch.float64).reshape(-1, 2).to(b_t.device)
    elif classes is None and min_object_size is not None:
        table = torch.as_tensor(min_object_size, dtype=torch.float64).reshape(1, 2).to(b_t.device)
    outs = _ext.boxes_to_gaussian_params(b_t, tens(centers), tens(classes), tens(is_valid), sizes, scale_xy, table,
                                         int(num_classes), int(heatmap_hw[0]), int(heatmap_hw[1]),
                                         float(radius_scaling_factor), float(min_radius), float(max_radius),
                                         float(min_fraction_area))
    names = ("bboxes_clipped", "center", "center_offset", "radii", "is_active", "hw_clipped", "fraction_area")
    return {k: RaggedBatch(v, sample_sizes=sizes) for k, v in zip(names, outs)}


def gaussian_radius(det_size: torch.Tensor, min_overlap: float = 0.5, variant: str = "cornernet", min_radius: int = 0,
                    return_float: bool = False):
    """CenterNet / CenterPoint ``gaussian_radius(det_size, min_overlap)`` for a float32 CUDA tensor ``[..., 2]`` of
    (height, width) in heatmap pixels -> int32 radii ``max(min_radius, int(r))`` (what CenterPoint's target assigner
    feeds to its ``draw_heatmap_gaussian``), ready for :func:`render_heatmap_batched` / ``draw_heatmap_batched``.
    ``variant``: ``"cornernet"`` (CornerNet / CenterNet / CenterPoint) or ``"mmdet"`` (corrected roots). Not part of the
    reference (its example helper uses the minimum border distance, :func:`radius_center`)."""
    v = {"cornernet": 0, "mmdet": 1}[variant]
    ri, rf = _ext.gaussian_radius(det_size.contiguous(), float(min_overlap), v, int(min_radius))
    return (ri, rf) if return_float else ri


class _GaussianFocalLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, centers, radii, sizes, labels, factor, k_scale, alpha, gamma, eps, fast_math):
        out2, grad = _ext.gaussian_focal_loss(pred.contiguous(), centers, radii, sizes, labels, factor, k_scale, alpha, gamma, eps,
                                              pred.requires_grad, fast_math)
        if pred.requires_grad:
            ctx.save_for_backward(grad)
        ctx.mark_non_differentiable(out2)
        return out2[0].to(torch.float32), out2

    @staticmethod
    def backward(ctx, g_loss, _g_out2):
        (grad,) = ctx.saved_tensors
        if g_loss.is_cuda and g_loss.dtype == torch.float32 and g_loss.numel() == 1:
            out = _ext.scale_by(grad, g_loss.contiguous(), False)  # streaming kernel, the scalar is read on the device
        else:
            out = grad * g_loss
        return (out,) + (None,) * 10


def gaussian_focal_loss(pred: torch.Tensor, centers, radii, diameter_to_sigma_factor: float = 6.0, k_scale: float = 1.0,
                        labels=None, alpha: float = 2.0, gamma: float = 4.0, eps: float = 1e-12, fast_math: bool = False):
    """Fused consumer of the heatmap: ``sum(GaussianFocalLoss(pred, target))`` where ``target`` is
    the heatmap ``draw_heatmap_batched`` would draw from ``centers`` / ``radii`` (/ ``labels``) into zeros — without
    ever writing the target to memory (formula: reference docs/intro.rst:7-26; mmdet's ``gaussian_focal_loss`` with
    ``reduction='sum'``). Returns ``(loss_sum, num_pos)``: a differentiable float32 scalar (the gradient w.r.t. ``pred``
    is produced by the same launch) and the number of ``target == 1`` pixels (the usual ``avg_factor``)."""
    c_t, sizes = _tensor_and_sizes(centers)
    r_t, _ = _tensor_and_sizes(radii)
    l_t = None if labels is None else _tensor_and_sizes(labels)[0]
    if sizes is None:
        sizes = torch.full((r_t.shape[0],), r_t.shape[1], dtype=torch.int64, device=r_t.device)
    loss, out2 = _GaussianFocalLoss.apply(pred, c_t, r_t, sizes, l_t, float(diameter_to_sigma_factor), float(k_scale),
                                          float(alpha), float(gamma), float(eps), bool(fast_math))
    return loss, out2[1].to(torch.int64)


def _flatten(data) -> list:
    """Depth-first flattening of a (nested) sequence of tensors (batched_processing_py.py:373-397)."""
    if isinstance(data, torch.Tensor):
        return [data]
    if isinstance(data, Sequence) and not isinstance(data, str):
        out = []
        for el in data:
            out += _flatten(el)
        return out
    raise AssertionError(f"The data to combine must be a tensor or a (nested) sequence of tensors. Got {type(data)}.")


def combine_data(data_list: Sequence[Union[Sequence, torch.Tensor]], other_with_same_sample_sizes: Optional[RaggedBatch] = None,
                 device: Optional[Union[torch.device, str]] = None, flatten_batch_dims: bool = True,
                 max_sample_size: Optional[int] = None) -> RaggedBatch:
    """``accvlab.batching_helpers.combine_data`` for CUDA inputs (batched_processing_py.py:275-438, flattening mode):
    a (nested) sequence of per-sample tensors ``[n_i, *data]`` -> ``RaggedBatch`` with ``tensor [B, max n_i, *data]``
    (zero padded) and int64 ``sample_sizes``. One kernel launch + one small H2D copy instead of one slice-assign per
    sample. ``max_sample_size`` (extension) pads to a fixed length. There is no CPU path: every element must live on
    the target CUDA device."""
    assert isinstance(data_list, Sequence), "`data_list` must be a sequence"
    assert len(data_list) > 0, "`data_list` must not be empty"
    if not flatten_batch_dims:
        raise NotImplementedError("only flatten_batch_dims=True (one batch dimension) feeds draw_heatmap")
    flat = _flatten(data_list)
    if device is not None:
        dev = torch.device(device)
        if dev.type == "cuda" and dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        for el in flat:
            if el.device != dev:
                raise RuntimeError(f"combine_data: element on {el.device}, expected {dev} (no CPU path, no implicit copies)")
    tensor, sizes = _ext.combine_data([el.contiguous() for el in flat], -1 if max_sample_size is None else int(max_sample_size))
    if other_with_same_sample_sizes is not None:
        assert len(flat) == other_with_same_sample_sizes.sample_sizes.shape[0], \
            "Number of samples does not match `other_with_same_sample_sizes`"
        assert tensor.shape[1] == other_with_same_sample_sizes.tensor.shape[1], \
            "Needed mask dimension does not match `other_with_same_sample_sizes`"
        return RaggedBatch(tensor, mask=other_with_same_sample_sizes._mask, sample_sizes=other_with_same_sample_sizes.sample_sizes)
    return RaggedBatch(tensor, sample_sizes=sizes)


def pack_radius_center(centers_list: Sequence[torch.Tensor], bboxes_list: Sequence[torch.Tensor], out_size_factor: float,
                       max_sample_size: Optional[int] = None) -> Tuple[RaggedBatch, RaggedBatch]:
    """Per-sample float centres ``[n_i,2]`` and boxes ``[n_i,4]`` -> (``centers`` int32 ``[B,T,2]``, ``radii`` int32
    ``[B,T]``) RaggedBatches sharing one int64 ``sample_sizes``: combine_data + get_centers_and_radii
    (DH/tests/_test_helpers.py:20-28) in one launch, ready for ``draw_heatmap_batched``."""
    c, r, sizes = _ext.pack_radius_center([t.contiguous() for t in centers_list], [t.contiguous() for t in bboxes_list],
                                          float(out_size_factor), -1 if max_sample_size is None else int(max_sample_size))
    return RaggedBatch(c, sample_sizes=sizes), RaggedBatch(r, sample_sizes=sizes)


def shard_range(batch_size: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous slice ``[lo, hi)`` of the (flattened batch x camera) dimension owned by ``rank``."""
    if not (0 <= rank < world_size):
        raise ValueError("rank must be in [0, world_size)")
    return batch_size * rank // world_size, batch_size * (rank + 1) // world_size


def launch_count() -> int:
    return int(_ext.launch_count())


def fill_probe(t: torch.Tensor) -> None:
    _ext.fill_probe(t)
