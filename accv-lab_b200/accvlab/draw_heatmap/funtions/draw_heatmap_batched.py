"""draw_heatmap_batched — RaggedBatch front end.

Same signature, defaults, assertions and in-place semantics as the reference wrapper
(/root/reference/packages/draw_heatmap/accvlab/draw_heatmap/funtions/draw_heatmap_batched.py:27-84).
Differences: ``sample_sizes`` is handed to the extension as it is (int32 or int64); the reference casts it to int32 on
every call (:62-63, which carries a TODO asking for exactly this). ``assume_zero_init`` is a keyword-only addition
(default False = the reference's in-place max-blend).

The signature, docstring structure and the four assertion messages follow the upstream wrapper, which is
Copyright (c) 2025, NVIDIA CORPORATION & AFFILIATES and licensed under the Apache License, Version 2.0
(http://www.apache.org/licenses/LICENSE-2.0); see NOTICE at the repository root.
"""
from __future__ import annotations

from typing import TYPE_CHECKING

import torch

from accvlab.draw_heatmap.draw_heatmap_ext import draw_heatmap_batched_classwise_impl, draw_heatmap_batched_impl

if TYPE_CHECKING:  # duck-typed: anything with .tensor and .sample_sizes works
    from accvlab.batching_helpers import RaggedBatch


def _sizes(sample_sizes: torch.Tensor) -> torch.Tensor:
    if sample_sizes.dtype in (torch.int32, torch.int64):
        return sample_sizes
    return sample_sizes.to(torch.int32)  # what the reference does for every dtype


def draw_heatmap_batched(
    heatmap: torch.Tensor,
    centers: "RaggedBatch",
    radii: "RaggedBatch",
    diameter_to_sigma_factor: float = 6.0,
    k_scale: float = 1.0,
    labels: "RaggedBatch" = None,
    *,
    assume_zero_init: bool = False,
):
    """Draw heatmaps for a batch of samples, in place (max-blend with the existing contents).

    Args:
        heatmap: float32 ``(batch_size, height, width)`` when ``labels`` is None, otherwise
            ``(batch_size, max_num_classes, height, width)``. Modified in place.
        centers: RaggedBatch of int32 ``(batch_size, max_num_targets, 2)`` — (x, y) centres.
        radii: RaggedBatch of int32 ``(batch_size, max_num_targets)``.
        diameter_to_sigma_factor: ``sigma = (2*radius+1) / diameter_to_sigma_factor``.
        k_scale: scale of the Gaussian peak.
        labels: optional RaggedBatch of int32 ``(batch_size, max_num_targets)`` class indices; if None all objects of
            a sample are drawn into one plane.
        assume_zero_init: keyword-only, not in the reference. True = the caller promises the heatmap holds zeros (or
            nothing worth keeping): every pixel is written exactly once with the zero-init fused into the render
            (no read of the old contents, no separate ``torch.zeros`` needed).
    """
    centers_tensor = centers.tensor
    radii_tensor = radii.tensor
    assert centers_tensor.shape[0] == radii_tensor.shape[0], "centers and radii must have the same size batch size"
    assert (
        centers_tensor.shape[1] == radii_tensor.shape[1]
    ), "centers and radii must have the same maximum number of objects"
    nums_targets = _sizes(centers.sample_sizes)
    if labels is None:
        draw_heatmap_batched_impl(heatmap, centers_tensor, radii_tensor, nums_targets, diameter_to_sigma_factor, k_scale,
                                  assume_zero_init=assume_zero_init)
    else:
        labels_tensor = labels.tensor
        assert (
            centers_tensor.shape[0] == labels_tensor.shape[0]
        ), "centers and labels must have the same size batch size"
        assert (
            centers_tensor.shape[1] == labels_tensor.shape[1]
        ), "centers and labels must have the same maximum number of objects"
        draw_heatmap_batched_classwise_impl(
            heatmap, centers_tensor, radii_tensor, nums_targets, labels_tensor, diameter_to_sigma_factor, k_scale,
            assume_zero_init=assume_zero_init,
        )
