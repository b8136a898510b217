"""accvlab.draw_heatmap — B200-native (sm_100a) drop-in for ACCV-Lab's Gaussian heatmap drawing.

Public surface identical to the reference package
(/root/reference/packages/draw_heatmap/accvlab/draw_heatmap/__init__.py:15-24):
``draw_heatmap``, ``draw_heatmap_batched``, ``__version__``.  Additions that are not part of the reference
(fresh/overwrite mode, fused bbox->radius, sharding helpers) live in :mod:`accvlab.draw_heatmap.b200`.

There is no CPU fallback: importing this package without the compiled extension raises ImportError.
"""
__version__ = "0.1.0"

from .funtions import draw_heatmap, draw_heatmap_batched

__all__ = ["__version__", "draw_heatmap", "draw_heatmap_batched"]
