"""Mirror of the reference sub-package (its name is misspelt upstream and kept for drop-in imports):
/root/reference/packages/draw_heatmap/accvlab/draw_heatmap/funtions/__init__.py:15-30."""
try:
    from accvlab.draw_heatmap.draw_heatmap_ext import draw_heatmap
except ImportError as e:  # fail loudly: the CUDA extension IS the product
    raise ImportError(
        "accvlab.draw_heatmap: the compiled extension 'draw_heatmap_ext' (and libdh_b200.so) is missing or cannot be "
        "loaded. Build it in-tree with `python accv-lab_b200/build.py` (or __graft_entry__.build()). "
        "There is no CPU/PyTorch fallback. Original error: " + str(e)
    ) from e

from .draw_heatmap_batched import draw_heatmap_batched

__all__ = ["draw_heatmap", "draw_heatmap_batched"]
