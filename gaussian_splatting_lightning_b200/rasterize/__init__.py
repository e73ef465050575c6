"""B200 twin of the reference's `gs_lightning.rasterize` Python entry points
(gs_lightning/rasterize/__init__.py:1-2): same function names, argument names and return
shapes as rasterize_gaussian (rasterize.py:28-47,125-127) and markVisible (rasterize.py:23-26),
computed by the CUDA rasterizer (libgsr_b200.so) instead of the pure-PyTorch loops.
"""
from .rasterize import markVisible, rasterize_gaussian  # noqa: F401
