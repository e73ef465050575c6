"""gaussian_splatting_lightning_b200 -- B200-native 3D Gaussian Splatting rasterizer.

Drop-in for the rasterizer hot path of pomelyu/gaussian_splatting_lightning:
`diff_gaussian_rasterization.{GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians}`
(see rasterizer.py) and the `gs_lightning.rasterize` entry points (see rasterize/).
Compute lives in libgsr_b200.so (csrc/, hand-written CUDA for sm_100a) behind the C-ABI of
include/gsr_b200.h.  Importing this package does not need a GPU; calling it does.
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                         _RasterizeGaussians, rasterize_gaussians)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians"]
