"""Import-name shim: `from diff_gaussian_rasterization import GaussianRasterizationSettings,
GaussianRasterizer` (gs_lightning/lightning/gs_lightning_module.py:8-9,
scripts/render_trained_image.py:10-11, tests/rasterizer_python/test_cases.py:2) resolves to the
B200 rasterizer when the repository root is on sys.path."""
from gaussian_splatting_lightning_b200.rasterizer import (GaussianRasterizationSettings,  # noqa: F401
                                                          GaussianRasterizer, _RasterizeGaussians,
                                                          rasterize_gaussians)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians"]
