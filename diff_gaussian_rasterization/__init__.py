"""Import-name shim: `from diff_gaussian_rasterization import GaussianRasterizationSettings,
GaussianRasterizer` (gs_lightning/lightning/gs_lightning_module.py:8-9,
scripts/render_trained_image.py:10-11, tests/rasterizer_python/test_cases.py:2) resolves to the
B200 rasterizer when the repository root is on sys.path."""
from gaussian_splatting_lightning_b200.rasterizer import (GaussianRasterizationSettings,  # noqa: F401
                                                          GaussianRasterizer, _RasterizeGaussians,
                                                          rasterize_gaussians)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians"]


def __getattr__(name):
    # `diff_gaussian_rasterization._C`: the torch C++ extension with upstream's entry points (rasterize_gaussians,
    # rasterize_gaussians_backward, mark_visible), loaded on first use (csrc/torch_ext.cpp)
    if name == "_C":
        from gaussian_splatting_lightning_b200 import _native
        m = _native.ext()
        if m is None:
            raise ImportError("diff_gaussian_rasterization._C is not available (GSR_FFI=ctypes or _C.so not built)")
        return m
    raise AttributeError(name)
