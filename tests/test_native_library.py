"""CPU-side checks of the boundary: the C-ABI library builds, loads and exports every symbol that
include/gsr_b200.h declares (no compute calls -- there is no GPU here)."""
import os
import re

from conftest import ROOT


def test_library_exports_every_declared_symbol():
    from gaussian_splatting_lightning_b200 import _native as N
    from gaussian_splatting_lightning_b200 import build
    build.build()
    L = N.lib()
    hdr = open(os.path.join(ROOT, "include", "gsr_b200.h")).read()
    declared = set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", hdr)) - {"gsr_alloc_fn"}
    assert declared == set(N.EXPORTED_SYMBOLS)
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in L.gsr_version()


def test_scratch_size_queries_are_consistent():
    from gaussian_splatting_lightning_b200 import _native as N
    L = N.lib()
    assert L.gsr_geom_bytes(1_000_000) >= 1_000_000 * (48 + 24 + 1 + 4 + 4)
    assert L.gsr_img_bytes(1920, 1080) >= 1920 * 1080 * 8 + 8160 * 8
    assert L.gsr_binning_bytes(3_000_000, 1920, 1080) >= 3_000_000 * 24
    assert L.gsr_binning_bytes(0, 16, 16) > 0
    assert L.gsr_backward_scratch_bytes(10) >= 480


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gaussian_splatting_lightning_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "libgsr_oracle" not in src, f


def test_cpu_tensors_are_rejected_loudly():
    import pytest
    import torch
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    z = torch.zeros
    rs = GaussianRasterizationSettings(8, 8, 1.0, 1.0, z(3), 1.0, torch.eye(4), torch.eye(4), 0, z(3), False, False, False)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        GaussianRasterizer(rs)(means3D=z(4, 3), means2D=None, opacities=z(4, 1), shs=z(4, 1, 3), scales=z(4, 3),
                               rotations=z(4, 4))


def test_torch_extension_builds_loads_and_has_upstreams_entry_points():
    """`_C` (csrc/torch_ext.cpp): the thin torch C++ extension north_star names as the FFI layer, with the entry points
    of upstream's diff_gaussian_rasterization._C; reachable under the reference's import name too."""
    import pytest
    import torch
    from gaussian_splatting_lightning_b200 import _native as N
    from gaussian_splatting_lightning_b200 import build
    build.build()
    build.build_torch_ext()
    C_ = N.ext()
    assert C_ is not None and os.path.exists(N.EXT_PATH)
    for fn in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible", "rasterize_gaussians_ex",
               "rasterize_gaussians_backward_ex"):
        assert callable(getattr(C_, fn)), fn
    assert "sm_100a" in C_.version()
    import diff_gaussian_rasterization
    assert diff_gaussian_rasterization._C is C_
    # upstream's positional order: 20 arguments, num_rendered first in the returned tuple
    doc = C_.rasterize_gaussians.__doc__
    assert doc.count("arg") == 20 and "tuple[int, torch.Tensor" in doc.replace("typing.", "")
    assert C_.rasterize_gaussians_backward.__doc__.count("arg") == 24
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        C_.mark_visible(torch.zeros(4, 3), torch.eye(4), torch.eye(4))
    z = torch.zeros
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        C_.rasterize_gaussians(z(3), z(4, 3), None, z(4, 1), z(4, 3), z(4, 4), 1.0, None, torch.eye(4), torch.eye(4), 1.0,
                               1.0, 8, 8, z(4, 1, 3), 0, z(3), False, False, False)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        C_.rasterize_gaussians(z(3), z(4, 2), None, z(4, 1), z(4, 3), z(4, 4), 1.0, None, torch.eye(4), torch.eye(4), 1.0,
                               1.0, 8, 8, z(4, 1, 3), 0, z(3), False, False, False)
