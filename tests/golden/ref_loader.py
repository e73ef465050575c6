"""Import the reference's Python rasterizer from /root/reference WITHOUT copying it.

Only usable in the build container (the GPU box has no /root/reference).  Used by
make_golden.py to mint the committed fixtures and by tests that are skipped when the
reference tree is absent.

Two things stand between `import gs_lightning.rasterize` and a clean import:
  * gs_lightning/__init__.py:1-4 pulls in pytorch_lightning / pycolmap / mlconfig, none of
    which exist here -> we register a stub parent package whose __path__ points at the
    reference directory, so only the sub-modules we ask for are executed;
  * render_tools.py:3,61 needs kornia solely for Quaternion(q).matrix() -> a stand-in with
    kornia's behaviour (normalise, then the (w,x,y,z) matrix of
    third_party/gaussian_splatting/utils/general_utils.py:78-99).
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("GSR_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "gs_lightning", "rasterize"))


def _install_kornia_shim():
    if "kornia" in sys.modules:
        return
    import torch

    class Quaternion:
        def __init__(self, data):
            self.data = data

        def matrix(self):
            q = torch.nn.functional.normalize(self.data, dim=-1)
            r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
            R = torch.stack([
                1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1)
            return R.reshape(-1, 3, 3)

    kornia = types.ModuleType("kornia")
    kornia.geometry = types.ModuleType("kornia.geometry")
    kornia.geometry.Quaternion = Quaternion
    sys.modules["kornia"] = kornia
    sys.modules["kornia.geometry"] = kornia.geometry


def load():
    """Returns the reference module gs_lightning.rasterize (rasterize_gaussian, markVisible)."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    _install_kornia_shim()
    for name in ("gs_lightning", "gs_lightning.utils"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = [os.path.join(REF_ROOT, *name.split("."))]
            sys.modules[name] = m
    # silence tqdm (rasterize.py:107)
    import tqdm as _tqdm

    class _Quiet:
        def __init__(self, *a, **k): pass
        def update(self, *a): pass
        def close(self): pass
    mod = importlib.import_module("gs_lightning.rasterize.rasterize")
    mod.tqdm = _Quiet
    return importlib.import_module("gs_lightning.rasterize")


def load_vendored_eval_sh():
    """third_party/gaussian_splatting/utils/sh_utils.py:57-112 (correct degree-3 SH)."""
    import importlib.util
    p = os.path.join(REF_ROOT, "gs_lightning", "third_party", "gaussian_splatting", "utils", "sh_utils.py")
    spec = importlib.util.spec_from_file_location("_ref_sh_utils", p)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.eval_sh
