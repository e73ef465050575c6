#!/usr/bin/env python
"""Mints tests/golden/sh_eval_vendored.npz: colours of the reference's OWN correct spherical-harmonics evaluation --
the vendored `eval_sh` of gs_lightning/third_party/gaussian_splatting/utils/sh_utils.py:57-112 (the in-tree
`utils/sh.py:41-98 apply_sh` reads degree-4 coefficients at degree 3: SURVEY.md Appendix B) -- for degrees 0..3, in
float64, on random coefficients and view directions.  The oracle's preprocess stage is pinned against them on the CPU
(tests/test_oracle_sh_golden.py); the CUDA path is compared with the oracle.

Run in the build container (imports /root/reference through ref_loader):  python tests/golden/make_sh_golden.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from ref_loader import load_vendored_eval_sh  # noqa: E402

from gaussian_splatting_lightning_b200 import synthetic  # noqa: E402


def main():
    eval_sh = load_vendored_eval_sh()
    W, H, P = 96, 64, 300
    cam = synthetic.make_camera(W, H)
    rng = np.random.default_rng(20261105)
    # Gaussians in front of the camera, well inside the frustum (every one visible), small and opaque
    view = cam["viewmatrix"].numpy().astype(np.float64)          # row-vector convention: p_view = [p, 1] @ view
    z = rng.uniform(2.0, 6.0, P)
    x = rng.uniform(-0.5, 0.5, P) * z * cam["tanfovx"]
    y = rng.uniform(-0.5, 0.5, P) * z * cam["tanfovy"]
    p_view = np.stack([x, y, z, np.ones(P)], 1)
    means = (p_view @ np.linalg.inv(view))[:, :3].astype(np.float32)
    shs = (rng.standard_normal((P, 16, 3)) * 0.6).astype(np.float32)     # about a third of the channels clamp at 0
    campos = cam["campos"].numpy().astype(np.float64)
    d = means.astype(np.float64) - campos
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    out = dict(image_width=W, image_height=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
               viewmatrix=cam["viewmatrix"].numpy(), projmatrix=cam["projmatrix"].numpy(), campos=cam["campos"].numpy(),
               means3D=means, shs=shs, scales=np.full((P, 3), 0.02, np.float32),
               rotations=np.tile(np.array([1, 0, 0, 0], np.float32), (P, 1)), opacities=np.full(P, 0.9, np.float32))
    for deg in range(4):
        n = (deg + 1) ** 2
        sh_t = torch.tensor(shs[:, :n, :], dtype=torch.float64).transpose(1, 2)      # (P, 3, n) as eval_sh wants it
        raw = eval_sh(deg, sh_t, torch.tensor(d)).numpy() + 0.5                      # gaussian_renderer convention
        out[f"raw_deg{deg}"] = raw
    np.savez_compressed(os.path.join(HERE, "sh_eval_vendored.npz"), **out)
    print("wrote sh_eval_vendored.npz;", {k: float(np.mean(out[f"raw_deg{k}"] < 0)) for k in range(4)})


if __name__ == "__main__":
    main()
