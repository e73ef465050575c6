"""Pin the oracle's analytic backward with float64 torch autograd.

The reference has no backward for this path in-tree (the Python rasterizer is forward-only
and the CUDA submodule is missing), so the analytic gradients of oracle/gsr_oracle.cpp are
checked against autograd through tests/torch_restatement.py, a float64 restatement of the
reference's Python forward.  Tolerance: 1e-4 relative plus 2e-5 of the tensor's largest
gradient (FP32 per-pixel arithmetic against an FP64 truth; sums of O(100) terms cancel).
"""
import numpy as np
import pytest
import torch

import torch_restatement as tr
from conftest import load_golden

PAIRS = [("means3D", "dL_dmeans3D"), ("scales", "dL_dscales"), ("rotations", "dL_drotations"),
         ("opacities", "dL_dopacity"), ("shs", "dL_dsh")]


def _check(oracle, name, aa=False, scale_modifier=1.0, with_invdepth=True):
    d, cam = load_golden(name)
    cam.antialiasing = aa
    cam.scale_modifier = scale_modifier
    st = oracle.forward(cam, d["means3D"], d["scales"], d["rotations"], d["opacities"], d["shs"])
    rng = np.random.default_rng(0)
    H, W = cam.image_height, cam.image_width
    dLc = rng.standard_normal((3, H, W)).astype(np.float32)
    dLd = rng.standard_normal((1, H, W)).astype(np.float32) if with_invdepth else None
    g = oracle.backward(st, dLc, dLd)

    t = {k: torch.tensor(d[k], dtype=torch.float64, requires_grad=True)
         for k in ["means3D", "scales", "rotations", "opacities", "shs"]}
    delta = torch.zeros(len(d["means3D"]), 3, dtype=torch.float64, requires_grad=True)
    col, dep = tr.render(cam, st, t["means3D"], t["scales"], t["rotations"], t["opacities"], t["shs"],
                         means2D_delta=delta)
    assert np.abs(col.detach().numpy() - st["color"]).max() < 2e-5
    loss = (col * torch.tensor(dLc, dtype=torch.float64)).sum()
    if dLd is not None:
        loss = loss + (dep * torch.tensor(dLd, dtype=torch.float64)).sum()
    loss.backward()
    pairs = [(t[k].grad.numpy(), g[gk].astype(np.float64), k) for k, gk in PAIRS]
    pairs.append((delta.grad.numpy()[:, :2], g["dL_dmeans2D"][:, :2].astype(np.float64), "means2D"))
    for a, b, k in pairs:
        tol = 1e-4 * np.abs(a) + 2e-5 * np.abs(a).max()
        assert (np.abs(a - b) <= tol).all(), k
        assert np.abs(a).max() > 0, k


@pytest.mark.parametrize("name", ["g2_deg3_100x70", "g4_saturated_64x48"])
def test_backward_matches_autograd(oracle_lib, name):
    _check(oracle_lib, name)


def test_backward_antialiasing_and_scale_modifier(oracle_lib):
    _check(oracle_lib, "g2_deg3_100x70", aa=True, scale_modifier=1.3)


def test_backward_without_invdepth_grad(oracle_lib):
    _check(oracle_lib, "g1_deg0_128", with_invdepth=False)
