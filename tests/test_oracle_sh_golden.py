"""The oracle's SH -> RGB stage (degrees 0..3, the +0.5 offset and the clamp at 0 with its clamp bits) against colours
computed by the REFERENCE'S OWN correct evaluator: the vendored `eval_sh` of
gs_lightning/third_party/gaussian_splatting/utils/sh_utils.py:57-112, run in float64 by tests/golden/make_sh_golden.py
(the in-tree `utils/sh.py apply_sh` reads degree-4 coefficients at degree 3, SURVEY.md Appendix B, which is why the
image goldens of degree 3 are zero padded; this fixture pins degree 3 without that detour)."""
import os

import numpy as np
import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
G = np.load(os.path.join(HERE, "golden", "sh_eval_vendored.npz"))


def _cam(deg):
    return oracle.Camera(int(G["image_height"]), int(G["image_width"]), float(G["tanfovx"]), float(G["tanfovy"]),
                         G["viewmatrix"], G["projmatrix"], G["campos"], bg=np.zeros(3, np.float32), sh_degree=deg)


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_oracle_sh_colours_equal_the_references_vendored_eval_sh(deg):
    st = oracle.forward(_cam(deg), G["means3D"], G["scales"], G["rotations"], G["opacities"], shs=G["shs"],
                        stop_after="preprocess")
    vis = st["radii"] > 0
    assert vis.sum() >= 0.95 * len(vis)                    # the fixture keeps every Gaussian inside the frustum
    raw = G[f"raw_deg{deg}"]                               # float64, before the clamp
    want = np.maximum(raw, 0.0)
    got = st["rgb"].astype(np.float64)
    # float32 evaluation of up to 16 terms of magnitude <= ~2: 1e-6 absolute covers it with margin
    assert np.abs(got[vis] - want[vis]).max() <= 2e-6, np.abs(got[vis] - want[vis]).max()
    # clamp bits: wherever the float64 value is clearly on one side of zero
    clear = np.abs(raw) > 1e-5
    assert np.array_equal((st["clamped"] != 0)[vis][clear[vis]], (raw < 0)[vis][clear[vis]])
    assert (raw < 0)[vis].mean() > (0.03 if deg else -1)   # the clamp is exercised (not at degree 0: 0.5 + C0 * sh)


@pytest.mark.gpu
@pytest.mark.parametrize("deg", [0, 3])
def test_cuda_sh_colours_equal_the_references_vendored_eval_sh(deg):
    """the same comparison for the CUDA preprocess kernel (colours exported from the geometry buffer)"""
    import torch

    import gpu_utils as U
    arrays = dict(means3D=G["means3D"], scales=G["scales"], rotations=G["rotations"], opacities=G["opacities"], shs=G["shs"])
    out = U.run_cuda_forward(_cam(deg), arrays)
    st = U.export(out)
    vis = out["radii"].cpu().numpy() > 0
    want = np.maximum(G[f"raw_deg{deg}"], 0.0)
    assert np.abs(st["rgb"].astype(np.float64)[vis] - want[vis]).max() <= 2e-6
    assert torch.isfinite(out["color"]).all()
