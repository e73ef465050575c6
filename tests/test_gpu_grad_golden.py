"""The CUDA backward (through the C-ABI) against the REFERENCE-DERIVED gradient goldens: float64 central differences
of a fixed linear loss through the reference's own Python rasterizer (tests/golden/*_grads.npz, generator
tests/golden/make_grad_golden.py).  Tolerance 1e-4 relative + 1e-4 of the tensor's largest derivative
(tests/grad_golden.py)."""
import numpy as np
import pytest

import gpu_utils as U
import grad_golden as GG
from conftest import load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GG.GRAD_GOLDENS)
def test_cuda_backward_matches_reference_finite_differences(name):
    g, d, w_img, w_dep = GG.load(name)
    _, cam = load_golden(name)
    arrays = dict(means3D=d["means3D"], scales=d["scales"], rotations=d["rotations"], opacities=d["opacities"],
                  shs=d["shs"])
    out = U.run_cuda_forward(cam, arrays)
    st = U.export(out)
    assert ((st["final_T"] >= 1e-3) == g["mask"]).mean() > 0.9999
    got = U.run_cuda_backward(out, w_img, w_dep)
    wd_img, wd_dep = GG.load_dir_weights(name)
    GG.compare(name, got, U.run_cuda_backward(out, wd_img, wd_dep))
