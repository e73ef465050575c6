"""The drop-in Python API on a GPU: GaussianRasterizationSettings / GaussianRasterizer /
rasterize_gaussians autograd Function used the way the reference's call sites use them
(gs_lightning/lightning/gs_lightning_module.py:316-348, scripts/render_trained_image.py:80-124,
tests/rasterizer_python/test_mark_visible.py:11-21)."""
import numpy as np
import pytest
import torch

import gpu_utils as U

pytestmark = pytest.mark.gpu


def _scene(P=6000, deg=3, W=160, H=96, seed=11):
    cam, arrays = U.synthetic_case(P, deg, W, H, seed=seed, bg=(0.1, 0.2, 0.3))
    return cam, arrays, U.to_dev(arrays)


def test_training_call_site_autograd(oracle_lib):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    cam, arrays, d = _scene()
    rs = U.settings_from_camera(cam)
    # tanfov arrive as 0-d float64 tensors at the training call site
    rs = rs._replace(tanfovx=torch.tensor(cam.tanfovx, dtype=torch.float64),
                     tanfovy=torch.tensor(cam.tanfovy, dtype=torch.float64))
    assert isinstance(rs, GaussianRasterizationSettings)
    params = {k: d[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    screenspace_points = torch.zeros_like(params["means3D"], requires_grad=True)
    rasterizer = GaussianRasterizer(raster_settings=rs)
    color, radii, invdepth = rasterizer(means3D=params["means3D"], means2D=screenspace_points, shs=params["shs"],
                                        colors_precomp=None, opacities=params["opacities"], scales=params["scales"],
                                        rotations=params["rotations"], cov3D_precomp=None)
    assert color.shape == (3, cam.image_height, cam.image_width) and invdepth.shape == (1, cam.image_height, cam.image_width)
    assert radii.dtype == torch.int32 and radii.shape == (6000,) and not radii.requires_grad
    rng = np.random.default_rng(0)
    wC = rng.standard_normal(color.shape).astype(np.float32)
    wD = rng.standard_normal(invdepth.shape).astype(np.float32)
    loss = (color * torch.tensor(wC, device=U.DEV)).sum() + (invdepth * torch.tensor(wD, device=U.DEV)).sum()
    loss.backward()
    ref = U.run_oracle_forward(cam, arrays)
    want = oracle_lib.backward(ref, wC, wD[0])
    pairs = [("means3D", "dL_dmeans3D"), ("shs", "dL_dsh"), ("opacities", "dL_dopacity"), ("scales", "dL_dscales"),
             ("rotations", "dL_drotations")]
    for p, k in pairs:
        w = want[k]
        g = params[p].grad.cpu().numpy().reshape(w.shape)
        scale = max(1.0, float(np.abs(w).max()))
        frac, mx = U.close_fraction(g / scale, w / scale, 1e-4, 1e-3)
        assert frac <= 2e-3 and mx <= 5e-2, (k, frac, mx)
    # the dummy means2D leaf receives dL_dmean2D; densification reads [:, :2] (gaussian_model.py:180)
    w = want["dL_dmeans2D"]
    g = screenspace_points.grad.cpu().numpy()
    scale = max(1.0, float(np.abs(w).max()))
    frac, mx = U.close_fraction(g / scale, w / scale, 1e-4, 1e-3)
    assert frac <= 2e-3 and mx <= 5e-2
    assert float(np.abs(g[:, 2]).max()) == 0.0


def test_color_only_loss_and_no_grad_render():
    from diff_gaussian_rasterization import GaussianRasterizer
    cam, arrays, d = _scene(P=2000, deg=1, W=64, H=48)
    rs = U.settings_from_camera(cam)
    m = d["means3D"].clone().requires_grad_(True)
    color, radii, invdepth = GaussianRasterizer(rs)(means3D=m, means2D=None, shs=d["shs"], colors_precomp=None,
                                                    opacities=d["opacities"], scales=d["scales"],
                                                    rotations=d["rotations"], cov3D_precomp=None)
    color.mean().backward()  # invdepth gets no gradient: dL_dinvdepth = NULL path
    assert torch.isfinite(m.grad).all() and float(m.grad.abs().sum()) > 0
    with torch.no_grad():  # scripts/render_trained_image.py:113-124
        c2, r2, d2 = GaussianRasterizer(rs)(means3D=d["means3D"], means2D=None, shs=d["shs"], colors_precomp=None,
                                            opacities=d["opacities"], scales=d["scales"], rotations=d["rotations"],
                                            cov3D_precomp=None)
    assert torch.equal(c2, color.detach()) and torch.equal(r2, radii)


def test_argument_validation_errors():
    from diff_gaussian_rasterization import GaussianRasterizer
    cam, arrays, d = _scene(P=100, deg=0, W=32, H=32)
    r = GaussianRasterizer(U.settings_from_camera(cam))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=d["means3D"], means2D=None, opacities=d["opacities"], scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=d["means3D"], means2D=None, opacities=d["opacities"], shs=d["shs"])
    with pytest.raises(RuntimeError, match="num_points, 3"):
        r(means3D=d["means3D"][:, :2], means2D=None, opacities=d["opacities"], shs=d["shs"], scales=d["scales"],
          rotations=d["rotations"])
    with pytest.raises(RuntimeError, match="CUDA"):
        r(means3D=d["means3D"].cpu(), means2D=None, opacities=d["opacities"], shs=d["shs"], scales=d["scales"],
          rotations=d["rotations"])


def test_gs_lightning_rasterize_entry_points(oracle_lib):
    from gaussian_splatting_lightning_b200.rasterize import markVisible, rasterize_gaussian
    cam, arrays, d = _scene(P=4000, deg=2, W=100, H=60)
    rs = U.settings_from_camera(cam)
    img, radius, depth = rasterize_gaussian(
        means3D=d["means3D"], opacities=d["opacities"], scales=d["scales"], rotations=d["rotations"], shs=d["shs"],
        scale_modifier=1.0, image_width=cam.image_width, image_height=cam.image_height, tanfovx=cam.tanfovx,
        tanfovy=cam.tanfovy, viewmatrix=rs.viewmatrix, projmatrix=rs.projmatrix, campos=rs.campos,
        background=rs.bg, sh_degree=2)
    ref = U.run_oracle_forward(cam, arrays)
    assert radius.dtype == torch.float32
    assert np.array_equal(radius.cpu().numpy().astype(np.int32), ref["radii"])
    frac, mx = U.close_fraction(img.cpu().numpy(), ref["color"])
    assert frac <= 2e-4 and mx <= 2e-2
    vis = markVisible(d["means3D"], rs.viewmatrix, rs.projmatrix)
    assert np.array_equal(vis.cpu().numpy(), oracle_lib.mark_visible(arrays["means3D"], cam.viewmatrix))
