from __future__ import annotations

import torch

from ..rasterizer import GaussianRasterizationSettings, forward_raw, mark_visible


def markVisible(means3D: torch.Tensor, viewmatrix: torch.Tensor, projmatrix: torch.Tensor) -> torch.Tensor:
    """visible[i] = ([p_i,1] @ viewmatrix).z > 0.2   (gs_lightning/rasterize/rasterize.py:23-26)."""
    return mark_visible(means3D, viewmatrix, projmatrix)


@torch.no_grad()
def rasterize_gaussian(
    means3D: torch.Tensor,
    opacities: torch.Tensor,
    scales: torch.Tensor,
    rotations: torch.Tensor,
    shs: torch.Tensor,
    scale_modifier: float,
    image_width: int,
    image_height: int,
    tanfovx: float,
    tanfovy: float,
    viewmatrix: torch.Tensor,
    projmatrix: torch.Tensor,
    campos: torch.Tensor,
    background: torch.Tensor,
    sh_degree: int,
):
    """Forward-only render with the reference's keyword signature
    (gs_lightning/rasterize/rasterize.py:28-47; called at scripts/render_trained_image.py:80-96).

    Returns (image (3,H,W), radius (P,) float, depth (1,H,W)) like rasterize.py:125-127; radius is
    float there, so it is cast here (0 for culled Gaussians).
    """
    rs = GaussianRasterizationSettings(
        image_height=int(image_height), image_width=int(image_width), tanfovx=float(tanfovx),
        tanfovy=float(tanfovy), bg=background, scale_modifier=float(scale_modifier), viewmatrix=viewmatrix,
        projmatrix=projmatrix, sh_degree=int(sh_degree), campos=campos, prefiltered=False, debug=False,
        antialiasing=False)
    _, color, radii, _, _, _, invdepth, _ = forward_raw(rs, means3D, shs, None, opacities, scales, rotations, None)
    return color, radii.to(torch.float32), invdepth
