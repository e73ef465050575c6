"""Synthetic scenes and cameras for the benchmark and the parity tests.

The generator is the one fixed in SURVEY.md section 8(d) so that builder, tests, bench
and judge all talk about the same workload.  The camera follows the reference's own
conventions: projection matrix from gs_lightning/utils/camera.py:26-41, row-vector
view / full-projection matrices as built in gs_lightning/datasets/colmap_dataset.py:85-98.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
import torch

ZNEAR, ZFAR = 0.01, 100.0
SCENE_CENTRE = (0.0, 0.0, 8.0)


def projection_matrix(fx: float, fy: float, w: int, h: int, znear: float = ZNEAR, zfar: float = ZFAR) -> np.ndarray:
    """Column-vector perspective matrix, z mapped to [0,1] (utils/camera.py:26-41)."""
    right = (w * 0.5) * (znear / fx)
    top = (h * 0.5) * (znear / fy)
    m = np.zeros((4, 4))
    m[0, 0] = (2 * znear) / (2 * right)
    m[1, 1] = (2 * znear) / (2 * top)
    m[3, 2] = 1.0
    m[2, 2] = (zfar + znear) / (zfar - znear)
    m[2, 3] = -(zfar * znear) / (zfar - znear)
    return m


def make_camera(width: int, height: int, k: int = 0, num_views: int = 1, fov_x_deg: float = 60.0) -> Dict:
    """View k of num_views: the camera orbits SCENE_CENTRE about the y axis by 2*pi*k/V.

    Returns a dict with the keys of the reference's per-view data dict
    (colmap_dataset.py:100-108): tanfovx, tanfovy, viewmatrix, projmatrix, campos (+ sizes).
    """
    tanfovx = math.tan(math.radians(fov_x_deg) / 2)
    tanfovy = tanfovx * height / width
    fx = width / (2 * tanfovx)
    fy = height / (2 * tanfovy)
    th = 2 * math.pi * k / max(num_views, 1)
    c, s = math.cos(th), math.sin(th)
    # camera-to-world rotation about y, camera position orbits the centre
    R_c2w = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)
    centre = np.array(SCENE_CENTRE)
    campos = centre - R_c2w @ centre  # k=0 -> origin
    R_w2c = R_c2w.T
    t = -R_w2c @ campos
    view = np.eye(4)
    view[:3, :3] = R_w2c.T          # row-vector convention: [p,1] @ view
    view[3, :3] = t
    proj = projection_matrix(fx, fy, width, height).T
    full = view @ proj
    return dict(
        image_height=height, image_width=width, tanfovx=tanfovx, tanfovy=tanfovy,
        viewmatrix=torch.tensor(view, dtype=torch.float32),
        projmatrix=torch.tensor(full, dtype=torch.float32),
        campos=torch.tensor(campos, dtype=torch.float32),
    )


def make_scene(num_gaussians: int, sh_degree: int, width: int, height: int, seed: int = 0,
               fov_x_deg: float = 60.0) -> Dict[str, torch.Tensor]:
    """P Gaussians back-projected from uniformly drawn pixels of view 0 (CPU tensors)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    P = num_gaussians
    tanfovx = math.tan(math.radians(fov_x_deg) / 2)
    focal = width / (2 * tanfovx)

    def U(n, lo, hi):
        return torch.rand(n, generator=g) * (hi - lo) + lo

    u = U(P, 0.0, float(width))
    v = U(P, 0.0, float(height))
    z = torch.exp(U(P, math.log(2.0), math.log(20.0)))
    x = (u - width / 2) / focal * z
    y = (v - height / 2) / focal * z
    means3D = torch.stack([x, y, z], -1)

    sigma_px = torch.exp(torch.randn(P, generator=g) * 0.6 + math.log(1.5))
    aniso = torch.exp(torch.randn(P, 3, generator=g) * 0.3)
    scales = (z / focal * sigma_px)[:, None] * aniso
    rotations = torch.nn.functional.normalize(torch.randn(P, 4, generator=g), dim=-1)
    opacities = U(P, 0.02, 0.6)[:, None]
    M = (sh_degree + 1) ** 2
    shs = torch.randn(P, M, 3, generator=g)
    shs[:, 1:] *= 0.1
    return dict(means3D=means3D.float().contiguous(), scales=scales.float().contiguous(),
                rotations=rotations.float().contiguous(), opacities=opacities.float().contiguous(),
                shs=shs.float().contiguous())


CONFIGS = {
    # name: (P, sh_degree, W, H)  -- BASELINE.json configs, SURVEY.md section 8
    "C1": (10_000, 0, 128, 128),
    "C2": (100_000, 3, 800, 800),
    "C3": (1_000_000, 3, 1920, 1080),
    "C4": (3_000_000, 3, 1600, 1200),
    "C5": (6_000_000, 3, 3840, 2160),
}


def config(name: str) -> Tuple[int, int, int, int]:
    return CONFIGS[name]
