"""Camera contract on the caller side of the rasterizer.

What ColmapDataset.build_item hands to GSLightningModule.render for one view
(gs_lightning/datasets/colmap_dataset.py:77-109, gs_lightning/utils/camera.py:4-41,
gs_lightning/lightning/gs_lightning_module.py:321-336), and its batched form: a DataLoader with batch_size V
collates the per-view dicts into tensors with a leading V, which `settings_from_batch` turns into the V
GaussianRasterizationSettings of one view-parallel step (view_parallel.render_step).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np
import torch

from .rasterizer import GaussianRasterizationSettings


def get_projection_matrix(fx: float, fy: float, w: int, h: int, znear: float, zfar: float) -> np.ndarray:
    """gs_lightning/utils/camera.py:4-41: column-vector perspective matrix mapping z to [0, 1]."""
    right = (w * 0.5) * (znear / fx)
    top = (h * 0.5) * (znear / fy)
    m = np.zeros((4, 4))
    m[0, 0] = (2 * znear) / (right - (-right))
    m[1, 1] = (2 * znear) / (top - (-top))
    m[3, 2] = 1.0
    m[2, 2] = (zfar + znear) / (zfar - znear)
    m[2, 3] = -(zfar * znear) / (zfar - znear)
    return m


def camera_from_colmap(cam_from_world: np.ndarray, fx: float, fy: float, width: int, height: int,
                       z_near: float = 0.01, z_far: float = 100.0) -> Dict[str, object]:
    """The camera entries of ColmapDataset.build_item (colmap_dataset.py:84-107) from a 3x4 cam_from_world matrix:
    row-vector viewmatrix (4,4) float32, projmatrix = viewmatrix @ P^T, campos = inverse(viewmatrix)[3, :3],
    tanfov = (size / 2) / focal as Python floats."""
    wv = np.eye(4)
    wv[:, :3] = np.asarray(cam_from_world, dtype=np.float64).reshape(3, 4).T
    world_view = torch.Tensor(wv)                                   # float32, as torch.Tensor(np.float64 array) gives
    campos = world_view.inverse()[3, :3]
    proj = torch.Tensor(get_projection_matrix(fx, fy, width, height, z_near, z_far).T)
    return dict(tanfovx=(width * 0.5) / fx, tanfovy=(height * 0.5) / fy, viewmatrix=world_view,
                projmatrix=world_view @ proj, campos=campos)


def settings_from_item(data_dict: Dict[str, object], sh_degree: int, image_hw: Sequence[int] = None
                       ) -> GaussianRasterizationSettings:
    """GSLightningModule.render's settings for one (un-batched) dataset item (gs_lightning_module.py:321-336)."""
    H, W = (data_dict["image"].shape[-2:] if image_hw is None else image_hw)
    return GaussianRasterizationSettings(
        image_height=int(H), image_width=int(W), tanfovx=data_dict["tanfovx"], tanfovy=data_dict["tanfovy"],
        bg=data_dict["background"], scale_modifier=1.0, viewmatrix=data_dict["viewmatrix"],
        projmatrix=data_dict["projmatrix"], sh_degree=int(sh_degree), campos=data_dict["campos"], prefiltered=False,
        debug=False, antialiasing=False)


def settings_from_batch(batch: Dict[str, object], sh_degree: int, device=None) -> List[GaussianRasterizationSettings]:
    """One GaussianRasterizationSettings per view of a collated batch: image (V,3,H,W), tanfovx / tanfovy (V,) (float64
    after default_collate), background (V,3), viewmatrix / projmatrix (V,4,4), campos (V,3).  The matrices of all
    views move to `device` in ONE copy each; the per-view settings hold views of those tensors."""
    V = int(batch["viewmatrix"].shape[0])
    H, W = batch["image"].shape[-2:]

    def dev(t):
        t = torch.as_tensor(t)
        return t.to(device, non_blocking=True) if device is not None else t

    view, proj, campos, bg = (dev(batch[k]).float() for k in ("viewmatrix", "projmatrix", "campos", "background"))
    tanx = [float(v) for v in torch.as_tensor(batch["tanfovx"]).reshape(-1).tolist()]
    tany = [float(v) for v in torch.as_tensor(batch["tanfovy"]).reshape(-1).tolist()]
    if not (view.shape == (V, 4, 4) and proj.shape == (V, 4, 4) and campos.shape == (V, 3) and len(tanx) == V):
        raise ValueError("settings_from_batch: expected a default_collate'd batch of ColmapDataset items")
    bg = bg.reshape(-1, 3).expand(V, 3) if bg.numel() in (3, 3 * V) else bg
    return [GaussianRasterizationSettings(
        image_height=int(H), image_width=int(W), tanfovx=tanx[k], tanfovy=tany[k], bg=bg[k], scale_modifier=1.0,
        viewmatrix=view[k], projmatrix=proj[k], sh_degree=int(sh_degree), campos=campos[k], prefiltered=False,
        debug=False, antialiasing=False) for k in range(V)]
