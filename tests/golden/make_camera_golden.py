"""Generates tests/golden/cameras.npz by calling the reference's own get_projection_matrix
(/root/reference/gs_lightning/utils/camera.py:4-41, imported from the reference tree) and following
ColmapDataset.build_item (gs_lightning/datasets/colmap_dataset.py:84-107) line by line for the rest.
Run in the build container (the reference tree is not on the GPU box): python tests/golden/make_camera_golden.py"""
import importlib.util
import os

import numpy as np
import torch

spec = importlib.util.spec_from_file_location("ref_camera", "/root/reference/gs_lightning/utils/camera.py")
ref_camera = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_camera)

rng = np.random.default_rng(0)
out = {}
for i in range(6):
    q = rng.standard_normal(4); q /= np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    t = rng.standard_normal(3) * 3
    c2w = np.concatenate([R, t[:, None]], 1)                     # image_info.cam_from_world().matrix(): 3x4
    fx, fy = float(rng.uniform(400, 2000)), float(rng.uniform(400, 2000))
    W, H = int(rng.integers(300, 4000)), int(rng.integers(200, 2200))
    zn, zf = 0.01, 100.0
    world_view_transform = np.eye(4)                             # colmap_dataset.py:84-87
    world_view_transform[:, :3] = c2w.T
    world_view_transform = torch.Tensor(world_view_transform)
    camera_center = world_view_transform.inverse()[3, :3]
    projection_matrix = torch.Tensor(ref_camera.get_projection_matrix(fx, fy, W, H, zn, zf).T)   # :88-96
    full_proj_transform = world_view_transform @ projection_matrix                                 # :97
    out[f"in_{i}"] = np.concatenate([c2w.reshape(-1), [fx, fy, W, H, zn, zf]])
    out[f"proj_{i}"] = ref_camera.get_projection_matrix(fx, fy, W, H, zn, zf)
    out[f"view_{i}"] = world_view_transform.numpy()
    out[f"full_{i}"] = full_proj_transform.numpy()
    out[f"campos_{i}"] = camera_center.numpy()
    out[f"tan_{i}"] = np.array([(W * 0.5) / fx, (H * 0.5) / fy])
np.savez(os.path.join(os.path.dirname(__file__), "cameras.npz"), **out)
print("wrote cameras.npz")
