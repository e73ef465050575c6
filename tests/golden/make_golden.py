"""Mint the committed golden fixtures by RUNNING THE REFERENCE's own Python rasterizer.

    python tests/golden/make_golden.py          # build container only (/root/reference)

Writes tests/golden/*.npz: scene inputs + the outputs of
gs_lightning.rasterize.rasterize_gaussian / markVisible (rasterize.py:23-127), unchanged
reference code imported through ref_loader.py.  Scenes are kept inside the validity
envelope of SURVEY.md Appendix B (every Gaussian has z_view > 0.2, normalised
quaternions, SH zero-padded to 25 coefficients for degree 3 because of sh.py:83).
The cameras of tests/rasterizer_python/test_cases.py:20-72 and its 53 points (:75-129)
are read from the reference file at generation time and stored in the fixture.
"""
from __future__ import annotations

import os
import sys
import types
from typing import NamedTuple

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_loader  # noqa: E402
from gaussian_splatting_lightning_b200 import synthetic  # noqa: E402


def _load_reference_test_cases():
    """tests/rasterizer_python/test_cases.py needs `diff_gaussian_rasterization`; give it a
    field-compatible NamedTuple so the file can be executed as is."""
    class GaussianRasterizationSettings(NamedTuple):
        image_height: int
        image_width: int
        tanfovx: float
        tanfovy: float
        bg: torch.Tensor
        scale_modifier: float
        viewmatrix: torch.Tensor
        projmatrix: torch.Tensor
        sh_degree: int
        campos: torch.Tensor
        prefiltered: bool
        debug: bool
        antialiasing: bool

    stub = types.ModuleType("diff_gaussian_rasterization")
    stub.GaussianRasterizationSettings = GaussianRasterizationSettings
    saved = sys.modules.get("diff_gaussian_rasterization")
    sys.modules["diff_gaussian_rasterization"] = stub
    try:
        import importlib.util
        p = os.path.join(ref_loader.REF_ROOT, "tests", "rasterizer_python", "test_cases.py")
        spec = importlib.util.spec_from_file_location("_ref_test_cases", p)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
    finally:
        if saved is None:
            del sys.modules["diff_gaussian_rasterization"]
        else:
            sys.modules["diff_gaussian_rasterization"] = saved
    return m.settings, m.points_3d.cpu()


def _filter_in_front(scene, view, zmin=0.3):
    p = scene["means3D"]
    z = (torch.cat([p, torch.ones(len(p), 1)], -1) @ view)[:, 2]
    keep = z > zmin
    return {k: v[keep].contiguous() for k, v in scene.items()}


def _run_reference(ref, scene, cam, sh_degree, bg):
    shs = scene["shs"]
    if sh_degree == 3:  # sh.py:83 runs the degree-4 block for degree 3: pad with zeros (exact)
        shs = torch.cat([shs, torch.zeros(len(shs), 25 - shs.shape[1], 3)], 1)
    with torch.no_grad():
        img, radius, depth = ref.rasterize_gaussian(
            means3D=scene["means3D"], opacities=scene["opacities"], scales=scene["scales"],
            rotations=scene["rotations"], shs=shs, scale_modifier=1.0,
            image_width=cam["image_width"], image_height=cam["image_height"],
            tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], viewmatrix=cam["viewmatrix"],
            projmatrix=cam["projmatrix"], campos=cam["campos"], background=bg, sh_degree=sh_degree)
    return img.numpy(), radius.numpy(), depth.numpy()


def _save(name, scene, cam, sh_degree, bg, out):
    img, radius, depth = out
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(
        path,
        means3D=scene["means3D"].numpy(), scales=scene["scales"].numpy(),
        rotations=scene["rotations"].numpy(), opacities=scene["opacities"].numpy(),
        shs=scene["shs"].numpy(), sh_degree=np.int32(sh_degree), bg=bg.numpy(),
        image_height=np.int32(cam["image_height"]), image_width=np.int32(cam["image_width"]),
        tanfovx=np.float64(cam["tanfovx"]), tanfovy=np.float64(cam["tanfovy"]),
        viewmatrix=cam["viewmatrix"].numpy(), projmatrix=cam["projmatrix"].numpy(),
        campos=cam["campos"].numpy(),
        ref_image=img, ref_radius=radius, ref_depth=depth)
    print(f"{name}: P={len(scene['means3D'])} {cam['image_width']}x{cam['image_height']} "
          f"deg{sh_degree} -> {os.path.getsize(path) / 1e3:.0f} kB")


def main():
    torch.manual_seed(0)
    ref = ref_loader.load()

    # G1: BASELINE config-1 shape, scaled down (SH degree 0, 128x128)
    W, H, deg = 128, 128, 0
    scene = synthetic.make_scene(2000, deg, W, H, seed=1)
    cam = synthetic.make_camera(W, H)
    bg = torch.zeros(3)
    _save("g1_deg0_128", scene, cam, deg, bg, _run_reference(ref, scene, cam, deg, bg))

    # G2: degree 3, image size not a multiple of 16, rotated camera, non-zero background
    W, H, deg = 100, 70, 3
    scene = synthetic.make_scene(1500, deg, W, H, seed=2)
    cam = synthetic.make_camera(W, H, k=1, num_views=12)
    scene = _filter_in_front(scene, cam["viewmatrix"])
    bg = torch.tensor([0.2, 0.5, 0.8])
    _save("g2_deg3_100x70", scene, cam, deg, bg, _run_reference(ref, scene, cam, deg, bg))

    # G3: the reference's own camera fixture (treehill view 0, test_cases.py:22-38) at half
    # resolution; Gaussians scattered around the reference's 53 fixture points (:75-129)
    settings, points = _load_reference_test_cases()
    s0 = settings[0]
    W, H, deg = s0.image_width // 2, s0.image_height // 2, 3
    cam = dict(image_width=W, image_height=H, tanfovx=s0.tanfovx, tanfovy=s0.tanfovy,
               viewmatrix=s0.viewmatrix.cpu(), projmatrix=s0.projmatrix.cpu(), campos=s0.campos.cpu())
    g = torch.Generator().manual_seed(3)
    extra = points[torch.randint(0, len(points), (700,), generator=g)] + torch.randn(700, 3, generator=g) * 0.8
    means = torch.cat([points, extra])
    P = len(means)
    scene = dict(
        means3D=means.float().contiguous(),
        scales=torch.exp(torch.randn(P, 3, generator=g) * 0.5 + np.log(0.03)).float(),
        rotations=torch.nn.functional.normalize(torch.randn(P, 4, generator=g), dim=-1),
        opacities=(torch.rand(P, 1, generator=g) * 0.5 + 0.05),
        shs=torch.cat([torch.randn(P, 1, 3, generator=g), torch.randn(P, 15, 3, generator=g) * 0.1], 1))
    scene = _filter_in_front(scene, cam["viewmatrix"])
    bg = torch.zeros(3)
    _save("g3_treehill0_half", scene, cam, deg, bg, _run_reference(ref, scene, cam, deg, bg))

    # G4: saturating scene (large opaque splats): exercises the T<1e-4 early-out where the
    # Python path is known to keep one extra Gaussian (rasterize.py:253-256, Appendix B3)
    W, H, deg = 64, 48, 2
    scene = synthetic.make_scene(900, deg, W, H, seed=4)
    scene["opacities"] = torch.full_like(scene["opacities"], 0.9)
    scene["scales"] = scene["scales"] * 3.0
    cam = synthetic.make_camera(W, H)
    bg = torch.tensor([1.0, 1.0, 1.0])
    _save("g4_saturated_64x48", scene, cam, deg, bg, _run_reference(ref, scene, cam, deg, bg))

    # markVisible on the reference's fixture: 53 points x 3 cameras (test_mark_visible.py:11-21)
    vis = []
    for s in settings:
        vis.append(ref.markVisible(points, viewmatrix=s.viewmatrix.cpu(), projmatrix=s.projmatrix.cpu()).numpy())
    np.savez_compressed(
        os.path.join(HERE, "mark_visible_treehill.npz"), points_3d=points.numpy(),
        viewmatrices=np.stack([s.viewmatrix.cpu().numpy() for s in settings]),
        projmatrices=np.stack([s.projmatrix.cpu().numpy() for s in settings]),
        camposes=np.stack([s.campos.cpu().numpy() for s in settings]),
        image_height=np.int32(settings[0].image_height), image_width=np.int32(settings[0].image_width),
        tanfovx=np.float64(settings[0].tanfovx), tanfovy=np.float64(settings[0].tanfovy),
        ref_visible=np.stack(vis))
    print("mark_visible_treehill: visible fractions", [float(v.mean()) for v in vis])


if __name__ == "__main__":
    main()
