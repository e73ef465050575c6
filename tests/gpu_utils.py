"""Shared helpers of the GPU parity tests: run the CUDA path through the C-ABI binding and the CPU
oracle on the same inputs."""
from __future__ import annotations

import numpy as np
import torch

import oracle
from gaussian_splatting_lightning_b200 import synthetic
from gaussian_splatting_lightning_b200.rasterizer import (GaussianRasterizationSettings, backward_raw, export_state,
                                                          forward_raw)

DEV = "cuda:0"


def settings_from_camera(cam: oracle.Camera, device=DEV, debug=False) -> GaussianRasterizationSettings:
    t = lambda a, s: torch.tensor(np.asarray(a, np.float32).reshape(s), device=device)
    return GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=t(cam.bg, (3,)), scale_modifier=cam.scale_modifier, viewmatrix=t(cam.viewmatrix, (4, 4)),
        projmatrix=t(cam.projmatrix, (4, 4)), sh_degree=cam.sh_degree, campos=t(cam.campos, (3,)),
        prefiltered=False, debug=debug, antialiasing=cam.antialiasing)


def synthetic_case(P, deg, W, H, seed=0, view_k=0, num_views=1, bg=(0.0, 0.0, 0.0), antialiasing=False,
                   scale_modifier=1.0):
    scene = synthetic.make_scene(P, deg, W, H, seed=seed)
    c = synthetic.make_camera(W, H, k=view_k, num_views=num_views)
    cam = oracle.Camera(H, W, c["tanfovx"], c["tanfovy"], c["viewmatrix"].numpy(), c["projmatrix"].numpy(),
                        c["campos"].numpy(), bg=np.asarray(bg, np.float32), sh_degree=deg,
                        antialiasing=antialiasing, scale_modifier=scale_modifier)
    arrays = {k: v.numpy() for k, v in scene.items()}
    return cam, arrays


def to_dev(arrays, device=DEV):
    return {k: (None if v is None else torch.tensor(np.ascontiguousarray(v), device=device)) for k, v in arrays.items()}


def run_cuda_forward(cam, arrays, hint=0, device=DEV, debug=False):
    rs = settings_from_camera(cam, device, debug)
    d = to_dev(arrays, device)
    R, color, radii, geom, binning, img, invdepth, cap = forward_raw(
        rs, d["means3D"], d.get("shs"), d.get("colors_precomp"), d["opacities"], d.get("scales"), d.get("rotations"),
        d.get("cov3D_precomp"), binning_hint=hint)
    torch.cuda.synchronize()
    return dict(rs=rs, d=d, R=R, color=color, radii=radii, geom=geom, binning=binning, img=img, invdepth=invdepth,
                cap=cap)


def export(out):
    P = out["d"]["means3D"].shape[0]
    st = export_state(out["rs"], P, out["radii"], out["geom"], out["binning"], out["img"], out["R"], out["cap"])
    return {k: v.cpu().numpy() for k, v in st.items()}


def run_cuda_backward(out, dL_dcolor, dL_dinvdepth):
    d = out["d"]
    g = backward_raw(out["rs"], d["means3D"], out["radii"], d.get("shs"), d.get("colors_precomp"), d["opacities"],
                     d.get("scales"), d.get("rotations"), d.get("cov3D_precomp"),
                     torch.tensor(dL_dcolor, device=out["color"].device),
                     None if dL_dinvdepth is None else torch.tensor(dL_dinvdepth, device=out["color"].device),
                     out["geom"], out["binning"], out["img"], out["R"])
    torch.cuda.synchronize()
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations"]
    return {n: t.cpu().numpy() for n, t in zip(names, g)}


def run_oracle_forward(cam, arrays):
    return oracle.forward(cam, arrays["means3D"], arrays.get("scales"), arrays.get("rotations"), arrays["opacities"],
                          shs=arrays.get("shs"), colors_precomp=arrays.get("colors_precomp"),
                          cov3D_precomp=arrays.get("cov3D_precomp"))


def close_fraction(a, b, atol=1e-4, rtol=1e-4):
    """fraction of elements violating |a-b| <= atol + rtol*|b| and the max abs error."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    err = np.abs(a - b)
    bad = err > atol + rtol * np.abs(b)
    return float(bad.mean()) if bad.size else 0.0, float(err.max()) if err.size else 0.0


# ---- the gradient bar (north_star: 1e-4 relative / abs) -------------------------------------------------------------
# Every tensor is compared on its own scale s = max(1, max|ref|): |got - ref| <= GRAD_ATOL * s + GRAD_RTOL * |ref|.
# A (pixel, splat) pair whose alpha sits within ulps of the 1/255 floor (ex2.approx vs expf) flips a decision and moves
# a few elements by more than that: at most GRAD_OUTLIER_FRAC of the elements may miss the bound (observed on the B200:
# at most 5e-5, profiles/r2_parity_observed.txt), by at most
# GRAD_OUTLIER_MAX * s.  (Round 1 allowed rtol 1e-3, 2e-3 of the elements and 5e-2: ten times looser.)
GRAD_ATOL = GRAD_RTOL = 1e-4
GRAD_OUTLIER_FRAC = 2e-4
GRAD_OUTLIER_MAX = 5e-3
GRAD_NAMES = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
              "dL_drotations")


def assert_grads_close(got: dict, want: dict, names=GRAD_NAMES, label=""):
    report = {}
    for k in names:
        w = want.get(k)
        if w is None:
            continue
        w = np.asarray(w, np.float64)
        g = np.asarray(got[k], np.float64).reshape(w.shape)
        s = max(1.0, float(np.abs(w).max())) if w.size else 1.0
        err = np.abs(g - w)
        bad = err > GRAD_ATOL * s + GRAD_RTOL * np.abs(w)
        frac = float(bad.mean()) if bad.size else 0.0
        mx = float(err.max() / s) if err.size else 0.0
        # the unscaled rule, reported for the record (elements far below the tensor's scale carry the float32 noise of
        # the large ones they were summed with)
        frac_unscaled = float((err > 1e-4 + 1e-4 * np.abs(w)).mean()) if err.size else 0.0
        report[k] = (s, frac, mx, frac_unscaled)
        print(f"{label} {k}: |ref|max {s:.3e}  outliers {frac:.2e} (bar {GRAD_OUTLIER_FRAC:.0e})  max err/scale {mx:.2e} "
              f"(bar {GRAD_OUTLIER_MAX:.0e})  unscaled-rule outliers {frac_unscaled:.2e}")
    for k, (s, frac, mx, _) in report.items():
        assert frac <= GRAD_OUTLIER_FRAC, (label, k, "outlier fraction", frac)
        assert mx <= GRAD_OUTLIER_MAX, (label, k, "max err / scale", mx)
    return report
