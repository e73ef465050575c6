"""Comparison of analytic gradients with the REFERENCE-DERIVED gradient goldens (tests/golden/*_grads.npz:
float64 central differences through the reference's own Python rasterizer, minted by
tests/golden/make_grad_golden.py).  Used by the CPU test of the oracle and the GPU test of the CUDA backward.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_grad_golden import TENSORS, direction, loss_weights  # noqa: E402  (pure numpy helpers; no reference needed)

GRAD_GOLDENS = ["g1_deg0_128", "g2_deg3_100x70", "g3_treehill0_half"]
GRAD_KEY = dict(means3D="dL_dmeans3D", scales="dL_dscales", rotations="dL_drotations", opacities="dL_dopacity",
                shs="dL_dsh")
# north_star: 1e-4 relative / abs.  The analytic side runs in float32 (sums over thousands of pixels), so the absolute
# part is taken relative to the largest derivative of the tensor among the golden's samples.
RTOL = 1e-4
ATOL_OF_TENSOR_MAX = 1e-4


def load(name):
    g = np.load(os.path.join(HERE, "golden", name + "_grads.npz"))
    d = np.load(os.path.join(HERE, "golden", name + ".npz"))
    H, W = int(d["image_height"]), int(d["image_width"])
    w_img, w_dep = loss_weights(H, W)
    mask = g["mask"]
    return g, d, (w_img * mask[None]).astype(np.float32), (w_dep * mask[None]).astype(np.float32)


def load_dir_weights(name):
    """Loss weights of the dense directions: `dir_mask` also leaves out the pixels on the alpha floor's contour."""
    g = np.load(os.path.join(HERE, "golden", name + "_grads.npz"))
    d = np.load(os.path.join(HERE, "golden", name + ".npz"))
    w_img, w_dep = loss_weights(int(d["image_height"]), int(d["image_width"]))
    m = g["dir_mask"]
    return (w_img * m[None]).astype(np.float32), (w_dep * m[None]).astype(np.float32)


def tangential(grads: dict, rotations: np.ndarray) -> dict:
    """The reference normalises q inside computeConv3D (render_tools.py:61): its derivative with respect to a unit
    quaternion is the tangential part (I - q q^T) g of the derivative g the CUDA semantics give."""
    out = dict(grads)
    q = rotations.astype(np.float64)
    q = q / np.linalg.norm(q, axis=-1, keepdims=True)
    gq = grads["dL_drotations"].astype(np.float64)
    out["dL_drotations"] = gq - q * (q * gq).sum(-1, keepdims=True)
    return out


def compare(name: str, grads: dict, grads_dir: dict = None, verbose=True):
    """grads: name -> array, the analytic gradients for dL_dcolor = w_img, dL_dinvdepth = w_dep of load(name);
    grads_dir: the same for the weights of load_dir_weights(name) (the dense directions' loss).
    Returns a list of (what, worst ratio err / tol) and asserts every sample within tolerance."""
    g, d, _, _ = load(name)
    grads = tangential(grads, d["rotations"])
    report = []
    for ti, tname in enumerate(TENSORS):
        sel = g["tensor"] == ti
        if not sel.any():
            continue
        fd = g["fd"][sel]
        got = np.asarray(grads[GRAD_KEY[tname]], np.float64).reshape(-1)[g["index"][sel]]
        tol = ATOL_OF_TENSOR_MAX * np.abs(fd).max() + RTOL * np.abs(fd)
        err = np.abs(got - fd)
        worst = float((err / tol).max())
        report.append((tname, worst, int(sel.sum()), float(np.abs(fd).max())))
        if verbose:
            print(f"{name} {tname}: {int(sel.sum())} scalars, |fd|max {np.abs(fd).max():.3e}, worst err/tol {worst:.3f}, "
                  f"max abs err {err.max():.3e}")
        assert worst <= 1.0, (name, tname, worst)
    # dense directions: sum_t h_t <g_t, d_t>
    if grads_dir is None:
        return report
    grads = tangential(grads_dir, d["rotations"])
    deg = int(d["sh_degree"])
    fdd = g["dirs_fd"]
    for j, k in enumerate(g["dir_k"]):
        tot, mag = 0.0, 0.0
        for ti, tname in enumerate(TENSORS):
            a = np.asarray(grads[GRAD_KEY[tname]], np.float64).reshape(d[tname].shape)
            dv = direction(int(k), tname, d[tname].shape)
            if tname == "shs" and deg < 3:
                dv[:, (deg + 1) ** 2:, :] = 0.0
            dv[~g["dir_gaussians"]] = 0.0     # Gaussians outside 1.3 x FOV are held still (make_grad_golden.py)
            tot += g["dir_steps"][ti] * float((a * dv).sum())
            mag += g["dir_steps"][ti] * float(np.abs(a * dv).sum())
        # the dot product cancels heavily (mag >> |tot|): float32 rounding of the analytic side is judged against mag
        tol = RTOL * abs(fdd[j]) + 2e-6 * mag
        ratio = abs(tot - fdd[j]) / tol
        report.append((f"dir{k}", ratio, 0, float(fdd[j])))
        if verbose:
            print(f"{name} direction {k}: fd {fdd[j]:.6e} analytic {tot:.6e} err/tol {ratio:.3f}")
        assert ratio <= 1.0, (name, k, tot, fdd[j])
    return report
