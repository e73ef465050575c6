"""Differentiable float64 torch restatement of the rasterizer forward (test helper).

Follows the reference's Python rasterizer (gs_lightning/rasterize/rasterize.py:28-127,
render_tools.py:13-144, camera_tools.py:5-32, utils/sh.py:41-81) with the CUDA semantics
of SURVEY.md Appendix B, written so that torch autograd yields the gradients the CUDA
backward is specified to return:

  * alpha = min(0.99, o*G) is treated straight-through (the CUDA backward does not gate
    the gradient on the clamp),
  * the 1.3*tanfov clamp on x/z, y/z gates the gradient through t.x / t.y and the clamped
    value is a constant (no d/dz term),
  * max(colour, 0) has zero gradient where clamped (the `clamped` mask).

The discrete decisions (sort order, tile ranges) are taken from an oracle forward state so
both sides differentiate the same piecewise-smooth function.  Used to pin the analytic
backward of oracle/gsr_oracle.cpp; O(entries x 256) memory per tile, small scenes only.
"""
from __future__ import annotations

import numpy as np
import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
      -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, d):
    x, y, z = d[:, 0:1], d[:, 1:2], d[:, 2:3]
    res = C0 * sh[:, 0]
    if deg > 0:
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6]
               + C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
    if deg > 2:
        res = (res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10]
               + C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12]
               + C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14]
               + C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return res


def quat_to_R(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)


def render(cam, state, means3D, scales, rotations, opacities, shs=None, colors_precomp=None,
           cov3D_precomp=None, means2D_delta=None):
    """cam: oracle.Camera; state: oracle.forward() dict (for point_list / ranges / radii).
    All tensor args float64 (requires_grad as wanted). Returns color (3,H,W), invdepth (1,H,W)."""
    dt = torch.float64
    W, H = cam.image_width, cam.image_height
    V = torch.tensor(cam.viewmatrix.reshape(4, 4), dtype=dt)
    PV = torch.tensor(cam.projmatrix.reshape(4, 4), dtype=dt)
    campos = torch.tensor(cam.campos, dtype=dt)
    bg = torch.tensor(cam.bg, dtype=dt)
    P = means3D.shape[0]
    ones = torch.ones(P, 1, dtype=dt)
    hom = torch.cat([means3D, ones], -1)
    pview = hom @ V
    ph = hom @ PV
    pw = 1.0 / (ph[:, 3:4] + 1e-7)
    ndc = ph[:, :2] * pw
    if means2D_delta is not None:
        ndc = ndc + means2D_delta[:, :2]
    pix = torch.stack([((ndc[:, 0] + 1.0) * W - 1.0) * 0.5, ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5], -1)

    if cov3D_precomp is None:
        R = quat_to_R(rotations)
        L = R * (scales * cam.scale_modifier)[:, None, :]
        Sigma = L @ L.transpose(1, 2)
    else:
        c = cov3D_precomp
        Sigma = torch.stack([c[:, 0], c[:, 1], c[:, 2], c[:, 1], c[:, 3], c[:, 4],
                             c[:, 2], c[:, 4], c[:, 5]], -1).reshape(P, 3, 3)

    fx, fy = W / (2 * cam.tanfovx), H / (2 * cam.tanfovy)
    tx, ty, tz = pview[:, 0], pview[:, 1], pview[:, 2]
    limx, limy = 1.3 * cam.tanfovx, 1.3 * cam.tanfovy
    txtz, tytz = tx / tz, ty / tz
    in_x = (txtz >= -limx) & (txtz <= limx)
    in_y = (tytz >= -limy) & (tytz <= limy)
    txc = torch.where(in_x, tx, (txtz.clamp(-limx, limx) * tz).detach())
    tyc = torch.where(in_y, ty, (tytz.clamp(-limy, limy) * tz).detach())
    J = torch.zeros(P, 2, 3, dtype=dt)
    J[:, 0, 0] = fx / tz
    J[:, 0, 2] = -(fx * txc) / (tz * tz)
    J[:, 1, 1] = fy / tz
    J[:, 1, 2] = -(fy * tyc) / (tz * tz)
    Rcw = V[:3, :3].T
    M = J @ Rcw
    cov = M @ Sigma @ M.transpose(1, 2)
    a0, b, c0 = cov[:, 0, 0], cov[:, 0, 1], cov[:, 1, 1]
    det0 = a0 * c0 - b * b
    a, c = a0 + 0.3, c0 + 0.3
    det = a * c - b * b
    conic = torch.stack([c / det, -b / det, a / det], -1)
    op = opacities.reshape(P)
    if cam.antialiasing:
        op = op * torch.sqrt(torch.clamp_min(det0 / det, 0.000025))

    if colors_precomp is None:
        d = means3D - campos[None]
        d = d / torch.linalg.norm(d, dim=-1, keepdim=True)
        rgb = torch.clamp_min(eval_sh(cam.sh_degree, shs, d) + 0.5, 0.0)
    else:
        rgb = colors_precomp
    invz = 1.0 / tz

    gx, gy = cam.grid
    ranges = state["ranges"].astype(np.int64)
    plist = torch.from_numpy(state["point_list"].astype(np.int64))
    color = torch.zeros(H, W, 3, dtype=dt)
    invdepth = torch.zeros(H, W, dtype=dt)
    for t in range(gx * gy):
        lo, hi = int(ranges[t, 0]), int(ranges[t, 1])
        x0, y0 = (t % gx) * 16, (t // gx) * 16
        x1, y1 = min(x0 + 16, W), min(y0 + 16, H)
        yy, xx = torch.meshgrid(torch.arange(y0, y1, dtype=dt), torch.arange(x0, x1, dtype=dt), indexing="ij")
        npix = yy.numel()
        if hi <= lo:
            color[y0:y1, x0:x1] = bg
            continue
        ids = plist[lo:hi]
        dx = pix[ids, 0:1] - xx.reshape(1, -1)
        dy = pix[ids, 1:2] - yy.reshape(1, -1)
        cn = conic[ids]
        power = -0.5 * (cn[:, 0:1] * dx * dx + cn[:, 2:3] * dy * dy) - cn[:, 1:2] * dx * dy
        raw = op[ids][:, None] * torch.exp(power)
        alpha = raw + (torch.clamp_max(raw, 0.99) - raw).detach()   # straight-through clamp
        keep = (power <= 0) & (alpha >= 1.0 / 255.0)
        alpha = torch.where(keep, alpha, torch.zeros_like(alpha))
        one_m = 1.0 - alpha
        T_incl = torch.cumprod(one_m, 0)
        T_excl = torch.cat([torch.ones(1, npix, dtype=dt), T_incl[:-1]], 0)
        stop = keep & (T_incl < 0.0001)                     # test_T < 1e-4 on a contributing entry
        stopped = torch.cumsum(stop.to(torch.int64), 0) > 0  # this entry and everything after
        wgt = torch.where(stopped, torch.zeros_like(alpha), alpha * T_excl)
        T_final = torch.where(stopped, torch.zeros_like(alpha), one_m).masked_fill(stopped, 1.0).prod(0)
        col = (wgt[:, :, None] * rgb[ids][:, None, :]).sum(0) + T_final[:, None] * bg[None]
        dep = (wgt * invz[ids][:, None]).sum(0)
        color[y0:y1, x0:x1] = col.reshape(y1 - y0, x1 - x0, 3)
        invdepth[y0:y1, x0:x1] = dep.reshape(y1 - y0, x1 - x0)
    return color.permute(2, 0, 1), invdepth[None]
