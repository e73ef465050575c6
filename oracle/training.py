"""CPU restatement (numpy) of the steps either side of the rasterizer in training_step.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): imported by tests/ and bench.py's checker legs, never by
the product package.

Pinned: the reference's implementations of these steps are torch's own CPU operators (nn.L1Loss,
torch.optim.Adam, boolean-mask updates), which run in this container AND on the GPU box, so
tests/test_oracle_training.py checks every function here directly against them (no fixtures needed).

    l1_loss        nn.L1Loss()                      gs_lightning/lightning/gs_lightning_module.py:99,278
    adam_step      torch.optim.Adam                 gs_lightning_module.py:111-127 (configs: eps 1e-15, default betas)
    densify_fold   update_max_radii2D / update_xyz_gradient over the views of a step
                                                    gs_lightning/modules/gaussian_model.py:174-181
"""
from __future__ import annotations

import math

import numpy as np

f32 = np.float32


def l1_loss(image: np.ndarray, target: np.ndarray):
    """(mean |image - target| accumulated in float64, unit gradient sign(image - target) / n in float32)."""
    a, b = np.asarray(image, f32), np.asarray(target, f32)
    d = a - b
    n = d.size
    loss = f32(np.abs(d).astype(np.float64).sum() / n)
    grad = (np.sign(d) * (f32(1.0) / f32(n))).astype(f32)
    return loss, grad


def adam_step(param, grad, exp_avg, exp_avg_sq, step: int, lr: float, beta1: float = 0.9, beta2: float = 0.999,
              eps: float = 1e-15):
    """One torch.optim.Adam update (no weight decay / amsgrad), every operation rounded to float32 exactly
    where the CUDA kernel rounds (optimizer.cu adam_one: no fused multiply-add).  Returns new (p, m, v)."""
    p, g, m, v = (np.asarray(t, f32) for t in (param, grad, exp_avg, exp_avg_sq))
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    nss = f32(-(lr / bc1))                  # Python-double scalars folded to float32 once, as torch does
    w1 = f32(1.0 - beta1)
    b2 = f32(beta2)
    w2 = f32(1.0 - beta2)
    bc2_sqrt = f32(math.sqrt(bc2))
    e = f32(eps)
    m = m + w1 * (g - m)
    v = v * b2 + (w2 * g) * g
    denom = np.sqrt(v) / bc2_sqrt + e
    p = p + nss * (m / denom)
    return p.astype(f32), m.astype(f32), v.astype(f32)


def densify_fold(d_max_radii, d_accum, d_count, max_radii2D, xyz_grad_accum, xyz_grad_count):
    """Running statistics after folding one step's deltas; returns new (max_radii2D, accum, count)."""
    seen = np.asarray(d_count) > 0
    mr, ac, ct = (np.array(t, f32, copy=True) for t in (max_radii2D, xyz_grad_accum, xyz_grad_count))
    mr[seen] = np.maximum(mr[seen], np.asarray(d_max_radii, f32)[seen])
    ac[seen] = ac[seen] + np.asarray(d_accum, f32)[seen]
    ct[seen] = ct[seen] + np.asarray(d_count, f32)[seen]
    return mr, ac, ct


def _gauss11() -> np.ndarray:
    """The 11-tap window of submodules/fused-ssim/tests/test.py:14-16 (sigma 1.5, normalised)."""
    g = np.exp(-((np.arange(11) - 5) ** 2) / (2 * 1.5 ** 2))
    return g / g.sum()


def _blur(a: np.ndarray) -> np.ndarray:
    """Separable 11x11 Gaussian filter with zero padding over the last two axes (float64)."""
    g = _gauss11()
    H, W = a.shape[-2:]
    p = np.pad(a, [(0, 0)] * (a.ndim - 2) + [(5, 5), (5, 5)])
    h = sum(g[k] * p[..., :, k:k + W] for k in range(11))
    return sum(g[k] * h[..., k:k + H, :] for k in range(11))


def fused_ssim(img1, img2, padding: str = "same", C1: float = 0.01 ** 2, C2: float = 0.03 ** 2):
    """(mean SSIM, d mean / d img1) in float64, the arithmetic of submodules/fused-ssim/ssim.cu:187-366
    (forward map + the three partial derivatives, backward = three filtered products) with dL_dmap = 1/N
    (fused_ssim/__init__.py:33-41; padding 'valid' drops a 5-pixel border from the mean)."""
    x, y = np.asarray(img1, np.float64), np.asarray(img2, np.float64)
    mu1, mu2 = _blur(x), _blur(y)
    s11, s22, s12 = _blur(x * x) - mu1 * mu1, _blur(y * y) - mu2 * mu2, _blur(x * y) - mu1 * mu2
    A, B = mu1 * mu1 + mu2 * mu2 + C1, s11 + s22 + C2
    C, D = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    m = C * D / (A * B)
    w = np.zeros_like(m)
    if padding == "valid":
        w[..., 5:-5, 5:-5] = 1.0
    else:
        w[...] = 1.0
    w /= w.sum()
    d_mu1 = 2 * mu2 * (D - C) / (A * B) - 2 * mu1 * C * D / (A * A * B) + 2 * mu1 * C * D / (A * B * B)
    d_s11 = -C * D / (A * B * B)
    d_s12 = 2 * C / (A * B)
    grad = _blur(w * d_mu1) + 2 * x * _blur(w * d_s11) + y * _blur(w * d_s12)
    return float((w * m).sum()), grad
