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
