"""The two steps either side of the rasterizer in `GSLightningModule.training_step`,
as hand-written sm_100a kernels of libgsr_b200.so behind the interfaces the reference uses:

    l1_loss(image, target)      == nn.L1Loss()                 gs_lightning/lightning/gs_lightning_module.py:99,278
    FusedAdam(param_groups,...) == torch.optim.Adam            gs_lightning_module.py:111-127, configs (eps 1e-15)
    densify_fold(...)           == GaussianModel.update_max_radii2D / update_xyz_gradient over the views of a step
                                                               gs_lightning/modules/gaussian_model.py:174-181

CUDA only: CPU tensors raise, a missing library raises ImportError (no fallback).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, Optional

import torch

from . import _native as N


def _need_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{what}: gaussian_splatting_lightning_b200 runs on CUDA devices only (no CPU fallback)")


class _L1Loss(torch.autograd.Function):
    """mean |image - target|; value and unit gradient come out of ONE pass over the two images."""

    @staticmethod
    def forward(ctx, image, target):
        _need_cuda(image, "l1_loss")
        if image.shape != target.shape:
            raise RuntimeError(f"l1_loss: shapes differ: {tuple(image.shape)} vs {tuple(target.shape)}")
        dev = image.device
        a, b = N.f32c(image), N.f32c(target, dev)
        n = a.numel()
        out = torch.empty((), dtype=torch.float32, device=dev)
        want_grad = ctx.needs_input_grad[0]
        unit = torch.empty_like(a) if want_grad else None
        with torch.cuda.device(dev):
            scratch = torch.zeros(N.lib().gsr_l1_loss_scratch_bytes(), dtype=torch.uint8, device=dev)
            N.check(N.lib().gsr_l1_loss(N.ptr(a), N.ptr(b), n, N.ptr(unit), out.data_ptr(), N.ptr(scratch),
                                        N.stream_ptr(dev)), "gsr_l1_loss")
        ctx.unit = unit
        ctx.shape = image.shape
        return out

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.unit is None:
            return None, None
        return (ctx.unit * grad_out).view(ctx.shape), None


def l1_loss(image: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """nn.L1Loss()(image, target) (reduction 'mean'); `target` gets no gradient."""
    if image.numel() == 0:
        raise RuntimeError("l1_loss: empty input")
    return _L1Loss.apply(image, target.detach())


def densify_fold(d_max_radii: torch.Tensor, d_accum: torch.Tensor, d_count: torch.Tensor,
                 max_radii2D: Optional[torch.Tensor], xyz_grad_accum: Optional[torch.Tensor],
                 xyz_grad_count: Optional[torch.Tensor]) -> None:
    """Fold a step's statistics deltas (StepBuffers, after the exchange) into the model's running statistics
    in one launch and without the host synchronisation boolean-mask indexing costs."""
    _need_cuda(d_count, "densify_fold")
    dev, P = d_count.device, d_count.numel()
    for name, t in (("d_max_radii", d_max_radii), ("d_accum", d_accum), ("d_count", d_count),
                    ("max_radii2D", max_radii2D), ("xyz_grad_accum", xyz_grad_accum),
                    ("xyz_grad_count", xyz_grad_count)):
        if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == P):
            raise RuntimeError(f"densify_fold: {name} must be a contiguous float32 CUDA tensor of {P} elements")
    with torch.cuda.device(dev):
        N.check(N.lib().gsr_densify_fold(P, N.ptr(d_max_radii), N.ptr(d_accum), N.ptr(d_count), N.ptr(max_radii2D),
                                         N.ptr(xyz_grad_accum), N.ptr(xyz_grad_count), N.stream_ptr(dev)),
                "gsr_densify_fold")


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam (no weight decay, no amsgrad, not maximize) with every parameter group updated by ONE
    kernel launch.  Same constructor arguments, `param_groups` (LR schedulers keep working:
    GSWarmUpExponentialDecayScheduler rewrites group["lr"]) and per-parameter state names
    (`step`, `exp_avg`, `exp_avg_sq`) as torch's, so state dicts are interchangeable.

    step(zero_grad=True) also zeroes the gradients in the same pass (optimizer.zero_grad(set_to_none=False)):
    with the gradients living in the flat all-reduce buffer (view_parallel.StepBuffers) that removes the
    buffer's memset from the next step.
    """

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps))

    @torch.no_grad()
    def step(self, closure=None, zero_grad: bool = False):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # (device, betas, eps, step) -> list of (param, grad, exp_avg, exp_avg_sq, lr)
        batches: Dict[tuple, list] = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                _need_cuda(p, "FusedAdam")
                g = p.grad
                if g.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                if not (p.dtype == torch.float32 and p.is_contiguous() and g.dtype == torch.float32
                        and g.is_contiguous()):
                    raise RuntimeError("FusedAdam needs contiguous float32 parameters and gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] = int(st["step"]) + 1
                key = (p.device, float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]), st["step"])
                batches.setdefault(key, []).append((p, g, st["exp_avg"], st["exp_avg_sq"], float(group["lr"])))
        for (dev, b1, b2, eps, t), items in batches.items():
            with torch.cuda.device(dev):
                for i in range(0, len(items), N.GSR_ADAM_MAX_GROUPS):
                    part = items[i:i + N.GSR_ADAM_MAX_GROUPS]
                    arr = (N.GsrAdamGroup * len(part))(*[
                        N.GsrAdamGroup(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), lr)
                        for p, g, m, v, lr in part if p.numel() > 0])
                    N.check(N.lib().gsr_adam_step(arr, len(arr), t, b1, b2, eps,
                                                  N.GSR_ADAM_ZERO_GRAD if zero_grad else 0, N.stream_ptr(dev)),
                            "gsr_adam_step")
        return loss


def parameter_groups_on(buffers, params: Dict[str, torch.Tensor]) -> None:
    """Point `.grad` of the model's parameters at their slices of the step's flat gradient buffer
    (view_parallel.StepBuffers), so that the rasterizer backward's fused `+=`, the one all-reduce and
    FusedAdam all work on the same memory: no per-parameter gradient tensors, no pack pass.

    params: name -> leaf tensor with names among means3D (P,3), shs (P,M,3) [or features_dc (P,1,3) +
    features_rest (P,M-1,3) when the SH block is stored split, gaussian_model.py:317-330 -- not
    representable as views of one (P,M,3) block, so only `shs` is accepted], opacities, scales, rotations.
    """
    names = dict(means3D="dL_dmeans3D", shs="dL_dsh", opacities="dL_dopacity", scales="dL_dscales",
                 rotations="dL_drotations")
    views = buffers.grads()
    for k, p in params.items():
        if k not in names:
            raise KeyError(f"unknown parameter {k!r}; expected one of {sorted(names)}")
        p.grad = views[names[k]].view(p.shape)


class _FusedSSIM(torch.autograd.Function):
    """mean SSIM map and its gradient w.r.t. img1 from ONE launch (gsr_fused_ssim)."""

    @staticmethod
    def forward(ctx, img1, img2, valid, train):
        _need_cuda(img1, "fused_ssim")
        if img1.dim() != 4 or img1.shape != img2.shape:
            raise RuntimeError(f"fused_ssim: expected two (B, C, H, W) tensors of one shape, got "
                               f"{tuple(img1.shape)} and {tuple(img2.shape)}")
        dev = img1.device
        a, b = N.f32c(img1), N.f32c(img2, dev)
        B, CH, H, W = a.shape
        out = torch.empty((), dtype=torch.float32, device=dev)
        want_grad = bool(train) and ctx.needs_input_grad[0]
        unit = torch.empty_like(a) if want_grad else None
        with torch.cuda.device(dev):
            scratch = torch.zeros(N.lib().gsr_fused_ssim_scratch_bytes(B * CH, H, W), dtype=torch.uint8, device=dev)
            N.check(N.lib().gsr_fused_ssim(N.ptr(a), N.ptr(b), B * CH, H, W, 0.01 ** 2, 0.03 ** 2, int(valid),
                                           N.ptr(unit), out.data_ptr(), N.ptr(scratch), N.stream_ptr(dev)),
                    "gsr_fused_ssim")
        ctx.unit = unit
        return out

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.unit is None:
            return None, None, None, None
        return ctx.unit * grad_out, None, None, None


def fused_ssim(img1: torch.Tensor, img2: torch.Tensor, padding: str = "same", train: bool = True) -> torch.Tensor:
    """fused_ssim.fused_ssim (submodules/fused-ssim/fused_ssim/__init__.py:33-41): mean SSIM of two (B, C, H, W)
    images, 11x11 Gaussian window, sigma 1.5, C1 = 0.01^2, C2 = 0.03^2; gradient flows to img1 only, as upstream.
    The reference's loss term is `1 - fused_ssim(rendered.unsqueeze(0), gt.unsqueeze(0))`
    (gs_lightning/lightning/gs_lightning_module.py:100,279)."""
    assert padding in ("same", "valid")
    return _FusedSSIM.apply(img1, img2.detach(), padding == "valid", train)


class _PhotometricLoss(torch.autograd.Function):
    """(1 - w) L1 + w (1 - ssim) and its gradient from ONE launch (gsr_photometric_loss)."""

    @staticmethod
    def forward(ctx, image, target, w_ssim):
        _need_cuda(image, "photometric_loss")
        if image.dim() != 4 or image.shape != target.shape:
            raise RuntimeError(f"photometric_loss: expected two (B, C, H, W) tensors of one shape, got "
                               f"{tuple(image.shape)} and {tuple(target.shape)}")
        dev = image.device
        a, b = N.f32c(image), N.f32c(target, dev)
        B, CH, H, W = a.shape
        out = torch.empty(3, dtype=torch.float32, device=dev)          # loss, L1, mean SSIM
        unit = torch.empty_like(a) if ctx.needs_input_grad[0] else None
        with torch.cuda.device(dev):
            scratch = torch.zeros(N.lib().gsr_fused_ssim_scratch_bytes(B * CH, H, W), dtype=torch.uint8, device=dev)
            p = out.data_ptr()
            N.check(N.lib().gsr_photometric_loss(N.ptr(a), N.ptr(b), B * CH, H, W, float(w_ssim), N.ptr(unit), p, p + 4,
                                                 p + 8, N.ptr(scratch), N.stream_ptr(dev)), "gsr_photometric_loss")
        ctx.unit = unit
        loss, l1, ssim = out[0], out[1], out[2]
        ctx.mark_non_differentiable(l1, ssim)
        return loss, l1, ssim

    @staticmethod
    def backward(ctx, grad_out, _g1, _g2):
        if ctx.unit is None:
            return None, None, None
        return ctx.unit * grad_out, None, None


def photometric_loss(image: torch.Tensor, target: torch.Tensor, w_ssim: float = 0.2):
    """GSLightningModule.calculate_loss's loss (gs_lightning/lightning/gs_lightning_module.py:278-291) for (B, C, H, W)
    images: returns (loss, recon, ssim) with loss = recon * (1 - w_ssim) + ssim * w_ssim, recon = nn.L1Loss(),
    ssim = 1 - fused_ssim(image, target) -- the three entries of the reference's loss_log.  One kernel reads the two
    images once and writes the combined gradient."""
    loss, l1, mean_ssim = _PhotometricLoss.apply(image, target.detach(), float(w_ssim))
    return loss, l1, 1.0 - mean_ssim
