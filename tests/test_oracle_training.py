"""The oracle of the training-step neighbours (oracle/training.py) pinned against the reference's own
implementations, which are torch CPU operators: nn.L1Loss (gs_lightning_module.py:99), torch.optim.Adam
(gs_lightning_module.py:111-127, configs eps 1e-15) and the boolean-mask statistics updates of
gaussian_model.py:174-181."""
import numpy as np
import torch

from oracle import training as OT


def test_l1_loss_matches_torch():
    g = torch.Generator().manual_seed(0)
    a = torch.rand(3, 37, 53, generator=g, requires_grad=True)
    b = torch.rand(3, 37, 53, generator=g)
    b[0, :4] = a.detach()[0, :4]                       # exact zeros in the difference: sign(0) = 0
    loss = torch.nn.L1Loss()(a.unsqueeze(0), b.unsqueeze(0))
    loss.backward()
    got_loss, got_grad = OT.l1_loss(a.detach().numpy(), b.numpy())
    assert abs(float(got_loss) - float(loss.detach())) <= 1e-7
    np.testing.assert_allclose(got_grad, a.grad.numpy(), rtol=0, atol=1e-12)


def test_adam_matches_torch_optim_adam_over_steps():
    g = torch.Generator().manual_seed(1)
    lrs = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 5e-2, 5e-3, 1e-3]       # the six groups of gs_lightning_module.py:120-127
    shapes = [(501, 3), (501, 1, 3), (501, 15, 3), (501, 1), (501, 3), (501, 4)]
    params = [torch.randn(s, generator=g).requires_grad_(True) for s in shapes]
    opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(params, lrs)], lr=0.0, eps=1e-15,
                           foreach=False)
    mine = [(p.detach().numpy().copy(), np.zeros(s, np.float32), np.zeros(s, np.float32)) for p, s in zip(params, shapes)]
    for step in range(1, 8):
        grads = [torch.randn(s, generator=g) * (10.0 ** float(torch.randint(-6, 1, (1,), generator=g))) for s in shapes]
        for p, gr in zip(params, grads):
            p.grad = gr.clone()
        opt.param_groups[0]["lr"] = lrs[0] * 0.97 ** step       # the xyz scheduler rewrites group["lr"] every step
        opt.step()
        for i, gr in enumerate(grads):
            lr = opt.param_groups[i]["lr"]
            mine[i] = OT.adam_step(mine[i][0], gr.numpy(), mine[i][1], mine[i][2], step, lr)
    for i, p in enumerate(params):
        st = opt.state[p]
        # torch's CPU kernels may contract a*b+c; agreement is to a few ulps of the operands (the lerp
        # cancels), not bit for bit: tolerance 2e-6 relative + 5e-7 of the array's largest magnitude
        for mine_t, ref_t in ((mine[i][0], p.detach()), (mine[i][1], st["exp_avg"]), (mine[i][2], st["exp_avg_sq"])):
            ref_np = ref_t.numpy()
            np.testing.assert_allclose(mine_t, ref_np, rtol=2e-6, atol=5e-7 * float(np.abs(ref_np).max()))


def test_densify_fold_matches_view_by_view_updates():
    g = torch.Generator().manual_seed(2)
    P, V = 1000, 5
    max_r, acc, cnt = torch.rand(P, generator=g) * 3, torch.rand(P, generator=g), torch.randint(0, 4, (P,), generator=g).float()
    ref = [max_r.clone(), acc.clone(), cnt.clone()]
    d_max, d_acc, d_cnt = torch.zeros(P), torch.zeros(P), torch.zeros(P)
    for _ in range(V):
        radii = torch.randint(0, 9, (P,), generator=g).float() * (torch.rand(P, generator=g) > 0.4)
        grad2d = torch.randn(P, 3, generator=g)
        vis = radii > 0
        # the reference, view by view (gaussian_model.py:174-181)
        ref[0][vis] = torch.max(ref[0][vis], radii[vis])
        ref[1][vis] += torch.norm(grad2d[vis, :2], dim=1)
        ref[2][vis] += 1
        # what the backward kernel leaves in the step's delta buffers
        d_max[vis] = torch.max(d_max[vis], radii[vis])
        d_acc[vis] += torch.norm(grad2d[vis, :2], dim=1)
        d_cnt[vis] += 1
    got = OT.densify_fold(d_max.numpy(), d_acc.numpy(), d_cnt.numpy(), max_r.numpy(), acc.numpy(), cnt.numpy())
    np.testing.assert_array_equal(got[0], ref[0].numpy())
    np.testing.assert_allclose(got[1], ref[1].numpy(), rtol=1e-6)
    np.testing.assert_array_equal(got[2], ref[2].numpy())


def _torch_ssim(img1, img2, padding):
    """The PyTorch SSIM the reference's own test compares fused-ssim with (submodules/fused-ssim/tests/test.py:14-54):
    grouped conv2d with the 11x11 Gaussian window, zero padding; 'valid' crops 5 pixels before the mean."""
    import torch.nn.functional as F
    ch = img1.shape[1]
    g = torch.tensor([np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)) for x in range(11)], dtype=img1.dtype)
    g = g / g.sum()
    win = (g[:, None] @ g[None, :]).expand(ch, 1, 11, 11).contiguous()
    blur = lambda t: F.conv2d(t, win, padding=5, groups=ch)
    mu1, mu2 = blur(img1), blur(img2)
    s11, s22, s12 = blur(img1 * img1) - mu1 ** 2, blur(img2 * img2) - mu2 ** 2, blur(img1 * img2) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2)) / ((mu1 ** 2 + mu2 ** 2 + 0.01 ** 2) * (s11 + s22 + 0.03 ** 2))
    if padding == "valid":
        m = m[:, :, 5:-5, 5:-5]
    return m.mean()


def test_ssim_oracle_matches_torch_autograd():
    g = torch.Generator().manual_seed(7)
    for shape, padding in (((2, 3, 37, 45), "same"), ((1, 3, 40, 33), "valid"), ((1, 1, 12, 11), "valid")):
        a = torch.rand(shape, generator=g, dtype=torch.float64, requires_grad=True)
        b = (a.detach() + 0.2 * torch.rand(shape, generator=g, dtype=torch.float64)).clamp(0, 1)
        val = _torch_ssim(a, b, padding)
        val.backward()
        got_val, got_grad = OT.fused_ssim(a.detach().numpy(), b.numpy(), padding)
        assert abs(got_val - float(val.detach())) <= 1e-12
        np.testing.assert_allclose(got_grad, a.grad.numpy(), rtol=1e-9, atol=1e-13)
