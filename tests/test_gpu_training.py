"""GPU parity of the training-step neighbours against oracle/training.py and the
reference's own operators (torch.optim.Adam, nn.L1Loss), all through the C-ABI of libgsr_b200.so:
fused L1 loss, multi-tensor Adam, densification fold, and the fused gradient accumulation of the autograd API."""
import contextlib
import copy

import numpy as np
import pytest
import torch

import gpu_utils as U
from oracle import training as OT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(3, 1080, 1920), (3, 37, 53), (1, 1, 5), (7,)])
def test_l1_loss_value_and_gradient(shape):
    from gaussian_splatting_lightning_b200.training import l1_loss
    g = torch.Generator().manual_seed(0)
    a = torch.rand(shape, generator=g)
    b = torch.rand(shape, generator=g)
    b.view(-1)[:3] = a.view(-1)[:3]                       # sign(0) = 0
    want_loss, want_grad = OT.l1_loss(a.numpy(), b.numpy())
    x = a.to(U.DEV).requires_grad_(True)
    loss = l1_loss(x, b.to(U.DEV))
    (loss * 0.75).backward()
    assert abs(float(loss.detach()) - float(want_loss)) <= 2e-7 * max(1.0, abs(float(want_loss)))
    # gradient: exactly +-(0.75/n) or 0 -- one float32 product, bit-exact against the oracle
    np.testing.assert_array_equal(x.grad.cpu().numpy(), (want_grad * np.float32(0.75)).astype(np.float32))
    # and against the reference's operator
    y = a.to(U.DEV).requires_grad_(True)
    ref = torch.nn.L1Loss()(y.unsqueeze(0), b.to(U.DEV).unsqueeze(0))
    ref.backward()
    assert abs(float(ref.detach()) - float(loss.detach())) <= 1e-6
    np.testing.assert_allclose(x.grad.cpu().numpy(), 0.75 * y.grad.cpu().numpy(), rtol=1e-6, atol=0)


def test_l1_loss_is_deterministic_and_rejects_cpu():
    from gaussian_splatting_lightning_b200.training import l1_loss
    a = torch.rand(3, 300, 500, device=U.DEV)
    b = torch.rand(3, 300, 500, device=U.DEV)
    vals = {float(l1_loss(a, b)) for _ in range(5)}
    assert len(vals) == 1
    with pytest.raises(RuntimeError):
        l1_loss(a.cpu(), b.cpu())


def _groups(P, M, g):
    shapes = [(P, 3), (P, M, 3), (P, 1), (P, 3), (P, 4)]
    lrs = [1.6e-4, 2.5e-3, 5e-2, 5e-3, 1e-3]
    return shapes, lrs, [torch.randn(s, generator=g) for s in shapes]


@pytest.mark.parametrize("P", [1, 333, 40_000])
def test_fused_adam_bit_exact_against_oracle_and_close_to_torch(P):
    from gaussian_splatting_lightning_b200.training import FusedAdam
    g = torch.Generator().manual_seed(3)
    shapes, lrs, init = _groups(P, 16, g)
    mine = [t.clone().to(U.DEV).requires_grad_(True) for t in init]
    ref = [t.clone().to(U.DEV).requires_grad_(True) for t in init]
    opt = FusedAdam([{"params": [p], "lr": lr} for p, lr in zip(mine, lrs)], lr=0.0, eps=1e-15)
    topt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(ref, lrs)], lr=0.0, eps=1e-15)
    orc = [(t.numpy().copy(), np.zeros(s, np.float32), np.zeros(s, np.float32)) for t, s in zip(init, shapes)]
    for step in range(1, 6):
        grads = [torch.randn(s, generator=g) * 10.0 ** float(torch.randint(-5, 1, (1,), generator=g)) for s in shapes]
        for p, q, gr in zip(mine, ref, grads):
            p.grad = gr.clone().to(U.DEV)
            q.grad = gr.clone().to(U.DEV)
        lr0 = lrs[0] * 0.9 ** step
        opt.param_groups[0]["lr"] = lr0                      # scheduler rewrites group["lr"]
        topt.param_groups[0]["lr"] = lr0
        opt.step(zero_grad=(step % 2 == 0))
        topt.step()
        for i, gr in enumerate(grads):
            orc[i] = OT.adam_step(orc[i][0], gr.numpy(), orc[i][1], orc[i][2], step, opt.param_groups[i]["lr"])
            if step % 2 == 0:
                assert float(mine[i].grad.abs().max()) == 0.0     # GSR_ADAM_ZERO_GRAD
            else:
                assert torch.equal(mine[i].grad.cpu(), gr)
    for i, p in enumerate(mine):
        st = opt.state[p]
        np.testing.assert_array_equal(p.detach().cpu().numpy(), orc[i][0])
        np.testing.assert_array_equal(st["exp_avg"].cpu().numpy(), orc[i][1])
        np.testing.assert_array_equal(st["exp_avg_sq"].cpu().numpy(), orc[i][2])
        tst = topt.state[ref[i]]
        for a, b in ((p.detach(), ref[i].detach()), (st["exp_avg"], tst["exp_avg"]), (st["exp_avg_sq"], tst["exp_avg_sq"])):
            b = b.cpu().numpy()
            np.testing.assert_allclose(a.cpu().numpy(), b, rtol=2e-6, atol=5e-7 * float(np.abs(b).max()))
        assert int(st["step"]) == 5


def test_fused_adam_unaligned_views_and_state_dict_roundtrip():
    from gaussian_splatting_lightning_b200.training import FusedAdam
    g = torch.Generator().manual_seed(4)
    base = torch.randn(10_001, generator=g).to(U.DEV)
    p = base[1:].detach().requires_grad_(True)            # 4-byte aligned only: scalar path
    q = p.detach().clone().requires_grad_(True)
    gr = torch.randn(10_000, generator=g).to(U.DEV)
    p.grad, q.grad = gr.clone(), gr.clone()
    a = FusedAdam([p], lr=1e-2, eps=1e-15)
    b = FusedAdam([q], lr=1e-2, eps=1e-15)
    a.step(); b.step()
    assert torch.equal(p.detach(), q.detach())
    sd = copy.deepcopy(a.state_dict())                    # load_state_dict would otherwise alias a's moment tensors
    c = torch.optim.Adam([q], lr=1e-2, eps=1e-15)
    sd["state"][0]["step"] = torch.tensor(float(sd["state"][0]["step"]))
    c.load_state_dict({"state": sd["state"], "param_groups": c.state_dict()["param_groups"]})
    p.grad, q.grad = gr.clone(), gr.clone()
    a.step(); c.step()
    np.testing.assert_allclose(p.detach().cpu().numpy(), q.detach().cpu().numpy(), rtol=2e-6, atol=1e-8)


def test_densify_fold_bit_exact():
    from gaussian_splatting_lightning_b200.training import densify_fold
    rng = np.random.default_rng(5)
    P = 70_001
    d_cnt = rng.integers(0, 4, P).astype(np.float32)
    d_max = (rng.integers(1, 50, P) * (d_cnt > 0)).astype(np.float32)
    d_acc = (rng.random(P) * (d_cnt > 0)).astype(np.float32)
    mr, ac, ct = rng.random(P).astype(np.float32) * 30, rng.random(P).astype(np.float32), rng.integers(0, 9, P).astype(np.float32)
    want = OT.densify_fold(d_max, d_acc, d_cnt, mr, ac, ct)
    t = [torch.tensor(x, device=U.DEV) for x in (d_max, d_acc, d_cnt, mr, ac, ct)]
    densify_fold(*t)
    for got, w in zip(t[3:], want):
        np.testing.assert_array_equal(got.cpu().numpy(), w)


def test_fused_gradient_accumulation_equals_autograd_accumulation():
    """accumulate_gradients_into: V views through the autograd API with the fused `+=` give the same sums
    (and the same densification statistics) as autograd's own .grad accumulation + the reference's updates."""
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer, accumulate_gradients_into
    from gaussian_splatting_lightning_b200.training import l1_loss, parameter_groups_on
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers
    P, deg, W, H, V = 5000, 3, 160, 96, 3
    names = ("means3D", "shs", "opacities", "scales", "rotations")
    cams, d = [], None
    for k in range(V):
        cam, arrays = U.synthetic_case(P, deg, W, H, seed=21, view_k=k, num_views=V, bg=(0.1, 0.2, 0.3))
        cams.append(cam)
        d = d or U.to_dev(arrays)
    gts = [torch.rand(3, H, W, generator=torch.Generator().manual_seed(k)).to(U.DEV) for k in range(V)]

    def run(fused):
        params = {k: d[k].clone().requires_grad_(True) for k in names}
        bufs = StepBuffers(P, 16, U.DEV)
        sps, radii_all = [], []
        if fused:
            parameter_groups_on(bufs, params)
        for k in range(V):
            sp = torch.zeros_like(params["means3D"], requires_grad=True)
            rs = U.settings_from_camera(cams[k])
            ctx = accumulate_gradients_into(bufs.grads(), bufs.densify_stats()) if fused else contextlib.nullcontext()
            with ctx:
                color, radii, _ = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=sp, shs=params["shs"],
                                                         colors_precomp=None, opacities=params["opacities"],
                                                         scales=params["scales"], rotations=params["rotations"],
                                                         cov3D_precomp=None)
            l1_loss(color, gts[k]).backward()          # backward outside the context: the sink travels with the graph
            sps.append(sp); radii_all.append(radii)
        return params, bufs, sps, radii_all

    pa, ba, spa, ra = run(True)
    pb, bb, spb, rb = run(False)
    for k in names:
        a, b = pa[k].grad, pb[k].grad
        assert a.data_ptr() == ba.grads()[dict(means3D="dL_dmeans3D", shs="dL_dsh", opacities="dL_dopacity",
                                               scales="dL_dscales", rotations="dL_drotations")[k]].data_ptr()
        s = max(1e-12, float(b.abs().max()))
        assert float((a - b).abs().max()) <= 2e-5 * s, k       # same kernels; only the summation order differs
    mr, ac, ct = torch.zeros(P, device=U.DEV), torch.zeros(P, device=U.DEV), torch.zeros(P, device=U.DEV)
    for sp_f, sp, radii in zip(spa, spb, rb):
        # the dummy means2D leaf still gets its gradient (two runs differ by the order of the L2 reductions)
        assert float((sp_f.grad - sp.grad).abs().max()) <= 2e-5 * float(sp.grad.abs().max())
        vis = radii > 0
        mr[vis] = torch.max(mr[vis], radii[vis].float())
        ac[vis] += torch.norm(sp.grad[vis, :2], dim=1)
        ct[vis] += 1
    st = ba.densify_stats()
    assert torch.equal(st["max_radii2D"], mr) and torch.equal(st["xyz_grad_count"], ct)
    np.testing.assert_allclose(st["xyz_grad_accum"].cpu().numpy(), ac.cpu().numpy(), rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("shape,padding", [((1, 3, 200, 320), "same"), ((2, 3, 67, 45), "same"), ((1, 3, 64, 96), "valid"),
                                           ((1, 1, 12, 11), "valid"), ((1, 3, 33, 31), "same")])
def test_fused_ssim_matches_oracle(shape, padding):
    from gaussian_splatting_lightning_b200.training import fused_ssim
    g = torch.Generator().manual_seed(11)
    a = torch.rand(shape, generator=g)
    b = (a + 0.2 * torch.rand(shape, generator=g)).clamp(0, 1)
    want_val, want_grad = OT.fused_ssim(a.numpy(), b.numpy(), padding)
    x = a.to(U.DEV).requires_grad_(True)
    val = fused_ssim(x, b.to(U.DEV), padding=padding)
    (1.0 - val).backward()                      # the reference's loss term: 1 - fused_ssim(...)
    assert abs(float(val.detach()) - want_val) <= 2e-6
    s = float(np.abs(want_grad).max())
    # tolerance: FP32 against the float64 oracle, 1e-4 of the largest gradient entry
    assert float(np.abs(-x.grad.cpu().numpy() - want_grad).max()) <= 1e-4 * s
    assert float(fused_ssim(x.detach(), b.to(U.DEV), padding=padding, train=False)) == float(val.detach())


def _reference_fused_ssim(mod, img1, img2, padding="same"):
    """fused_ssim.fused_ssim restated over the reference's compiled kernels (fused_ssim/__init__.py:8-41)."""
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = mod.fusedssim(C1, C2, img1, img2, True)
    m = ssim_map[:, :, 5:-5, 5:-5] if padding == "valid" else ssim_map
    dL_dmap = torch.full_like(m, 1.0 / m.numel())
    if padding == "valid":
        full = torch.zeros_like(img1)
        full[:, :, 5:-5, 5:-5] = dL_dmap
        dL_dmap = full
    grad = mod.fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
    return m.mean(), grad


@pytest.mark.parametrize("shape,padding", [((1, 3, 1080, 1920), "same"), ((2, 3, 211, 173), "same"),
                                           ((1, 3, 300, 200), "valid")])
def test_fused_ssim_matches_the_reference_cuda_kernels(shape, padding):
    """oracle/_ref/fused_ssim_cuda.so is the reference's own fused-ssim extension compiled for sm_100a by
    oracle/build_ref.py from /root/reference/submodules/fused-ssim; it travels to the GPU box with the snapshot."""
    from oracle import build_ref
    if not build_ref.available():
        pytest.skip("oracle/_ref/fused_ssim_cuda.so has not been built (needs /root/reference)")
    from gaussian_splatting_lightning_b200.training import fused_ssim
    mod = build_ref.load()
    g = torch.Generator().manual_seed(12)
    a = torch.rand(shape, generator=g).to(U.DEV)
    b = (a + 0.3 * torch.rand(shape, generator=g).to(U.DEV)).clamp(0, 1)
    ref_val, ref_grad = _reference_fused_ssim(mod, a, b, padding)
    x = a.clone().requires_grad_(True)
    val = fused_ssim(x, b, padding=padding)
    val.backward()
    assert abs(float(val.detach()) - float(ref_val)) <= 2e-6
    s = float(ref_grad.abs().max())
    assert float((x.grad - ref_grad).abs().max()) <= 1e-4 * s


@pytest.mark.parametrize("shape,w", [((1, 3, 200, 320), 0.2), ((2, 3, 67, 45), 0.5), ((1, 3, 33, 31), 0.0)])
def test_photometric_loss_equals_the_two_terms(shape, w):
    """(1 - w) L1 + w (1 - ssim) from one launch against the oracle's two terms (and the reference's operators)."""
    from gaussian_splatting_lightning_b200.training import photometric_loss
    g = torch.Generator().manual_seed(13)
    a = torch.rand(shape, generator=g)
    b = (a + 0.2 * torch.rand(shape, generator=g)).clamp(0, 1)
    l1_val, l1_grad = OT.l1_loss(a.numpy(), b.numpy())
    ss_val, ss_grad = OT.fused_ssim(a.numpy(), b.numpy(), "same")
    want = (1 - w) * float(l1_val) + w * (1 - ss_val)
    want_grad = (1 - w) * l1_grad.astype(np.float64) - w * ss_grad
    x = a.to(U.DEV).requires_grad_(True)
    loss, recon, ssim = photometric_loss(x, b.to(U.DEV), w)
    (2.0 * loss).backward()
    assert abs(float(loss.detach()) - want) <= 3e-6
    assert abs(float(recon) - float(l1_val)) <= 2e-7 and abs(float(ssim) - (1 - ss_val)) <= 2e-6
    assert not recon.requires_grad and not ssim.requires_grad
    s = float(np.abs(want_grad).max())
    assert float(np.abs(x.grad.cpu().numpy() - 2.0 * want_grad).max()) <= 2e-4 * s
