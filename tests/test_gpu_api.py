"""The drop-in Python API on a GPU: GaussianRasterizationSettings / GaussianRasterizer /
rasterize_gaussians autograd Function used the way the reference's call sites use them
(gs_lightning/lightning/gs_lightning_module.py:316-348, scripts/render_trained_image.py:80-124,
tests/rasterizer_python/test_mark_visible.py:11-21)."""
import numpy as np
import pytest
import torch

import gpu_utils as U

pytestmark = pytest.mark.gpu


def _scene(P=6000, deg=3, W=160, H=96, seed=11):
    cam, arrays = U.synthetic_case(P, deg, W, H, seed=seed, bg=(0.1, 0.2, 0.3))
    return cam, arrays, U.to_dev(arrays)


def test_training_call_site_autograd(oracle_lib):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    cam, arrays, d = _scene()
    rs = U.settings_from_camera(cam)
    # tanfov arrive as 0-d float64 tensors at the training call site
    rs = rs._replace(tanfovx=torch.tensor(cam.tanfovx, dtype=torch.float64),
                     tanfovy=torch.tensor(cam.tanfovy, dtype=torch.float64))
    assert isinstance(rs, GaussianRasterizationSettings)
    params = {k: d[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    screenspace_points = torch.zeros_like(params["means3D"], requires_grad=True)
    rasterizer = GaussianRasterizer(raster_settings=rs)
    color, radii, invdepth = rasterizer(means3D=params["means3D"], means2D=screenspace_points, shs=params["shs"],
                                        colors_precomp=None, opacities=params["opacities"], scales=params["scales"],
                                        rotations=params["rotations"], cov3D_precomp=None)
    assert color.shape == (3, cam.image_height, cam.image_width) and invdepth.shape == (1, cam.image_height, cam.image_width)
    assert radii.dtype == torch.int32 and radii.shape == (6000,) and not radii.requires_grad
    rng = np.random.default_rng(0)
    wC = rng.standard_normal(color.shape).astype(np.float32)
    wD = rng.standard_normal(invdepth.shape).astype(np.float32)
    loss = (color * torch.tensor(wC, device=U.DEV)).sum() + (invdepth * torch.tensor(wD, device=U.DEV)).sum()
    loss.backward()
    ref = U.run_oracle_forward(cam, arrays)
    want = oracle_lib.backward(ref, wC, wD[0])
    pairs = [("means3D", "dL_dmeans3D"), ("shs", "dL_dsh"), ("opacities", "dL_dopacity"), ("scales", "dL_dscales"),
             ("rotations", "dL_drotations")]
    got = {k: params[p].grad.cpu().numpy() for p, k in pairs}
    # the dummy means2D leaf receives dL_dmean2D; densification reads [:, :2] (gaussian_model.py:180)
    g = got["dL_dmeans2D"] = screenspace_points.grad.cpu().numpy()
    U.assert_grads_close(got, want, names=[k for _, k in pairs] + ["dL_dmeans2D"], label="autograd call site")
    assert float(np.abs(g[:, 2]).max()) == 0.0


def test_color_only_loss_and_no_grad_render():
    from diff_gaussian_rasterization import GaussianRasterizer
    cam, arrays, d = _scene(P=2000, deg=1, W=64, H=48)
    rs = U.settings_from_camera(cam)
    m = d["means3D"].clone().requires_grad_(True)
    color, radii, invdepth = GaussianRasterizer(rs)(means3D=m, means2D=None, shs=d["shs"], colors_precomp=None,
                                                    opacities=d["opacities"], scales=d["scales"],
                                                    rotations=d["rotations"], cov3D_precomp=None)
    color.mean().backward()  # invdepth gets no gradient: dL_dinvdepth = NULL path
    assert torch.isfinite(m.grad).all() and float(m.grad.abs().sum()) > 0
    with torch.no_grad():  # scripts/render_trained_image.py:113-124
        c2, r2, d2 = GaussianRasterizer(rs)(means3D=d["means3D"], means2D=None, shs=d["shs"], colors_precomp=None,
                                            opacities=d["opacities"], scales=d["scales"], rotations=d["rotations"],
                                            cov3D_precomp=None)
    assert torch.equal(c2, color.detach()) and torch.equal(r2, radii)


def test_argument_validation_errors():
    from diff_gaussian_rasterization import GaussianRasterizer
    cam, arrays, d = _scene(P=100, deg=0, W=32, H=32)
    r = GaussianRasterizer(U.settings_from_camera(cam))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=d["means3D"], means2D=None, opacities=d["opacities"], scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=d["means3D"], means2D=None, opacities=d["opacities"], shs=d["shs"])
    with pytest.raises(RuntimeError, match="num_points, 3"):
        r(means3D=d["means3D"][:, :2], means2D=None, opacities=d["opacities"], shs=d["shs"], scales=d["scales"],
          rotations=d["rotations"])
    with pytest.raises(RuntimeError, match="CUDA"):
        r(means3D=d["means3D"].cpu(), means2D=None, opacities=d["opacities"], shs=d["shs"], scales=d["scales"],
          rotations=d["rotations"])


def test_gs_lightning_rasterize_entry_points(oracle_lib):
    from gaussian_splatting_lightning_b200.rasterize import markVisible, rasterize_gaussian
    cam, arrays, d = _scene(P=4000, deg=2, W=100, H=60)
    rs = U.settings_from_camera(cam)
    img, radius, depth = rasterize_gaussian(
        means3D=d["means3D"], opacities=d["opacities"], scales=d["scales"], rotations=d["rotations"], shs=d["shs"],
        scale_modifier=1.0, image_width=cam.image_width, image_height=cam.image_height, tanfovx=cam.tanfovx,
        tanfovy=cam.tanfovy, viewmatrix=rs.viewmatrix, projmatrix=rs.projmatrix, campos=rs.campos,
        background=rs.bg, sh_degree=2)
    ref = U.run_oracle_forward(cam, arrays)
    assert radius.dtype == torch.float32
    assert np.array_equal(radius.cpu().numpy().astype(np.int32), ref["radii"])
    frac, mx = U.close_fraction(img.cpu().numpy(), ref["color"])
    assert frac <= 2e-4 and mx <= 2e-2
    vis = markVisible(d["means3D"], rs.viewmatrix, rs.projmatrix)
    assert np.array_equal(vis.cpu().numpy(), oracle_lib.mark_visible(arrays["means3D"], cam.viewmatrix))


def test_backward_accumulate_stats_and_optional_outputs():
    """GSR_GRADS_ACCUMULATE sums views into one buffer; the densification statistics are taken per view
    (gaussian_model.py:174-181, visible_mask = radii > 0); NULL outputs are not produced."""
    from gaussian_splatting_lightning_b200.rasterizer import backward_raw
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")
    P = 9000
    scene = U.synthetic.make_scene(P, 3, 192, 128, seed=21)
    arrays = {k: v.numpy() for k, v in scene.items()}
    rng = np.random.default_rng(5)
    flat = {n: None for n in names}
    max_radii = torch.zeros(P, device=U.DEV)
    accum = torch.zeros(P, device=U.DEV)
    count = torch.zeros(P, device=U.DEV)
    want_sum, want_max, want_accum, want_count = None, np.zeros(P, np.float32), np.zeros(P, np.float64), np.zeros(P)
    for k in (0, 3, 5):  # three views of an 8-view orbit; view 3/5 cull part of the scene
        c = U.synthetic.make_camera(192, 128, k=k, num_views=8)
        import oracle
        cam = oracle.Camera(128, 192, c["tanfovx"], c["tanfovy"], c["viewmatrix"].numpy(), c["projmatrix"].numpy(),
                            c["campos"].numpy(), sh_degree=3)
        out = U.run_cuda_forward(cam, arrays)
        dC = torch.tensor(rng.standard_normal((3, 128, 192)).astype(np.float32), device=U.DEV)
        d = out["d"]
        args = (out["rs"], d["means3D"], out["radii"], d["shs"], None, d["opacities"], d["scales"], d["rotations"],
                None, dC, None, out["geom"], out["binning"], out["img"], out["R"])
        plain = dict(zip(names, backward_raw(*args)))
        if flat["dL_dsh"] is None:
            for n in ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations"):
                flat[n] = torch.zeros_like(plain[n])
        acc = backward_raw(*args, accumulate_into={n: t for n, t in flat.items() if t is not None},
                           stats=dict(max_radii2D=max_radii, xyz_grad_accum=accum, xyz_grad_count=count))
        assert acc[0] is None and acc[1] is None and acc[4] is None  # dL_dmeans2D / dL_dcolors / dL_dcov3D not wanted
        vis = (out["radii"] > 0).cpu().numpy()
        assert 0 < vis.sum()
        want_sum = {n: plain[n].double().cpu().numpy() + (0 if want_sum is None else want_sum[n]) for n in flat if flat[n] is not None}
        want_max = np.where(vis, np.maximum(want_max, out["radii"].cpu().numpy().astype(np.float32)), want_max)
        want_accum += np.where(vis, np.linalg.norm(plain["dL_dmeans2D"].double().cpu().numpy()[:, :2], axis=1), 0.0)
        want_count += vis
        # only-what-is-asked-for path of the non-accumulating call
        only = backward_raw(*args, want=("dL_dopacity", "dL_dsh"))
        assert [t is not None for t in only] == [False, False, True, False, False, True, False, False]
        assert torch.equal(only[2], plain["dL_dopacity"]) or torch.allclose(only[2], plain["dL_dopacity"], rtol=1e-4, atol=1e-6)
    torch.cuda.synchronize()
    for n, w in want_sum.items():
        g = flat[n].cpu().numpy().astype(np.float64)
        s = max(1.0, float(np.abs(w).max()))
        assert np.abs(g - w).max() / s <= 2e-4, n  # atomics make each backward run-to-run different in the last bits
    assert np.array_equal(max_radii.cpu().numpy(), want_max)
    assert np.array_equal(count.cpu().numpy(), want_count.astype(np.float32))
    assert np.abs(accum.cpu().numpy() - want_accum).max() <= 2e-4 * max(1.0, want_accum.max())


def test_ply_checkpoint_and_collated_cameras_through_the_render_path(tmp_path):
    """SURVEY 8(f) rows 3-4 end to end: Gaussians saved to a PLY, loaded straight to the device, rendered for a
    default_collate'd batch of ColmapDataset-style items; same pixels as rendering the original tensors per view."""
    from torch.utils.data import default_collate
    from gaussian_splatting_lightning_b200 import cameras, io, synthetic
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer
    P, deg, W, H, V = 4000, 3, 160, 96, 3
    scene = synthetic.make_scene(P, deg, W, H, seed=5)
    path = str(tmp_path / "point_cloud.ply")
    io.save_ply(path, scene["means3D"], scene["shs"][:, :1], scene["shs"][:, 1:], scene["opacities"], scene["scales"],
                scene["rotations"])
    ck = io.load_ply(path, device=U.DEV, canonical=True)
    shs = torch.cat([ck["features_dc"], ck["features_rest"]], 1)
    assert ck["xyz"].is_cuda and torch.equal(shs.cpu(), scene["shs"]) and ck["active_sh_degree"] == 3
    items = []
    for k in range(V):
        c = synthetic.make_camera(W, H, k=k, num_views=V)
        items.append(dict(image=torch.zeros(3, H, W), background=torch.tensor([0.1, 0.2, 0.3]), tanfovx=c["tanfovx"],
                          tanfovy=c["tanfovy"], viewmatrix=c["viewmatrix"], projmatrix=c["projmatrix"], campos=c["campos"]))
    sets = cameras.settings_from_batch(default_collate(items), sh_degree=deg, device=U.DEV)
    d = {k: v.to(U.DEV) for k, v in scene.items()}
    for s, it in zip(sets, items):
        a = GaussianRasterizer(s)(means3D=ck["xyz"], means2D=None, shs=shs, opacities=ck["opacity"],
                                  scales=ck["scaling"], rotations=ck["rotation"])[0]
        b = GaussianRasterizer(cameras.settings_from_item(it, deg))(means3D=d["means3D"], means2D=None, shs=d["shs"],
                                                                   opacities=d["opacities"], scales=d["scales"],
                                                                   rotations=d["rotations"])[0]
        assert torch.equal(a, b)


@pytest.mark.parametrize("P", [5001, 5002, 5003])
def test_step_buffers_with_P_not_a_multiple_of_4(P):
    """after a densify / prune P % 4 != 0; every section of the step buffer must still satisfy
    gsr_backward's alignment (dL_drotations 16 bytes) -- through render_step and accumulate_gradients_into."""
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer, accumulate_gradients_into, backward_raw
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers, render_step
    W, H, deg = 160, 96, 3
    scene = {k: v.to(U.DEV) for k, v in U.synthetic.make_scene(P, deg, W, H, seed=P).items()}
    cams = [U.settings_from_camera(U.synthetic_case(8, deg, W, H, view_k=k, num_views=8)[0]) for k in (0, 1)]
    g = torch.Generator().manual_seed(3)
    dC = torch.randn(3, H, W, generator=g).to(U.DEV)
    bufs = StepBuffers(P, (deg + 1) ** 2, U.DEV)
    for v in bufs.grads().values():
        assert v.data_ptr() % 128 == 0
    render_step(cams, scene, dC, None, bufs, exchange=False)
    step_sum = {n: t.clone() for n, t in bufs.grads().items()}
    # the same two views one at a time through the plain (non-accumulating) backward
    want = None
    for rs in cams:
        from gaussian_splatting_lightning_b200.rasterizer import forward_raw
        R, color, radii, geom, binning, img, invd, cap = forward_raw(rs, scene["means3D"], scene["shs"], None,
                                                                    scene["opacities"], scene["scales"], scene["rotations"], None)
        gr = backward_raw(rs, scene["means3D"], radii, scene["shs"], None, scene["opacities"], scene["scales"],
                          scene["rotations"], None, dC, None, geom, binning, img, R)
        named = dict(dL_dopacity=gr[2], dL_dmeans3D=gr[3], dL_dsh=gr[5], dL_dscales=gr[6], dL_drotations=gr[7])
        want = named if want is None else {n: want[n] + named[n] for n in named}
    for n, w in want.items():
        s = max(1.0, float(w.abs().max()))
        assert float((step_sum[n] - w).abs().max()) / s <= 2e-5, n
    # and through the autograd API with the fused accumulation
    bufs.zero_()
    params = {k: scene[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    for rs in cams:
        with accumulate_gradients_into(bufs.grads(), bufs.densify_stats()):
            color, radii, invd = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=None, shs=params["shs"],
                                                        opacities=params["opacities"], scales=params["scales"],
                                                        rotations=params["rotations"])
        (color * dC).sum().backward()
    for n, w in want.items():
        s = max(1.0, float(w.abs().max()))
        assert float((bufs.grads()[n] - w).abs().max()) / s <= 2e-5, n
    assert all(p.grad is None for p in params.values())      # leaves: the sink took their gradients


def test_sink_does_not_cut_the_chain_through_activations():
    """the reference passes ACTIVATIONS to the rasterizer (exp(_scaling), sigmoid(_opacity),
    normalize(_rotation), cat(dc, rest): gs_lightning_module.py:337-348).  Inside accumulate_gradients_into the raw
    parameters must still receive their gradients through autograd -- the sink only takes over leaf inputs."""
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer, accumulate_gradients_into
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers
    P, W, H, deg = 6000, 160, 96, 3
    scene = {k: v.to(U.DEV) for k, v in U.synthetic.make_scene(P, deg, W, H, seed=77).items()}
    rs = U.settings_from_camera(U.synthetic_case(8, deg, W, H)[0])
    g = torch.Generator().manual_seed(5)
    dC = torch.randn(3, H, W, generator=g).to(U.DEV)

    def raw_params():
        return dict(xyz=scene["means3D"].clone().requires_grad_(True),
                    scaling=scene["scales"].log().requires_grad_(True),
                    opacity=torch.logit(scene["opacities"]).requires_grad_(True),
                    rotation=(scene["rotations"] * 1.7).requires_grad_(True),
                    dc=scene["shs"][:, :1].clone().requires_grad_(True),
                    rest=scene["shs"][:, 1:].clone().requires_grad_(True))

    def render(p):
        return GaussianRasterizer(rs)(means3D=p["xyz"], means2D=None, shs=torch.cat([p["dc"], p["rest"]], 1),
                                      opacities=torch.sigmoid(p["opacity"]), scales=torch.exp(p["scaling"]),
                                      rotations=torch.nn.functional.normalize(p["rotation"]))[0]

    a = raw_params()
    (render(a) * dC).sum().backward()
    b = raw_params()
    bufs = StepBuffers(P, (deg + 1) ** 2, U.DEV)
    with accumulate_gradients_into(bufs.grads(), bufs.densify_stats()):
        img = render(b)
    (img * dC).sum().backward()
    for k in a:
        # xyz IS a leaf input of the rasterizer: its gradient went into the sink instead
        got = bufs.grads()["dL_dmeans3D"] if k == "xyz" else b[k].grad
        assert got is not None, k
        s = max(1.0, float(a[k].grad.abs().max()))
        assert float((a[k].grad - got).abs().max()) / s <= 2e-5, k
    assert b["xyz"].grad is None and float(bufs.densify_stats()["xyz_grad_count"].sum()) > 0


def test_both_bindings_of_the_c_abi_agree(monkeypatch):
    """`_C` (torch C++ extension, the default FFI) and the ctypes binding (GSR_FFI=ctypes) drive the same C-ABI:
    same pixels, same integers; gradients equal up to the order of the atomics."""
    from gaussian_splatting_lightning_b200 import _native as N
    cam, arrays = U.synthetic_case(12_000, 3, 224, 128, seed=4, bg=(0.2, 0.1, 0.4))
    rng = np.random.default_rng(2)
    dC = rng.standard_normal((3, 128, 224)).astype(np.float32)
    dD = rng.standard_normal((1, 128, 224)).astype(np.float32)
    assert N.ext() is not None, "_C.so has not been built"
    a = U.run_cuda_forward(cam, arrays)
    ga = U.run_cuda_backward(a, dC, dD)
    monkeypatch.setenv("GSR_FFI", "ctypes")
    assert N.ext() is None
    b = U.run_cuda_forward(cam, arrays)
    gb = U.run_cuda_backward(b, dC, dD)
    monkeypatch.delenv("GSR_FFI")
    assert a["R"] == b["R"] and torch.equal(a["color"], b["color"]) and torch.equal(a["radii"], b["radii"])
    assert torch.equal(a["invdepth"], b["invdepth"])
    for k in ga:
        s = max(1.0, float(np.abs(ga[k]).max()))
        assert np.abs(ga[k] - gb[k]).max() / s <= 1e-5, k


def test_upstream_signature_of_the_extension():
    """diff_gaussian_rasterization._C.rasterize_gaussians / rasterize_gaussians_backward / mark_visible called the way
    upstream's Python front end calls them (positional, empty tensors for absent inputs): 7- and 8-tuples."""
    import diff_gaussian_rasterization as dgr
    C_ = dgr._C
    cam, arrays = U.synthetic_case(6000, 2, 160, 96, seed=8)
    rs = U.settings_from_camera(cam)
    d = U.to_dev(arrays)
    e = torch.Tensor([]).to(U.DEV)
    args = (rs.bg, d["means3D"], e, d["opacities"], d["scales"], d["rotations"], rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, d["shs"], rs.sh_degree, rs.campos,
            rs.prefiltered, rs.antialiasing, rs.debug)
    outs = [C_.rasterize_gaussians(*args) for _ in range(2)]     # the second call runs on the stored binning hint
    num_rendered, color, radii, geom, binning, img, invdepth = outs[1]
    assert isinstance(num_rendered, int) and geom.dtype == torch.uint8 and binning.dtype == torch.uint8
    assert torch.equal(color, outs[0][1]) and num_rendered == outs[0][0]
    want = U.run_cuda_forward(cam, arrays)
    assert torch.equal(color, want["color"]) and torch.equal(radii, want["radii"]) and num_rendered == want["R"]
    g = torch.Generator().manual_seed(0)
    dC = torch.randn(3, 96, 160, generator=g).to(U.DEV)
    dD = torch.randn(1, 96, 160, generator=g).to(U.DEV)
    grads = C_.rasterize_gaussians_backward(rs.bg, d["means3D"], radii, e, d["opacities"], d["scales"], d["rotations"],
                                            rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dC,
                                            dD, d["shs"], rs.sh_degree, rs.campos, geom, num_rendered, binning, img,
                                            rs.antialiasing, rs.debug)
    assert len(grads) == 8
    ref = U.run_cuda_backward(want, dC.cpu().numpy(), dD.cpu().numpy())
    for t, k in zip(grads, ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                            "dL_drotations")):
        s = max(1.0, float(np.abs(ref[k]).max()))
        assert np.abs(t.cpu().numpy().reshape(ref[k].shape) - ref[k]).max() / s <= 1e-5, k
    vis = C_.mark_visible(d["means3D"], rs.viewmatrix, rs.projmatrix)
    assert vis.dtype == torch.bool and vis.shape == (6000,)


@pytest.mark.parametrize("P,deg,V", [(9001, 3, 3), (5000, 0, 3), (3000, 2, 20)])
def test_fused_multi_view_preprocess_backward_equals_the_per_view_sum(P, deg, V):
    """gsr_preprocess_backward_views (all views of a step in one pass over the Gaussians) against the per-view backward:
    gradients equal up to summation order, the densification statistics (per-view norms, counts, max radii) too.
    deg 0 takes the library's per-view fallback (SH rows of 12 bytes), V = 20 two groups of views."""
    from gaussian_splatting_lightning_b200.rasterizer import (backward_raw, backward_render_only, forward_raw,
                                                              preprocess_backward_views)
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers, ViewStreams, render_step
    W, H = 176, 112
    scene = {k: v.to(U.DEV) for k, v in U.synthetic.make_scene(P, deg, W, H, seed=P).items()}
    cams = [U.settings_from_camera(U.synthetic_case(8, deg, W, H, view_k=k % 8, num_views=8)[0]) for k in range(V)]
    g = torch.Generator().manual_seed(9)
    dC = [torch.randn(3, H, W, generator=g).to(U.DEV) for _ in range(V)]
    dD = [torch.randn(1, H, W, generator=g).to(U.DEV) for _ in range(V)]
    M = (deg + 1) ** 2
    a, b = StepBuffers(P, M, U.DEV), StepBuffers(P, M, U.DEV)
    render_step(cams, scene, lambda k, c, d: dC[k], lambda k, c, d: dD[k], a, exchange=False, fused=False)
    render_step(cams, scene, lambda k, c, d: dC[k], lambda k, c, d: dD[k], b, exchange=False, fused=True,
                streams=ViewStreams(U.DEV, 3))
    torch.cuda.synchronize()
    assert float(a.grads()["dL_dmeans3D"].abs().max()) > 0
    for n in a.grads():
        w, got = a.grads()[n], b.grads()[n]
        s = max(1.0, float(w.abs().max()))
        assert float((got - w).abs().max()) / s <= 2e-5, n
    assert torch.equal(a.max_radii, b.max_radii) and torch.equal(a.stats[1], b.stats[1])
    assert float((a.stats[0] - b.stats[0]).abs().max()) <= 2e-5 * max(1.0, float(a.stats[0].max()))
    # the storing (non-accumulating) form with a selection of outputs
    states, want = [], None
    for k in range(min(V, 3)):
        R, color, radii, geom, binning, img, invd, cap = forward_raw(cams[k], scene["means3D"], scene["shs"], None,
                                                                    scene["opacities"], scene["scales"], scene["rotations"], None)
        args = (cams[k], scene["means3D"], radii, scene["shs"], None, scene["opacities"], scene["scales"], scene["rotations"],
                None, dC[k], dD[k], geom, binning, img, R)
        states.append((cams[k], radii, geom, backward_render_only(*args)))
        one = backward_raw(*args)
        want = one if want is None else tuple(x + y for x, y in zip(want, one))
    got = preprocess_backward_views(states, scene["means3D"], scene["shs"], None, scene["opacities"], scene["scales"],
                                    scene["rotations"], None, want=("dL_dmeans3D", "dL_dsh", "dL_drotations", "dL_dcov3D"))
    assert [t is not None for t in got] == [False, False, False, True, True, True, False, True]
    for i in (3, 4, 5, 7):
        s = max(1.0, float(want[i].abs().max()))
        assert float((got[i] - want[i]).abs().max()) / s <= 2e-5, i


def test_step_sink_defers_the_autograd_backward_to_one_fused_pass():
    """rasterizer.StepSink: inside it the autograd backward of a view runs the compositing backward only and parks the
    view; flush() runs the preprocess backward of all parked views in one pass.  Same gradient sum and statistics as the
    immediate accumulation (accumulate_gradients_into); a view that asks for dL_dmeans2D is not parked."""
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer, StepSink, accumulate_gradients_into
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers
    P, W, H, deg = 7000, 160, 96, 3
    scene = {k: v.to(U.DEV) for k, v in U.synthetic.make_scene(P, deg, W, H, seed=31).items()}
    cams = [U.settings_from_camera(U.synthetic_case(8, deg, W, H, view_k=k, num_views=8)[0]) for k in (0, 2, 5)]
    g = torch.Generator().manual_seed(4)
    dC = [torch.randn(3, H, W, generator=g).to(U.DEV) for _ in cams]

    def params():
        return {k: scene[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}

    def render(p, rs, sp=None):
        return GaussianRasterizer(rs)(means3D=p["means3D"], means2D=sp, shs=p["shs"], opacities=p["opacities"],
                                      scales=p["scales"], rotations=p["rotations"])[0]

    a, pa = StepBuffers(P, 16, U.DEV), params()
    for rs, w in zip(cams, dC):
        with accumulate_gradients_into(a.grads(), a.densify_stats()):
            img = render(pa, rs)
        (img * w).sum().backward()
    b, pb = StepBuffers(P, 16, U.DEV), params()
    sink = StepSink(b.grads(), b.densify_stats())
    for rs, w in zip(cams, dC):
        with sink:
            img = render(pb, rs)
        (img * w).sum().backward()
    assert len(sink.pending) == 3 and float(b.grads()["dL_dmeans3D"].abs().max()) == 0.0     # nothing written yet
    sink.flush()
    assert not sink.pending
    for n in a.grads():
        s = max(1.0, float(a.grads()[n].abs().max()))
        assert float((a.grads()[n] - b.grads()[n]).abs().max()) / s <= 2e-5, n
    assert torch.equal(a.max_radii, b.max_radii) and torch.equal(a.stats[1], b.stats[1])
    assert float((a.stats[0] - b.stats[0]).abs().max()) <= 2e-5 * max(1.0, float(a.stats[0].max()))
    # a view whose dummy screen-space leaf wants its gradient takes the immediate path
    sp = torch.zeros_like(pb["means3D"], requires_grad=True)
    with sink:
        img = render(pb, cams[0], sp)
    (img * dC[0]).sum().backward()
    assert not sink.pending and sp.grad is not None and float(sp.grad.abs().max()) > 0


def test_chunked_fused_step_equals_the_unchunked_one():
    """fused_backward_and_exchange over a chunk-major StepBuffers (the layout the pipelined exchange uses): the same
    gradient rows and statistics as the whole-P pass, range by range (gsr_preprocess_backward_views_range)."""
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers, render_step
    P, deg, W, H, V = 9001, 3, 176, 112, 5
    scene = {k: v.to(U.DEV) for k, v in U.synthetic.make_scene(P, deg, W, H, seed=12).items()}
    cams = [U.settings_from_camera(U.synthetic_case(8, deg, W, H, view_k=k, num_views=8)[0]) for k in range(V)]
    g = torch.Generator().manual_seed(2)
    dC = [torch.randn(3, H, W, generator=g).to(U.DEV) for _ in range(V)]
    a, b = StepBuffers(P, 16, U.DEV), StepBuffers(P, 16, U.DEV, chunks=4)
    assert b.num_chunks == 3 and b.chunks[-1][0] + b.chunks[-1][1] == P     # boundaries on multiples of 1024 Gaussians
    b._store.fill_(123.0)        # the pass overwrites every gradient row: stale contents must not survive
    render_step(cams, scene, lambda k, c, d: dC[k], None, a, exchange=False)
    render_step(cams, scene, lambda k, c, d: dC[k], None, b, exchange=False)
    torch.cuda.synchronize()
    for n in a.grads():
        got = torch.cat([b.grads(c)[n] for c in range(b.num_chunks)])
        s = max(1.0, float(a.grads()[n].abs().max()))
        assert float((got - a.grads()[n]).abs().max()) / s <= 2e-5, n
    assert torch.equal(torch.cat([b.densify_stats(c)["xyz_grad_count"] for c in range(b.num_chunks)]), a.stats[1])
    assert torch.equal(b.max_radii, a.max_radii)
