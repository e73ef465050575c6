"""GPU parity tests: the CUDA path (through the C-ABI of include/gsr_b200.h) against the CPU oracle
and the committed golden fixtures of the reference's Python rasterizer.

Bars (BASELINE.json north_star):
  * integer / index work -- radii, tiles_touched, offsets, unsorted and sorted (key, value) pairs,
    tile ranges -- BIT-EXACT against the oracle;
  * images, inverse depth, final_T and gradients within 1e-4 abs + 1e-4 rel.  A (pixel, splat) pair
    whose alpha sits within a few ulp of the 1/255 floor or whose T sits at the 1e-4 stop can fall on
    the other side of the test under ex2.approx vs. glibc expf; such a flip moves a pixel by up to
    ~4e-3.  The tests therefore allow a fraction OUTLIER_FRAC of elements outside the tolerance and
    bound them by OUTLIER_MAX.  The observed fractions are printed (-s) and kept in
    profiles/r2_parity_observed.txt.
"""
import numpy as np
import pytest
import torch

from conftest import GOLDEN_SCENES, load_golden
import gpu_utils as U

pytestmark = pytest.mark.gpu

ATOL = RTOL = 1e-4
OUTLIER_FRAC = 1e-4   # observed on the B200: at most 1.6e-5 (profiles/r2_parity_observed.txt)
OUTLIER_MAX = 1e-2    # observed: at most 1.7e-3

SYNTH_CASES = {
    # name: (P, deg, W, H, kwargs)
    "c1_10k_deg0_128": (10_000, 0, 128, 128, {}),
    "deg3_20k_320x200": (20_000, 3, 320, 200, dict(bg=(0.3, 0.1, 0.7))),
    "deg1_ragged_203x117_view3of8": (15_000, 1, 203, 117, dict(view_k=3, num_views=8, bg=(1.0, 1.0, 1.0))),
    "deg2_aa_160x160": (8_000, 2, 160, 160, dict(antialiasing=True, scale_modifier=1.7)),
    "deg3_behind_camera_view4of8": (12_000, 3, 256, 144, dict(view_k=4, num_views=8)),
}


def _case(name):
    if name in GOLDEN_SCENES:
        d, cam = load_golden(name)
        arrays = dict(means3D=d["means3D"], scales=d["scales"], rotations=d["rotations"], opacities=d["opacities"],
                      shs=d["shs"])
        return cam, arrays
    P, deg, W, H, kw = SYNTH_CASES[name]
    return U.synthetic_case(P, deg, W, H, **kw)


ALL_CASES = GOLDEN_SCENES + list(SYNTH_CASES)


@pytest.fixture(scope="module")
def results(oracle_lib):
    cache = {}

    def get(name):
        if name not in cache:
            cam, arrays = _case(name)
            ref = U.run_oracle_forward(cam, arrays)
            out = U.run_cuda_forward(cam, arrays)
            cache[name] = (cam, arrays, ref, out, U.export(out))
        return cache[name]
    return get


@pytest.mark.parametrize("name", ALL_CASES)
def test_preprocess_bit_exact_integers(results, name):
    cam, arrays, ref, out, st = results(name)
    radii = out["radii"].cpu().numpy()
    assert np.array_equal(radii, ref["radii"])
    assert np.array_equal(st["tiles_touched"].view(np.uint32), ref["tiles_touched"])
    assert np.array_equal(st["point_offsets"].view(np.uint32), ref["point_offsets"])
    assert out["R"] == ref["num_rendered"]
    # depth and pixel position feed the keys and rects: bit-exact
    assert np.array_equal(st["depths"].view(np.uint32), ref["depths"].view(np.uint32))
    assert np.array_equal(st["means2D"].view(np.uint32), ref["means2D"].view(np.uint32))
    vis = radii > 0
    assert np.array_equal(st["clamped"][vis], ref["clamped"][vis])
    for k in ("cov3D", "conic_opacity", "rgb"):
        np.testing.assert_allclose(st[k], ref[k], rtol=1e-6, atol=1e-7, err_msg=k)


@pytest.mark.parametrize("name", ALL_CASES)
def test_keys_sort_ranges_bit_exact(results, name):
    cam, arrays, ref, out, st = results(name)
    assert np.array_equal(st["keys_unsorted"].view(np.uint64), ref["keys_unsorted"])
    assert np.array_equal(st["vals_unsorted"].view(np.uint32), ref["vals_unsorted"])
    assert np.array_equal(st["keys_sorted"].view(np.uint64), ref["keys"])
    assert np.array_equal(st["vals_sorted"].view(np.uint32), ref["point_list"])
    assert np.array_equal(st["ranges"].view(np.uint32), ref["ranges"])


@pytest.mark.parametrize("name", ALL_CASES)
def test_forward_image(results, name):
    cam, arrays, ref, out, st = results(name)
    for got, want, what in ((out["color"].cpu().numpy(), ref["color"], "color"),
                            (out["invdepth"].cpu().numpy(), ref["invdepth"], "invdepth"),
                            (st["final_T"], ref["final_T"], "final_T")):
        frac, mx = U.close_fraction(got, want, ATOL, RTOL)
        print(f"{name} {what}: outlier fraction {frac:.2e}, max abs err {mx:.2e}")
        assert frac <= OUTLIER_FRAC, (what, frac)
        assert mx <= OUTLIER_MAX, (what, mx)
    mism = float((st["n_contrib"].view(np.uint32) != ref["n_contrib"]).mean())
    print(f"{name} n_contrib mismatch fraction {mism:.2e}")
    assert mism <= OUTLIER_FRAC


@pytest.mark.parametrize("name", GOLDEN_SCENES)
def test_forward_matches_reference_python_golden(results, name):
    """Directly against the stored output of gs_lightning.rasterize.rasterize_gaussian (tests/golden)."""
    d, cam = load_golden(name)
    _, _, ref, out, st = results(name)
    unsat = st["final_T"] >= 1e-3
    img = out["color"].cpu().numpy()
    err = np.abs(img - d["ref_image"])
    tol = 1e-4 + 1e-4 * np.abs(d["ref_image"])
    if unsat.any():
        assert ((err > tol)[:, unsat]).mean() <= OUTLIER_FRAC
    assert err.max() <= 2e-2  # saturated pixels: rasterize.py:253-256 keeps one more splat (Appendix B3)
    radii = out["radii"].cpu().numpy()
    vis = radii > 0
    assert np.array_equal(radii[vis], d["ref_radius"][vis].astype(np.int32))


@pytest.mark.parametrize("name", ["g2_deg3_100x70", "g4_saturated_64x48", "deg3_20k_320x200",
                                  "deg1_ragged_203x117_view3of8", "deg2_aa_160x160", "c1_10k_deg0_128"])
def test_backward_matches_oracle(results, oracle_lib, name):
    cam, arrays, ref, out, st = results(name)
    rng = np.random.default_rng(7)
    H, W = cam.image_height, cam.image_width
    dC = rng.standard_normal((3, H, W)).astype(np.float32)
    dD = rng.standard_normal((1, H, W)).astype(np.float32)
    want = oracle_lib.backward(ref, dC, dD[0])
    got = U.run_cuda_backward(out, dC, dD)
    U.assert_grads_close(got, want, label=name)


def test_backward_without_invdepth_grad_and_precomp(oracle_lib):
    """colors_precomp + cov3D_precomp inputs (the other branch of both argument pairs)."""
    cam, arrays = U.synthetic_case(5000, 0, 96, 64, seed=3)
    ref0 = U.run_oracle_forward(cam, arrays)
    rng = np.random.default_rng(1)
    arrays2 = dict(means3D=arrays["means3D"], opacities=arrays["opacities"],
                   colors_precomp=rng.random((5000, 3)).astype(np.float32), cov3D_precomp=ref0["cov3D"].copy())
    # cov3D of culled Gaussians is zero in the oracle dump: recompute for all from a second pass
    cam_all = cam
    ref = U.run_oracle_forward(cam_all, arrays2)
    out = U.run_cuda_forward(cam_all, arrays2)
    assert np.array_equal(out["radii"].cpu().numpy(), ref["radii"])
    frac, mx = U.close_fraction(out["color"].cpu().numpy(), ref["color"], ATOL, RTOL)
    assert frac <= OUTLIER_FRAC and mx <= OUTLIER_MAX
    dC = rng.standard_normal((3, 64, 96)).astype(np.float32)
    want = oracle_lib.backward(ref, dC, None)
    got = U.run_cuda_backward(out, dC, None)
    U.assert_grads_close(got, want, names=("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D"),
                         label="precomp")


def test_mark_visible_matches_reference_golden():
    import os
    from conftest import GOLDEN_DIR
    from gaussian_splatting_lightning_b200.rasterizer import mark_visible
    d = np.load(os.path.join(GOLDEN_DIR, "mark_visible_treehill.npz"))
    pts = torch.tensor(d["points_3d"], device=U.DEV)
    for v, ref in zip(d["viewmatrices"], d["ref_visible"]):
        got = mark_visible(pts, torch.tensor(v, device=U.DEV)).cpu().numpy()
        assert not np.logical_xor(got, ref).any()


def test_binning_hint_paths_give_identical_results():
    """Enough, exact and too-small binning hints (the re-run path) must all equal the synchronous path."""
    cam, arrays = U.synthetic_case(20_000, 1, 320, 200, seed=5)
    base = U.run_cuda_forward(cam, arrays, hint=0)
    R = base["R"]
    for hint in (R, R + 12345, max(1, R // 3), 1):
        o = U.run_cuda_forward(cam, arrays, hint=hint)
        assert o["R"] == R
        assert torch.equal(o["color"], base["color"])
        assert torch.equal(o["radii"], base["radii"])
        a, b = U.export(o), U.export(base)
        for k in ("keys_sorted", "vals_sorted", "ranges", "n_contrib"):
            assert np.array_equal(a[k], b[k]), (hint, k)


def test_empty_and_all_culled_inputs():
    cam, arrays = U.synthetic_case(64, 0, 50, 30, seed=2, bg=(0.2, 0.4, 0.6))
    # everything behind the camera
    arrays["means3D"] = arrays["means3D"].copy()
    arrays["means3D"][:, 2] = -5.0
    o = U.run_cuda_forward(cam, arrays)
    assert o["R"] == 0 and int(o["radii"].abs().sum()) == 0
    img = o["color"].cpu().numpy()
    assert np.allclose(img, np.array([0.2, 0.4, 0.6], np.float32)[:, None, None])
    assert float(o["invdepth"].abs().max()) == 0.0
    g = U.run_cuda_backward(o, np.ones((3, 30, 50), np.float32), None)
    assert all(float(np.abs(v).max()) == 0.0 for v in g.values() if v.size)
    # P == 0
    empty = {k: v[:0] for k, v in arrays.items()}
    o = U.run_cuda_forward(cam, empty)
    assert o["R"] == 0 and np.allclose(o["color"].cpu().numpy()[:, 0, 0], [0.2, 0.4, 0.6])


def test_debug_flag_runs_stage_by_stage():
    cam, arrays = U.synthetic_case(3000, 3, 64, 64, seed=9)
    a = U.run_cuda_forward(cam, arrays)
    b = U.run_cuda_forward(cam, arrays, debug=True)
    assert torch.equal(a["color"], b["color"])
