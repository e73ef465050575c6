"""Tile-segmented binning (csrc/tile_binning.cu) against the oracle's global stable sort, on scenes built to reach
every path of the per-tile sort: segments of a few instances, of 1 000 - 8 000 (the 64 KB shared-memory launch),
above 8 192 (the in-place network through L2), and depths with many exact ties (the Gaussian index must break
them the way the reference's stable radix sort does).  Bit-exact: sorted keys, sorted point list, ranges."""
import numpy as np
import pytest
import torch

import gpu_utils as U

pytestmark = pytest.mark.gpu

CASES = {
    # name: (P, W, H, number of distinct depths or 0)
    "sparse_tiles": (300, 256, 256, 0),
    "segments_of_thousands": (30_000, 64, 48, 0),
    "segments_above_8192": (60_000, 32, 32, 0),
    "ties_everywhere": (20_000, 64, 64, 5),
    "ties_and_huge_segments": (50_000, 32, 16, 3),
}


def _case(name):
    P, W, H, levels = CASES[name]
    cam, arrays = U.synthetic_case(P, 1, W, H, seed=7)
    if levels:
        m = arrays["means3D"].copy()
        z = m[:, 2]
        q = np.round((np.log(z) - np.log(2.0)) / (np.log(20.0) - np.log(2.0)) * (levels - 1))
        znew = (2.0 * (10.0 ** (q / max(levels - 1, 1)))).astype(np.float32)
        m[:, 0] *= znew / z
        m[:, 1] *= znew / z
        m[:, 2] = znew
        arrays["means3D"] = np.ascontiguousarray(m.astype(np.float32))
    return cam, arrays


@pytest.mark.parametrize("name", list(CASES))
def test_segmented_binning_equals_global_stable_sort(oracle_lib, name):
    cam, arrays = _case(name)
    ref = U.run_oracle_forward(cam, arrays)
    out = U.run_cuda_forward(cam, arrays)
    st = U.export(out)
    assert out["R"] == ref["num_rendered"]
    counts = ref["ranges"][:, 1].astype(np.int64) - ref["ranges"][:, 0]
    print(f"{name}: R = {out['R']}, largest segment {counts.max()}, distinct depths "
          f"{np.unique(ref['depths'][ref['radii'] > 0]).size}")
    assert np.array_equal(st["ranges"].view(np.uint32), ref["ranges"])
    assert np.array_equal(st["vals_sorted"].view(np.uint32), ref["point_list"])
    assert np.array_equal(st["keys_sorted"].view(np.uint64), ref["keys"])
    frac, mx = U.close_fraction(out["color"].cpu().numpy(), ref["color"], 1e-4, 1e-4)
    assert frac <= 2e-4 and mx <= 2e-2, (frac, mx)


@pytest.mark.parametrize("name", ["segments_of_thousands", "ties_and_huge_segments"])
def test_too_small_hint_reruns_to_the_same_lists(name):
    cam, arrays = _case(name)
    base = U.export(U.run_cuda_forward(cam, arrays, hint=0))
    for hint in (1, 1000):
        o = U.export(U.run_cuda_forward(cam, arrays, hint=hint))
        for k in ("ranges", "vals_sorted", "keys_sorted"):
            assert np.array_equal(o[k], base[k]), (hint, k)


def test_global_onesweep_binning_gives_the_same_lists():
    """gsr_set_binning(GSR_BINNING_ONESWEEP) selects the global radix-sort binning at run time: the same sorted lists,
    ranges, images and gradients as the tile-segmented binning, on this file's crowded / tied scenes and the parity
    scenes, in ONE process (round 1 needed a subprocess: the mode was a process-static)."""
    from gaussian_splatting_lightning_b200 import _native as N
    import test_gpu_parity as TP
    assert N.get_binning() == "tile"
    names = list(CASES) if "CASES" in globals() else []
    cases = [(_case(n)) for n in names] + [TP._case(n) for n in ("g2_deg3_100x70", "deg3_20k_320x200", "deg1_ragged_203x117_view3of8")]
    rng = np.random.default_rng(0)
    for cam, arrays in cases:
        H, W = cam.image_height, cam.image_width
        dC = rng.standard_normal((3, H, W)).astype(np.float32)
        a = U.run_cuda_forward(cam, arrays)
        sa, ga = U.export(a), U.run_cuda_backward(a, dC, None)
        try:
            N.set_binning("onesweep")
            assert N.get_binning() == "onesweep"
            b = U.run_cuda_forward(cam, arrays)
            sb, gb = U.export(b), U.run_cuda_backward(b, dC, None)
        finally:
            N.set_binning("tile")
        for k in ("keys_sorted", "vals_sorted", "ranges", "n_contrib", "keys_unsorted", "vals_unsorted"):
            assert np.array_equal(sa[k], sb[k]), k
        assert torch.equal(a["color"], b["color"]) and a["R"] == b["R"]
        for k in ga:
            s_ = max(1.0, float(np.abs(ga[k]).max()))
            assert np.abs(ga[k] - gb[k]).max() / s_ <= 1e-5, k


@pytest.mark.parametrize("n", [1, 15, 16, 17, 255, 511, 512, 513, 1023, 1024, 1025, 2047, 2048, 2049, 4097, 8191, 8192,
                               8193, 20000])
def test_one_segment_of_exactly_n_instances(n):
    """Every Gaussian lands in the same single tile, so that tile's segment holds exactly n instances: the size-class
    boundaries of the per-tile sort (512 / 2048 / 8192), the power-of-two paddings on either side of them and the
    skipped merges.  Expected order: ascending depth bits, ties by Gaussian index (numpy lexsort)."""
    from gaussian_splatting_lightning_b200 import synthetic
    import oracle
    W = H = 64
    c = synthetic.make_camera(W, H)
    cam = oracle.Camera(H, W, c["tanfovx"], c["tanfovy"], c["viewmatrix"].numpy(), c["projmatrix"].numpy(),
                        c["campos"].numpy(), bg=np.zeros(3, np.float32), sh_degree=0)
    rng = np.random.default_rng(n)
    focal = W / (2 * c["tanfovx"])
    levels = max(2, n // 3)                                  # about three Gaussians per distinct depth
    z = (2.0 + 18.0 * rng.integers(0, levels, n) / levels).astype(np.float32)
    u, v = 24.0 + rng.uniform(-1, 1, n), 24.0 + rng.uniform(-1, 1, n)      # well inside tile (1, 1)
    means = np.stack([(u - W / 2 + 0.5) / focal * z, (v - H / 2 + 0.5) / focal * z, z], -1).astype(np.float32)
    arrays = dict(means3D=means, scales=np.full((n, 3), 1e-4, np.float32),
                  rotations=np.tile(np.array([1, 0, 0, 0], np.float32), (n, 1)),
                  opacities=np.full((n, 1), 0.01, np.float32), shs=rng.normal(size=(n, 1, 3)).astype(np.float32))
    out = U.run_cuda_forward(cam, arrays)
    st = U.export(out)
    assert out["R"] == n
    ranges = st["ranges"].view(np.uint32).reshape(-1, 2)
    assert ranges[5].tolist() == [0, n] and np.count_nonzero(ranges) == 1
    depth_bits = st["depths"].view(np.uint32)
    want = np.lexsort((np.arange(n), depth_bits)).astype(np.uint32)
    assert np.array_equal(st["vals_sorted"].view(np.uint32), want)
    assert np.array_equal(st["keys_sorted"].view(np.uint64), (np.uint64(5) << np.uint64(32)) | depth_bits[want].astype(np.uint64))


def test_priority_lanes_give_the_same_frames_on_several_streams():
    """gsr_set_priority_lanes(1): the stages before compositing run on a high-priority stream per caller stream (fork /
    join by events).  Same lists, images and gradients as on the caller's stream alone -- also with several views in
    flight on several streams, with a too-small binning hint (the retry forks again) and on reused streams."""
    from gaussian_splatting_lightning_b200 import _native as N
    import test_gpu_parity as TP
    cases = [TP._case(n) for n in ("g2_deg3_100x70", "deg3_20k_320x200", "deg1_ragged_203x117_view3of8")]
    rng = np.random.default_rng(1)
    before = N.get_priority_lanes()
    try:
        results = {}
        for on in (False, True):
            N.set_priority_lanes(on)
            assert N.get_priority_lanes() == on
            streams = [torch.cuda.Stream(device=U.DEV) for _ in range(3)]
            outs = []
            prepared = [(U.settings_from_camera(cam), U.to_dev(arrays)) for cam, arrays in cases]
            torch.cuda.synchronize()
            for rep in range(2):                 # second round: the lanes' streams and events are reused
                for k, (rs, d) in enumerate(prepared):
                    with torch.cuda.stream(streams[k % 3]):      # no synchronisation between the views
                        R, color, radii, geom, binning, img, invdepth, cap = U.forward_raw(
                            rs, d["means3D"], d.get("shs"), d.get("colors_precomp"), d["opacities"], d.get("scales"),
                            d.get("rotations"), d.get("cov3D_precomp"), binning_hint=(0 if rep == 0 else 7))
                    outs.append(dict(rs=rs, d=d, R=R, color=color, radii=radii, geom=geom, binning=binning, img=img,
                                     invdepth=invdepth, cap=cap))
            torch.cuda.synchronize()
            results[on] = outs
        for a, b, (cam, arrays) in zip(results[False], results[True], cases + cases):
            sa, sb = U.export(a), U.export(b)
            for k in ("keys_sorted", "vals_sorted", "ranges", "n_contrib"):
                assert np.array_equal(sa[k], sb[k]), k
            assert torch.equal(a["color"], b["color"]) and torch.equal(a["invdepth"], b["invdepth"]) and a["R"] == b["R"]
            dC = rng.standard_normal((3, cam.image_height, cam.image_width)).astype(np.float32)
            ga, gb = U.run_cuda_backward(a, dC, None), U.run_cuda_backward(b, dC, None)
            for k in ga:
                s_ = max(1.0, float(np.abs(ga[k]).max()))
                assert np.abs(ga[k] - gb[k]).max() / s_ <= 1e-5, k
    finally:
        N.set_priority_lanes(before)
