"""Reference-derived GRADIENT goldens: float64 central differences through the reference's own
Python rasterizer, unchanged (gs_lightning/rasterize/rasterize.py:28-127, render_tools.py:13-144,
utils/sh.py:41-98, imported by ref_loader.py).

    python tests/golden/make_grad_golden.py [g1_deg0_128 ...]      # build container only (/root/reference)

The reference has no backward (its Python path is forward-only and autograd trips on the in-place
dilation at render_tools.py:79-80), so its derivative is taken numerically.  The reference code runs in
float64 end to end when torch's default dtype is float64 (its canvases are `torch.zeros(...)`,
rasterize.py:104-105) -- nothing of it is modified.

For each forward golden <name>.npz this writes <name>_grads.npz:

  loss      L = sum(w_img * image) + sum(w_dep * depth) with the closed-form weights of loss_weights()
            (no stored random tensors), ZEROED on pixels whose final transmittance is below 1e-3:
            there the Python path keeps one more Gaussian than the CUDA semantics (rasterize.py:253-256,
            SURVEY.md Appendix B3).  The transmittance comes from the reference itself:
            T = image(bg = 1) - image(bg = 0).  The mask is stored.
  scalars   `tensor` (id into TENSORS), `index` (flat index into that tensor), `fd` = central difference
            dL/dparam with step h, `fd2` the same with step 2h.  A parameter is kept only if the two agree
            to 2e-6 of max(|fd|, the tensor's median |fd|): a finite step that crosses one of the rasterizer's
            discontinuities (ceil of the radius, the 1/255 alpha floor, a depth-order swap) does not.
  dirs      directional derivatives dL/dt of L(theta + t * d_k) for the closed-form dense directions of
            direction(): every parameter of the scene takes part.  Their loss uses `dir_mask`, which also leaves out
            the pixels holding a (pixel, splat) pair within 2e-3 of the 1/255 alpha floor (see main()).

Conventions the comparison has to respect (tests/test_oracle_grad_golden.py, tests/test_gpu_grad_golden.py):
  * the reference normalises the quaternion inside computeConv3D (kornia Quaternion.matrix(),
    render_tools.py:61); the CUDA semantics take it as given.  For unit quaternions (all goldens) the
    reference's derivative is the tangential part: (I - q q^T) dL/dq.
  * degree-3 SH are zero-padded to 25 coefficients (sh.py:83); only the first 16 are differentiated.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_loader  # noqa: E402

TENSORS = ("means3D", "scales", "rotations", "opacities", "shs")
SCALARS_PER_TENSOR = {"means3D": 48, "scales": 36, "rotations": 36, "opacities": 24, "shs": 64}
NUM_DIRS = 12
REL_STEP = 2e-5
DIR_STEP = 0.01   # dense directions move by DIR_STEP * (per-tensor scalar step) * d


def loss_weights(H: int, W: int):
    """Closed-form, platform-independent loss weights (float64): w_img (3,H,W), w_dep (1,H,W)."""
    y, x = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64), indexing="ij")
    w_img = np.stack([np.sin(0.37 * x + 0.61 * y + 1.3 * c) + 0.5 * np.cos(1.93 * x - 2.71 * y + 0.7 * c) for c in range(3)])
    w_dep = (np.cos(0.53 * x - 0.29 * y) + 0.5 * np.sin(2.39 * x + 1.61 * y))[None]
    return w_img, w_dep


def direction(k: int, name: str, shape):
    """Closed-form dense direction k for tensor `name` (float64), O(1) entries."""
    n = int(np.prod(shape))
    i = np.arange(n, dtype=np.float64)
    phase = {"means3D": 0.1, "scales": 0.9, "rotations": 1.7, "opacities": 2.3, "shs": 3.1}[name]
    return np.sin(i * (0.7548776662466927 + 0.1 * k) + phase + k).reshape(shape)


def tensor_step(name: str, arr: np.ndarray) -> float:
    """Finite-difference step of a tensor: REL_STEP of its typical magnitude."""
    return REL_STEP * max(float(np.abs(arr).mean()), 1e-3)


class Scene:
    def __init__(self, path: str):
        d = np.load(path)
        self.d = d
        self.deg = int(d["sh_degree"])
        self.H, self.W = int(d["image_height"]), int(d["image_width"])
        self.params = {k: d[k].astype(np.float64) for k in TENSORS}
        self.cam = dict(image_width=self.W, image_height=self.H, tanfovx=float(d["tanfovx"]), tanfovy=float(d["tanfovy"]),
                        viewmatrix=torch.tensor(d["viewmatrix"], dtype=torch.float64),
                        projmatrix=torch.tensor(d["projmatrix"], dtype=torch.float64),
                        campos=torch.tensor(d["campos"], dtype=torch.float64))
        self.bg = torch.tensor(d["bg"], dtype=torch.float64)

    def render(self, ref, params, bg=None):
        shs = torch.tensor(params["shs"])
        if self.deg == 3:
            shs = torch.cat([shs, torch.zeros(len(shs), 25 - shs.shape[1], 3)], 1)
        with torch.no_grad():
            img, _radius, depth = ref.rasterize_gaussian(
                means3D=torch.tensor(params["means3D"]), opacities=torch.tensor(params["opacities"]),
                scales=torch.tensor(params["scales"]), rotations=torch.tensor(params["rotations"]), shs=shs,
                scale_modifier=1.0, background=self.bg if bg is None else bg, sh_degree=self.deg, **self.cam)
        return img.numpy(), depth.numpy()


def main():
    dirs_only = "--dirs-only" in sys.argv   # keep the scalar samples of the existing fixture, redo the directions
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["g1_deg0_128", "g2_deg3_100x70", "g3_treehill0_half"]
    torch.set_default_dtype(torch.float64)   # the reference's canvases follow the default dtype
    torch.set_num_threads(max(1, (os.cpu_count() or 2)))
    ref = ref_loader.load()
    for name in names:
        t_start = time.time()
        sc = Scene(os.path.join(HERE, name + ".npz"))
        H, W = sc.H, sc.W
        one, zero = torch.ones(3, dtype=torch.float64), torch.zeros(3, dtype=torch.float64)
        T = sc.render(ref, sc.params, one)[0][0] - sc.render(ref, sc.params, zero)[0][0]
        mask = T >= 1e-3
        w_img, w_dep = loss_weights(H, W)
        w_img = w_img * mask[None]
        w_dep = w_dep * mask[None]
        evals = [0]

        def loss(params):
            evals[0] += 1
            img, dep = sc.render(ref, params)
            return float((w_img * img).sum() + (w_dep * dep).sum())

        def central(make, h):
            return (loss(make(+h)) - loss(make(-h))) / (2 * h)

        rng = np.random.default_rng(20260 + len(name))
        rec = dict(tensor=[], index=[], fd=[], fd2=[], step=[])
        rejected = 0
        if dirs_only:
            old = np.load(os.path.join(HERE, name + "_grads.npz"))
            rec = dict(tensor=list(old["tensor"]), index=list(old["index"]), fd=list(old["fd"]), fd2=list(old["fd2"]),
                       step=list(old["step"]))
            rejected = int(old["rejected"])
        for ti, tname in enumerate(() if dirs_only else TENSORS):
            base = sc.params[tname]
            h = tensor_step(tname, base)
            n_coef = (sc.deg + 1) ** 2
            cand = [int(f) for f in rng.permutation(base.size)
                    if not (tname == "shs" and (f // 3) % base.shape[1] >= n_coef)]   # unused SH bands: zero derivative
            cand = cand[:SCALARS_PER_TENSOR[tname] + SCALARS_PER_TENSOR[tname] // 4]
            f1s, f2s = [], []
            for flat in cand:
                def make(step, flat=flat):
                    p = dict(sc.params)
                    a = base.copy().reshape(-1)
                    a[flat] += step
                    p[tname] = a.reshape(base.shape)
                    return p
                f1s.append(central(make, h)); f2s.append(central(make, 2 * h))
            f1s, f2s = np.array(f1s), np.array(f2s)
            # h vs 2h must agree to 2e-6 of the value (or of the tensor's typical derivative for the small ones)
            floor = float(np.median(np.abs(f1s)))
            ok = np.abs(f1s - f2s) <= 2e-6 * np.maximum(np.abs(f1s), floor)
            rejected += int((~ok).sum())
            kept = 0
            for flat, f1, f2, good in zip(cand, f1s, f2s, ok):
                if good and kept < SCALARS_PER_TENSOR[tname]:
                    rec["tensor"].append(ti); rec["index"].append(flat); rec["fd"].append(float(f1))
                    rec["fd2"].append(float(f2)); rec["step"].append(h)
                    kept += 1
            print(f"  {name} {tname}: {kept} scalars, step {h:.2e}, {evals[0]} reference renders so far", flush=True)

        # dense directions: every parameter moves, so the step is 100x smaller than the scalar one (the chance that
        # SOME pixel/splat pair crosses the alpha floor grows with the number of moving Gaussians); float64 leaves room
        # ... and the loss of the directions leaves out every pixel that holds a (pixel, splat) pair whose alpha lies
        # within 2e-3 (relative) of the 1/255 floor: with ~1e5 pairs on the floor's contour, some pair sits within 1e-6
        # of it (g3: alpha * 255 = 0.999999 at pixel (113, 195)) and flips under ANY dense step.  Found through the
        # reference itself: scaling all opacities by (1 +- 2e-3) moves a smooth pixel by O(1e-6) in second difference
        # and a pixel with a flip by O(1e-3).
        dir_k, dirs_fd, dirs_fd2 = [], [], []
        hs = {t: tensor_step(t, sc.params[t]) for t in TENSORS}
        up = sc.render(ref, {**sc.params, "opacities": sc.params["opacities"] * (1 + 2e-3)})
        dn = sc.render(ref, {**sc.params, "opacities": sc.params["opacities"] * (1 - 2e-3)})
        mid = sc.render(ref, sc.params)
        second = np.abs(up[0] + dn[0] - 2 * mid[0]).max(0) + np.abs(up[1] + dn[1] - 2 * mid[1])[0]
        dir_mask = mask & (second < 2e-5)
        w_img_d, w_dep_d = loss_weights(H, W)
        w_img_d, w_dep_d = w_img_d * dir_mask[None], w_dep_d * dir_mask[None]

        def loss_d(params):
            evals[0] += 1
            img, dep = sc.render(ref, params)
            return float((w_img_d * img).sum() + (w_dep_d * dep).sum())

        def central_d(make, h):
            return (loss_d(make(+h)) - loss_d(make(-h))) / (2 * h)
        # Gaussians outside 1.3 x the field of view do not move: there the reference clamps x/z, y/z before building the
        # EWA Jacobian (render_tools.py:26-31) and its exact derivative sees x = lim * z, while the CUDA semantics the
        # oracle and the kernels follow gate dL/dx to zero and treat the clamped x as a constant in dJ/dz -- a known
        # difference of the two backward conventions for splats that sit outside the image (g2: 296 of 1500 Gaussians;
        # with them moving, the directional derivatives differed by 1e-4 .. 4e-4 relative, without by < 5e-6)
        import importlib
        ct = importlib.import_module("gs_lightning.rasterize.camera_tools")
        pv = ct.apply_extrinsic_matrix(torch.tensor(sc.params["means3D"]), sc.cam["viewmatrix"]).numpy()
        free = (np.abs(pv[:, 0] / pv[:, 2]) <= 1.3 * sc.cam["tanfovx"]) & (np.abs(pv[:, 1] / pv[:, 2]) <= 1.3 * sc.cam["tanfovy"])
        cand_fd = []
        for k in range(3 * NUM_DIRS):
            dvec = {t: direction(k, t, sc.params[t].shape) for t in TENSORS}
            if sc.deg < 3:   # no movement in unused SH coefficients
                dvec["shs"][:, (sc.deg + 1) ** 2:, :] = 0.0
            for t in TENSORS:
                dvec[t][~free] = 0.0

            def make(step):
                return {t: sc.params[t] + step * hs[t] * dvec[t] for t in TENSORS}
            # L(theta + s * (h_t d_t)): derivative in s = sum_t h_t <g_t, d_t>
            cand_fd.append((k, central_d(make, DIR_STEP), central_d(make, 2 * DIR_STEP)))
        floor = float(np.median([abs(f1) for _, f1, _ in cand_fd]))
        for k, f1, f2 in cand_fd:
            if abs(f1 - f2) <= 2e-6 * max(abs(f1), floor) and len(dir_k) < NUM_DIRS:
                dir_k.append(k); dirs_fd.append(f1); dirs_fd2.append(f2)
        print(f"  {name}: {len(dir_k)} of {len(cand_fd)} candidate directions pass the h/2h check; "
              f"{int((mask & ~dir_mask).sum())} pixels near the alpha floor left out of their loss, "
              f"{int((~free).sum())} Gaussians outside 1.3 x FOV held still", flush=True)
        out = os.path.join(HERE, name + "_grads.npz")
        np.savez_compressed(
            out, mask=mask, dir_mask=dir_mask, dir_gaussians=free, tensor=np.array(rec["tensor"], np.int32), index=np.array(rec["index"], np.int64),
            fd=np.array(rec["fd"], np.float64), fd2=np.array(rec["fd2"], np.float64), step=np.array(rec["step"], np.float64),
            dir_k=np.array(dir_k, np.int32), dirs_fd=np.array(dirs_fd, np.float64), dirs_fd2=np.array(dirs_fd2, np.float64),
            dir_steps=np.array([hs[t] for t in TENSORS], np.float64), dir_step=np.float64(DIR_STEP),
            rel_step=np.float64(REL_STEP), masked_pixels=np.int64((~mask).sum()), rejected=np.int64(rejected))
        print(f"{name}: {len(rec['fd'])} scalar derivatives ({rejected} rejected by the h/2h check), {len(dir_k)} directions, "
              f"{int((~mask).sum())} of {mask.size} pixels masked, {evals[0]} reference renders, "
              f"{time.time() - t_start:.0f} s -> {os.path.getsize(out) / 1e3:.0f} kB", flush=True)


if __name__ == "__main__":
    main()
