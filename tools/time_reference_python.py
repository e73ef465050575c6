#!/usr/bin/env python
"""north_star's reported CPU baseline: the reference's OWN gs_lightning Python rasterizer
(/root/reference/gs_lightning/rasterize/rasterize.py:28-127, unchanged; only the kornia stand-in of
tests/golden/ref_loader.py and a silenced progress bar) timed on BASELINE.json configs[0] = C1:
10k synthetic Gaussians, SH degree 0, 128x128, one view, forward only -- the only configuration it can hold
(its (tiles x P) mask and (Ng, 256, 256) intermediates cap it).

    python tools/time_reference_python.py            # build container only: /root/reference does not travel

1 warm-up + median of 5, torch threads = all host cores; writes profiles/reference_python_c1.json, which bench.py
copies into its line (`reference_python_rasterizer_c1`).  The CUDA-semantics oracle port is timed beside it on the
same frame for scale."""
import json
import os
import platform
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import ref_loader  # noqa: E402
from gaussian_splatting_lightning_b200 import synthetic  # noqa: E402


def cpu_model():
    try:
        for l in open("/proc/cpuinfo"):
            if l.startswith("model name"):
                return l.split(":", 1)[1].strip()
    except OSError:
        pass
    return platform.processor()


def main():
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    ref = ref_loader.load()
    P, deg, W, H = synthetic.config("C1")
    scene = synthetic.make_scene(P, deg, W, H, seed=0)
    cam = synthetic.make_camera(W, H)
    bg = torch.zeros(3)

    def frame():
        with torch.no_grad():
            return ref.rasterize_gaussian(
                means3D=scene["means3D"], opacities=scene["opacities"], scales=scene["scales"], rotations=scene["rotations"],
                shs=scene["shs"], scale_modifier=1.0, image_width=W, image_height=H, tanfovx=cam["tanfovx"],
                tanfovy=cam["tanfovy"], viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"], campos=cam["campos"],
                background=bg, sh_degree=deg)
    frame()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); img, radius, depth = frame(); ts.append(time.perf_counter() - t0)
    med = float(np.median(ts))
    import oracle
    oracle.build()
    oracle.set_num_threads(cores)
    ocam = oracle.Camera(H, W, cam["tanfovx"], cam["tanfovy"], cam["viewmatrix"].numpy(), cam["projmatrix"].numpy(),
                         cam["campos"].numpy(), sh_degree=deg)
    a = {k: v.numpy() for k, v in scene.items()}
    to = []
    for _ in range(6):
        t0 = time.perf_counter()
        st = oracle.forward(ocam, a["means3D"], a["scales"], a["rotations"], a["opacities"], shs=a["shs"])
        to.append(time.perf_counter() - t0)
    # the two CPU paths render the same frame (Appendix B differences aside): a sanity check of what was timed
    err = float(np.abs(img.numpy() - st["color"]).max())
    out = {"what": "gs_lightning.rasterize.rasterize_gaussian (reference, unchanged) on C1: 10k Gaussians, SH degree 0, "
                   "128x128, one view, forward only; 1 warm-up + median of 5",
           "frames_per_s": 1.0 / med, "seconds_per_frame": med, "all_seconds": [round(t, 4) for t in ts],
           "cores": cores, "cpu": cpu_model(), "torch_threads": torch.get_num_threads(), "where": "build container",
           "oracle_port_same_frame": {"frames_per_s": 1.0 / float(np.median(to[1:])), "seconds_per_frame": float(np.median(to[1:])),
                                      "threads": oracle.num_threads(), "max_abs_image_difference": err}}
    json.dump(out, open(os.path.join(ROOT, "profiles", "reference_python_c1.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
