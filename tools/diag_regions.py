#!/usr/bin/env python
"""Why are the timed regions of bench.py bimodal (12.1 / 15.7 / 12.1 / 14.8 ms per step)?

Times REGIONS regions of STEPS C3 steps (the bench's step: 8 views on `--streams` streams, fwd+bwd into the step
buffers) and records per region: device ms per step, host enqueue ms per step, binning-hint retries, caching-allocator
cudaMalloc calls, SM clock before/after (nvidia-smi one-shot, outside the region).  Variants on the command line.
    python tools/diag_regions.py [--streams 3] [--regions 10] [--steps 20] [--poller] [--sync-r]
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from gaussian_splatting_lightning_b200 import rasterizer as RZ, synthetic  # noqa: E402
from gaussian_splatting_lightning_b200.view_parallel import StepBuffers, ViewStreams  # noqa: E402


def smi():
    try:
        o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,temperature.gpu", "--format=csv,noheader,nounits",
                            "-i", "0"], capture_output=True, text=True, timeout=20).stdout.strip()
        return o
    except Exception as e:
        return str(e)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=3)
    ap.add_argument("--regions", type=int, default=10)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--poller", action="store_true")
    ap.add_argument("--sync-r", action="store_true", help="binning_hint=0: blocking read of num_rendered")
    ap.add_argument("--workload", default="C3")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    P, deg, W, H = synthetic.config(a.workload)
    M = (deg + 1) ** 2
    scene = {k: v.to(dev) for k, v in synthetic.make_scene(P, deg, W, H, seed=0).items()}
    cams = bench.make_cameras(W, H, 0, a.views, dev)
    g = torch.Generator(device="cpu").manual_seed(1234)
    dC = torch.randn(3, H, W, generator=g).to(dev)
    dD = torch.randn(1, H, W, generator=g).to(dev)
    bufs = StepBuffers(P, M, dev)
    acc, st = bufs.grads(), bufs.densify_stats()
    vs = ViewStreams(dev, a.streams)

    def one_view(c):
        rs = c["settings"]
        R, color, radii, geom, binning, img, invd, cap = RZ.forward_raw(
            rs, scene["means3D"], scene["shs"], None, scene["opacities"], scene["scales"], scene["rotations"], None,
            binning_hint=0 if a.sync_r else None)
        RZ.backward_raw(rs, scene["means3D"], radii, scene["shs"], None, scene["opacities"], scene["scales"],
                        scene["rotations"], None, dC, dD, geom, binning, img, R, accumulate_into=acc, stats=st)

    def step():
        bufs.zero_()
        vs.run(len(cams), lambda j: one_view(cams[j]))

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    poller = None
    if a.poller:
        poller = bench.ClockSampler(0)
        poller.start()
    rows = []
    for r in range(a.regions):
        ms0 = torch.cuda.memory_stats(dev)
        c0 = dict(RZ.counters)
        clk0 = smi() if not a.poller else ""
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        ms1 = torch.cuda.memory_stats(dev)
        rows.append(dict(region=r, dev_ms_per_step=round(e0.elapsed_time(e1) / a.steps, 3),
                         host_enqueue_ms_per_step=round((t1 - t0) * 1e3 / a.steps, 3),
                         wall_ms_per_step=round((t2 - t0) * 1e3 / a.steps, 3),
                         hint_retries=RZ.counters["hint_retries"] - c0["hint_retries"],
                         cuda_mallocs=ms1.get("num_device_alloc", 0) - ms0.get("num_device_alloc", 0),
                         alloc_retries=ms1.get("num_alloc_retries", 0) - ms0.get("num_alloc_retries", 0),
                         reserved_mb=ms1.get("reserved_bytes.all.current", 0) >> 20, smi_before=clk0))
        print(json.dumps(rows[-1]), flush=True)
    if poller:
        print(json.dumps(poller.stop(0, time.time() + 1)))
    ms = sorted(x["dev_ms_per_step"] for x in rows)
    print(json.dumps(dict(args=vars(a), min=ms[0], median=ms[len(ms) // 2], max=ms[-1])))


if __name__ == "__main__":
    main()
