#!/usr/bin/env python
"""Where does the HOST time of an end-to-end step go?  cProfile of bench.E2E.step on a small workload (C2: the GPU
finishes a view in 0.18 ms, so the step is bound by the Python / launch path).
    python tools/profile_host.py [--workload C2] [--views 8] [--streams 3] [--steps 60]
"""
import argparse
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--streams", type=int, default=3)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--top", type=int, default=45)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    rig = bench.Rig(a.workload, a.views, a.streams, dev, 0, 1, "block")
    e = bench.E2E(rig)
    for _ in range(10):
        e.step()
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(a.steps):
        e.step()
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(a.top)
    print("per step: %.3f ms host" % (1e3 * st.total_tt / a.steps))


if __name__ == "__main__":
    main()
