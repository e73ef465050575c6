#!/usr/bin/env python
"""bench.py -- rendered frames/sec of the rasterizer hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3]

Workload (config.workload): BASELINE.json configs[2] = "C3": 1M synthetic Gaussians, SH degree 3,
1920x1080, forward + backward training step with densification statistics.  One STEP = one batch of
`--views-per-rank` (default 8) camera views per GPU, each view a full forward + backward of the hot
path with gradients accumulated into one flat per-Gaussian buffer; with N > 1 ranks every rank renders
its own views against the replicated Gaussian set and the step ends with one NCCL all-reduce of the
flat gradient buffer and of the densification statistics (view-parallel, weak scaling).
value = frames (views) per second over all ranks, device-timed, max over ranks.

Printed JSON keys follow the driver's contract; see DESIGN.md section "Measurement".
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from gaussian_splatting_lightning_b200 import synthetic  # noqa: E402

METRIC = "rendered frames/sec fwd+bwd @1M Gaussians 1080p SH3"
UNIT = "frames/s"
ORBIT_VIEWS = 256  # bench views orbit the scene centre in steps of 2*pi/256 around view 0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=list(synthetic.CONFIGS))
    ap.add_argument("--views-per-rank", type=int, default=8)
    ap.add_argument("--streams", type=int, default=3, help="CUDA streams the views of a step are spread over")
    ap.add_argument("--partition", default="block", choices=["block", "interleaved"],
                    help="which of the step's N*views-per-rank views a rank owns (see view_index)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--per-view-backward", action="store_true",
                    help="run every view's preprocess backward on its own (default: one fused pass per step over all views)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the secondary blocks (single view per step, C1, C4 strong scaling, C5 render only)")
    return ap.parse_args()


def view_index(rank: int, j: int, vpr: int, world: int = 1, partition: str = "block") -> int:
    """global view g -> orbit slot: 0, +1, -1, +2, -2, ... around view 0.

    The views get lighter away from view 0 (fewer Gaussians in the frustum).  "block" (default): rank r owns
    g = r*vpr .. r*vpr + vpr - 1, so rank 0 renders exactly the N = 1 views whatever N and the max-over-ranks
    timing of the weak-scaling bench is never flattered by lighter views.  "interleaved": g = j*N + r, the
    assignment of view_parallel.partition_views (rank r owns k = r mod N) -- the balanced split of a FIXED
    set of N*vpr views, for the strong-scaling runs of C4 (64 cameras per step)."""
    g = rank * vpr + j if partition == "block" else j * world + rank
    return ((g + 1) // 2) * (1 if g % 2 else -1) % ORBIT_VIEWS


def make_cameras(W, H, rank, vpr, device, world=1, partition="block"):
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizationSettings
    cams = []
    bg = torch.zeros(3, device=device)
    for j in range(vpr):
        c = synthetic.make_camera(W, H, k=view_index(rank, j, vpr, world, partition), num_views=ORBIT_VIEWS)
        cams.append(dict(host=c, settings=GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=c["tanfovx"], tanfovy=c["tanfovy"], bg=bg, scale_modifier=1.0,
            viewmatrix=c["viewmatrix"].to(device), projmatrix=c["projmatrix"].to(device), sh_degree=3,
            campos=c["campos"].to(device), prefiltered=False, debug=False, antialiasing=False)))
    return cams


class ClockSampler:
    """SM clock + throttle reasons during the timed region (B200_PROFILING.md clocks line).

    Default: one `nvidia-smi -lms 200` child process (the recipe's query list).  start() BLOCKS until the
    first sample arrives: on a fresh box the child's driver/NVML initialisation takes seconds and, when it
    overlapped the timed region, stalled this process's launches (measured: 90.7 ms per step instead of
    24.4 ms).  An in-process NVML polling thread (GSR_BENCH_SAMPLER=nvml) was measured slower still
    (254-304 vs 328 frames/s: it competes with the launching thread for the GIL), so it is not the default.
    """
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("sw_power_cap", 0x4))
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, mode=None, period=0.25):
        self.gpu = gpu_index
        self.mode = mode or os.environ.get("GSR_BENCH_SAMPLER", "smi")
        self.period = period
        self.samples = []      # (time, sm_mhz, sm_max_mhz, reasons bitmask)
        self.proc = None
        self._stop = threading.Event()
        self.t = None

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x.strip() for x in vis.split(",") if x.strip()]
            if self.gpu < len(ids) and ids[self.gpu].isdigit():
                return int(ids[self.gpu])
        return self.gpu

    def start(self):
        if self.mode == "none":
            return
        if self.mode == "nvml":
            try:
                import pynvml
                pynvml.nvmlInit()
                self.nv = pynvml
                self.h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index())
                self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
                self.t = threading.Thread(target=self._poll_nvml, daemon=True)
                self.t.start()
                return
            except Exception:
                self.mode = "smi"
        try:
            ms = int(os.environ.get("GSR_BENCH_SMI_MS", "200"))
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self._physical_index()), "-lms", str(ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read_smi, daemon=True)
            self.t.start()
            t_end = time.time() + 30.0
            while not self.samples and time.time() < t_end and self.proc.poll() is None:
                time.sleep(0.05)
        except Exception:
            self.proc = None

    def _poll_nvml(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    rs = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((time.time(), sm, self.sm_max, rs))
            except Exception:
                pass
            self._stop.wait(self.period)

    def _read_smi(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.strip().split(",")]
            try:
                rs = 0
                for (name, bit), v in zip(self.REASONS, f[5:9]):
                    if v.lower().startswith("active"):
                        rs |= bit
                self.samples.append((time.time(), float(f[1]), float(f[2]), rs))
            except Exception:
                continue

    def stop(self, t0, t1):
        if self.mode == "none":
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["sampler disabled"]}
        self._stop.set()
        if self.proc is not None:
            time.sleep(0.1)
            self.proc.terminate()
        if self.t is not None:
            self.t.join(timeout=1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        rows = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples
        bits = 0
        for s in rows:
            bits |= s[3]
        return {"sm_mhz": float(np.median([s[1] for s in rows])), "sm_max_mhz": max(s[2] for s in rows),
                "reasons": sorted(n for n, b in self.REASONS if bits & b), "samples": len(rows),
                "sampler": self.mode}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)", 1965.0


def stage_model(P, M, Pvis, R, L, A, HW, T, passes):
    """Algorithmic bytes / flops per frame of every stage.  The binning stages are
    those of the build: tile-segmented (csrc/tile_binning.cu) unless GSR_BINNING=onesweep."""
    if os.environ.get("GSR_BINNING") == "onesweep":
        binning = {"scan": dict(bytes=8 * P), "duplicate": dict(bytes=8 * P + 12 * Pvis + 12 * R),
                   "sort": dict(bytes=(8 + 24 * passes) * R), "tile_ranges": dict(bytes=8 * R + 8 * T)}
    else:
        binning = {"scan": dict(bytes=4 * T + 8 * T),                  # tile scan: counts in, ranges out
                   "duplicate": dict(bytes=4 * P + 36 * Pvis + 12 * R),  # scatter: radii, record, 4 B cursor + 8 B composite
                   "sort": dict(bytes=8 * T + 12 * R),                  # per-tile sort: composite in, point list out
                   "tile_ranges": dict(bytes=12 * T)}                   # tile order (LPT): ranges in, order out
    return {
        "preprocess": dict(bytes=52 * P + (12 * M + 67) * Pvis),
        **binning,
        "render_forward": dict(bytes=44 * L + 24 * HW, flops=24 * A, ex2=A),
        "render_backward": dict(bytes=84 * L + 24 * HW, flops=75 * A, ex2=A),
        "preprocess_backward": dict(bytes=(187 + 24 * M) * Pvis),
    }


# which roofline bounds a stage: the compositing kernels are bound by FP32 instruction issue, the
# per-Gaussian kernels by HBM bytes, the binning stages by latency / L2 transactions (reported against HBM bytes)
STAGE_BOUND = {"render_forward": "fp32_issue", "render_backward": "fp32_issue"}
EXCHANGE_CHUNKS = int(os.environ.get("GSR_EXCHANGE_CHUNKS", "1"))   # Gaussian ranges the fused step backward + all-reduce are pipelined over (N > 1).
# Measured on 8 B200 (C3, 248 MB per step): 1 chunk 11.29 ms per step, 2: 11.39, 4: 11.49, 8: 11.59 -- the smaller collectives lose
# more than the overlap with the fused pass (0.77 ms) wins; on 2 GPUs 4 chunks won 0.08 ms.  Default: one collective.
DENSIFY_INTERVAL = 100   # gs_lightning/lightning/gs_lightning_module.py:32,198: densification reads max_radii2D every 100 steps
NCU_COUNTS = os.path.join(ROOT, "profiles", "ncu_counts.json")   # per-launch counters of the committed ncu capture


class Rig:
    """One workload on one rank: replicated scene, this rank's cameras, step buffers, view streams."""

    def __init__(self, workload, vpr, streams, device, rank, world, partition, seed_grad=1234, fused=True):
        from gaussian_splatting_lightning_b200.view_parallel import StepBuffers, ViewStreams
        self.workload, self.vpr, self.world, self.rank, self.device = workload, vpr, world, rank, device
        self.P, self.deg, self.W, self.H = synthetic.config(workload)
        self.M = (self.deg + 1) ** 2
        self.scene_cpu = synthetic.make_scene(self.P, self.deg, self.W, self.H, seed=0)
        self.scene = {k: v.to(device) for k, v in self.scene_cpu.items()}
        self.cams = make_cameras(self.W, self.H, rank, vpr, device, world, partition)
        for c in self.cams:
            c["settings"] = c["settings"]._replace(sh_degree=self.deg)
        g = torch.Generator(device="cpu").manual_seed(seed_grad + rank)
        self.dL_dcolor = torch.randn(3, self.H, self.W, generator=g).to(device)
        self.dL_dinv = torch.randn(1, self.H, self.W, generator=g).to(device)
        # per-step exchange buffers: flat gradient buffer [means3D 3 | sh 3M | opacity 1 | scales 3 | rot 4]
        # (59 floats per Gaussian at M = 16) + densification-statistics deltas (view_parallel.py)
        # (chunk-major with N > 1 ranks: the exchange then follows the fused preprocess backward range by range)
        self.bufs = StepBuffers(self.P, self.M, device, chunks=(EXCHANGE_CHUNKS if world > 1 and fused else 1))
        one = self.bufs.num_chunks == 1
        self.acc_into, self.stat_into = (self.bufs.grads(), self.bufs.densify_stats()) if one else (None, None)
        self.n_streams = streams
        self.steps_done = 0
        self.fused, self.states = fused, []
        self.vstreams = ViewStreams(device, streams)   # independent views of a step round-robin over CUDA streams

    def one_view(self, c, fwd_only=False):
        from gaussian_splatting_lightning_b200.rasterizer import backward_raw, backward_render_only, forward_raw
        rs, sc = c["settings"], self.scene
        R, color, radii, geom, binning, img, invd, cap = forward_raw(
            rs, sc["means3D"], sc["shs"], None, sc["opacities"], sc["scales"], sc["rotations"], None)
        if fwd_only:
            return color
        if self.fused:
            # the compositing backward of the view; its preprocess backward runs with the other views' in ONE pass
            # over the Gaussians at the end of the step (rasterizer.preprocess_backward_views)
            scratch = backward_render_only(rs, sc["means3D"], radii, sc["shs"], None, sc["opacities"], sc["scales"],
                                           sc["rotations"], None, self.dL_dcolor, self.dL_dinv, geom, binning, img, R)
            self.states.append((rs, radii, geom, scratch))
            return color
        # gradients += into the flat all-reduce buffer and the view's densification statistics
        # (gaussian_model.py:174-181), both fused into the preprocess backward kernel
        backward_raw(rs, sc["means3D"], radii, sc["shs"], None, sc["opacities"], sc["scales"], sc["rotations"], None,
                     self.dL_dcolor, self.dL_dinv, geom, binning, img, R, accumulate_into=self.acc_into,
                     stats=self.stat_into)
        return color

    def step(self, fwd_only=False):
        if not fwd_only:
            # per-step buffers: gradient sum and statistics DELTAS of this step's views; the radii candidates stay
            # (max commutes with the steps: their exchange is deferred to densification time, StepBuffers.all_reduce)
            # (the fused pass overwrites every gradient row: only the statistics need clearing)
            self.bufs.zero_(keep_max_radii=self.world > 1, stats_only=self.fused)
        self.states = []
        self.vstreams.run(len(self.cams), lambda j: self.one_view(self.cams[j], fwd_only))
        if fwd_only:
            return
        if self.fused:
            # ONE pass over the Gaussians for all views of the step, Gaussian range by Gaussian range; each range's
            # all-reduce (sum of gradients + statistics: one collective) runs under the pass over the next range
            from gaussian_splatting_lightning_b200.view_parallel import fused_backward_and_exchange
            fused_backward_and_exchange(self.states, self.scene, self.bufs, exchange=self.world > 1, max_radii=False)
            self.states = []
        elif self.world > 1:
            self.bufs.all_reduce(max_radii=False)   # sum(flat + stats): ONE collective per step
        if self.world > 1:
            self.steps_done += 1
            if self.steps_done % DENSIFY_INTERVAL == 0:
                self.bufs.all_reduce_max_radii()    # the deferred max, when densification would read max_radii2D

    def free(self):
        for k in list(self.__dict__):
            delattr(self, k)
        torch.cuda.empty_cache()


def timed(fn, steps, warmup, device, world):
    """`steps` calls of fn between barrier + synchronize, CUDA events, max over ranks -> (ms, t0, t1)."""
    import torch.distributed as dist
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t1 = time.time()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), t0, t1


def regions(fn, steps, n, device, world, warmup=0):
    """n timed regions of `steps` steps each -> (median ms, sorted list of ms, [(t0, t1)])."""
    out = [timed(fn, steps, warmup if i == 0 else 0, device, world) for i in range(n)]
    ms = sorted(o[0] for o in out)
    return ms[len(ms) // 2] if len(ms) % 2 else 0.5 * (ms[len(ms) // 2 - 1] + ms[len(ms) // 2]), ms, [(o[1], o[2]) for o in out]


def workload_counters(rig):
    """P_vis, R, L, A averaged over this rank's views, from radii / n_contrib of the CUDA forward (the oracle's
    definitions: L = sum over tiles of the max n_contrib, A = sum over pixels), and what the backward compositing
    kernel really evaluates, counted from the forward's hit words (rasterizer.hit_word_stats)."""
    from gaussian_splatting_lightning_b200.rasterizer import export_state, forward_raw, hit_word_stats
    acc = dict(Pvis=0.0, R=0.0, L=0.0, A=0.0, visits_both=0.0, visits_single=0.0, pairs_evaluated=0.0)
    first = None
    sc = rig.scene
    for c in rig.cams:
        rs = c["settings"]
        R, color, radii, geom, binning, img, invd, cap = forward_raw(
            rs, sc["means3D"], sc["shs"], None, sc["opacities"], sc["scales"], sc["rotations"], None, binning_hint=0)
        st = export_state(rs, rig.P, radii, geom, binning, img, R, cap)
        nc = st["n_contrib"].to(torch.int64)
        H, W = nc.shape
        gy, gx = (H + 15) // 16, (W + 15) // 16
        pad = torch.zeros((gy * 16, gx * 16), dtype=torch.int64, device=rig.device)
        pad[:H, :W] = nc
        tile_max = pad.view(gy, 16, gx, 16).amax(dim=(1, 3))
        hw = hit_word_stats(rs, binning, st["n_contrib"], st["ranges"], cap)
        acc["Pvis"] += float((radii > 0).sum()); acc["R"] += float(R)
        acc["L"] += float(tile_max.sum()); acc["A"] += float(nc.sum())
        for k in ("visits_both", "visits_single", "pairs_evaluated"):
            acc[k] += float(hw[k])
        if first is None:
            first = dict(hw, R=float(R))
        del st
    out = {k: v / len(rig.cams) for k, v in acc.items()}
    out["view0"] = first
    return out


class E2E:
    """The same step through the public autograd API with HOST inputs: per view a pinned-host ground-truth image and
    the camera are copied host->device (copy stream, a ring of streams+1 slots like a prefetching loader, inside the
    timed region) and the loss is read back device->host.  The parameters are leaves whose .grad are views of the
    step's flat gradient buffer (training.parameter_groups_on); the rasterizer backward adds into it in-kernel
    (rasterizer.accumulate_gradients_into).  The activations and optimizer.step of GSLightningModule.training_step are
    NOT part of this region (they are torch elementwise kernels + one fused Adam launch: e2e.neighbours times Adam)."""

    def __init__(self, rig):
        from gaussian_splatting_lightning_b200.training import l1_loss, parameter_groups_on, photometric_loss
        self.rig = rig
        dev, H, W, vpr = rig.device, rig.H, rig.W, rig.vpr
        self.params = {k: rig.scene[k].clone().requires_grad_(True)
                       for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        self.gt_host = [torch.rand(3, H, W).pin_memory() for _ in rig.cams]
        # the same images as a dataset stores them: 8 bits per channel (uploaded as such when self.u8 is set)
        self.gt_host_u8 = [(t * 255).round().to(torch.uint8).pin_memory() for t in self.gt_host]
        self.u8 = False
        self.cam_host = [torch.cat([c["host"]["viewmatrix"].reshape(-1), c["host"]["projmatrix"].reshape(-1),
                                    c["host"]["campos"].reshape(-1), torch.zeros(3)]).pin_memory() for c in rig.cams]
        self.loss_host = torch.zeros(2, vpr).pin_memory()    # double buffered: step s is read while step s + 1 runs
        self.done = [torch.cuda.Event(), torch.cuda.Event()]
        self.steps, self.uploaded = 0, 0                     # steps run; global views uploaded so far
        if rig.bufs.num_chunks == 1:
            parameter_groups_on(rig.bufs, self.params)     # .grad = the parameter's rows of the step buffer
        self.copy_stream = torch.cuda.Stream(dev)
        self.cam_stream = torch.cuda.Stream(dev)
        self.n_slots = rig.n_streams + 2
        self.slots = [dict(cam=torch.empty(38, device=dev), gt=torch.empty(3, H, W, device=dev),
                           gt_u8=torch.empty(3, H, W, dtype=torch.uint8, device=dev), cam_ready=torch.cuda.Event(),
                           ready=torch.cuda.Event(), free=torch.cuda.Event()) for _ in range(self.n_slots)]
        self.loss_fns = {"l1": l1_loss, "full": lambda c_, g_: photometric_loss(c_[None], g_[None], 0.2)[0]}
        self.loss_kind = "l1"
        from gaussian_splatting_lightning_b200.rasterizer import StepSink
        one = rig.bufs.num_chunks == 1
        # chunk-major buffer (N > 1): the sink only parks the views; names of what it covers are all it needs
        self.sink = StepSink(rig.bufs.grads(0), rig.bufs.densify_stats(0)) if not one else StepSink(rig.acc_into, rig.stat_into)
        self.h2d_bytes_per_step = vpr * (3 * H * W * 4 + 38 * 4)
        self.d2h_bytes_per_step = vpr * 4

    def upload_until(self, g_end):
        """host inputs of the global views [uploaded, g_end) -> device, on the copy stream, a ring of n_slots slots the
        way a prefetching loader hands batches over (global view g = step * views_per_rank + j: the ring runs across
        steps); every byte of every step is copied inside the timed region"""
        vpr = self.rig.vpr
        while self.uploaded < g_end:
            g = self.uploaded
            sl, j = self.slots[g % self.n_slots], g % vpr
            with torch.cuda.stream(self.cam_stream):              # the forward needs only the camera (152 bytes): its
                self.cam_stream.wait_event(sl["free"])            # copy does not queue behind the previous view's image
                sl["cam"].copy_(self.cam_host[j], non_blocking=True)
                sl["cam_ready"].record(self.cam_stream)
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(sl["free"])           # the view that used this slot last is done with it
                if self.u8:
                    sl["gt_u8"].copy_(self.gt_host_u8[j], non_blocking=True)
                else:
                    sl["gt"].copy_(self.gt_host[j], non_blocking=True)
                sl["ready"].record(self.copy_stream)              # the loss needs the image
            self.uploaded += 1

    def view(self, j):
        from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer, accumulate_gradients_into
        rig, p = self.rig, self.params
        st = torch.cuda.current_stream()                 # ViewStreams.run: view j on stream j mod n
        g = self.steps * rig.vpr + j
        self.upload_until(g + self.n_slots - 1)          # the slot view g - 1 just released; overlaps the rendering
        sl = self.slots[g % self.n_slots]
        st.wait_event(sl["cam_ready"])
        cam_d, gt = sl["cam"], sl["gt"]
        rs = rig.cams[j]["settings"]._replace(viewmatrix=cam_d[0:16].view(4, 4), projmatrix=cam_d[16:32].view(4, 4),
                                              campos=cam_d[32:35], bg=cam_d[35:38])
        if rig.fused:
            # StepSink: the view's backward runs the compositing backward and parks the view; the preprocess backward
            # of all views is ONE pass at the end of the step.  No dummy screen-space leaf: it exists in the reference
            # (gs_lightning_module.py:316-317) to carry dL_dmeans2D to the densification statistics, which the fused
            # pass produces itself.
            with self.sink:
                color, radii, invd = GaussianRasterizer(rs)(
                    means3D=p["means3D"], means2D=None, shs=p["shs"], colors_precomp=None, opacities=p["opacities"],
                    scales=p["scales"], rotations=p["rotations"], cov3D_precomp=None)
        else:
            sp = torch.zeros_like(p["means3D"], requires_grad=True)
            with accumulate_gradients_into(rig.acc_into, rig.stat_into):
                color, radii, invd = GaussianRasterizer(rs)(
                    means3D=p["means3D"], means2D=sp, shs=p["shs"], colors_precomp=None, opacities=p["opacities"],
                    scales=p["scales"], rotations=p["rotations"], cov3D_precomp=None)
        st.wait_event(sl["ready"])
        if self.u8:
            torch.mul(sl["gt_u8"], 1.0 / 255.0, out=gt)      # 8-bit image -> float32 on the device (one kernel)
        loss = self.loss_fns[self.loss_kind](color, gt)
        loss.backward()
        self.loss_host[self.steps & 1, j].copy_(loss.detach(), non_blocking=True)
        sl["free"].record(st)

    def step(self):
        rig = self.rig
        t_host = time.perf_counter()
        main = torch.cuda.current_stream()
        rig.bufs.zero_(keep_max_radii=rig.world > 1, stats_only=rig.fused)
        self.upload_until(self.steps * rig.vpr + self.n_slots - 1)
        rig.vstreams.run(rig.vpr, self.view)
        if rig.fused:
            from gaussian_splatting_lightning_b200.view_parallel import fused_backward_and_exchange
            assert len(self.sink.pending) == rig.vpr      # every view parked: the pass may overwrite the buffer
            p = self.params
            fused_backward_and_exchange(self.sink.pending, dict(means3D=p["means3D"], shs=p["shs"], opacities=p["opacities"],
                                                                scales=p["scales"], rotations=p["rotations"]),
                                        rig.bufs, exchange=rig.world > 1, max_radii=False)
            self.sink.pending = []
        elif rig.world > 1:
            rig.bufs.all_reduce(max_radii=False)
        # the step's result (its losses) goes device -> host every step; the host reads it one step late, the way a
        # training loop logs, so that it never stalls the GPU between steps (the last step is waited for by the caller's
        # synchronize)
        self.done[self.steps & 1].record(main)
        self.steps += 1
        # host time inside the step before the loss read: enqueueing + every forward's wait for its view's
        # num_rendered (the mailbox read of gsr_forward, which ties the host to the GPU about one view ahead)
        self.host_busy_ms = (time.perf_counter() - t_host) * 1e3
        if self.steps >= 2:
            self.done[self.steps & 1].synchronize()      # step s - 1
            return float(self.loss_host[self.steps & 1].sum())
        return 0.0


def ev_ms(fn, reps=20):
    for _ in range(3):
        fn()
    a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a_.record()
    for _ in range(reps):
        fn()
    b_.record(); torch.cuda.synchronize()
    return a_.elapsed_time(b_) / reps


def neighbours(rig, e2e):
    """The training-step neighbours of the path, each one launch, HBM bound."""
    from gaussian_splatting_lightning_b200.training import FusedAdam, densify_fold, fused_ssim, l1_loss
    device, P, H, W = rig.device, rig.P, rig.H, rig.W
    params, bufs = e2e.params, rig.bufs
    if bufs.num_chunks > 1:      # the single-kernel measurements want whole-P gradient tensors
        from gaussian_splatting_lightning_b200.training import parameter_groups_on
        from gaussian_splatting_lightning_b200.view_parallel import StepBuffers
        bufs = StepBuffers(P, rig.M, device)
        parameter_groups_on(bufs, params)
    hbm_pk = peaks()[0]
    opt = FusedAdam([{"params": [params[k]], "lr": lr} for k, lr in
                     (("means3D", 1.6e-4), ("shs", 2.5e-3), ("opacities", 5e-2), ("scales", 5e-3), ("rotations", 1e-3))],
                    lr=0.0, eps=1e-15)
    n_par = sum(p_.numel() for p_ in params.values())
    img_a, img_b = torch.rand(3, H, W, device=device, requires_grad=True), torch.rand(3, H, W, device=device)
    run_stats = [torch.zeros(P, device=device) for _ in range(3)]
    out = []
    for name, fn, nbytes in (
            ("adam (5 groups, one launch, zero_grad fused)", lambda: opt.step(zero_grad=True), 32 * n_par),
            ("l1_loss value + unit gradient", lambda: l1_loss(img_a, img_b), 12 * 3 * H * W),
            ("densify_fold", lambda: densify_fold(bufs.max_radii, bufs.stats[0], bufs.stats[1], *run_stats), 32 * P)):
        ms_ = ev_ms(fn)
        out.append({"kernel": name, "ms": ms_, "algorithmic_bytes": nbytes, "GBps": nbytes / (ms_ * 1e6),
                    "hbm_frac": nbytes / (ms_ * 1e6) / hbm_pk})
    # fused SSIM (value + gradient, one launch) next to the reference's own CUDA kernels (oracle/_ref, compiled from
    # /root/reference/submodules/fused-ssim by oracle/build_ref.py) on the same image pair: fusedssim + fusedssim_backward
    im_a, im_b = img_a.detach()[None].clone().requires_grad_(True), img_b[None]
    ms_ = ev_ms(lambda: fused_ssim(im_a, im_b))
    ent = {"kernel": "fused_ssim value + gradient (one launch)", "ms": ms_, "algorithmic_bytes": 12 * 3 * H * W,
           "GBps": 12 * 3 * H * W / (ms_ * 1e6), "hbm_frac": 12 * 3 * H * W / (ms_ * 1e6) / hbm_pk}
    try:
        from oracle import build_ref
        if build_ref.available():
            ref_mod = build_ref.load()
            dmap = torch.full_like(im_b, 1.0 / im_b.numel())

            def ref_ssim():
                _m, d1, d2, d3 = ref_mod.fusedssim(0.01 ** 2, 0.03 ** 2, im_a.detach(), im_b, True)
                ref_mod.fusedssim_backward(0.01 ** 2, 0.03 ** 2, im_a.detach(), im_b, dmap, d1, d2, d3)
            ent["reference_cuda_ms"] = ev_ms(ref_ssim)
            ent["reference"] = "submodules/fused-ssim fusedssim + fusedssim_backward (oracle/_ref), same GPU"
    except Exception as ex:  # the checker is optional on the box
        ent["reference"] = f"unavailable: {ex}"
    out.append(ent)
    return out


def extra_block(workload, vpr, streams, partition, steps, device, rank, world, fwd_only=False, with_e2e=False, n_regions=3):
    """A secondary measurement of the same path (BASELINE.json's other configurations / usages): median of n_regions
    regions of `steps` steps, same timing rules as the headline."""
    rig = Rig(workload, vpr, streams, device, rank, world, partition)
    for _ in range(3):
        rig.step(fwd_only)
    med, ms, _ = regions(lambda: rig.step(fwd_only), steps, n_regions, device, world)
    frames = steps * vpr * world
    out = {"workload": f"{workload}: {rig.P} Gaussians, SH degree {rig.deg}, {rig.W}x{rig.H}", "views_per_rank": vpr,
           "streams_per_rank": streams, "partition": partition, "n_gpus": world, "steps": steps,
           "what": "render only (forward)" if fwd_only else "forward + backward + densification stats"
                   + (", one NCCL all-reduce of the step buffer per step" if world > 1 else ""),
           "value": frames / (med / 1e3), "unit": UNIT, "ms_per_step": med / steps,
           "regions_ms_per_step": [round(m / steps, 3) for m in ms]}
    if not fwd_only:
        medf, _, _ = regions(lambda: rig.step(True), steps, n_regions, device, world, warmup=2)
        out["forward_only"] = {"value": frames / (medf / 1e3), "unit": UNIT, "ms_per_step": medf / steps}
    if with_e2e:
        e = E2E(rig)
        mede, _, _ = regions(e.step, steps, n_regions, device, world, warmup=2)
        out["e2e"] = {"value": frames / (mede / 1e3), "unit": UNIT, "h2d_bytes_per_step": e.h2d_bytes_per_step,
                      "d2h_bytes_per_step": e.d2h_bytes_per_step}
        del e
    rig.free()
    return out


def run_ours(args):
    from gaussian_splatting_lightning_b200 import _native as N
    from gaussian_splatting_lightning_b200 import rasterizer as RZ
    from gaussian_splatting_lightning_b200.view_parallel import ViewStreams
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    vpr = args.views_per_rank
    rig = Rig(args.workload, vpr, args.streams, device, rank, world, args.partition, fused=not args.per_view_backward)
    P, deg, W, H, M = rig.P, rig.deg, rig.W, rig.H, rig.M
    counters = workload_counters(rig)
    K, Wm = args.steps, max(args.warmup, 3)
    REGIONS = 5

    # ---- headline: fwd+bwd, inputs resident in HBM -------------------------------------------------
    # REGIONS timed regions of exactly K steps each (barrier + synchronize on both sides, CUDA events, max over
    # ranks) with nvidia-smi sampling clocks concurrently; value = the MEDIAN region, min / max listed beside it.
    # One more region runs first with the poller already up and is not counted: the poller is another driver
    # client and its start was measured to stall this process's launches (82 ms per step for one region,
    # tools/diag_regions.py --poller).  Instrumented per region: binning-hint retries, cudaMalloc calls.
    for _ in range(Wm):
        rig.step()
    torch.cuda.synchronize()
    l0 = N.launch_count()
    ms_unsampled, _, _ = timed(rig.step, K, 0, device, world)
    launches = N.launch_count() - l0
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    timed(rig.step, K, 0, device, world)
    per_region = []
    for _ in range(REGIONS):
        c0, m0 = dict(RZ.counters), torch.cuda.memory_stats(device).get("num_device_alloc", 0)
        ms_, t0_, t1_ = timed(rig.step, K, 0, device, world)
        per_region.append(dict(ms=ms_, t0=t0_, t1=t1_, hint_retries=RZ.counters["hint_retries"] - c0["hint_retries"],
                               cuda_mallocs=torch.cuda.memory_stats(device).get("num_device_alloc", 0) - m0))
    order = sorted(per_region, key=lambda r: r["ms"])
    ms_total = order[len(order) // 2]["ms"]
    clocks = sampler.stop(min(r["t0"] for r in per_region), max(r["t1"] for r in per_region)) if rank == 0 else None
    if clocks is not None:
        clocks["regions_ms_per_step"] = [round(r["ms"] / K, 3) for r in per_region]
        clocks["unsampled_ms_per_step"] = round(ms_unsampled / K, 3)
        clocks["value_from"] = f"median of {REGIONS} sampled regions of {K} steps"
        clocks["min_ms_per_step"], clocks["max_ms_per_step"] = round(order[0]["ms"] / K, 3), round(order[-1]["ms"] / K, 3)
        clocks["hint_retries_per_region"] = [r["hint_retries"] for r in per_region]
        clocks["cuda_mallocs_per_region"] = [r["cuda_mallocs"] for r in per_region]
    # per-stage CUDA events (on the launching stream, inside the C library) in a second pass of the same steps
    # (views one after another on ONE stream here, so that a stage's events bracket that stage alone)
    vstreams_timed, rig.vstreams = rig.vstreams, ViewStreams(device, 1)
    N.profile_reset(); N.profile_enable(True)
    timed(rig.step, K, 0, device, world)
    prof = N.profile_read(); N.profile_enable(False)
    rig.vstreams = vstreams_timed
    ms_fwd, fwd_regions, _ = regions(lambda: rig.step(True), K, 3, device, world, warmup=2)

    frames = K * vpr * world
    value = frames / (ms_total / 1e3)
    fwd_value = frames / (ms_fwd / 1e3)

    # ---- e2e: public API, host buffers (pinned GT image + camera) -> loss on the host -------------
    e2e = None
    if not args.no_e2e:
        E = E2E(rig)
        ms_e2e, e2e_regions, _ = regions(E.step, K, 3, device, world, warmup=2)
        e2e_host_ms = E.host_busy_ms
        E.loss_kind = "full"     # the reference's training loss: 0.8 L1 + 0.2 (1 - SSIM), one fused kernel
        ms_e2e_full, _, _ = regions(E.step, K, 3, device, world, warmup=2)
        E.loss_kind = "l1"
        E.u8 = True              # ground truth uploaded as 8-bit images (4x fewer H2D bytes), converted on the device
        ms_e2e_u8, _, _ = regions(E.step, K, 3, device, world, warmup=2)
        E.u8 = False
        e2e = {"value": frames / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": E.h2d_bytes_per_step,
               "d2h_bytes_per_step": E.d2h_bytes_per_step, "regions_ms_per_step": [round(m / K, 3) for m in e2e_regions],
               "host_ms_per_step_before_loss_read": round(e2e_host_ms, 3),
               "what": "GaussianRasterizer autograd API (through the `_C` torch extension) + fused L1 loss (training.l1_loss) "
                       "with fused gradient accumulation (accumulate_gradients_into) on the same CUDA streams as `value`; "
                       "per view: pinned GT image and camera H2D (copy stream, ring of streams+1 slots, inside the timed "
                       "region), loss D2H.  Parameters are leaf tensors; the activations and optimizer.step of the "
                       "reference's training_step are outside this region (Adam is timed in `neighbours`)",
               "with_uint8_targets": {"value": frames / (ms_e2e_u8 / 1e3), "unit": UNIT,
                                      "h2d_bytes_per_step": vpr * (3 * H * W + 38 * 4), "what": "same, the ground-truth images "
                                      "uploaded as 8 bits per channel (the form a dataset stores) and converted to float32 by one "
                                      "device kernel per view"},
               "with_reference_training_loss": {"value": frames / (ms_e2e_full / 1e3), "unit": UNIT, "what": "same, with "
                   "the loss of GSLightningModule.calculate_loss: 0.8 L1 + 0.2 (1 - fused_ssim) (training.photometric_loss)"}}
        e2e["neighbours"] = neighbours(rig, E)
        del E

    # ---- per-stage roofline (rank 0 computes; the other ranks only take part in the extras) --------
    line = None
    if rank == 0:
        line = headline(args, rig, counters, prof, clocks, value, fwd_value, ms_total, ms_fwd, e2e, launches, K, Wm, world)
    cpu_scene, cam0 = rig.scene_cpu, rig.cams[0]["host"]
    rig.free()

    # ---- the other configurations and usages north_star / BASELINE.json name, carried in the same line ----------
    if not args.no_extras:
        extras = {}
        S = max(2, min(K, 3))
        # the reference trains ONE view per step (gs_lightning_module.py:139-140): single-view latency, one stream
        extras["single_view_per_step"] = extra_block(args.workload, 1, 1, "block", 8 * S, device, rank, world, with_e2e=True)
        # BASELINE.json configs[0]: the reference Python rasterizer's CPU-runnable case, on the GPU path
        extras["c1"] = extra_block("C1", 1, 1, "block", 50, device, rank, world)
        # configs[3]: 3M Gaussians, 64 cameras per step at 1600x1200 split over the ranks (STRONG scaling, the
        # library's own interleaved view assignment), NCCL all-reduce of the step buffer
        if 64 % world == 0:
            extras["c4_strong"] = dict(extra_block("C4", 64 // world, args.streams, "interleaved", S, device, rank, world),
                                       scaling="strong", frames_per_step=64)
        # configs[4]: 6M Gaussians, 3840x2160, 8 views per GPU, render only (replicas, no collective)
        extras["c5_render_only"] = extra_block("C5", 8, args.streams, "block", S, device, rank, world, fwd_only=True)
        if line is not None:
            line["extras"] = extras
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_oracle_frames_per_s(args.workload, cpu_scene, cam0, 6)   # ~11 s of CPU work
        line["reference_python_rasterizer_c1"] = reference_python_c1()
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def ffi_name():
    from gaussian_splatting_lightning_b200 import _native
    return "torch C++ extension _C" if _native.ext() is not None else "ctypes"


def reference_python_c1():
    """north_star's reported baseline: the reference's own gs_lightning Python rasterizer on BASELINE.json configs[0]
    (C1).  /root/reference does not travel to the GPU box, so the number is the one measured in the build container by
    tools/time_reference_python.py (committed with its output, profiles/reference_python_c1.json)."""
    p = os.path.join(ROOT, "profiles", "reference_python_c1.json")
    if os.path.exists(p):
        return json.load(open(p))
    return None


def headline(args, rig, counters, prof, clocks, value, fwd_value, ms_total, ms_fwd, e2e, launches, K, Wm, world):
    """The JSON line of rank 0: per-stage roofline from the CUDA-event stage times and the workload counters."""
    P, deg, W, H, M, vpr = rig.P, rig.deg, rig.W, rig.H, rig.M, rig.vpr
    hbm_peak, peak_src, sm_max = peaks()
    T = ((W + 15) // 16) * ((H + 15) // 16)
    passes = (32 + max(1, (T - 1).bit_length()) + 7) // 8
    model = stage_model(P, M, counters["Pvis"], counters["R"], counters["L"], counters["A"], H * W, T, passes)
    ncu = json.load(open(NCU_COUNTS)) if os.path.exists(NCU_COUNTS) else {"kernels": {}}
    sm_mhz = (clocks or {}).get("sm_mhz") or sm_max
    issue_peak = 148 * 4 * sm_mhz * 1e6            # warp instructions per second: 4 schedulers per SM, one issue per clock
    v0 = counters["view0"]
    visit_scale = ((counters["visits_both"] + counters["visits_single"]) / max(1.0, v0["visits_both"] + v0["visits_single"]))
    stages, stage_sum = [], 0.0
    if rig.fused:
        # one fused pass per STEP over all views: inputs in once (44 + 12 M), the gradient row + statistics out once
        # (44 + 12 M + 12), per view the radius of every Gaussian (4) and 49 B per visible one (raw sums 48, clamp bits 1);
        # reported per frame (1 / views_per_rank of the pass)
        model["preprocess_backward"] = dict(bytes=(P * (2 * (44 + 12 * M) + 12)) / vpr + 4 * P + 49 * counters["Pvis"])
    for name, (tot_ms, calls) in prof.items():
        if calls == 0:
            continue
        if rig.fused and name == "preprocess_backward":
            calls *= vpr          # per-frame share of the per-step pass
        avg_ms = tot_ms / calls
        stage_sum += tot_ms
        m = model[name]
        s = {"stage": name, "bound": STAGE_BOUND.get(name, "hbm"), "ms_per_frame": avg_ms, "calls": calls,
             "algorithmic_bytes": m["bytes"], "GBps": m["bytes"] / (avg_ms * 1e6),
             "hbm_frac": m["bytes"] / (avg_ms * 1e6) / hbm_peak}
        k = ncu["kernels"].get(name)
        if k:
            # ncu counters of ONE launch on view 0 (profiles/): DRAM traffic against the algorithmic bytes, and -- for
            # the issue-bound kernels -- warp instructions per launch, scaled from view 0 to the mean view of the step
            # by the number of (warp, entry) visits counted from the hit words
            s["dram_traffic_bytes_view0"] = k.get("dram_bytes")
            if k.get("dram_bytes") and k.get("algorithmic_bytes_view0"):
                s["traffic_over_algorithmic"] = k["dram_bytes"] / k["algorithmic_bytes_view0"]
            if k.get("inst") and s["bound"] == "fp32_issue":
                inst = k["inst"] * visit_scale
                s["warp_inst_per_launch"] = inst
                s["Gwarp_inst_per_s"] = inst / (avg_ms * 1e6)
                s["issue_frac"] = inst / (avg_ms * 1e-3) / issue_peak
                s["ncu_issue_active_pct"] = k.get("issue_active_pct")
                s["ncu_fma_pipe_cycles_pct"] = k.get("fma_pipe_cycles_pct")
        if name == "render_backward":
            # FP32 work actually done: 75 flop (SURVEY 8(d) convention) per (pixel, splat) pair the kernel EVALUATES
            # (hit words), not per pair of A = sum(n_contrib), which also counts what the culling skipped
            s["pairs_evaluated"] = counters["pairs_evaluated"]
            s["pairs_A_convention"] = counters["A"]
            s["TFLOPs_fp32_evaluated"] = 75 * counters["pairs_evaluated"] / (avg_ms * 1e9)
        stages.append(s)
    for s in stages:
        s["share_of_kernel_time"] = s["ms_per_frame"] * s["calls"] / stage_sum
    dom = max(stages, key=lambda s: s["ms_per_frame"] * s["calls"])
    fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
    if dom["bound"] == "fp32_issue" and "issue_frac" in dom:
        roofline = {"kernel": dom["stage"], "bound": "fp32_issue", "achieved": dom["Gwarp_inst_per_s"],
                    "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s", "frac": dom["issue_frac"],
                    "traffic": dom.get("dram_traffic_bytes_view0"),
                    "peak_source": f"148 SMs x 4 schedulers x {sm_mhz:.0f} MHz (observed SM clock)",
                    "inst_source": ncu.get("_source"), "ms_per_launch": dom["ms_per_frame"],
                    "warp_inst_per_launch": dom["warp_inst_per_launch"],
                    "hbm": {"achieved_GBps": dom["GBps"], "peak_GBps": hbm_peak, "frac": dom["hbm_frac"],
                            "algorithmic_bytes_per_launch": dom["algorithmic_bytes"], "peak_source": peak_src},
                    "note": "the dominant kernel is bound by FP32 instruction issue (SURVEY.md 8(d)), so `frac` is its "
                            "issue-slot utilisation: warp instructions per launch (ncu smsp__inst_executed.sum of the "
                            "committed capture, scaled by visits) / CUDA-event duration / issue peak; its HBM fraction "
                            "is kept under `hbm`; the HBM-bound stages carry hbm_frac in stages[]"}
        if "TFLOPs_fp32_evaluated" in dom:
            roofline["fp32"] = {"achieved_TFLOPs": dom["TFLOPs_fp32_evaluated"], "peak_TFLOPs_at_observed_clock": fp32_peak,
                                "frac": dom["TFLOPs_fp32_evaluated"] / fp32_peak,
                                "note": "75 flop x pairs EVALUATED (hit words); the A-based figure of round 1 read 1.135"}
    else:
        roofline = {"kernel": dom["stage"], "bound": "hbm", "achieved": dom["GBps"], "peak": hbm_peak, "unit": "GB/s",
                    "frac": dom["hbm_frac"], "traffic": dom.get("dram_traffic_bytes_view0"), "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": dom["algorithmic_bytes"], "ms_per_launch": dom["ms_per_frame"]}
    hbm_stages = [s for s in stages if s["bound"] == "hbm" and s["stage"] in ("preprocess", "preprocess_backward")]
    if hbm_stages:
        top = max(hbm_stages, key=lambda s: s["ms_per_frame"])
        roofline["hbm_bound_kernel"] = {"kernel": top["stage"], "achieved_GBps": top["GBps"], "peak_GBps": hbm_peak,
                                        "frac": top["hbm_frac"], "traffic_over_algorithmic": top.get("traffic_over_algorithmic")}
    return {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {P} Gaussians, SH degree {deg}, {W}x{H}, fwd+bwd + densification "
                               f"stats, {vpr} views per GPU per step" + (", NCCL all-reduce of 59 floats/Gaussian + stats" if world > 1 else ""),
                   "views_per_rank": vpr, "frames_per_step": vpr * world, "streams_per_rank": args.streams,
                   "parallelism": f"view-parallel x{world}", "partition": args.partition,
                   "ffi": ffi_name(),
                   "preprocess_backward": "one fused pass per step over all views" if rig.fused else "per view",
                   "l2": (lambda mb: f"inputs ({mb:.0f} MB of Gaussian parameters + {64 * P / 1e6:.0f} MB records) "
                          + ("exceed the 126 MB L2; no flush" if mb > 126 else
                             "fit the 126 MB L2: a parity-test size, not a bench line"))((44 + 12 * M) * P / 1e6),
                   "num_rendered_mean": counters["R"], "visible_mean": counters["Pvis"],
                   "pairs_A_mean": counters["A"], "list_entries_L_mean": counters["L"],
                   "pairs_evaluated_mean": counters["pairs_evaluated"], "view0": counters["view0"]},
        "forward_only": {"value": fwd_value, "unit": UNIT, "ms_per_step": ms_fwd / K},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "stages": stages,
        "cpu_baseline": None,
    }


def cpu_oracle_frames_per_s(workload, scene_cpu, cam_host, repeats):
    """The CPU restatement (oracle/, C++/OpenMP, 'port') timed on the host cores: full frames of the
    same workload, forward + backward.  This is the one place bench.py executes oracle/."""
    import oracle
    oracle.build()
    P, deg, W, H = synthetic.config(workload)
    cam = oracle.Camera(H, W, cam_host["tanfovx"], cam_host["tanfovy"], cam_host["viewmatrix"].numpy(),
                        cam_host["projmatrix"].numpy(), cam_host["campos"].numpy(), sh_degree=deg)
    a = {k: v.numpy() for k, v in scene_cpu.items()}
    rng = np.random.default_rng(0)
    dC = rng.standard_normal((3, H, W)).astype(np.float32)
    dD = rng.standard_normal((H, W)).astype(np.float32)
    # all the host cores this process may use, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1)
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    oracle.set_num_threads(max(1, avail))
    cores = oracle.num_threads()
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        st = oracle.forward(cam, a["means3D"], a["scales"], a["rotations"], a["opacities"], shs=a["shs"])
        t1 = time.perf_counter()
        oracle.backward(st, dC, dD)
        t2 = time.perf_counter()
        times.append((t2 - t0, t1 - t0))
    tot = float(np.median([t[0] for t in times]))
    fwd = float(np.median([t[1] for t in times]))
    return {"value": 1.0 / tot, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{repeats} full frame(s) of {workload} (view 0), forward + backward, oracle/gsr_oracle.cpp with OpenMP "
                      f"on {cores} threads; forward alone {fwd:.2f} s, forward+backward {tot:.2f} s per frame",
            "forward_only_value": 1.0 / fwd}


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path.  Its CUDA rasterizer source is
    absent (un-vendored submodule) and its Python rasterizer is forward-only and cannot hold this
    workload (O(tiles x P) mask), so the arm times the CPU restatement under oracle/ ('port')."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    P, deg, W, H = synthetic.config(args.workload)
    scene_cpu = synthetic.make_scene(P, deg, W, H, seed=0)
    cam = synthetic.make_camera(W, H, k=0, num_views=ORBIT_VIEWS)
    K = max(1, min(args.steps, 20))
    for _ in range(min(args.warmup, 1)):
        cpu_oracle_frames_per_s(args.workload, scene_cpu, cam, 1)
    t0 = time.perf_counter()
    cb = cpu_oracle_frames_per_s(args.workload, scene_cpu, cam, K)
    wall = time.perf_counter() - t0
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": min(args.warmup, 1), "ms_per_step": 1e3 / cb["value"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {P} Gaussians, SH degree {deg}, {W}x{H}, fwd+bwd; one frame per step",
                       "note": "one step = one full frame on the CPU; steps capped at 20 so the run ends within a minute; wall %.1f s" % wall},
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


_JSON_OUT = None


def emit(line: dict) -> None:
    """The ONE JSON line goes to the process's original stdout; everything else that libraries write to fd 1 during
    the run (NCCL prints its version banner there) was redirected to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_OUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_OUT, data)


def main():
    global _JSON_OUT
    args = parse_args()
    sys.stdout.flush()
    _JSON_OUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
