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
                   "tile_ranges": dict(bytes=0)}
    return {
        "preprocess": dict(bytes=52 * P + (12 * M + 67) * Pvis),
        **binning,
        "render_forward": dict(bytes=44 * L + 24 * HW, flops=24 * A, ex2=A),
        "render_backward": dict(bytes=84 * L + 24 * HW, flops=75 * A, ex2=A),
        "preprocess_backward": dict(bytes=(187 + 24 * M) * Pvis),
    }


def workload_counters(cams, scene, device):
    """P_vis, R, L, A averaged over this rank's views, from radii / n_contrib of the CUDA forward
    (the oracle's definitions: L = sum over tiles of the max n_contrib, A = sum over pixels)."""
    from gaussian_splatting_lightning_b200.rasterizer import export_state, forward_raw
    acc = dict(Pvis=0.0, R=0.0, L=0.0, A=0.0)
    for c in cams:
        rs = c["settings"]
        R, color, radii, geom, binning, img, invd, cap = forward_raw(
            rs, scene["means3D"], scene["shs"], None, scene["opacities"], scene["scales"], scene["rotations"], None,
            binning_hint=0)
        st = export_state(rs, scene["means3D"].shape[0], radii, geom, binning, img, R, cap)
        nc = st["n_contrib"].to(torch.int64)
        H, W = nc.shape
        gy, gx = (H + 15) // 16, (W + 15) // 16
        pad = torch.zeros((gy * 16, gx * 16), dtype=torch.int64, device=device)
        pad[:H, :W] = nc
        tile_max = pad.view(gy, 16, gx, 16).amax(dim=(1, 3))
        acc["Pvis"] += float((radii > 0).sum()); acc["R"] += float(R)
        acc["L"] += float(tile_max.sum()); acc["A"] += float(nc.sum())
        del st
    return {k: v / len(cams) for k, v in acc.items()}


def run_ours(args):
    from gaussian_splatting_lightning_b200 import _native as N
    from gaussian_splatting_lightning_b200.rasterizer import GaussianRasterizer, backward_raw, forward_raw
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    P, deg, W, H = synthetic.config(args.workload)
    M = (deg + 1) ** 2
    vpr = args.views_per_rank
    scene_cpu = synthetic.make_scene(P, deg, W, H, seed=0)
    scene = {k: v.to(device) for k, v in scene_cpu.items()}
    cams = make_cameras(W, H, rank, vpr, device, world, args.partition)
    for c in cams:
        c["settings"] = c["settings"]._replace(sh_degree=deg)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    dL_dcolor = torch.randn(3, H, W, generator=g).to(device)
    dL_dinv = torch.randn(1, H, W, generator=g).to(device)

    # per-step exchange buffers: flat gradient buffer [means3D 3 | sh 3M | opacity 1 | scales 3 | rot 4]
    # (59 floats per Gaussian at M = 16) + densification-statistics deltas (view_parallel.py)
    from gaussian_splatting_lightning_b200.view_parallel import StepBuffers
    bufs = StepBuffers(P, M, device)
    acc_into, stat_into = bufs.grads(), bufs.densify_stats()

    def one_view(c, fwd_only=False):
        rs = c["settings"]
        R, color, radii, geom, binning, img, invd, cap = forward_raw(
            rs, scene["means3D"], scene["shs"], None, scene["opacities"], scene["scales"], scene["rotations"], None)
        if fwd_only:
            return color
        # gradients += into the flat all-reduce buffer and the view's densification statistics
        # (gaussian_model.py:174-181), both fused into the preprocess backward kernel
        backward_raw(rs, scene["means3D"], radii, scene["shs"], None, scene["opacities"], scene["scales"],
                     scene["rotations"], None, dL_dcolor, dL_dinv, geom, binning, img, R,
                     accumulate_into=acc_into, stats=stat_into)
        return color

    from gaussian_splatting_lightning_b200.view_parallel import ViewStreams
    vstreams = ViewStreams(device, args.streams)   # independent views of a step round-robin over CUDA streams

    def step(fwd_only=False):
        if not fwd_only:
            bufs.zero_()        # per-step buffers: gradient sum and the statistics DELTAS of this step's views
        vstreams.run(len(cams), lambda j: one_view(cams[j], fwd_only))
        if world > 1 and not fwd_only:
            bufs.all_reduce()   # sum(flat + stats) in one collective, max(max_radii) over the ranks

    def timed(fn, steps, warmup):
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

    counters = workload_counters(cams, scene, device)
    K, Wm = args.steps, max(args.warmup, 3)

    # ---- headline: fwd+bwd, inputs resident in HBM -------------------------------------------------
    # The timed region (exactly K steps, barrier + synchronize on both sides, CUDA events, max over
    # ranks) is measured ATTEMPTS times with nvidia-smi sampling clocks concurrently, and once more
    # with no poller at all.  The poller is another driver client: on some boxes its queries were
    # measured to stall this process's launches for whole regions (43.8 and 90.7 ms per step against
    # 19.4 ms).  value = the fastest sampled attempt; if even that is > 10 % slower than the unsampled
    # region the unsampled one is reported and `clocks` says so.  All attempts are listed.
    ATTEMPTS = 3
    for _ in range(Wm):
        step()
    torch.cuda.synchronize()
    l0 = N.launch_count()
    ms_clean, _, _ = timed(step, K, 0)
    launches = N.launch_count() - l0
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    attempts = [timed(step, K, 0) for _ in range(ATTEMPTS)]
    best = min(attempts, key=lambda a: a[0])
    clocks = sampler.stop(best[1], best[2]) if rank == 0 else None
    ms_total = best[0]
    if clocks is not None:
        clocks["sampled_attempts_ms_per_step"] = [round(a[0] / K, 3) for a in attempts]
        clocks["unsampled_ms_per_step"] = round(ms_clean / K, 3)
        clocks["value_from"] = "fastest sampled attempt"
    if ms_total > 1.10 * ms_clean:
        ms_total = ms_clean
        if clocks is not None:
            clocks["value_from"] = ("unsampled region (every sampled attempt was > 10 % slower: the nvidia-smi poller "
                                    "stalled launches); clocks are those of the adjacent sampled attempts")
    # per-stage CUDA events (on the launching stream, inside the C library) in a second pass of the same steps
    # (views one after another on ONE stream here, so that a stage's events bracket that stage alone)
    vstreams_timed, vstreams = vstreams, ViewStreams(device, 1)
    N.profile_reset(); N.profile_enable(True)
    timed(step, K, 0)
    prof = N.profile_read(); N.profile_enable(False)
    vstreams = vstreams_timed
    ms_fwd = min(timed(lambda: step(True), K, 2 if a == 0 else 0)[0] for a in range(ATTEMPTS))  # fastest of the attempts, like value

    frames = K * vpr * world
    value = frames / (ms_total / 1e3)
    fwd_value = frames / (ms_fwd / 1e3)

    # ---- e2e: public API, host buffers (pinned GT image + camera) -> loss on the host -------------
    e2e = None
    if not args.no_e2e:
        params = {k: scene[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        gt_host = [torch.rand(3, H, W).pin_memory() for _ in cams]
        cam_host = [torch.cat([c["host"]["viewmatrix"].reshape(-1), c["host"]["projmatrix"].reshape(-1),
                               c["host"]["campos"].reshape(-1), torch.zeros(3)]).pin_memory() for c in cams]
        loss_host = torch.zeros(vpr).pin_memory()

        # .grad of the parameters = their slices of the step's flat gradient buffer: the backward kernel adds
        # into it (accumulate_gradients_into), the one all-reduce sends it, FusedAdam would read it
        from gaussian_splatting_lightning_b200.rasterizer import accumulate_gradients_into
        from gaussian_splatting_lightning_b200.training import l1_loss, parameter_groups_on
        parameter_groups_on(bufs, params)
        copy_stream = torch.cuda.Stream(device)
        n_slots = args.streams + 1
        slots = [dict(cam=torch.empty(38, device=device), gt=torch.empty(3, H, W, device=device),
                      ready=torch.cuda.Event(), free=torch.cuda.Event()) for _ in range(n_slots)]

        def upload(j):
            """this step's host inputs of view j -> device, on the copy stream (a ring of streams+1 slots, the
            way a prefetching data loader hands batches over); every byte is copied inside the timed region"""
            sl = slots[j % n_slots]
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(sl["free"])           # the view that used this slot last is done with it
                sl["cam"].copy_(cam_host[j], non_blocking=True)
                sl["gt"].copy_(gt_host[j], non_blocking=True)
                sl["ready"].record(copy_stream)

        from gaussian_splatting_lightning_b200.training import photometric_loss
        loss_fns = {"l1": l1_loss, "full": lambda c_, g_: photometric_loss(c_[None], g_[None], 0.2)[0]}

        def e2e_view(j):
            loss_fn = loss_fns[e2e_loss[0]]
            st = torch.cuda.current_stream()                 # ViewStreams.run: view j on stream j mod n
            if j + n_slots - 1 < vpr:
                upload(j + n_slots - 1)                      # its slot was freed by view j-1; overlaps the rendering
            sl = slots[j % n_slots]
            st.wait_event(sl["ready"])
            cam_d, gt = sl["cam"], sl["gt"]
            rs = cams[j]["settings"]._replace(viewmatrix=cam_d[0:16].view(4, 4), projmatrix=cam_d[16:32].view(4, 4),
                                              campos=cam_d[32:35], bg=cam_d[35:38])
            sp = torch.zeros_like(params["means3D"], requires_grad=True)
            with accumulate_gradients_into(acc_into, stat_into):
                color, radii, invd = GaussianRasterizer(rs)(
                    means3D=params["means3D"], means2D=sp, shs=params["shs"], colors_precomp=None,
                    opacities=params["opacities"], scales=params["scales"], rotations=params["rotations"],
                    cov3D_precomp=None)
            loss = loss_fn(color, gt)
            loss.backward()
            loss_host[j].copy_(loss.detach(), non_blocking=True)
            sl["free"].record(st)

        def e2e_step():
            main = torch.cuda.current_stream()
            bufs.zero_()
            for j in range(min(n_slots - 1, vpr)):
                upload(j)
            vstreams.run(vpr, e2e_view)
            if world > 1:
                bufs.all_reduce()
            main.synchronize()                               # the step's result (losses) is on the host
            return float(loss_host.sum())

        e2e_loss = ["l1"]
        ms_e2e, _, _ = timed(e2e_step, K, 2)
        e2e_loss[0] = "full"     # the reference's training loss: 0.8 L1 + 0.2 (1 - SSIM), one fused kernel
        ms_e2e_full, _, _ = timed(e2e_step, K, 2)
        e2e_loss[0] = "l1"
        h2d = vpr * (3 * H * W * 4 + 38 * 4)
        e2e = {"value": frames / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": vpr * 4,
               "what": "GaussianRasterizer autograd API + fused L1 loss (training.l1_loss) with fused gradient accumulation "
                       "(accumulate_gradients_into) on the same CUDA streams as `value`; per view: pinned GT image and "
                       "camera H2D (copy stream, ring of streams+1 slots, inside the timed region), loss D2H",
               "with_reference_training_loss": {"value": frames / (ms_e2e_full / 1e3), "unit": UNIT, "what": "same, with "
                   "the loss of GSLightningModule.calculate_loss: 0.8 L1 + 0.2 (1 - fused_ssim) (training.photometric_loss)"}}
        # ---- the training-step neighbours of the path, each one launch, HBM bound --------
        from gaussian_splatting_lightning_b200.training import FusedAdam, densify_fold

        def ev_ms(fn, reps=20):
            for _ in range(3):
                fn()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); a_.record()
            for _ in range(reps):
                fn()
            b_.record(); torch.cuda.synchronize()
            return a_.elapsed_time(b_) / reps

        hbm_pk = peaks()[0]
        opt = FusedAdam([{"params": [params[k]], "lr": lr} for k, lr in
                         (("means3D", 1.6e-4), ("shs", 2.5e-3), ("opacities", 5e-2), ("scales", 5e-3), ("rotations", 1e-3))],
                        lr=0.0, eps=1e-15)
        n_par = sum(p_.numel() for p_ in params.values())
        img_a, img_b = torch.rand(3, H, W, device=device, requires_grad=True), torch.rand(3, H, W, device=device)
        run_stats = [torch.zeros(P, device=device) for _ in range(3)]
        neighbours = []
        for name, fn, nbytes in (
                ("adam (6 groups, one launch, zero_grad fused)", lambda: opt.step(zero_grad=True), 32 * n_par),
                ("l1_loss value + unit gradient", lambda: l1_loss(img_a, img_b), 12 * 3 * H * W),
                ("densify_fold", lambda: densify_fold(bufs.max_radii, bufs.stats[0], bufs.stats[1], *run_stats), 32 * P)):
            ms_ = ev_ms(fn)
            neighbours.append({"kernel": name, "ms": ms_, "algorithmic_bytes": nbytes, "GBps": nbytes / (ms_ * 1e6),
                               "hbm_frac": nbytes / (ms_ * 1e6) / hbm_pk})
        # fused SSIM (value + gradient, one launch) next to the reference's own CUDA kernels (oracle/_ref, compiled from
        # /root/reference/submodules/fused-ssim by oracle/build_ref.py) on the same image pair: fusedssim + fusedssim_backward
        from gaussian_splatting_lightning_b200.training import fused_ssim
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
        neighbours.append(ent)
        e2e["neighbours"] = neighbours
        del params, opt

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-stage roofline ------------------------------------------------------------------------
    hbm_peak, peak_src, sm_max = peaks()
    T = ((W + 15) // 16) * ((H + 15) // 16)
    passes = (32 + max(1, (T - 1).bit_length()) + 7) // 8
    model = stage_model(P, M, counters["Pvis"], counters["R"], counters["L"], counters["A"], H * W, T, passes)
    stages, stage_sum = [], 0.0
    for name, (tot_ms, calls) in prof.items():
        if calls == 0:
            continue
        avg_ms = tot_ms / calls
        stage_sum += tot_ms
        m = model[name]
        s = {"stage": name, "ms_per_frame": avg_ms, "calls": calls, "algorithmic_bytes": m["bytes"],
             "GBps": m["bytes"] / (avg_ms * 1e6), "hbm_frac": m["bytes"] / (avg_ms * 1e6) / hbm_peak}
        if "flops" in m:
            s["TFLOPs_fp32"] = m["flops"] / (avg_ms * 1e9)
            s["Gex2_per_s"] = m["ex2"] / (avg_ms * 1e6)
        stages.append(s)
    for s in stages:
        s["share_of_kernel_time"] = s["ms_per_frame"] * s["calls"] / stage_sum
    dom = max(stages, key=lambda s: s["ms_per_frame"] * s["calls"])
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "r1_v7_traffic.json")      # dram bytes per launch from the ncu --set full capture
    if os.path.exists(tp):
        tj = json.load(open(tp))
        ent = tj.get(dom["stage"] + "_kernel") or tj.get({"sort": "onesweep_pass_kernel"}.get(dom["stage"], ""))
        if ent:
            traffic, traffic_src = ent["dram_bytes_per_launch"], "profiles/r1_v7_traffic.json: " + tj.get("_source", "")
    roofline = {"kernel": dom["stage"], "bound": "hbm", "achieved": dom["GBps"], "peak": hbm_peak, "unit": "GB/s",
                "frac": dom["hbm_frac"], "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": dom["algorithmic_bytes"], "ms_per_launch": dom["ms_per_frame"],
                "note": "the compositing kernels are FP32-issue bound, not HBM bound: see stages[].TFLOPs_fp32"}
    fp32_peak = 148 * 128 * 2 * (clocks["sm_mhz"] or sm_max) * 1e6 / 1e12 if clocks else None
    if "TFLOPs_fp32" in dom and fp32_peak:
        roofline["fp32"] = {"achieved_TFLOPs": dom["TFLOPs_fp32"], "peak_TFLOPs_at_observed_clock": fp32_peak,
                            "frac": dom["TFLOPs_fp32"] / fp32_peak}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_baseline = cpu_oracle_frames_per_s(args.workload, scene_cpu, cams[0]["host"], 6)   # ~11 s of CPU work

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {P} Gaussians, SH degree {deg}, {W}x{H}, fwd+bwd + densification "
                               f"stats, {vpr} views per GPU per step" + (", NCCL all-reduce of 59 floats/Gaussian + stats" if world > 1 else ""),
                   "views_per_rank": vpr, "frames_per_step": vpr * world, "streams_per_rank": args.streams, "parallelism": f"view-parallel x{world}", "partition": args.partition,
                   "l2": (lambda mb: f"inputs ({mb:.0f} MB of Gaussian parameters + {48 * P / 1e6:.0f} MB records) "
                          + ("exceed the 126 MB L2; no flush" if mb > 126 else
                             "fit the 126 MB L2: a parity-test size, not a bench line"))((44 + 12 * M) * P / 1e6),
                   "num_rendered_mean": counters["R"], "visible_mean": counters["Pvis"],
                   "pairs_A_mean": counters["A"], "list_entries_L_mean": counters["L"]},
        "forward_only": {"value": fwd_value, "unit": UNIT, "ms_per_step": ms_fwd / K},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "stages": stages,
        "cpu_baseline": cpu_baseline,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


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
