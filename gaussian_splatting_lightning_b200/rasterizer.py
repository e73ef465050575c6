"""Drop-in replacement of the `diff_gaussian_rasterization` Python package the reference imports
(gs_lightning/lightning/gs_lightning_module.py:8-9, scripts/render_trained_image.py:10-11,
tests/rasterizer_python/test_cases.py:2, tests/rasterizer_python/test_mark_visible.py:2).

Same names, argument meaning and error behaviour as the package those call sites use:
GaussianRasterizationSettings (13 fields, test_cases.py:7-18), GaussianRasterizer
(.forward / .markVisible), rasterize_gaussians and the autograd Function _RasterizeGaussians
returning (color, radii, invdepth) (gs_lightning_module.py:339, render_trained_image.py:115).
The compute is libgsr_b200.so (hand-written sm_100a CUDA) reached through _native.py.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
import os
import threading
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _native as N


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


# Last num_rendered per (device, W, H): sizes the binning buffer of the next frame up front so
# that the frame is enqueued without a device->host round trip (gsr_forward's binning_hint).
_hint = {}
_recent = {}
# How often the hint was too small and gsr_forward re-ran binning + compositing (bench.py reports it per region).
counters = {"forward_calls": 0, "hint_retries": 0}


def _bucket(n: int) -> int:
    """Binning capacity rounded up to 4-8 steps per octave.  Every view has its own R; sizing the binning buffer to
    R * 1.125 exactly made every request a different size, which the caching allocator answers with splits and, in
    the end, new cudaMalloc calls inside steady-state steps (measured: reserved memory grew from 18.4 to 20.6 GB over
    200 C3 steps and the regions that saw cudaMalloc calls ran at 13.4 ms per step instead of 12.1 --
    tools/diag_regions.py).  Bucketed sizes repeat, so freed blocks are reused as they are."""
    if n <= 65536:
        return 65536
    step = 1 << (n.bit_length() - 3)
    return (n + step - 1) // step * step


def _sync_mode() -> bool:
    return os.environ.get("GSR_SYNC_NUM_RENDERED", "0") == "1"


class _Frame:
    """Device tensors + ctypes structs of one call; keeps every tensor alive while the C call runs."""

    def __init__(self, rs: GaussianRasterizationSettings, means3D, sh, colors_precomp, opacities, scales,
                 rotations, cov3Ds_precomp):
        if means3D.dim() != 2 or means3D.shape[1] != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        if not means3D.is_cuda:
            raise RuntimeError("gaussian_splatting_lightning_b200 rasterizes on CUDA devices only (no CPU fallback)")
        dev = means3D.device
        self.device = dev
        self.P = int(means3D.shape[0])
        self.means3D = N.f32c(means3D) if self.P else means3D
        self.sh = N.f32c(sh)
        self.colors = N.f32c(colors_precomp)
        self.opacities = N.f32c(opacities)
        self.scales = N.f32c(scales)
        self.rotations = N.f32c(rotations)
        self.cov3D = N.f32c(cov3Ds_precomp)
        self.M = int(self.sh.shape[1]) if self.sh is not None else 0
        self.bg = N.f32c(rs.bg, dev)
        self.view = N.f32c(rs.viewmatrix, dev)
        self.proj = N.f32c(rs.projmatrix, dev)
        self.campos = N.f32c(rs.campos, dev)
        self.H, self.W = int(rs.image_height), int(rs.image_width)
        self.settings = N.GsrSettings(
            self.H, self.W, float(rs.tanfovx), float(rs.tanfovy), float(rs.scale_modifier), int(rs.sh_degree),
            int(bool(rs.prefiltered)), int(bool(rs.debug)), int(bool(rs.antialiasing)),
            N.ptr(self.bg), N.ptr(self.view), N.ptr(self.proj), N.ptr(self.campos))
        self.gaussians = N.GsrGaussians(
            self.P, self.M, N.ptr(self.means3D), N.ptr(self.opacities), N.ptr(self.sh), N.ptr(self.colors),
            N.ptr(self.scales), N.ptr(self.rotations), N.ptr(self.cov3D))


def _next_hint(key, binning_hint, R):
    counters["forward_calls"] += 1
    if 0 < binning_hint < R:
        counters["hint_retries"] += 1
    # the largest of the last 8 frames at this resolution: the views of a step differ in R, and a hint that is too
    # small costs a second binning + compositing pass while one that is too large costs only memory
    recent = _recent.setdefault(key, [])
    recent.append(_bucket(int(R * 1.125) + 4096))
    del recent[:-8]
    _hint[key] = max(recent)


def forward_raw(rs: GaussianRasterizationSettings, means3D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, binning_hint: Optional[int] = None):
    """_C.rasterize_gaussians: returns (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer,
    invdepths, binning_capacity)."""
    ext = N.ext()
    if ext is not None:
        if not means3D.is_cuda:
            raise RuntimeError("gaussian_splatting_lightning_b200 rasterizes on CUDA devices only (no CPU fallback)")
        H, W = int(rs.image_height), int(rs.image_width)
        key = (means3D.device.index, W, H)
        if binning_hint is None:
            binning_hint = 0 if _sync_mode() else _hint.get(key, 0)
        try:
            R, color, radii, geom, binning, img, invdepth, cap = ext.rasterize_gaussians_ex(
                rs.bg, means3D, colors_precomp, opacities, scales, rotations, float(rs.scale_modifier), cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, float(rs.tanfovx), float(rs.tanfovy), H, W, sh, int(rs.sh_degree),
                rs.campos, bool(rs.prefiltered), bool(rs.antialiasing), bool(rs.debug), int(binning_hint))
        except RuntimeError as e:
            if "gsr_" in str(e):
                raise N.GsrError(str(e)) from None
            raise
        _next_hint(key, binning_hint, R)
        return R, color, radii, geom, binning, img, invdepth, cap
    fr = _Frame(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
    dev = fr.device
    with torch.cuda.device(dev):
        color = torch.empty((3, fr.H, fr.W), dtype=torch.float32, device=dev)
        invdepth = torch.empty((1, fr.H, fr.W), dtype=torch.float32, device=dev)
        radii = torch.empty((fr.P,), dtype=torch.int32, device=dev)
        alloc = N.Allocator(dev)
        key = (dev.index, fr.W, fr.H)
        if binning_hint is None:
            binning_hint = 0 if _sync_mode() else _hint.get(key, 0)
        R = C.c_int64(0)
        cap = C.c_int64(0)
        rc = N.lib().gsr_forward(C.byref(fr.settings), C.byref(fr.gaussians), color.data_ptr(),
                                 N.ptr(radii), invdepth.data_ptr(), alloc.cfunc, None,
                                 int(binning_hint), C.byref(R), C.byref(cap), N.stream_ptr(dev))
        if alloc.error is not None:
            raise alloc.error
        N.check(rc, "gsr_forward")
        _next_hint(key, binning_hint, R.value)
    return (int(R.value), color, radii, alloc.buffers[N.GSR_BUF_GEOM], alloc.buffers[N.GSR_BUF_BINNING],
            alloc.buffers[N.GSR_BUF_IMG], invdepth, int(cap.value))


_GRAD_NAMES = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
               "dL_drotations")


def backward_raw(rs: GaussianRasterizationSettings, means3D, radii, sh, colors_precomp, opacities, scales,
                 rotations, cov3Ds_precomp, dL_dcolor, dL_dinvdepth, geomBuffer, binningBuffer, imgBuffer,
                 num_rendered: int, want=None, accumulate_into=None, stats=None):
    """_C.rasterize_gaussians_backward: returns (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D,
    dL_dcov3D, dL_dsh, dL_dscales, dL_drotations).

    Extensions of the C-ABI (include/gsr_b200.h), all optional:
      want            names of the gradients to produce (others come back as None, not computed);
      accumulate_into dict name -> tensor: those gradients are ADDED into the given tensors
                      (GSR_GRADS_ACCUMULATE: several views into one flat gradient buffer);
      stats           dict with any of max_radii2D / xyz_grad_accum / xyz_grad_count, (P,) float32:
                      the view's densification statistics (gaussian_model.py:174-181), fused.
    """
    ext = N.ext()
    if ext is not None:
        if accumulate_into is not None:
            acc, mask = [accumulate_into.get(n) for n in _GRAD_NAMES], 0
        else:
            names = _GRAD_NAMES if want is None else tuple(want)
            acc, mask = [], sum(1 << i for i, n in enumerate(_GRAD_NAMES) if n in names)
        st = [stats.get("max_radii2D"), stats.get("xyz_grad_accum"), stats.get("xyz_grad_count")] if stats else []
        if dL_dcolor is None and dL_dinvdepth is None:
            dL_dcolor = torch.zeros((3, int(rs.image_height), int(rs.image_width)), dtype=torch.float32,
                                    device=means3D.device)
        try:
            return tuple(ext.rasterize_gaussians_backward_ex(
                rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, float(rs.scale_modifier),
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, float(rs.tanfovx), float(rs.tanfovy), dL_dcolor,
                dL_dinvdepth, sh, int(rs.sh_degree), rs.campos, geomBuffer, int(num_rendered), binningBuffer, imgBuffer,
                bool(rs.antialiasing), bool(rs.debug), mask, acc, st))
        except RuntimeError as e:
            if "gsr_" in str(e):
                raise N.GsrError(str(e)) from None
            raise
    fr = _Frame(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
    dev, P, M = fr.device, fr.P, fr.M
    shapes = dict(dL_dmeans2D=(P, 3), dL_dcolors=(P, 3), dL_dopacity=(P, 1), dL_dmeans3D=(P, 3), dL_dcov3D=(P, 6),
                  dL_dsh=(P, M, 3), dL_dscales=(P, 3), dL_drotations=(P, 4))
    with torch.cuda.device(dev):
        if accumulate_into is not None:
            g = {n: accumulate_into.get(n) for n in _GRAD_NAMES}
            for n, t in g.items():
                if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
                                          and t.numel() == math.prod(shapes[n])):
                    raise RuntimeError(f"accumulate_into[{n!r}] must be a contiguous float32 CUDA tensor of shape {shapes[n]}")
        else:
            names = _GRAD_NAMES if want is None else tuple(want)
            g = {n: (torch.empty(shapes[n], dtype=torch.float32, device=dev) if n in names else None)
                 for n in _GRAD_NAMES}
        if P == 0:
            return tuple(g.values())
        dL_dcolor = N.f32c(dL_dcolor, dev)
        if dL_dcolor is None:
            dL_dcolor = torch.zeros((3, fr.H, fr.W), dtype=torch.float32, device=dev)
        dL_dinvdepth = N.f32c(dL_dinvdepth, dev)
        scratch = torch.empty(N.lib().gsr_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        grads = N.GsrGrads(*[N.ptr(t) for t in g.values()])
        st = None
        if stats:
            for k, t in stats.items():
                if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == P):
                    raise RuntimeError(f"stats[{k!r}] must be a contiguous float32 CUDA tensor of {P} elements")
            st = N.GsrDensifyStats(N.ptr(stats.get("max_radii2D")), N.ptr(stats.get("xyz_grad_accum")),
                                   N.ptr(stats.get("xyz_grad_count")))
        rc = N.lib().gsr_backward(C.byref(fr.settings), C.byref(fr.gaussians), N.ptr(radii), N.ptr(dL_dcolor),
                                  N.ptr(dL_dinvdepth), N.ptr(geomBuffer), N.ptr(binningBuffer), N.ptr(imgBuffer),
                                  int(num_rendered), N.ptr(scratch), C.byref(grads),
                                  N.GSR_GRADS_ACCUMULATE if accumulate_into is not None else 0,
                                  C.byref(st) if st is not None else None, N.stream_ptr(dev))
        N.check(rc, "gsr_backward")
    return tuple(g.values())


def backward_render_only(rs: GaussianRasterizationSettings, means3D, radii, sh, colors_precomp, opacities, scales,
                         rotations, cov3Ds_precomp, dL_dcolor, dL_dinvdepth, geomBuffer, binningBuffer, imgBuffer,
                         num_rendered: int) -> torch.Tensor:
    """The compositing backward of one view alone (gsr_backward with GSR_BWD_RENDER_ONLY).  Returns the view's scratch
    buffer -- (P,12) raw pixel sums per Gaussian -- to hand to preprocess_backward_views together with the view's
    settings, radii and geometry buffer."""
    ext = N.ext()
    if dL_dcolor is None and dL_dinvdepth is None:
        dL_dcolor = torch.zeros((3, int(rs.image_height), int(rs.image_width)), dtype=torch.float32, device=means3D.device)
    if ext is not None:
        try:
            return ext.render_backward_only(
                rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, float(rs.scale_modifier),
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, float(rs.tanfovx), float(rs.tanfovy), dL_dcolor,
                dL_dinvdepth, sh, int(rs.sh_degree), rs.campos, geomBuffer, int(num_rendered), binningBuffer, imgBuffer,
                bool(rs.antialiasing), bool(rs.debug))
        except RuntimeError as e:
            if "gsr_" in str(e):
                raise N.GsrError(str(e)) from None
            raise
    fr = _Frame(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
    dev, P = fr.device, fr.P
    with torch.cuda.device(dev):
        scratch = torch.empty(N.lib().gsr_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        if P == 0:
            return scratch
        dC = N.f32c(dL_dcolor, dev)
        if dC is None:
            dC = torch.zeros((3, fr.H, fr.W), dtype=torch.float32, device=dev)
        dD = N.f32c(dL_dinvdepth, dev)
        N.check(N.lib().gsr_backward(C.byref(fr.settings), C.byref(fr.gaussians), N.ptr(radii), N.ptr(dC), N.ptr(dD),
                                     N.ptr(geomBuffer), N.ptr(binningBuffer), N.ptr(imgBuffer), int(num_rendered),
                                     N.ptr(scratch), None, N.GSR_BWD_RENDER_ONLY, None, N.stream_ptr(dev)), "gsr_backward")
    return scratch


def preprocess_backward_views(views, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                              want=None, accumulate_into=None, stats=None, overwrite=False, first=0, count=None):
    """gsr_preprocess_backward_views: the preprocess backward of ALL views of a training step in one pass over the
    Gaussians.  views: sequence of (raster_settings, radii, geomBuffer, scratch), scratch from backward_render_only.
    Returns the 8-tuple of backward_raw with dL_dmeans2D = None (a per-view output); `want`, `accumulate_into` and
    `stats` as in backward_raw; overwrite=True WRITES the tensors of accumulate_into (every row; zeros where no view
    saw the Gaussian) instead of adding to them, so a step whose views all go through one call needs no cleared buffer
    and no reductions at the L2.  first / count restrict the pass to the Gaussians [first, first + count): outputs,
    accumulate_into and stats then hold `count` rows (view_parallel.StepBuffers(chunks=...) exchanges range by range).  The result equals the sum of per-view backward_raw calls up to float summation order;
    the densification statistics are the per-view ones (norm taken per view, then summed)."""
    views = list(views)
    if not views:
        raise ValueError("preprocess_backward_views: no views")
    rs0 = views[0][0]
    ext = N.ext()
    if accumulate_into is not None:
        acc = [None if n == "dL_dmeans2D" else accumulate_into.get(n) for n in _GRAD_NAMES]
        mask = 0
    else:
        names = tuple(n for n in (_GRAD_NAMES if want is None else want) if n != "dL_dmeans2D")
        acc, mask = [], sum(1 << i for i, n in enumerate(_GRAD_NAMES) if n in names)
    st = [stats.get("max_radii2D"), stats.get("xyz_grad_accum"), stats.get("xyz_grad_count")] if stats else []
    if ext is not None:
        vt = [(rs.bg, rs.viewmatrix, rs.projmatrix, float(rs.tanfovx), float(rs.tanfovy), int(rs.image_height),
               int(rs.image_width), rs.campos, radii, geom, scratch) for rs, radii, geom, scratch in views]
        try:
            return tuple(ext.preprocess_backward_views(
                vt, means3D, colors_precomp, opacities, scales, rotations, float(rs0.scale_modifier), cov3Ds_precomp, sh,
                int(rs0.sh_degree), bool(rs0.antialiasing), bool(rs0.debug), mask, acc, st, bool(overwrite), int(first),
                -1 if count is None else int(count)))
        except RuntimeError as e:
            if "gsr_" in str(e):
                raise N.GsrError(str(e)) from None
            raise
    frames = [_Frame(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp) for rs, *_ in views]
    fr = frames[0]
    dev, Pall, M = fr.device, fr.P, fr.M
    P = Pall - first if count is None else int(count)
    shapes = dict(dL_dmeans2D=(P, 3), dL_dcolors=(P, 3), dL_dopacity=(P, 1), dL_dmeans3D=(P, 3), dL_dcov3D=(P, 6),
                  dL_dsh=(P, M, 3), dL_dscales=(P, 3), dL_drotations=(P, 4))
    with torch.cuda.device(dev):
        if accumulate_into is not None:
            g = dict(zip(_GRAD_NAMES, acc))
        else:
            g = {n: (torch.empty(shapes[n], dtype=torch.float32, device=dev) if mask >> i & 1 else None)
                 for i, n in enumerate(_GRAD_NAMES)}
        if P == 0:
            return tuple(g.values())
        arr = (N.GsrViewState * len(views))(*[
            N.GsrViewState(C.pointer(f.settings), N.ptr(radii), N.ptr(geom), N.ptr(scratch))
            for f, (_rs, radii, geom, scratch) in zip(frames, views)])
        grads = N.GsrGrads(*[N.ptr(t) for t in g.values()])
        ds = N.GsrDensifyStats(*[N.ptr(t) for t in st]) if st else None
        N.check(N.lib().gsr_preprocess_backward_views_range(
            arr, len(views), C.byref(fr.gaussians), C.byref(grads),
            N.GSR_GRADS_ACCUMULATE if (accumulate_into is not None and not overwrite) else 0,
            C.byref(ds) if ds is not None else None, int(first), P, N.stream_ptr(dev)), "gsr_preprocess_backward_views")
    return tuple(g.values())


def mark_visible(positions: torch.Tensor, viewmatrix: torch.Tensor, projmatrix: torch.Tensor = None) -> torch.Tensor:
    """_C.mark_visible (tests/rasterizer_python/test_mark_visible.py:13-14)."""
    if not positions.is_cuda:
        raise RuntimeError("mark_visible runs on CUDA devices only (no CPU fallback)")
    ext = N.ext()
    if ext is not None:
        return ext.mark_visible(positions, viewmatrix, projmatrix)
    dev = positions.device
    p = N.f32c(positions)
    P = int(positions.shape[0])
    present = torch.zeros((P,), dtype=torch.uint8, device=dev)
    if P:
        v = N.f32c(viewmatrix, dev)
        with torch.cuda.device(dev):
            N.check(N.lib().gsr_mark_visible(P, N.ptr(p), N.ptr(v), N.ptr(present), N.stream_ptr(dev)),
                    "gsr_mark_visible")
    return present.bool()


class _Sink(threading.local):
    grads = None
    stats = None
    step = None      # a StepSink: views of a step whose preprocess backward is deferred to ONE fused pass


class StepSink:
    """Collects the views of one training step for the fused preprocess backward (preprocess_backward_views).

        sink = StepSink(buffers.grads(), buffers.densify_stats())
        for view in views:
            with sink:                                    # = accumulate_gradients_into(...) with deferral
                color, radii, depth = rasterizer(...)     # parameters are LEAF tensors, means2D=None
            loss(color).backward()                        # runs the compositing backward only, parks the view here
        sink.flush()                                      # ONE pass over the Gaussians for all parked views

    A view is parked only if every gradient autograd wants from it is covered by the sink (leaf inputs, no dL_dmeans2D
    request: the dummy screen-space leaf of gs_lightning_module.py:316-317 exists for the densification statistics, which
    the fused pass produces itself); any other view takes the per-view path at once, adding into the same tensors."""

    def __init__(self, grads, stats=None):
        self.grads, self.stats = dict(grads), (dict(stats) if stats else None)
        self.pending = []        # (rs, radii, geom, scratch)
        self.inputs = None       # (means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
        self._prev = None

    def __enter__(self):
        self._prev = (_sink.grads, _sink.stats, _sink.step)
        _sink.grads, _sink.stats, _sink.step = self.grads, self.stats, self
        return self

    def __exit__(self, *exc):
        _sink.grads, _sink.stats, _sink.step = self._prev
        return False

    def flush(self, overwrite: bool = False):
        """Run the fused preprocess backward over the parked views (on the current stream) and forget them.
        overwrite=True writes the gradient tensors instead of adding to them (every view of the step was parked and the
        tensors were not cleared: preprocess_backward_views)."""
        if not self.pending:
            return
        views, self.pending = self.pending, []
        means3D, sh, colors, opac, scales, rot, cov = self.inputs
        cur = torch.cuda.current_stream(means3D.device)
        for _rs, radii, geom, scratch in views:      # possibly allocated under other streams' pools
            for t in (radii, geom, scratch):
                t.record_stream(cur)
        preprocess_backward_views(views, means3D, sh, colors, opac, scales, rot, cov,
                                  accumulate_into={n: t for n, t in self.grads.items() if n != "dL_dmeans2D"},
                                  stats=self.stats, overwrite=overwrite)


_sink = _Sink()


@contextlib.contextmanager
def accumulate_gradients_into(grads, stats=None):
    """Fused gradient accumulation for the autograd API.

    GaussianRasterizer calls made inside this context remember `grads` (name -> tensor, the names of
    backward_raw's `accumulate_into`, e.g. view_parallel.StepBuffers.grads()) and `stats`
    (StepBuffers.densify_stats()).  Their backward then ADDS the parameter gradients straight into
    those tensors inside the preprocess-backward kernel (GSR_GRADS_ACCUMULATE) and updates the
    densification statistics there, and hands autograd None for the inputs it covered -- so autograd
    neither allocates per-view gradient tensors nor runs its own accumulation kernels.  Only LEAF inputs
    are covered: an input that is the result of an activation (exp / sigmoid / normalize / cat, the way the
    unchanged GSLightningModule calls the rasterizer) gets its gradient through autograd as usual, so the
    chain to the raw parameters is never cut.  Point
    `param.grad` at the same tensors (training.parameter_groups_on) and the optimiser sees the sum over
    the views.  The dummy means2D leaf still receives its gradient through autograd when it requires one.
    """
    prev = (_sink.grads, _sink.stats)
    _sink.grads, _sink.stats = dict(grads), (dict(stats) if stats else None)
    try:
        yield
    finally:
        _sink.grads, _sink.stats = prev


_SINK_INPUT = dict(dL_dmeans3D=0, dL_dsh=2, dL_dcolors=3, dL_dopacity=4, dL_dscales=5, dL_drotations=6, dL_dcov3D=7)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        ctx.sink = (_sink.grads, _sink.stats) if _sink.grads is not None else None
        ctx.step_sink = _sink.step
        # The sink may take over the gradient of an input only when that input is a LEAF: for an activation
        # (the reference passes exp(_scaling), sigmoid(_opacity), normalize(_rotation), cat(dc, rest):
        # gs_lightning_module.py:337-348) autograd has to receive the gradient to carry it on to the parameters.
        ctx.leaf = tuple(t is None or t.is_leaf for t in (means3D, means2D, sh, colors_precomp, opacities, scales,
                                                          rotations, cov3Ds_precomp))
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths, _cap = forward_raw(
            raster_settings, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.has = tuple(t is not None and t.numel() > 0 for t in (sh, colors_precomp, scales, rotations, cov3Ds_precomp))
        z = means3D.new_empty(0)
        nz = lambda t: z if t is None else t
        ctx.save_for_backward(nz(colors_precomp), means3D, nz(scales), nz(rotations), nz(cov3Ds_precomp), radii,
                              nz(sh), opacities, geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)
        return color, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, grad_out_depth):
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, geomBuffer,
         binningBuffer, imgBuffer) = ctx.saved_tensors
        has_sh, has_col, has_scale, has_rot, has_cov = ctx.has
        # inputs in forward order: means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp
        need = ctx.needs_input_grad
        want = [n for n, ok in (("dL_dmeans3D", need[0]), ("dL_dmeans2D", need[1]), ("dL_dsh", has_sh and need[2]),
                                ("dL_dcolors", has_col and need[3]), ("dL_dopacity", need[4]),
                                ("dL_dscales", has_scale and need[5]), ("dL_drotations", has_rot and need[6]),
                                ("dL_dcov3D", has_cov and need[7])) if ok]
        if ctx.sink is not None:
            # fused accumulation (accumulate_gradients_into): wanted gradients the sink covers are added in
            # place by the kernel; autograd gets None for them.  Anything else wanted is produced fresh
            # (into a zeroed tensor, since the accumulate flag covers the whole call).
            sink_grads, sink_stats = ctx.sink
            P, M = means3D.shape[0], (sh.shape[1] if has_sh else 0)
            step = ctx.step_sink
            if step is not None and all(n in sink_grads and n != "dL_dmeans2D" and ctx.leaf[_SINK_INPUT[n]] for n in want):
                # park the view: compositing backward now, its preprocess backward in the step's ONE fused pass
                nz = lambda t, ok: t if ok else None
                inputs = (means3D, nz(sh, has_sh), nz(colors_precomp, has_col), opacities, nz(scales, has_scale),
                          nz(rotations, has_rot), nz(cov3Ds_precomp, has_cov))
                if step.inputs is None or not step.pending:
                    step.inputs = inputs
                same = lambda a, b: (a is None and b is None) or (a is not None and b is not None and
                                                               a.data_ptr() == b.data_ptr() and a.shape == b.shape)
                if all(same(a, b) for a, b in zip(step.inputs, inputs)):
                    scratch = backward_render_only(ctx.raster_settings, *inputs[:1], radii, *inputs[1:], grad_out_color,
                                                   grad_out_depth, geomBuffer, binningBuffer, imgBuffer, ctx.num_rendered)
                    step.pending.append((ctx.raster_settings, radii, geomBuffer, scratch))
                    return (None,) * 9
            shapes = dict(dL_dmeans2D=(P, 3), dL_dcolors=(P, 3), dL_dopacity=(P, 1), dL_dmeans3D=(P, 3),
                          dL_dcov3D=(P, 6), dL_dsh=(P, M, 3), dL_dscales=(P, 3), dL_drotations=(P, 4))
            acc, fresh = {}, {}
            for n in want:
                if n in sink_grads and ctx.leaf[1 if n == "dL_dmeans2D" else _SINK_INPUT[n]]:
                    acc[n] = sink_grads[n]
                else:
                    acc[n] = fresh[n] = torch.zeros(shapes[n], dtype=torch.float32, device=means3D.device)
            backward_raw(ctx.raster_settings, means3D, radii, sh, colors_precomp, opacities, scales, rotations,
                         cov3Ds_precomp, grad_out_color, grad_out_depth, geomBuffer, binningBuffer, imgBuffer,
                         ctx.num_rendered, accumulate_into=acc, stats=sink_stats)
            out = [None] * 9
            for n, t in fresh.items():
                out[1 if n == "dL_dmeans2D" else _SINK_INPUT[n]] = t
            return tuple(out)
        (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
         dL_drotations) = backward_raw(ctx.raster_settings, means3D, radii, sh, colors_precomp, opacities, scales,
                                       rotations, cov3Ds_precomp, grad_out_color, grad_out_depth, geomBuffer,
                                       binningBuffer, imgBuffer, ctx.num_rendered, want=want)
        return (dL_dmeans3D, dL_dmeans2D, dL_dsh if has_sh else None, dL_dcolors if has_col else None, dL_dopacity,
                dL_dscales if has_scale else None, dL_drotations if has_rot else None,
                dL_dcov3D if has_cov else None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if means2D is None:  # scripts/render_trained_image.py:117 passes None
            means2D = means3D.new_zeros(means3D.shape)
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)


def export_state(rs: GaussianRasterizationSettings, P: int, radii, geomBuffer, binningBuffer, imgBuffer,
                 num_rendered: int, binning_capacity: int = 0) -> dict:
    """Canonical-layout dump of the opaque scratch buffers as torch tensors."""
    dev = radii.device
    H, W = int(rs.image_height), int(rs.image_width)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    R = int(num_rendered)
    f = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
    out = dict(depths=f(P), means2D=f(P, 2), cov3D=f(P, 6), conic_opacity=f(P, 4), rgb=f(P, 3),
               clamped=torch.zeros((P, 3), dtype=torch.uint8, device=dev),
               tiles_touched=torch.zeros((P,), dtype=torch.int32, device=dev),
               point_offsets=torch.zeros((P,), dtype=torch.int32, device=dev),
               keys_unsorted=torch.zeros((R,), dtype=torch.int64, device=dev),
               vals_unsorted=torch.zeros((R,), dtype=torch.int32, device=dev),
               keys_sorted=torch.zeros((R,), dtype=torch.int64, device=dev),
               vals_sorted=torch.zeros((R,), dtype=torch.int32, device=dev),
               ranges=torch.zeros((T, 2), dtype=torch.int32, device=dev),
               final_T=f(H, W), n_contrib=torch.zeros((H, W), dtype=torch.int32, device=dev))
    bg, view, proj, campos = (N.f32c(t, dev) for t in (rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos))
    st = N.GsrSettings(H, W, float(rs.tanfovx), float(rs.tanfovy), float(rs.scale_modifier), int(rs.sh_degree), 0, 0,
                       int(bool(rs.antialiasing)), N.ptr(bg), N.ptr(view), N.ptr(proj), N.ptr(campos))
    ex = N.GsrStateExport(*[N.ptr(out[k]) for k in N._EXPORT_FIELDS])
    with torch.cuda.device(dev):
        N.check(N.lib().gsr_export_state(C.byref(st), P, N.ptr(radii), N.ptr(geomBuffer), N.ptr(binningBuffer),
                                         N.ptr(imgBuffer), R, int(binning_capacity), C.byref(ex),
                                         N.stream_ptr(dev)), "gsr_export_state")
        torch.cuda.synchronize(dev)
    return out


def sort_pairs(keys: torch.Tensor, vals: torch.Tensor, end_bit: int = 64):
    """Stand-alone entry to the onesweep radix sort: stable sort of (int64 key, int32 value) pairs on key
    bits [0, end_bit) (keys are treated as unsigned)."""
    assert keys.is_cuda and keys.dtype == torch.int64 and vals.dtype == torch.int32 and keys.numel() == vals.numel()
    dev, n = keys.device, keys.numel()
    keys, vals = keys.contiguous(), vals.contiguous()
    ko, vo, kt, vt = torch.empty_like(keys), torch.empty_like(vals), torch.empty_like(keys), torch.empty_like(vals)
    ws = torch.empty(N.lib().gsr_sort_workspace_bytes(n), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        N.check(N.lib().gsr_sort_pairs(N.ptr(keys), N.ptr(vals), N.ptr(ko), N.ptr(vo), N.ptr(kt), N.ptr(vt), n,
                                       int(end_bit), N.ptr(ws), N.stream_ptr(dev)), "gsr_sort_pairs")
    return ko, vo


def hit_word_stats(rs: GaussianRasterizationSettings, binningBuffer, n_contrib, ranges, binning_capacity: int) -> dict:
    """What the backward compositing kernel actually evaluates, counted from the forward's hit words (debug accessor,
    bench.py's roofline): per (warp, list entry) VISIT the kernel runs the packed two-pixel path when the entry reached
    both 8x4 sub-tiles of the warp's 8x8 block and the one-pixel path otherwise, so

        pairs_evaluated = 32 lanes * (2 * visits_both + visits_single)        (pixel, splat) evaluations,

    against A = sum(n_contrib), the convention of SURVEY.md 8(d) that also counts pairs the culling skipped.
    n_contrib (H,W) int32 and ranges (T,2) int32 as export_state returns them."""
    dev = n_contrib.device
    H, W = int(rs.image_height), int(rs.image_width)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    cap = int(binning_capacity)
    hits_off = 128 + (cap * 4 + 127) // 128 * 128                  # gsr_bin_hits_off (csrc/gsr_common.cuh)
    n_words = (cap // 32 + T + 2) * 8
    hits = binningBuffer[hits_off:hits_off + 4 * n_words].view(torch.int32).view(-1, 8)
    pad = torch.zeros((gy * 16, gx * 16), dtype=torch.int64, device=dev)
    pad[:H, :W] = n_contrib.to(torch.int64)
    # last contributor per 8x8 block = per warp; warp w of a tile owns block (w & 1, w >> 1)
    blk = pad.view(gy, 2, 8, gx, 2, 8).amax(dim=(2, 5))           # (gy, by, gx, bx)
    warp_last = blk.permute(0, 2, 1, 3).reshape(T, 4)             # [tile][warp = by * 2 + bx]
    nchunks = (warp_last + 31) // 32                               # chunks of 32 entries the backward walks
    slot0 = (ranges[:, 0].to(torch.int64) >> 5) + torch.arange(T, device=dev)
    both = single = 0
    for w in range(4):
        colA = (w >> 1) * 4 + (w & 1)
        n = nchunks[:, w]
        tile_of = torch.repeat_interleave(torch.arange(T, device=dev), n)
        first = torch.cumsum(n, 0) - n
        k = torch.arange(tile_of.numel(), device=dev) - first[tile_of]
        rows = hits[slot0[tile_of] + k]
        a, b = rows[:, colA], rows[:, colA + 2]
        pc = lambda x: sum(((x >> s) & 0x01010101) for s in range(8)).view(torch.uint8).to(torch.int64).sum()
        both += int(pc(a & b))
        single += int(pc(a ^ b))
    return dict(visits_both=both, visits_single=single, pairs_evaluated=32 * (2 * both + single))
