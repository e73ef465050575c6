"""View-parallel training step of the rasterizer hot path.

The path shards by camera view: every rank holds the full (replicated) Gaussian set, renders its own
views of the optimiser step and accumulates their gradients -- straight from the preprocess-backward
kernel, GSR_GRADS_ACCUMULATE -- into ONE flat fp32 buffer laid out for a single all-reduce:

    flat = [ dL_dmeans3D (P,3) | dL_dsh (P,M,3) | dL_dopacity (P,1) | dL_dscales (P,3) | dL_drotations (P,4) ]
           = 59 floats per Gaussian at SH degree 3, every section padded to a 128-byte boundary

plus the step's densification-statistics deltas, taken per view BEFORE the views are summed
(gs_lightning/modules/gaussian_model.py:174-181, gs_lightning/lightning/gs_lightning_module.py:192-195):

    stats[0] = sum over views of || dL_dmeans2D[:, :2] ||  where radii > 0     (xyz_grad_accum delta)
    stats[1] = number of views with radii > 0                                   (xyz_grad_count delta)
    max_radii = max over views of radii                                         (max_radii2D candidate)

One exchange per step: ONE all_reduce(sum) of `flat` + `stats` (adjacent sections of one storage), all_reduce(max) of
`max_radii`.  Every rank then applies the identical optimiser / densification update, so the
replicas stay bit-identical.  There is no collective inside the render path.

Only torch + torch.distributed here (works on any device and backend: NCCL on the GPUs, gloo in the
CPU tests of the exchange); rendering itself is CUDA-only (rasterizer.py).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

GRAD_FIELDS = ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations")


def partition_views(num_views: int, world_size: int, rank: int) -> List[int]:
    """Views of one optimiser step owned by `rank`: {k : k mod world_size == rank}."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of size {world_size}")
    return list(range(rank, num_views, world_size))


class StepBuffers:
    """The per-step exchange buffers of one rank.

    ONE storage [grad sections | stats | max_radii]; EVERY section starts on a 128-byte boundary whatever P is
    (after the first densify / prune P is almost never a multiple of 4, and gsr_backward needs a 16-byte aligned
    dL_drotations and takes the bulk-reduction path for dL_dsh only when that block is 16-byte aligned).  The
    padding floats stay zero and travel with the all-reduce (< 128 bytes per section).
    """

    ALIGN = 32   # floats: 128 bytes

    def __init__(self, num_gaussians: int, sh_coeffs: int, device, dtype=torch.float32):
        P, M = int(num_gaussians), int(sh_coeffs)
        self.P, self.M = P, M
        self.widths = dict(dL_dmeans3D=3, dL_dsh=3 * M, dL_dopacity=1, dL_dscales=3, dL_drotations=4)
        self.floats_per_gaussian = sum(self.widths.values())
        pad = lambda n: (n + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.offsets: Dict[str, int] = {}
        o = 0
        for name, w in self.widths.items():
            self.offsets[name] = o
            o += pad(P * w)
        n_flat, n_stats = o, 2 * pad(P)
        self._store = torch.zeros(n_flat + n_stats + pad(P), device=device, dtype=dtype)
        self._summed = self._store[:n_flat + n_stats]
        self.flat = self._store[:n_flat]                         # the gradient sections incl. their padding
        self._stats2 = self._store[n_flat:n_flat + n_stats].view(2, pad(P))
        self.stats = self._stats2[:, :P]                         # rows start 128-byte aligned
        self.max_radii = self._store[n_flat + n_stats:n_flat + n_stats + P]
        self._views: Dict[str, torch.Tensor] = {}
        for name, w in self.widths.items():
            shape = (P, M, 3) if name == "dL_dsh" else (P, w)
            self._views[name] = self.flat[self.offsets[name]:self.offsets[name] + P * w].view(shape)

    # -- what the rasterizer backward writes into -------------------------------------------------
    def grads(self) -> Dict[str, torch.Tensor]:
        """name -> view into `flat`; pass as backward_raw(..., accumulate_into=...)."""
        return dict(self._views)

    def densify_stats(self) -> Dict[str, torch.Tensor]:
        """pass as backward_raw(..., stats=...)."""
        return dict(max_radii2D=self.max_radii, xyz_grad_accum=self.stats[0], xyz_grad_count=self.stats[1])

    def zero_(self, keep_max_radii: bool = False, stats_only: bool = False) -> None:
        """Clear the step.  keep_max_radii: leave the radii candidates alone (deferred max exchange, see all_reduce).
        stats_only: leave the gradient sections alone too -- the fused preprocess backward of the whole step overwrites
        every row of them (preprocess_backward_views(..., overwrite=True))."""
        if stats_only:
            self._stats2.zero_()
            if not keep_max_radii:
                self.max_radii.zero_()
        else:
            (self._summed if keep_max_radii else self._store).zero_()

    # -- the one exchange step --------------------------------------------------------------------
    def all_reduce(self, group=None, async_op: bool = False, max_radii: bool = True):
        """Sum gradients and statistics deltas (ONE collective over adjacent sections) and max the radii over the
        ranks of `group`.

        max_radii=False defers the second collective: max commutes with the steps, so a rank may keep folding its OWN
        views' radii into `max_radii` step after step (zero_(keep_max_radii=True)) and call all_reduce_max_radii()
        only when densification reads max_radii2D (every `densify_interval` steps in the reference,
        gs_lightning_module.py:197-214) -- the result is the same maximum."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return []
        works = [dist.all_reduce(self._summed, op=dist.ReduceOp.SUM, group=group, async_op=async_op)]
        if max_radii:
            works.append(dist.all_reduce(self.max_radii, op=dist.ReduceOp.MAX, group=group, async_op=async_op))
        return [w for w in works if w is not None]

    def all_reduce_max_radii(self, group=None, async_op: bool = False):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.max_radii, op=dist.ReduceOp.MAX, group=group, async_op=async_op)

    def exchanged_bytes(self, max_radii: bool = True) -> int:
        return (self._summed.numel() + (self.max_radii.numel() if max_radii else 0)) * self.flat.element_size()

    # -- consumers ----------------------------------------------------------------------------------
    def apply_densify_stats(self, max_radii2D: torch.Tensor, xyz_grad_accum: torch.Tensor,
                            xyz_grad_count: torch.Tensor) -> None:
        """Fold the (all-reduced) deltas of this step into the model's running statistics, the way
        GaussianModel.update_max_radii2D / update_xyz_gradient do view by view."""
        seen = self.stats[1] > 0
        max_radii2D[seen] = torch.maximum(max_radii2D[seen], self.max_radii[seen].to(max_radii2D.dtype))
        xyz_grad_accum.add_(self.stats[0].to(xyz_grad_accum.dtype).view_as(xyz_grad_accum))
        xyz_grad_count.add_(self.stats[1].to(xyz_grad_count.dtype).view_as(xyz_grad_count))


class ViewStreams:
    """Round-robin CUDA streams for the independent views of one step.

    A frame is a chain of dependent kernels of which the binning stages (scan, key duplication, six
    sort passes, tile ranges) are latency-bound and leave most issue slots idle; running the views of a
    step on 2-3 streams lets those stages of one view overlap the issue-bound compositing kernels of
    another (measured on B200, C3: 470 -> 502 frames/s fwd+bwd, 1107 -> 1231 fwd with 3 streams).
    The gradient accumulation is an L2 reduction, so concurrent views may share the step's buffers.
    """

    def __init__(self, device, num_streams: int = 3):
        self.device = torch.device(device)
        self.side = [torch.cuda.Stream(self.device) for _ in range(max(1, int(num_streams)) - 1)]

    def run(self, num_views: int, fn) -> list:
        """fn(k) for k in range(num_views), view k on stream k mod n; joins back into the current stream."""
        main = torch.cuda.current_stream(self.device)
        if not self.side:
            return [fn(k) for k in range(num_views)]
        fork = torch.cuda.Event()
        fork.record(main)
        for st in self.side:
            st.wait_event(fork)
        n = len(self.side) + 1
        out = []
        for k in range(num_views):
            st = main if k % n == 0 else self.side[k % n - 1]
            with torch.cuda.stream(st):
                out.append(fn(k))
        for st in self.side:
            join = torch.cuda.Event()
            join.record(st)
            main.wait_event(join)
        # what fn returned was allocated under the side streams' pools and is consumed on `main` from here on: tell
        # the caching allocator, or a block freed while main-stream work still reads it could be reissued to a side stream
        for k, o in enumerate(out):
            if k % n:
                for t in (o if isinstance(o, (list, tuple)) else (o,)):
                    if isinstance(t, torch.Tensor) and t.is_cuda:
                        t.record_stream(main)
        return out


def render_step(settings_per_view: Sequence, scene: Dict[str, torch.Tensor], dL_dcolor, dL_dinvdepth,
                buffers: StepBuffers, group=None, exchange: bool = True,
                streams: Optional[ViewStreams] = None, fused: bool = True) -> List[torch.Tensor]:
    """Forward + backward of this rank's views into `buffers`, then the exchange.

    settings_per_view: GaussianRasterizationSettings of the views this rank owns;
    dL_dcolor / dL_dinvdepth: callables view_index, color, invdepth -> gradient tensors (or fixed tensors).
    streams: spread the views over CUDA streams (ViewStreams); None = the current stream only.
    fused (default): every view runs forward + the compositing backward; the preprocess backward of ALL views then
    runs as one pass over the Gaussians (rasterizer.preprocess_backward_views: inputs read once, the step's gradient row
    written once).  fused=False: view by view, each view's preprocess backward adding into the buffers.
    Returns the rendered colour images.
    """
    from .rasterizer import backward_raw, backward_render_only, forward_raw, preprocess_backward_views
    buffers.zero_(stats_only=fused and len(settings_per_view) > 0)
    grads, stats = buffers.grads(), buffers.densify_stats()
    states = [None] * len(settings_per_view)

    def one_view(k):
        rs = settings_per_view[k]
        R, color, radii, geom, binning, img, invd, _cap = forward_raw(
            rs, scene["means3D"], scene["shs"], None, scene["opacities"], scene["scales"], scene["rotations"], None)
        gC = dL_dcolor(k, color, invd) if callable(dL_dcolor) else dL_dcolor
        gD = dL_dinvdepth(k, color, invd) if callable(dL_dinvdepth) else dL_dinvdepth
        if fused:
            scratch = backward_render_only(rs, scene["means3D"], radii, scene["shs"], None, scene["opacities"],
                                           scene["scales"], scene["rotations"], None, gC, gD, geom, binning, img, R)
            states[k] = (rs, radii, geom, scratch)
        else:
            backward_raw(rs, scene["means3D"], radii, scene["shs"], None, scene["opacities"], scene["scales"],
                         scene["rotations"], None, gC, gD, geom, binning, img, R, accumulate_into=grads, stats=stats)
        return color

    if streams is None:
        images = [one_view(k) for k in range(len(settings_per_view))]
    else:
        images = streams.run(len(settings_per_view), one_view)
    if fused and states:
        if streams is not None:   # the views' buffers were allocated on the side streams and are read on this one
            cur = torch.cuda.current_stream(scene["means3D"].device)
            for _rs, radii, geom, scratch in states:
                for t in (radii, geom, scratch):
                    t.record_stream(cur)
        preprocess_backward_views(states, scene["means3D"], scene["shs"], None, scene["opacities"], scene["scales"],
                                  scene["rotations"], None, accumulate_into=grads, stats=stats, overwrite=True)
    if exchange:
        buffers.all_reduce(group)
    return images
