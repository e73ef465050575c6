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

    ONE storage; EVERY section starts on a 128-byte boundary whatever P is (after the first densify / prune P is almost
    never a multiple of 4, and gsr_backward needs a 16-byte aligned dL_drotations and takes the bulk-reduction path for
    dL_dsh only when that block is 16-byte aligned).  The padding floats stay zero and travel with the all-reduce
    (< 128 bytes per section).

    chunks = 1 (default):  [ grad sections of all P | stats (2, P) | max_radii (P) ]
    chunks = C:            [ chunk 0: grad sections of its Gaussians | its stats ] ... [ chunk C-1 ] [ max_radii (P) ]
    The chunk-major form lets the exchange follow the fused preprocess backward range by range: the pass over chunk c
    writes exactly chunk c's contiguous block, whose all-reduce then runs under the pass over chunk c + 1
    (fused_backward_and_exchange).  Whole-P views (`flat`, `stats`, `grads()` without a chunk) exist only for chunks = 1.
    """

    ALIGN = 32   # floats: 128 bytes

    def __init__(self, num_gaussians: int, sh_coeffs: int, device, dtype=torch.float32, chunks: int = 1):
        P, M = int(num_gaussians), int(sh_coeffs)
        self.P, self.M = P, M
        self.widths = dict(dL_dmeans3D=3, dL_dsh=3 * M, dL_dopacity=1, dL_dscales=3, dL_drotations=4)
        self.floats_per_gaussian = sum(self.widths.values())
        pad = lambda n: (n + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        C = max(1, min(int(chunks), max(1, P // 1024)))
        per = (P + C - 1) // C
        per = (per + 1023) // 1024 * 1024 if C > 1 else per        # chunk boundaries on multiples of 1024 Gaussians
        self.chunks: List[tuple] = [(f, min(per, P - f)) for f in range(0, max(P, 1), max(per, 1)) if f < P] or [(0, 0)]
        # layout
        lay, o = [], 0
        for f, n in self.chunks:
            sec = {}
            start = o
            for name, w in self.widths.items():
                sec[name] = o
                o += pad(n * w)
            sec["stats"] = o
            o += 2 * pad(n)
            lay.append((start, o, sec))
        n_summed = o
        self._store = torch.zeros(n_summed + pad(P), device=device, dtype=dtype)
        self._summed = self._store[:n_summed]
        self.max_radii = self._store[n_summed:n_summed + P]
        self._chunk_block = [self._store[a:b] for a, b, _ in lay]          # what chunk c's all-reduce sends
        self._chunk_views: List[Dict[str, torch.Tensor]] = []
        self._chunk_stats: List[torch.Tensor] = []
        for (f, n), (_a, _b, sec) in zip(self.chunks, lay):
            views = {}
            for name, w in self.widths.items():
                shape = (n, M, 3) if name == "dL_dsh" else (n, w)
                views[name] = self._store[sec[name]:sec[name] + n * w].view(shape)
            self._chunk_views.append(views)
            self._chunk_stats.append(self._store[sec["stats"]:sec["stats"] + 2 * pad(n)].view(2, pad(n))[:, :n])
        if len(self.chunks) == 1:
            a0, _b0, sec0 = lay[0]
            self.flat = self._store[a0:sec0["stats"]]               # the gradient sections incl. their padding
            self._stats2 = self._store[sec0["stats"]:n_summed].view(2, pad(P))
            self.stats = self._stats2[:, :P]                         # rows start 128-byte aligned
            self.offsets = {name: sec0[name] for name in self.widths}
            self._views = self._chunk_views[0]

    @property
    def num_chunks(self) -> int:
        return len(self.chunks)

    def _one(self, what):
        if len(self.chunks) != 1:
            raise ValueError(f"{what} without a chunk index needs StepBuffers(chunks=1); this one has {len(self.chunks)} chunks")

    # -- what the rasterizer backward writes into -------------------------------------------------
    def grads(self, chunk: Optional[int] = None) -> Dict[str, torch.Tensor]:
        """name -> view of the gradient rows (of chunk `chunk`); pass as backward_raw(..., accumulate_into=...)."""
        if chunk is None:
            self._one("grads()")
            chunk = 0
        return dict(self._chunk_views[chunk])

    def densify_stats(self, chunk: Optional[int] = None) -> Dict[str, torch.Tensor]:
        """pass as backward_raw(..., stats=...)."""
        if chunk is None:
            self._one("densify_stats()")
            chunk = 0
        f, n = self.chunks[chunk]
        st = self._chunk_stats[chunk]
        return dict(max_radii2D=self.max_radii[f:f + n], xyz_grad_accum=st[0], xyz_grad_count=st[1])

    def zero_(self, keep_max_radii: bool = False, stats_only: bool = False) -> None:
        """Clear the step.  keep_max_radii: leave the radii candidates alone (deferred max exchange, see all_reduce).
        stats_only: leave the gradient sections alone too -- the fused preprocess backward of the whole step overwrites
        every row of them (preprocess_backward_views(..., overwrite=True))."""
        if stats_only:
            for st in self._chunk_stats:
                st.zero_()
            if not keep_max_radii:
                self.max_radii.zero_()
        else:
            (self._summed if keep_max_radii else self._store).zero_()

    # -- the exchange -----------------------------------------------------------------------------
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

    def all_reduce_chunk(self, chunk: int, group=None, async_op: bool = True):
        """Sum chunk `chunk`'s gradients + statistics deltas over the ranks (its block is contiguous)."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self._chunk_block[chunk], op=dist.ReduceOp.SUM, group=group, async_op=async_op)

    def all_reduce_max_radii(self, group=None, async_op: bool = False):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.max_radii, op=dist.ReduceOp.MAX, group=group, async_op=async_op)

    def exchanged_bytes(self, max_radii: bool = True) -> int:
        return (self._summed.numel() + (self.max_radii.numel() if max_radii else 0)) * self._store.element_size()

    # -- consumers ----------------------------------------------------------------------------------
    def apply_densify_stats(self, max_radii2D: torch.Tensor, xyz_grad_accum: torch.Tensor,
                            xyz_grad_count: torch.Tensor) -> None:
        """Fold the (all-reduced) deltas of this step into the model's running statistics, the way
        GaussianModel.update_max_radii2D / update_xyz_gradient do view by view."""
        acc, cnt = xyz_grad_accum.view(-1), xyz_grad_count.view(-1)
        for (f, n), st in zip(self.chunks, self._chunk_stats):
            seen = st[1] > 0
            mr = max_radii2D[f:f + n]
            mr[seen] = torch.maximum(mr[seen], self.max_radii[f:f + n][seen].to(max_radii2D.dtype))
            acc[f:f + n].add_(st[0].to(acc.dtype))
            cnt[f:f + n].add_(st[1].to(cnt.dtype))


def fused_backward_and_exchange(states, scene: Dict[str, torch.Tensor], buffers: StepBuffers, group=None,
                                exchange: bool = True, max_radii: bool = True) -> None:
    """The end of a fused step: the preprocess backward of all parked views (`states`: (settings, radii, geom, scratch)
    per view, from rasterizer.backward_render_only) over the Gaussians, chunk by chunk of `buffers`, each chunk's
    all-reduce started as soon as its pass is enqueued -- it runs on NCCL's stream under the pass over the next chunk.
    The gradient rows are overwritten (no cleared buffer needed); the statistics are added to."""
    from .rasterizer import preprocess_backward_views
    dev = scene["means3D"].device
    cur = torch.cuda.current_stream(dev)
    for _rs, radii, geom, scratch in states:      # allocated under the view streams' pools, read on this stream
        for t in (radii, geom, scratch):
            t.record_stream(cur)
    works = []
    for c, (f, n) in enumerate(buffers.chunks):
        if states and n:
            preprocess_backward_views(states, scene["means3D"], scene.get("shs"), scene.get("colors_precomp"),
                                      scene["opacities"], scene.get("scales"), scene.get("rotations"),
                                      scene.get("cov3D_precomp"), accumulate_into=buffers.grads(c),
                                      stats=buffers.densify_stats(c), overwrite=True, first=f, count=n)
        if exchange:
            w = buffers.all_reduce_chunk(c, group, async_op=True)
            if w is not None:
                works.append(w)
    if exchange and max_radii:
        w = buffers.all_reduce_max_radii(group, async_op=True)
        if w is not None:
            works.append(w)
    for w in works:
        w.wait()


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
    from .rasterizer import backward_raw, backward_render_only, forward_raw
    buffers.zero_(stats_only=fused and len(settings_per_view) > 0)
    grads, stats = (None, None) if fused else (buffers.grads(), buffers.densify_stats())
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
    if fused:
        fused_backward_and_exchange(states, scene, buffers, group, exchange)
    elif exchange:
        buffers.all_reduce(group)
    return images
