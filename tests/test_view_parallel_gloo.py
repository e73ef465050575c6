"""World-size-2 gloo test (CPU) of the view-parallel exchange step: view partition,
flat gradient-buffer layout, sum / sum / max all-reduce and the per-view densification statistics.
The per-view gradients are synthetic here -- rendering is CUDA-only; the GPU tests cover the kernels
that write these buffers (tests/test_gpu_api.py::test_backward_accumulate_stats_and_optional_outputs)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gaussian_splatting_lightning_b200.view_parallel import GRAD_FIELDS, StepBuffers, partition_views

P, M, VIEWS, WORLD = 257, 16, 7, 2


def _view_contribution(k: int):
    """Deterministic fake result of rendering view k: gradients, radii, screen-space gradient norm."""
    g = torch.Generator().manual_seed(1000 + k)
    shapes = dict(dL_dmeans3D=(P, 3), dL_dsh=(P, M, 3), dL_dopacity=(P, 1), dL_dscales=(P, 3), dL_drotations=(P, 4))
    grads = {n: torch.randn(s, generator=g) for n, s in shapes.items()}
    radii = torch.randint(0, 40, (P,), generator=g)
    radii[torch.rand(P, generator=g) < 0.3] = 0          # culled in this view
    norm = torch.rand(P, generator=g)
    return grads, radii, norm


def _accumulate_like_the_kernel(buf: StepBuffers, k: int):
    grads, radii, norm = _view_contribution(k)
    vis = radii > 0
    for n, t in buf.grads().items():
        t[vis] += grads[n][vis]                            # GSR_GRADS_ACCUMULATE leaves culled rows alone
    st = buf.densify_stats()
    st["max_radii2D"][vis] = torch.maximum(st["max_radii2D"][vis], radii[vis].float())
    st["xyz_grad_accum"][vis] += norm[vis]
    st["xyz_grad_count"][vis] += 1


def _worker(rank: int, port: int, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        buf = StepBuffers(P, M, "cpu")
        buf.zero_()
        mine = partition_views(VIEWS, WORLD, rank)
        for k in mine:
            _accumulate_like_the_kernel(buf, k)
        buf.all_reduce()
        out[rank] = dict(flat=buf.flat.clone(), stats=buf.stats.clone(), max_radii=buf.max_radii.clone(), views=mine)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_partition_covers_every_view_once():
    for world in (1, 2, 3, 8):
        for views in (0, 1, 7, 64):
            got = sorted(k for r in range(world) for k in partition_views(views, world, r))
            assert got == list(range(views))
    assert partition_views(64, 8, 3) == list(range(3, 64, 8))


def test_flat_layout_is_one_contiguous_buffer():
    buf = StepBuffers(10, 16, "cpu")
    # sections padded to 32 floats: 30 -> 32, 480, 10 -> 32, 30 -> 32, 40 -> 64
    assert buf.floats_per_gaussian == 59 and buf.flat.numel() == 32 + 480 + 32 + 32 + 64
    off = 0
    for n in GRAD_FIELDS:
        v = buf.grads()[n]
        assert v.is_contiguous() and v.data_ptr() == buf.flat.data_ptr() + 4 * off
        off += (v.numel() + 31) // 32 * 32
    assert off == buf.flat.numel()
    assert buf.exchanged_bytes() == (640 + 64 + 10) * 4 and buf.exchanged_bytes(max_radii=False) == (640 + 64) * 4


@pytest.mark.parametrize("P", [1, 2, 3, 5001, 5002, 5003, 4096])
def test_every_section_is_128_byte_aligned_whatever_P(P):
    """After a densify / prune P is rarely a multiple of 4; gsr_backward needs dL_drotations 16-byte aligned
    (csrc/api.cu) and takes the bulk-reduction path for dL_dsh only when that block is aligned."""
    buf = StepBuffers(P, 16, "cpu")
    base = buf.flat.data_ptr()
    assert base % 64 == 0          # torch CPU allocations are 64-byte aligned; CUDA ones 512
    for n, v in buf.grads().items():
        assert (v.data_ptr() - base) % 128 == 0 and v.is_contiguous(), n
    for t in (buf.stats[0], buf.stats[1], buf.max_radii):
        assert (t.data_ptr() - base) % 128 == 0 and t.is_contiguous() and t.numel() == P
    # sections do not overlap: fill each with its own value, read the others back
    for i, v in enumerate(list(buf.grads().values()) + [buf.stats[0], buf.stats[1], buf.max_radii]):
        v.fill_(float(i + 1))
    for i, v in enumerate(list(buf.grads().values()) + [buf.stats[0], buf.stats[1], buf.max_radii]):
        assert bool((v == float(i + 1)).all())


def test_world2_exchange_matches_single_process_sum():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(_free_port(), out), nprocs=WORLD, join=True)
    ref = StepBuffers(P, M, "cpu")
    for k in range(VIEWS):
        _accumulate_like_the_kernel(ref, k)
    assert sorted(out[0]["views"] + out[1]["views"]) == list(range(VIEWS))
    for r in range(WORLD):
        assert torch.allclose(out[r]["flat"], ref.flat, rtol=1e-6, atol=1e-6)
        assert torch.allclose(out[r]["stats"][0], ref.stats[0], rtol=1e-6, atol=1e-6)
        assert torch.equal(out[r]["stats"][1], ref.stats[1])
        assert torch.equal(out[r]["max_radii"], ref.max_radii)
    assert torch.equal(out[0]["flat"], out[1]["flat"])      # replicas stay bit-identical

    # folding the deltas into the model's running statistics == the reference's view-by-view update
    max_r, acc, cnt = torch.zeros(P), torch.zeros(P, 1), torch.zeros(P, 1)
    ref.apply_densify_stats(max_r, acc, cnt)
    w_max, w_acc, w_cnt = torch.zeros(P), torch.zeros(P, 1), torch.zeros(P, 1)
    for k in range(VIEWS):
        _, radii, norm = _view_contribution(k)
        vis = radii > 0
        w_max[vis] = torch.max(w_max[vis], radii[vis].float())      # gaussian_model.py:174-176
        w_acc[vis] += norm[vis, None]                                # gaussian_model.py:178-181
        w_cnt[vis] += 1
    assert torch.equal(max_r, w_max) and torch.allclose(acc, w_acc, atol=1e-6) and torch.equal(cnt, w_cnt)


def test_step_buffers_share_one_storage():
    """flat | stats | max_radii are 128-byte aligned sections of one storage: one memset clears the step,
    the two summed sections are adjacent (one all-reduce)."""
    buf = StepBuffers(10, 16, "cpu")
    base = buf.flat.data_ptr()
    assert (buf.stats.data_ptr() - base) % 128 == 0 and (buf.max_radii.data_ptr() - base) % 128 == 0
    assert buf.flat.data_ptr() < buf.stats.data_ptr() < buf.max_radii.data_ptr()
    assert buf.stats.shape == (2, 10) and buf.stats[0].is_contiguous() and buf.max_radii.shape == (10,)
    assert buf._summed.data_ptr() == base
    assert buf._summed.data_ptr() + 4 * buf._summed.numel() == buf.max_radii.data_ptr()
    buf.flat.fill_(1.0); buf.stats.fill_(2.0); buf.max_radii.fill_(3.0)
    assert float(buf._summed.sum()) == 640 + 40 and float(buf.max_radii.sum()) == 30   # sections do not overlap
    buf.zero_(keep_max_radii=True)
    assert not buf.flat.any() and not buf.stats.any() and float(buf.max_radii.sum()) == 30
    buf.zero_()
    assert not buf.flat.any() and not buf.stats.any() and not buf.max_radii.any()


def test_bench_partitions_cover_the_same_views():
    """bench.py --partition block / interleaved split the SAME global views; interleaved is partition_views."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location(
        "gsr_bench", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    for world, vpr in ((1, 8), (2, 32), (4, 16), (8, 8)):
        block = [[bench.view_index(r, j, vpr, world, "block") for j in range(vpr)] for r in range(world)]
        inter = [[bench.view_index(r, j, vpr, world, "interleaved") for j in range(vpr)] for r in range(world)]
        assert sorted(sum(block, [])) == sorted(sum(inter, [])) and len(set(sum(block, []))) == world * vpr
        assert block[0] == [bench.view_index(0, j, vpr) for j in range(vpr)]          # rank 0 = the N = 1 views
        slot = [bench.view_index(0, g, world * vpr) for g in range(world * vpr)]       # global view g -> orbit slot
        for r in range(world):
            assert inter[r] == [slot[g] for g in partition_views(world * vpr, world, r)]


def _worker_chunked(rank: int, port: int, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        P2 = 5000
        buf = StepBuffers(P2, M, "cpu", chunks=3)
        g = torch.Generator().manual_seed(77 + rank)
        full = {n: torch.randn((P2, M, 3) if n == "dL_dsh" else (P2, w), generator=g) for n, w in buf.widths.items()}
        radii = torch.randint(0, 50, (P2,), generator=g).float()
        for c, (f, n) in enumerate(buf.chunks):
            for name, t in buf.grads(c).items():
                t.copy_(full[name][f:f + n])
            st = buf.densify_stats(c)
            st["xyz_grad_accum"].fill_(rank + 1.0)
            st["xyz_grad_count"].fill_(1.0)
            st["max_radii2D"].copy_(radii[f:f + n])
        works = [buf.all_reduce_chunk(c, async_op=True) for c in range(buf.num_chunks)]
        works.append(buf.all_reduce_max_radii(async_op=True))
        for w in works:
            w.wait()
        got = {n: torch.cat([buf.grads(c)[n] for c in range(buf.num_chunks)]) for n in buf.widths}
        out[rank] = dict(full=full, got=got, radii=radii, max_radii=buf.max_radii.clone(),
                         accum=torch.cat([buf.densify_stats(c)["xyz_grad_accum"] for c in range(buf.num_chunks)]))
    finally:
        dist.destroy_process_group()


def test_chunk_major_buffers_exchange_range_by_range():
    """StepBuffers(chunks=C): every chunk is one contiguous, 128-byte aligned block [its gradient rows | its statistics];
    all_reduce_chunk per chunk + the max of the radii give what the single all-reduce of the whole buffer gives."""
    buf = StepBuffers(5000, 16, "cpu", chunks=3)
    assert buf.chunks == [(0, 2048), (2048, 2048), (4096, 904)] and buf.num_chunks == 3
    base = buf._store.data_ptr()
    prev_end = 0
    for c, (f, n) in enumerate(buf.chunks):
        blk = buf._chunk_block[c]
        assert (blk.data_ptr() - base) == 4 * prev_end and (blk.data_ptr() - base) % 128 == 0
        prev_end += blk.numel()
        for t in list(buf.grads(c).values()) + list(buf.densify_stats(c).values())[1:]:
            assert t.is_contiguous() and (t.data_ptr() - base) % 128 == 0
            assert blk.data_ptr() <= t.data_ptr() and t.data_ptr() + 4 * t.numel() <= blk.data_ptr() + 4 * blk.numel()
        assert buf.grads(c)["dL_dsh"].shape == (n, 16, 3) and buf.densify_stats(c)["max_radii2D"].shape == (n,)
    with pytest.raises(ValueError):
        buf.grads()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_chunked, args=(_free_port(), out), nprocs=WORLD, join=True)
    for n in out[0]["full"]:
        want = out[0]["full"][n] + out[1]["full"][n]
        assert torch.allclose(out[0]["got"][n], want, atol=1e-6) and torch.equal(out[0]["got"][n], out[1]["got"][n])
    assert torch.equal(out[0]["max_radii"], torch.maximum(out[0]["radii"], out[1]["radii"]))
    assert torch.equal(out[0]["accum"], torch.full((5000,), 3.0))
    # folding the chunked deltas into running statistics
    mr, acc, cnt = torch.zeros(5000), torch.zeros(5000, 1), torch.zeros(5000, 1)
    for c in range(buf.num_chunks):
        st = buf.densify_stats(c)
        st["xyz_grad_accum"].fill_(0.5); st["xyz_grad_count"].fill_(2.0); st["max_radii2D"].fill_(7.0)
    buf.apply_densify_stats(mr, acc, cnt)
    assert float(mr.min()) == 7.0 and float(acc.min()) == 0.5 and float(cnt.max()) == 2.0
