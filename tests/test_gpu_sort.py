"""Stand-alone parity of the hand-written onesweep radix sort (gsr_sort_pairs) against numpy's stable
argsort: bit-exact keys AND values (stability), ragged sizes around the 4096-pair tile, empty input,
heavy ties, restricted bit ranges, and at BASELINE size (3M pairs, 45 bits) the size-independent
properties sortedness + permutation + stability."""
import numpy as np
import pytest
import torch

from gaussian_splatting_lightning_b200.rasterizer import sort_pairs

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ref(keys, end_bit):
    mask = np.uint64((1 << end_bit) - 1) if end_bit < 64 else np.uint64(0xFFFFFFFFFFFFFFFF)
    order = np.argsort(keys & mask, kind="stable")
    return order


@pytest.mark.parametrize("n", [0, 1, 2, 31, 33, 4095, 4096, 4097, 8191, 12289, 100_003])
@pytest.mark.parametrize("end_bit", [39, 45, 64])
def test_sort_matches_stable_argsort(n, end_bit):
    rng = np.random.default_rng(n * 131 + end_bit)
    keys = rng.integers(0, 1 << min(end_bit, 63), size=n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)[::-1].copy()
    ko, vo = sort_pairs(torch.tensor(keys.view(np.int64), device=DEV), torch.tensor(vals.view(np.int32), device=DEV),
                        end_bit)
    order = _ref(keys, end_bit)
    assert np.array_equal(ko.cpu().numpy().view(np.uint64), keys[order])
    assert np.array_equal(vo.cpu().numpy().view(np.uint32), vals[order])


def test_sort_heavy_ties_is_stable():
    rng = np.random.default_rng(0)
    n = 50_000
    keys = (rng.integers(0, 7, size=n, dtype=np.uint64) << np.uint64(32)) | rng.integers(0, 3, size=n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = sort_pairs(torch.tensor(keys.view(np.int64), device=DEV), torch.tensor(vals.view(np.int32), device=DEV), 40)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(vo.cpu().numpy().view(np.uint32), vals[order])
    assert np.array_equal(ko.cpu().numpy().view(np.uint64), keys[order])


def test_sort_ignores_bits_above_end_bit():
    rng = np.random.default_rng(1)
    n = 10_000
    keys = rng.integers(0, 1 << 62, size=n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = sort_pairs(torch.tensor(keys.view(np.int64), device=DEV), torch.tensor(vals.view(np.int32), device=DEV), 16)
    order = _ref(keys, 16)
    assert np.array_equal(vo.cpu().numpy().view(np.uint32), vals[order])


def test_sort_full_size_properties():
    """BASELINE C3 size: R = 3M (tile_id<<32 | depth bits) keys, 45 significant bits."""
    g = torch.Generator(device="cpu").manual_seed(0)
    n = 3_000_000
    tiles = torch.randint(0, 8160, (n,), generator=g, dtype=torch.int64)
    depth = (torch.rand(n, generator=g) * 18 + 2).float().view(torch.int32).to(torch.int64)
    keys = ((tiles << 32) | depth).to(DEV)
    vals = torch.arange(n, dtype=torch.int32, device=DEV)
    ko, vo = sort_pairs(keys, vals, 45)
    assert bool((ko[1:] >= ko[:-1]).all())                      # sortedness
    assert bool((keys[vo.long()] == ko).all())                   # values travelled with their keys
    assert int(torch.unique(vo).numel()) == n                    # a permutation
    same = ko[1:] == ko[:-1]
    assert bool((vo[1:][same] > vo[:-1][same]).all())            # stability
    ref = torch.sort(keys, stable=True)
    assert torch.equal(ref.values, ko) and torch.equal(ref.indices.to(torch.int32), vo)
