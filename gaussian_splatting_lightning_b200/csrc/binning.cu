// binning.cu -- K3 (prefix sum), K4 (key duplication) and K6 (tile ranges) for sm_100a.
//
// All three are HBM-bound integer kernels.  The number of instances R is only known on the
// device (GsrControl::num_rendered); every kernel after the scan reads it from there and
// clamps it to the capacity of the binning buffer, so a frame can be enqueued without a
// host round trip.
//
//   K3  single-pass chained scan with decoupled look-back (one 64-bit status word per
//       2048-item tile: flag in the top two bits, running sum below); tile id = blockIdx.x (a
//       global atomic counter for dynamic ids serialises the tiles).      8 B / Gaussian
//   K4  key = (tile_id << 32) | float_bits(z_view), value = Gaussian index, rect walked
//       y-outer / x-inner (gs_lightning/rasterize/render_tools.py:134-139 gives the rect,
//       rasterize.py:110 the tile id).                            12 B / instance written
//   K6  ranges[t] = [first, last) of tile t in the sorted key list.   8 B / instance read
#include "gsr_math.cuh"

namespace gsr {

namespace {
constexpr unsigned long long SCAN_FLAG_AGG = 1ull << 62;
constexpr unsigned long long SCAN_FLAG_INC = 2ull << 62;
constexpr unsigned long long SCAN_VAL_MASK = (1ull << 62) - 1;
}  // namespace

__global__ void __launch_bounds__(256) scan_kernel(int P, const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                   unsigned long long* state, GsrControl* ctrl) {
  constexpr int IPT = 8;
  __shared__ unsigned int s_warp[8];
  __shared__ unsigned long long s_excl;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned int tile = blockIdx.x;  // in-order CTA dispatch keeps every predecessor resident or done
  const long long base = (long long)tile * GeomState::SCAN_ITEMS + (long long)tid * IPT;

  uint32_t v[IPT];
  if (base + IPT <= P) {
    const uint4 lo = reinterpret_cast<const uint4*>(in + base)[0];
    const uint4 hi = reinterpret_cast<const uint4*>(in + base)[1];
    v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w; v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
  } else {
#pragma unroll
    for (int k = 0; k < IPT; ++k) v[k] = (base + k < P) ? in[base + k] : 0u;
  }
  uint32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < IPT; ++k) { tsum += v[k]; v[k] = tsum; }  // thread-local inclusive
  uint32_t wincl = tsum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, wincl, d);
    if (lane >= d) wincl += t;
  }
  if (lane == 31) s_warp[warp] = wincl;
  __syncthreads();
  uint32_t wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const uint32_t c = s_warp[w];
    if (w < warp) wbase += c;
    total += c;
  }
  if (tid == 0) {
    unsigned long long excl = 0;
    volatile unsigned long long* st = state;
    if (tile == 0) {
      st[0] = SCAN_FLAG_INC | (unsigned long long)total;
    } else {
      st[tile] = SCAN_FLAG_AGG | (unsigned long long)total;
      long long p = (long long)tile - 1;
      while (true) {
        const unsigned long long w = st[p];
        if (w & SCAN_FLAG_INC) { excl += w & SCAN_VAL_MASK; break; }
        if (w & SCAN_FLAG_AGG) { excl += w & SCAN_VAL_MASK; --p; }
      }
      st[tile] = SCAN_FLAG_INC | (excl + total);
    }
    s_excl = excl;
    const unsigned int last_tile = (unsigned int)((P + GeomState::SCAN_ITEMS - 1) / GeomState::SCAN_ITEMS) - 1u;
    if (tile == last_tile) {
      const unsigned long long R = excl + total;
      ctrl->num_rendered = R > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)R;
    }
  }
  __syncthreads();
  const uint32_t off = (uint32_t)s_excl + wbase + (wincl - tsum);
  if (base + IPT <= P) {
    reinterpret_cast<uint4*>(out + base)[0] = make_uint4(off + v[0], off + v[1], off + v[2], off + v[3]);
    reinterpret_cast<uint4*>(out + base)[1] = make_uint4(off + v[4], off + v[5], off + v[6], off + v[7]);
  } else {
#pragma unroll
    for (int k = 0; k < IPT; ++k)
      if (base + k < P) out[base + k] = off + v[k];
  }
}

cudaError_t launch_scan(int P, const GeomState& geom, GsrControl* ctrl, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  const int tiles = (P + GeomState::SCAN_ITEMS - 1) / GeomState::SCAN_ITEMS;
  scan_kernel<<<tiles, 256, 0, s>>>(P, geom.tiles_touched, geom.point_offsets, geom.scan_state, ctrl);
  count_launch();
  return cudaGetLastError();
}

// K4.  When `hist` is given the kernel also builds the digit histograms of ALL radix passes of the
// sort that follows (shared-memory privatised, one global atomic per non-empty bin and CTA), which
// saves the sort its own read of the keys: the depth digits (key bits 0..31) are the same for every
// instance of a Gaussian, so they cost one shared atomic per Gaussian and pass; only the tile-id
// digits (bits 32..) are counted per instance.  Only instances actually written (below `capacity`)
// are counted, so the histograms always describe exactly the keys the sort will see.
__global__ void __launch_bounds__(256) duplicate_kernel(int P, int gx, int gy, const int32_t* __restrict__ radii,
                                                        const GsrRec* __restrict__ recs,
                                                        const uint32_t* __restrict__ offsets,
                                                        unsigned long long* __restrict__ keys,
                                                        uint32_t* __restrict__ vals, uint32_t capacity,
                                                        GsrControl* ctrl, BinningHeader* header,
                                                        uint32_t* __restrict__ hist, int npass) {
  __shared__ uint32_t sh[gsr_sort::MAX_PASSES * gsr_sort::RADIX];
  constexpr int RB = gsr_sort::RADIX_BITS, RMASK = gsr_sort::RADIX - 1;
  if (hist) {
    for (int k = threadIdx.x; k < npass * gsr_sort::RADIX; k += 256) sh[k] = 0;
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && header) { header->capacity = capacity; header->num_tiles = (uint32_t)(gx * gy); header->segmented = 0u; }
  bool over = false;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
    const int r = radii[i];
    if (r <= 0) continue;
    uint32_t off = (i == 0) ? 0u : offsets[i - 1];
    const float4 a = recs[i].a;
    const float depth = recs[i].d.w;
    int x0, y0, x1, y1;
    tile_rect(a.x, a.y, (float)r, gx, gy, x0, y0, x1, y1);
    const uint32_t dbits32 = __float_as_uint(depth);
    const unsigned long long dbits = (unsigned long long)dbits32;
    if (hist) {
      const uint32_t total = (uint32_t)((x1 - x0) * (y1 - y0));
      const uint32_t written = min(off + total, capacity) - min(off, capacity);
      if (written) {
#pragma unroll
        for (int p = 0; p < 32 / RB; ++p)
          if (p < npass) atomicAdd(&sh[p * gsr_sort::RADIX + ((dbits32 >> (p * RB)) & RMASK)], written);
      }
    }
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        if (off < capacity) {
          const uint32_t t = (uint32_t)(y * gx + x);
          keys[off] = ((unsigned long long)t << 32) | dbits;
          vals[off] = (uint32_t)i;
          if (hist)
            for (int p = 32 / RB; p < npass; ++p) atomicAdd(&sh[p * gsr_sort::RADIX + ((t >> (p * RB - 32)) & RMASK)], 1u);
        } else {
          over = true;
        }
        ++off;
      }
  }
  if (over) ctrl->overflow = 1u;
  if (hist) {
    __syncthreads();
    for (int k = threadIdx.x; k < npass * gsr_sort::RADIX; k += 256) {
      const uint32_t c = sh[k];
      if (c) atomicAdd(&hist[k], c);
    }
  }
}

cudaError_t launch_duplicate(const GsrView& v, int P, const int32_t* radii, const GeomState& geom,
                             unsigned long long* keys, uint32_t* vals, uint32_t capacity, GsrControl* ctrl,
                             BinningHeader* header, uint32_t* hist, int npass, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  static_assert(32 % gsr_sort::RADIX_BITS == 0, "depth and tile-id digits must not share a radix digit");
  int blocks = (P + 255) / 256;
  if (hist && blocks > 148 * 8) blocks = 148 * 8;  // few, long-lived CTAs: the histogram flush is per CTA
  duplicate_kernel<<<blocks, 256, 0, s>>>(P, v.gx, v.gy, radii, geom.recs, geom.point_offsets, keys, vals, capacity,
                                          ctrl, header, hist, npass);
  count_launch();
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) tile_ranges_kernel(const unsigned long long* __restrict__ keys,
                                                          const unsigned int* __restrict__ d_n, uint32_t capacity,
                                                          uint2* __restrict__ ranges) {
  const uint32_t n = min(*d_n, capacity);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
      ranges[t].x = 0;
    } else {
      const uint32_t tp = (uint32_t)(keys[i - 1] >> 32);
      if (tp != t) { ranges[tp].y = i; ranges[t].x = i; }
    }
    if (i == n - 1) ranges[t].y = n;
  }
}

cudaError_t launch_tile_ranges(const unsigned long long* keys_sorted, const unsigned int* d_n, uint32_t capacity,
                               uint2* ranges, cudaStream_t s) {
  if (capacity == 0) return cudaSuccess;
  unsigned int blocks = (capacity + 255u) / 256u;
  if (blocks > 148u * 16u) blocks = 148u * 16u;
  tile_ranges_kernel<<<blocks, 256, 0, s>>>(keys_sorted, d_n, capacity, ranges);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
