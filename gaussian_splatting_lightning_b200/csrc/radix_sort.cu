// radix_sort.cu -- K5: hand-written onesweep LSD radix sort of (u64 key, u32 value) pairs.
//
// Sorts the duplicated (tile_id << 32 | depth_bits, gaussian_index) instances between key
// duplication and tile-range extraction.  Stable, 8-bit digits, only the significant key bits
// [0, 32 + bits(num_tiles)) are visited (6 passes at 1080p instead of 8).
//
// Structure ("onesweep", one read + one write of the data per digit):
//   1. one histogram kernel computes the 256-bin digit histograms of ALL passes in a single
//      read of the keys (shared-memory privatised, one global atomic per non-empty bin);
//   2. per pass, one kernel in which every CTA owns a 4096-pair tile: warp-level multi-split
//      ranking with ballots (stable inside the tile), a decoupled look-back over the
//      per-tile digit counts of the preceding tiles (status word = 2 flag bits + 30-bit count,
//      one per digit per tile) to get each digit's global base without a second pass, a
//      shared-memory exchange into digit order, and a coalesced scatter.
// The number of pairs is read from device memory (GsrControl::num_rendered).
//
// HBM traffic: 8R (histogram) + passes * 24R bytes, R = number of pairs.
#include "gsr_common.cuh"

namespace gsr {
using namespace gsr_sort;

namespace {
constexpr uint32_t LB_FLAG_AGG = 1u << 30;
constexpr uint32_t LB_FLAG_INC = 2u << 30;
constexpr uint32_t LB_VAL_MASK = (1u << 30) - 1u;

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Exclusive prefix sum over the first RADIX (= 256) threads of the CTA; every thread of the CTA
// must call it (threads >= RADIX pass 0 and ignore the result).
__device__ __forceinline__ uint32_t digit_exclusive_scan(uint32_t v, uint32_t* s_warp /*[8]*/, int lane, int warp) {
  uint32_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  __syncthreads();  // protects s_warp against the previous use
  if (lane == 31 && warp < RADIX / 32) s_warp[warp] = incl;
  __syncthreads();
  uint32_t wbase = 0;
#pragma unroll
  for (int w = 0; w < RADIX / 32; ++w)
    if (w < warp) wbase += s_warp[w];
  return wbase + incl - v;
}
}  // namespace

__global__ void __launch_bounds__(256) sort_histogram_kernel(const unsigned long long* __restrict__ keys,
                                                             const unsigned int* __restrict__ d_n, uint32_t capacity,
                                                             int npass, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[MAX_PASSES * RADIX];
  const int tid = threadIdx.x;
  for (int k = tid; k < npass * RADIX; k += 256) sh[k] = 0;
  __syncthreads();
  const uint32_t n = min(*d_n, capacity);
  for (uint32_t i = blockIdx.x * 256u + tid; i < n; i += gridDim.x * 256u) {
    const unsigned long long k = keys[i];
    for (int p = 0; p < npass; ++p) atomicAdd(&sh[p * RADIX + (int)((k >> (p * RADIX_BITS)) & (RADIX - 1))], 1u);
  }
  __syncthreads();
  for (int k = tid; k < npass * RADIX; k += 256) {
    const uint32_t c = sh[k];
    if (c) atomicAdd(&hist[k], c);
  }
}

__global__ void __launch_bounds__(THREADS, 2) onesweep_pass_kernel(
    const unsigned long long* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
    unsigned long long* __restrict__ keys_out, uint32_t* __restrict__ vals_out, const unsigned int* __restrict__ d_n,
    uint32_t capacity, int shift, const uint32_t* __restrict__ hist /* this pass, [RADIX] */,
    uint32_t* lookback /* this pass, [ntiles][RADIX] */, uint32_t* lookback_group /* [ngroups][RADIX] */) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  unsigned long long* keys_s = reinterpret_cast<unsigned long long*>(dyn_smem);
  uint32_t* vals_s = reinterpret_cast<uint32_t*>(dyn_smem + (size_t)TILE_ITEMS * 8);
  __shared__ uint32_t s_warp_hist[WARPS][RADIX];
  __shared__ uint32_t s_dig_off[RADIX];   // exclusive scan of this tile's digit counts
  __shared__ uint32_t s_bin_base[RADIX];  // global position of the tile's digit run minus its local offset
  __shared__ uint32_t s_scan[RADIX / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n = min(*d_n, capacity);
  const uint32_t ntiles = (n + TILE_ITEMS - 1) / TILE_ITEMS;
  // Tile id = blockIdx.x.  The look-back only needs every tile's predecessors to be resident or done,
  // which in-order CTA dispatch gives (the same reliance as cub::DeviceScan).  A dynamic id taken from
  // one global atomic counter serialised the whole pass instead: ~1000 same-address atomics with
  // return per pass cost 25 us of a 48 us pass (ablation measured on B200).
  const uint32_t tile = blockIdx.x;
  if (tile >= ntiles) return;
  for (int k = tid; k < WARPS * RADIX; k += THREADS) (&s_warp_hist[0][0])[k] = 0;
  __syncthreads();
  const uint32_t hist_d = tid < RADIX ? hist[tid] : 0u;  // needed late (global digit offsets): load it now
  const uint32_t tile_base = tile * TILE_ITEMS;
  const uint32_t tile_n = min((uint32_t)TILE_ITEMS, n - tile_base);
  const uint32_t warp_base = tile_base + warp * (32 * IPT);

  // ---- load (warp-striped: item order inside the tile = warp, i, lane) ----------------
  unsigned long long key[IPT];
  uint32_t val[IPT];
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const uint32_t idx = warp_base + i * 32 + lane;
    key[i] = idx < n ? keys_in[idx] : ~0ull;
  }
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const uint32_t idx = warp_base + i * 32 + lane;
    val[i] = idx < n ? vals_in[idx] : 0u;
  }

  // ---- stable rank inside the warp, per digit (ballot multi-split) -----------------------
  // peers = lanes of the warp holding the same digit, built from one ballot per digit bit
  // (MATCH.ANY computes the same mask in one instruction but was measured to bound the pass:
  // ~42 us per pass at 0.25 IPC with short-scoreboard stalls, whatever the occupancy).
  uint32_t rank2[IPT / 2];  // two 16-bit ranks per register (a rank is < TILE_ITEMS)
#pragma unroll
  for (int i = 0; i < IPT / 2; ++i) rank2[i] = 0;
  const uint32_t lt = lanemask_lt();
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const uint32_t d = (uint32_t)(key[i] >> shift) & (RADIX - 1);
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < RADIX_BITS; ++b) {
      const bool bit = (d >> b) & 1u;
      const uint32_t m = __ballot_sync(0xffffffffu, bit);
      peers &= bit ? m : ~m;
    }
    const int leader = __ffs(peers) - 1;
    uint32_t pre = 0;
    if (lane == leader) {
      pre = s_warp_hist[warp][d];
      s_warp_hist[warp][d] = pre + __popc(peers);
    }
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank2[i >> 1] |= (pre + __popc(peers & lt)) << (16 * (i & 1));
    __syncwarp();
  }
  __syncthreads();

  // ---- per digit (thread == digit, first 256 threads): prefix over warps, tile count ----
  const bool is_digit = tid < RADIX;
  uint32_t count = 0;
  volatile uint32_t* lb = lookback;
  if (is_digit) {
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = s_warp_hist[w][tid];
      s_warp_hist[w][tid] = count;
      count += c;
    }
    // publish the tile's digit count as early as possible
    if (tile == 0) lb[tid] = LB_FLAG_INC | count;
    else lb[(size_t)tile * RADIX + tid] = LB_FLAG_AGG | count;
  }
  const uint32_t dig_off = digit_exclusive_scan(count, s_scan, lane, warp);
  const uint32_t glob_excl = digit_exclusive_scan(hist_d, s_scan, lane, warp);

  // ---- two-level decoupled look-back ------------------------------------------------------
  // Tiles of one wave start together, so a tile's predecessors mostly hold AGGREGATE words when it
  // looks back and a flat walk is as long as its position in the wave (hundreds of dependent L2
  // round trips / LB_UNROLL).  Tiles are therefore grouped by LB_GROUP = 32: the LAST tile of a
  // group also publishes the group's status word (aggregate over the 32 tiles as soon as it has
  // walked its own group, inclusive prefix once it knows it).  A tile walks at most 31 tile words
  // inside its group, then group words (32 tiles per step).  Publishers are predecessors of their
  // readers (dynamic tile ids), so nobody waits on a tile that has not started.
  if (is_digit) {
    uint32_t excl = 0;
    constexpr int LB_UNROLL = 8;
    const bool publisher = (tile % LB_GROUP) == LB_GROUP - 1;
    volatile uint32_t* lbg = lookback_group;
    const uint32_t group = tile / LB_GROUP;
    bool fin = (tile == 0);
    if (!fin) {
      const long long g0 = (long long)group * LB_GROUP;
      long long p = (long long)tile - 1;
      while (!fin && p >= g0) {  // level 1: tile words of my own group
        uint32_t w[LB_UNROLL];
#pragma unroll
        for (int k = 0; k < LB_UNROLL; ++k)
          w[k] = (p - k >= g0) ? (uint32_t)lb[(size_t)(p - k) * RADIX + tid] : 0xFFFFFFFFu;  // group boundary
        bool stop = false;
        int consumed = 0;
#pragma unroll
        for (int k = 0; k < LB_UNROLL; ++k) {
          if (!fin && !stop) {
            const uint32_t f = w[k] >> 30;
            if (f == 0u || f == 3u) {
              stop = true;  // not published yet (poll again from here) or group boundary
            } else {
              excl += w[k] & LB_VAL_MASK;
              ++consumed;
              if (f & 2u) fin = true;
            }
          }
        }
        p -= consumed;
      }
      if (!fin && group == 0) fin = true;  // walked back to tile 0's predecessor: nothing before it
      if (publisher && !fin) lbg[(size_t)group * RADIX + tid] = LB_FLAG_AGG | (excl + count);
      long long g = (long long)group - 1;
      while (!fin) {  // level 2: group words
        uint32_t w[LB_UNROLL];
#pragma unroll
        for (int k = 0; k < LB_UNROLL; ++k)
          w[k] = (g - k >= 0) ? (uint32_t)lbg[(size_t)(g - k) * RADIX + tid] : (uint32_t)(2u << 30);  // before group 0
        bool stop = false;
        int consumed = 0;
#pragma unroll
        for (int k = 0; k < LB_UNROLL; ++k) {
          if (!fin && !stop) {
            const uint32_t f = w[k] >> 30;
            if (f == 0u) {
              stop = true;
            } else {
              excl += w[k] & LB_VAL_MASK;
              ++consumed;
              if (f & 2u) fin = true;
            }
          }
        }
        g -= consumed;
      }
      lb[(size_t)tile * RADIX + tid] = LB_FLAG_INC | (excl + count);
    }
    if (publisher) lbg[(size_t)group * RADIX + tid] = LB_FLAG_INC | (excl + count);
    s_dig_off[tid] = dig_off;
    s_bin_base[tid] = glob_excl + excl - dig_off;
  }
  __syncthreads();

  // ---- exchange through shared memory into digit order ---------------------------------
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const uint32_t d = (uint32_t)(key[i] >> shift) & (RADIX - 1);
    const uint32_t pos = s_dig_off[d] + s_warp_hist[warp][d] + ((rank2[i >> 1] >> (16 * (i & 1))) & 0xffffu);
    keys_s[pos] = key[i];
    vals_s[pos] = val[i];
  }
  __syncthreads();

  // ---- coalesced scatter (runs of equal digits are contiguous) --------------------------
#pragma unroll
  for (int k = 0; k < IPT; ++k) {
    const uint32_t p = tid + k * THREADS;
    if (p < tile_n) {
      const unsigned long long kk = keys_s[p];
      const uint32_t d = (uint32_t)(kk >> shift) & (RADIX - 1);
      const uint32_t dst = s_bin_base[d] + p;
      keys_out[dst] = kk;
      vals_out[dst] = vals_s[p];
    }
  }
}

cudaError_t launch_onesweep(unsigned long long* keys_a, uint32_t* vals_a, unsigned long long* keys_b,
                            uint32_t* vals_b, const unsigned int* d_n, uint32_t capacity, int end_bit,
                            const Workspace& ws, bool hist_ready, cudaStream_t s) {
  if (capacity == 0) return cudaSuccess;
  const int npass = gsr_sort_passes(end_bit);
  if (npass > MAX_PASSES || npass < 1) return cudaErrorInvalidValue;
  const uint32_t ntiles = (capacity + TILE_ITEMS - 1) / TILE_ITEMS;
  if (ntiles > ws.ntiles) return cudaErrorInvalidValue;
  const size_t dyn = (size_t)TILE_ITEMS * 12;
  cudaError_t e = cudaFuncSetAttribute(onesweep_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
  if (e != cudaSuccess) return e;

  if (!hist_ready) {
    unsigned int hblocks = (capacity + 256u * 8u - 1u) / (256u * 8u);
    if (hblocks > 148u * 4u) hblocks = 148u * 4u;
    sort_histogram_kernel<<<hblocks, 256, 0, s>>>(keys_a, d_n, capacity, npass, ws.hist);
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }

  unsigned long long* kin = keys_a; uint32_t* vin = vals_a;
  unsigned long long* kout = keys_b; uint32_t* vout = vals_b;
  for (int p = 0; p < npass; ++p) {
    onesweep_pass_kernel<<<ntiles, THREADS, dyn, s>>>(kin, vin, kout, vout, d_n, capacity, p * RADIX_BITS,
                                                      ws.hist + (size_t)p * RADIX,
                                                      ws.lookback + (size_t)p * ws.ntiles * RADIX,
                                                      ws.lookback_group + (size_t)p * ws.ngroups * RADIX);
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    unsigned long long* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  return cudaSuccess;
}

}  // namespace gsr
