// tile_binning.cu -- tile-segmented binning: K3' (tile scan), K4' (scatter) and K5' (per-tile sort).
//
// The reference pipeline sorts all R (tile << 32 | depth, index) instances with one global stable
// radix sort (a12: rasterize.py:227-234 sorts per tile; upstream CUDA uses cub::DeviceRadixSort) and
// then finds the tile boundaries.  Its RESULT is: per tile, the instances in ascending depth bits,
// ties in ascending Gaussian index (the duplicated list is emitted in Gaussian order and the radix
// sort is stable).  That result is produced here without a global sort:
//   * the preprocess kernel counts the instances of every tile (one RED per instance);
//   * K3' a chained scan over the T counts (1024 tiles per CTA): ranges[t] = [first, last) directly (no K6), R = total;
//   * K4' every Gaussian claims a slot in each of its tiles' segments with an atomic cursor (a second
//     counter array the scan set to the segment starts) and
//     stores the 64-bit composite (depth bits << 32 | Gaussian index) there, in arbitrary order;
//   * K5' one CTA per tile (a second, list-driven launch for the tiles above 1024 instances) sorts the segment's composites ascending -- depth first, index breaks
//     ties, which is exactly the stable order -- with a bitonic network in shared memory
//     (segments above 8192 instances: the same network in place in global memory) and writes the
//     sorted point list.
// HBM traffic: 8 B/instance written by K4', 8 read + 4 written by K5', against (8 + 24 * 6) for the
// radix sort at 1080p; the instruction count falls from ~1150 to ~330 per 32 instances
// (profiles/README.md, r1_v7).  Measured on C3: 0.342 -> 0.174 ms per frame for the whole binning.
#include "gsr_math.cuh"

namespace gsr {

// ---- K3': exclusive scan of the per-tile counts ---------------------------------------------------
// Chained scan: CTA b takes the tiles [1024 b, 1024 b + 1024) (four per thread; eight CTAs at 1080p, so that eight
// SMs share the strided counter loads and cursor stores), publishes its total in a 64-bit status word and walks
// back over its predecessors' words (decoupled look-back; tile = blockIdx.x, in-order CTA dispatch keeps every
// predecessor resident or done).  Empty tiles get [0, 0) like the reference's zero-initialised ranges.  Tiles above
// 512 instances go on the lists of the MID and LARGE sort launches (one array: MID from the front, LARGE from the
// back), one global atomic per CTA and list.
namespace {
constexpr unsigned long long TSCAN_AGG = 1ull << 62, TSCAN_INC = 2ull << 62, TSCAN_VAL = (1ull << 62) - 1;
constexpr int TSCAN_TILES = 1024;
}  // namespace

__global__ void __launch_bounds__(256) tile_scan_kernel(int T, int gx, const uint32_t* __restrict__ count,
                                                        uint2* __restrict__ ranges, uint32_t* __restrict__ cursor,
                                                        uint32_t* __restrict__ large_tiles, unsigned long long* state,
                                                        GsrControl* ctrl, uint32_t* __restrict__ order) {
  constexpr int IPT = TSCAN_TILES / 256;
  __shared__ unsigned long long s_warp[8];
  __shared__ unsigned long long s_excl;
  __shared__ unsigned int s_mid, s_large, s_mid_base, s_large_base;
  __shared__ unsigned int s_cls[64];   // tile order: histogram, then running base, of the 64 length classes
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned int blk = blockIdx.x;
  if (tid == 0) { s_mid = 0u; s_large = 0u; }
  if (tid < 64) s_cls[tid] = 0u;
  const int k0 = (int)blk * TSCAN_TILES + IPT * tid;
  uint32_t c[IPT];
#pragma unroll
  for (int q = 0; q < IPT; ++q) {  // the tile's half of its pair's 64-bit word
    const int t = k0 + q, y = t / gx, x = t - y * gx;
    c[q] = (t < T) ? count[gsr_tile_pair_word(x, y, gx) * 2 + (x & 1)] : 0u;
  }
  unsigned long long sum = 0;
#pragma unroll
  for (int q = 0; q < IPT; ++q) sum += c[q];
  unsigned long long incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned long long wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const unsigned long long x = s_warp[w];
    if (w < warp) wbase += x;
    total += x;
  }
  // list slots inside the CTA (c[q] = 0 beyond T): one shared atomic per warp and list
  const uint32_t lt = (1u << lane) - 1u;
  uint32_t slot[IPT];
#pragma unroll
  for (int q = 0; q < IPT; ++q) {
    const bool large = c[q] > 2048u, mid = c[q] > 512u && !large;
    const uint32_t mm = __ballot_sync(0xffffffffu, mid), ml = __ballot_sync(0xffffffffu, large);
    uint32_t bm = 0, bl = 0;
    if (mm && lane == __ffs(mm) - 1) bm = atomicAdd(&s_mid, __popc(mm));
    if (ml && lane == __ffs(ml) - 1) bl = atomicAdd(&s_large, __popc(ml));
    if (mm) bm = __shfl_sync(0xffffffffu, bm, __ffs(mm) - 1);
    if (ml) bl = __shfl_sync(0xffffffffu, bl, __ffs(ml) - 1);
    slot[q] = mid ? bm + __popc(mm & lt) : bl + __popc(ml & lt);
  }
  __syncthreads();
  if (tid == 0) {
    unsigned long long excl = 0;
    volatile unsigned long long* st = state;
    if (blk == 0) {
      st[0] = TSCAN_INC | total;
    } else {
      st[blk] = TSCAN_AGG | total;
      long long p = (long long)blk - 1;
      while (true) {
        const unsigned long long w = st[p];
        if (w & TSCAN_INC) { excl += w & TSCAN_VAL; break; }
        if (w & TSCAN_AGG) { excl += w & TSCAN_VAL; --p; }
      }
      st[blk] = TSCAN_INC | (excl + total);
    }
    s_excl = excl;
    if (blk == gridDim.x - 1) {
      const unsigned long long R = excl + total;
      ctrl->num_rendered = R > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)R;
    }
    s_mid_base = s_mid ? atomicAdd(&ctrl->num_mid, s_mid) : 0u;
    s_large_base = s_large ? atomicAdd(&ctrl->num_large, s_large) : 0u;
  }
  // ---- tile order of the compositing kernels (blockIdx -> tile) ------------------------------------------------
  // One CTA per tile, dispatched in blockIdx order: with the tiles in raster order the heavy tiles sit wherever the
  // scene puts them and the last wave runs at low occupancy (sm__warps_active 41 % / 52 % in the r1 captures).
  // Every scan CTA lists ITS tiles by descending list length (counting sort over 64 length classes about 1/8 of the
  // CTA's mean length wide) and the CTAs' lists are interleaved: order[rank * nCTA + cta].  Approximately
  // longest-processing-time-first for the in-order dispatch, at no extra launch and no global synchronisation.
  if (order) {
    const int ntile_cta = min(TSCAN_TILES, T - (int)blk * TSCAN_TILES);
    const uint32_t mean = (uint32_t)(total / (unsigned long long)(ntile_cta > 0 ? ntile_cta : 1));
    int shift = 0;
    while ((mean >> (shift + 3)) > 0u) ++shift;
    uint32_t cls[IPT];
#pragma unroll
    for (int q = 0; q < IPT; ++q) {
      cls[q] = 63u - min(63u, c[q] >> shift);
      if (k0 + q < T) atomicAdd(&s_cls[cls[q]], 1u);
    }
    __syncthreads();
    if (warp == 0) {  // exclusive scan of the 64 class counts: two per lane
      const uint32_t h0 = s_cls[2 * lane], h1 = s_cls[2 * lane + 1];
      uint32_t inc = h0 + h1;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
      }
      s_cls[2 * lane] = inc - h0 - h1;
      s_cls[2 * lane + 1] = inc - h1;
    }
    __syncthreads();
    const uint32_t ncta = gridDim.x;
    const uint32_t n_last = (uint32_t)(T - (int)(ncta - 1) * TSCAN_TILES);   // tiles of the last (partial) CTA
#pragma unroll
    for (int q = 0; q < IPT; ++q) {
      if (k0 + q < T) {
        const uint32_t r = atomicAdd(&s_cls[cls[q]], 1u);   // rank inside this CTA, heaviest class first
        // ranks below n_last exist in every CTA; above, only in the ncta - 1 full ones
        const uint32_t pos = r < n_last ? r * ncta + blk : n_last * ncta + (r - n_last) * (ncta - 1u) + blk;
        order[pos] = (uint32_t)(k0 + q);
      }
    }
  }
  __syncthreads();
  unsigned long long run = s_excl + wbase + incl - sum;
#pragma unroll
  for (int q = 0; q < IPT; ++q) {
    if (k0 + q < T) {
      const unsigned long long lo = run, hi = run + c[q];
      const uint32_t lo32 = lo > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)lo;
      const uint32_t hi32 = hi > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)hi;
      ranges[k0 + q] = c[q] ? make_uint2(lo32, hi32) : make_uint2(0u, 0u);
      {
        const int t = k0 + q, y = t / gx, x = t - y * gx;
        cursor[gsr_tile_pair_word(x, y, gx) * 2 + (x & 1)] = lo32;  // the scatter's cursor starts at the segment
      }
      if (c[q] > 2048u) large_tiles[T - 1 - (int)(s_large_base + slot[q])] = (uint32_t)(k0 + q);
      else if (c[q] > 512u) large_tiles[s_mid_base + slot[q]] = (uint32_t)(k0 + q);  // see launch_tile_sort
      run = hi;
    }
  }
}

// state: (T + 1023) / 1024 status words, zeroed; ctrl->num_mid / num_large zeroed.
cudaError_t launch_tile_scan(int gx, int gy, const uint32_t* tile_count, uint2* ranges, uint32_t* tile_cursor,
                             uint32_t* large_tiles, unsigned long long* state, GsrControl* ctrl, uint32_t* order,
                             cudaStream_t s) {
  const int T = gx * gy;
  if (T <= 0) return cudaSuccess;
  tile_scan_kernel<<<(T + TSCAN_TILES - 1) / TSCAN_TILES, 256, 0, s>>>(T, gx, tile_count, ranges, tile_cursor,
                                                                      large_tiles, state, ctrl, order);
  count_launch();
  return cudaGetLastError();
}

// ---- K4': scatter the composites into the tile segments ------------------------------------------
__global__ void __launch_bounds__(256) tile_scatter_kernel(int P, int gx, int gy, const int32_t* __restrict__ radii,
                                                           const GsrRec* __restrict__ recs, uint32_t* cursor,
                                                           unsigned long long* __restrict__ comp, uint32_t capacity,
                                                           GsrControl* ctrl, BinningHeader* header) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0 && header) { header->capacity = capacity; header->num_tiles = (uint32_t)(gx * gy); header->segmented = 1u; }
  if (i >= P) return;
  const int r = radii[i];
  const float4 a = recs[i].a;        // requested together with the radius (a culled Gaussian's row is stale, unused)
  const float depth = recs[i].d.w;
  if (r <= 0) return;
  int x0, y0, x1, y1;
  tile_rect(a.x, a.y, (float)r, gx, gy, x0, y0, x1, y1);
  const unsigned long long c = ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned long long)(uint32_t)i;
  // walk the rect by PAIRS of tiles (2k, 2k + 1): one 64-bit atomic on the pair's cursor word returns the slots of
  // both halves; four pairs at a time, so that four independent atomics are in flight before the first store
  unsigned long long* cur = reinterpret_cast<unsigned long long*>(cursor);
  const int xp0 = x0 >> 1, npx = ((x1 - 1) >> 1) - xp0 + 1, npairs = npx * (y1 - y0);
  bool over = false;
  int xp = xp0, y = y0;
  for (int k0 = 0; k0 < npairs; k0 += 4) {
    unsigned long long old[4];
    uint32_t use[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      use[q] = 0u;
      old[q] = 0ull;
      if (k0 + q < npairs) {
        use[q] = (2 * xp >= x0 ? 1u : 0u) | (2 * xp + 1 < x1 ? 2u : 0u);
        const unsigned long long inc = (unsigned long long)(use[q] & 1u) | ((unsigned long long)(use[q] >> 1) << 32);
        old[q] = atomicAdd(cur + gsr_tile_pair_word(2 * xp, y, gx), inc);  // cursors start at ranges[t].x
        if (++xp == xp0 + npx) { xp = xp0; ++y; }
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t ilo = (uint32_t)old[q], ihi = (uint32_t)(old[q] >> 32);
      if (use[q] & 1u) { if (ilo < capacity) comp[ilo] = c; else over = true; }
      if (use[q] & 2u) { if (ihi < capacity) comp[ihi] = c; else over = true; }
    }
  }
  if (over) ctrl->overflow = 1u;
}

cudaError_t launch_tile_scatter(const GsrView& v, int P, const int32_t* radii, const GeomState& geom, uint32_t* cursor,
                                unsigned long long* comp, uint32_t capacity, GsrControl* ctrl, BinningHeader* header,
                                cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  static const int threads = env_threads("GSR_SCATTER_THREADS", 32);
  tile_scatter_kernel<<<(P + threads - 1) / threads, threads, 0, s>>>(P, v.gx, v.gy, radii, geom.recs, cursor, comp,
                                                                      capacity, ctrl, header);
  count_launch();
  return cudaGetLastError();
}

// ---- K5': per-tile sort ---------------------------------------------------------------------------
namespace {
typedef unsigned long long u64;

// compare-exchange, ascending: one 64-bit compare (two ISETP), four 32-bit selects.  Written on the halves
// because ptxas turns the plain C form into min(x, y) and max(x, y) with a compare each.
__device__ __forceinline__ void cx(u64& x, u64& y) {
  unsigned xl, xh, yl, yh, al, ah, bl, bh;
  asm("mov.b64 {%0,%1}, %2;" : "=r"(xl), "=r"(xh) : "l"(x));
  asm("mov.b64 {%0,%1}, %2;" : "=r"(yl), "=r"(yh) : "l"(y));
  asm("{\n\t.reg .pred p;\n\tsetp.lt.u64 p, %8, %9;\n\tselp.b32 %0, %6, %4, p;\n\tselp.b32 %1, %7, %5, p;\n\t"
      "selp.b32 %2, %4, %6, p;\n\tselp.b32 %3, %5, %7, p;\n\t}"
      : "=r"(al), "=r"(ah), "=r"(bl), "=r"(bh) : "r"(xl), "r"(xh), "r"(yl), "r"(yh), "l"(y), "l"(x));
  asm("mov.b64 %0, {%1,%2};" : "=l"(x) : "r"(al), "r"(ah));
  asm("mov.b64 %0, {%1,%2};" : "=l"(y) : "r"(bl), "r"(bh));
}

// Shared memory holds the segment as 16-byte chunks of two composites; chunk c lives at c ^ ((c >> 3) & 7).
// Every access below is a thread reading 4 consecutive composites (two chunks) out of 16 consecutive ones, or
// out of 4 columns a power of two apart; the swizzle spreads the eight threads of a quarter warp over the
// eight 16-byte bank groups in both patterns.  The network is bound by shared-memory bandwidth (every step
// moves the whole segment), hence 16 composites per thread and up to ten steps per pass over the data.
__device__ __forceinline__ uint32_t swz(uint32_t c) { return c ^ ((c >> 3) & 7u); }

__device__ __forceinline__ void load4(const ulonglong2* s2, uint32_t e0 /* multiple of 4 */, u64* v) {
  const ulonglong2 a = s2[swz(e0 >> 1)], b = s2[swz((e0 >> 1) + 1u)];
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void store4(ulonglong2* s2, uint32_t e0, const u64* v) {
  s2[swz(e0 >> 1)] = make_ulonglong2(v[0], v[1]);
  s2[swz((e0 >> 1) + 1u)] = make_ulonglong2(v[2], v[3]);
}

// halving steps of distance 8, 4, 2, 1 on 16 composites in registers
__device__ __forceinline__ void tail16(u64* e) {
#pragma unroll
  for (int d = 8; d >= 1; d >>= 1)
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if ((i & d) == 0) cx(e[i], e[i + d]);
}
// the merges of size 2, 4, 8, 16 on 16 composites in registers
__device__ __forceinline__ void sort16(u64* e) {
#pragma unroll
  for (int k = 2; k <= 16; k <<= 1) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int j = (i & ~(k - 1)) + (k - 1) - (i & (k - 1));  // mirror partner inside the block of k
      if (i < j) cx(e[i], e[j]);
    }
#pragma unroll
    for (int d = k >> 2; d >= 1; d >>= 1)
#pragma unroll
      for (int i = 0; i < 16; ++i)
        if ((i & d) == 0) cx(e[i], e[i + d]);
  }
}

// A warp's 32 units of one round (unit = tid + m * THREADS, 16 composites each) cover one aligned block of 512
// composites in every pass whose partners lie inside aligned blocks of <= 512: between two such passes the
// warp only waits for itself.
template <bool ONE_WARP>  // ONE_WARP: the segment belongs to a single warp (<= 512 composites), never a CTA barrier
__device__ __forceinline__ void pass_sync(uint32_t span_done, uint32_t span_next) {
  if (!ONE_WARP && (span_done > 512u || span_next > 512u)) __syncthreads();
  else __syncwarp();
}

// Bitonic network in its all-ascending form (the first step of a merge mirrors the upper half) over pow2
// composites (pow2 >= 16, padded with ~0).  A merge of size k is: mirror(k), halve(k/4) ... halve(16), then
// tail16.  The steps before the tail are taken two per pass (four columns of 4 composites per thread); when
// their number is odd the mirror step goes alone.
// Only composites below n are real.  A merge of size k is skipped for a block that holds at most k/2 of them:
// its lower half is sorted and its upper half is all padding, so it is sorted already (a segment of 540 costs
// one sort of 512, a little work on the 28 and ONE merge of 1024, not a sort of 1024).
template <int THREADS>
__device__ __forceinline__ void bitonic_sort_shared(u64* s, uint32_t pow2, uint32_t n, int tid) {
  ulonglong2* s2 = reinterpret_cast<ulonglong2*>(s);
  const uint32_t units = pow2 >> 4;
  for (uint32_t u = tid; u < units; u += THREADS) {
    if (16u * u >= n) continue;
    u64 e[16];
#pragma unroll
    for (int g = 0; g < 4; ++g) load4(s2, 16u * u + 4u * g, e + 4 * g);
    sort16(e);
#pragma unroll
    for (int g = 0; g < 4; ++g) store4(s2, 16u * u + 4u * g, e + 4 * g);
  }
  pass_sync<THREADS == 32>(16u, 32u);
  for (uint32_t k = 32; k <= pow2; k <<= 1) {
    const uint32_t nsteps = (uint32_t)(31 - __clz((int)k)) - 4u;  // mirror(k), halve(k/4) .. halve(16)
    uint32_t jj = k >> 2;                                        // distance of the next halving step
    if (nsteps & 1u) {  // mirror(k) alone: 8 pairs per thread
      const uint32_t half = k >> 1;
      for (uint32_t u = tid; u < units; u += THREADS) {
        const uint32_t p0 = 8u * u, r = p0 & (half - 1u), base = (p0 - r) << 1;
        const uint32_t lo = base + r, hi = base + k - 8u - r;
        if (n <= base + half) continue;
        u64 L[8], H[8];
        load4(s2, lo, L); load4(s2, lo + 4u, L + 4); load4(s2, hi, H); load4(s2, hi + 4u, H + 4);
#pragma unroll
        for (int q = 0; q < 8; ++q) cx(L[q], H[7 - q]);
        store4(s2, lo, L); store4(s2, lo + 4u, L + 4); store4(s2, hi, H); store4(s2, hi + 4u, H + 4);
      }
      pass_sync<THREADS == 32>(k, 16u);
    } else {  // mirror(k) + halve(k/4): columns A, B at offset r, C, D at the mirrored offset
      const uint32_t j2 = jj, upb = j2 >> 2;  // units per block of 4 * j2 = k
      for (uint32_t u = tid; u < units; u += THREADS) {
        const uint32_t r = 4u * (u & (upb - 1u)), b = (u / upb) * k, rm = j2 - 4u - r;
        const uint32_t iA = b + r, iB = iA + j2, iC = b + 2u * j2 + rm, iD = iC + j2;
        if (n <= b + (k >> 1)) continue;
        u64 A[4], B[4], C[4], D[4];
        load4(s2, iA, A); load4(s2, iB, B); load4(s2, iC, C); load4(s2, iD, D);
#pragma unroll
        for (int q = 0; q < 4; ++q) { cx(A[q], D[3 - q]); cx(B[q], C[3 - q]); }
#pragma unroll
        for (int q = 0; q < 4; ++q) { cx(A[q], B[q]); cx(C[q], D[q]); }
        store4(s2, iA, A); store4(s2, iB, B); store4(s2, iC, C); store4(s2, iD, D);
      }
      pass_sync<THREADS == 32>(k, 16u);
      jj >>= 1;
    }
    for (; jj >= 32u; jj >>= 2) {  // halve(jj) + halve(jj / 2)
      const uint32_t j2 = jj >> 1, upb = j2 >> 2, blk = 2u * jj;
      for (uint32_t u = tid; u < units; u += THREADS) {
        const uint32_t r = 4u * (u & (upb - 1u)), iA = (u / upb) * blk + r, iB = iA + j2, iC = iB + j2, iD = iC + j2;
        if (n <= (iA & ~(k - 1u)) + (k >> 1)) continue;
        u64 A[4], B[4], C[4], D[4];
        load4(s2, iA, A); load4(s2, iB, B); load4(s2, iC, C); load4(s2, iD, D);
#pragma unroll
        for (int q = 0; q < 4; ++q) { cx(A[q], C[q]); cx(B[q], D[q]); }
#pragma unroll
        for (int q = 0; q < 4; ++q) { cx(A[q], B[q]); cx(C[q], D[q]); }
        store4(s2, iA, A); store4(s2, iB, B); store4(s2, iC, C); store4(s2, iD, D);
      }
      pass_sync<THREADS == 32>(blk, 16u);
    }
    for (uint32_t u = tid; u < units; u += THREADS) {
      if (n <= ((16u * u) & ~(k - 1u)) + (k >> 1)) continue;
      u64 e[16];
#pragma unroll
      for (int g = 0; g < 4; ++g) load4(s2, 16u * u + 4u * g, e + 4 * g);
      tail16(e);
#pragma unroll
      for (int g = 0; g < 4; ++g) store4(s2, 16u * u + 4u * g, e + 4 * g);
    }
    pass_sync<THREADS == 32>(16u, 2u * k);
  }
}

// The same network for a segment that does not fit shared memory: in place through L2, one pair per
// thread and step, virtual padding (a pair whose upper index is >= n is skipped).
template <int THREADS>
__device__ __forceinline__ void bitonic_sort_global(volatile u64* a, uint32_t n, int tid) {
  const uint32_t pow2 = 1u << (32 - __clz((int)(n - 1u)));
  const uint32_t npairs = pow2 >> 1;
  for (uint32_t k = 2; k <= pow2; k <<= 1) {
    const uint32_t half = k >> 1;
    for (uint32_t p = tid; p < npairs; p += THREADS) {
      const uint32_t r = p & (half - 1u), base = (p - r) << 1, i = base + r, j = base + k - 1u - r;
      if (j < n) { const u64 x = a[i], y = a[j]; if (y < x) { a[i] = y; a[j] = x; } }
    }
    __syncthreads();
    for (uint32_t jj = half >> 1; jj; jj >>= 1) {
      for (uint32_t p = tid; p < npairs; p += THREADS) {
        const uint32_t r = p & (jj - 1u), i = 2u * p - r, j = i + jj;
        if (j < n) { const u64 x = a[i], y = a[j]; if (y < x) { a[i] = y; a[j] = x; } }
      }
      __syncthreads();
    }
  }
}
}  // namespace

// Size classes, one warp per 512 composites in each:
//   SMALL  a warp per tile: segments of 1..512 instances, 4 KB of shared memory;
//   MID    CTAs of 128 threads walking the list of tiles with 513..2048 instances (16 KB), same launch as SMALL;
//   LARGE  CTAs of 512 threads walking the list of tiles above 2048 (64 KB up to 8192 instances, above that
//          the global-memory network), a second launch.
// The tile scan makes the two lists.  All decide on the count clamped to the capacity, so a truncated frame
// (binning hint too small) stays in bounds: whatever a clamp moved below 513 the SMALL warps take.
constexpr uint32_t TS_SMALL_CAP = 512, TS_MID_CAP = 2048, TS_LARGE_CAP = 8192;

template <int CAP, int THREADS>
__device__ __forceinline__ void sort_segment(u64* s_comp, u64* comp, uint32_t* vals, uint32_t lo, uint32_t n, int tid) {
  ulonglong2* s2 = reinterpret_cast<ulonglong2*>(s_comp);
  u64* seg = comp + lo;
  if (n <= (uint32_t)CAP) {
    const uint32_t pow2 = n <= 16u ? 16u : 1u << (32 - __clz((int)(n - 1u)));
    for (uint32_t c = tid; c < (pow2 >> 1); c += THREADS) {
      const uint32_t i = 2u * c;
      s2[swz(c)] = make_ulonglong2(i < n ? seg[i] : ~0ull, i + 1u < n ? seg[i + 1u] : ~0ull);
    }
    if (THREADS == 32) __syncwarp(); else __syncthreads();
    bitonic_sort_shared<THREADS>(s_comp, pow2, n, tid);
    if (THREADS == 32) __syncwarp(); else __syncthreads();
    for (uint32_t i = tid; i < n; i += THREADS) {
      const ulonglong2 v = s2[swz(i >> 1)];
      vals[lo + i] = (uint32_t)((i & 1u) ? v.y : v.x);
    }
  } else {
    bitonic_sort_global<THREADS>((volatile u64*)seg, n, tid);
    for (uint32_t i = tid; i < n; i += THREADS) vals[lo + i] = (uint32_t)((volatile u64*)seg)[i];
  }
}

// SMALL and MID in one launch (CTAs of 128 threads, 16 KB): the first `mid_ctas` CTAs walk the MID list, one tile
// at a time; every other CTA takes four consecutive tiles, one per warp (a warp sorts up to 512 composites with
// no CTA barrier at all).  The few, long MID sorts start first and run under the many short ones.
__global__ void __launch_bounds__(128) tile_sort_kernel(uint2* ranges, const uint32_t* __restrict__ mid_tiles, int T,
                                                        const GsrControl* __restrict__ ctrl, u64* comp,
                                                        uint32_t* __restrict__ vals, uint32_t capacity,
                                                        uint32_t mid_ctas) {
  __shared__ __align__(16) u64 s_comp[TS_MID_CAP];
  if (blockIdx.x >= mid_ctas) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t t = (blockIdx.x - mid_ctas) * 4u + (uint32_t)warp;
    if (t >= (uint32_t)T) return;
    const uint2 r = ranges[t];
    const uint32_t lo = min(r.x, capacity), hi = min(r.y, capacity), n = hi - lo;
    if ((lo != r.x || hi != r.y) && lane == 0) ranges[t] = make_uint2(lo, hi);  // truncated frame (hint too small)
    if (n == 0u || n > TS_SMALL_CAP) return;
    sort_segment<TS_SMALL_CAP, 32>(s_comp + TS_SMALL_CAP * warp, comp, vals, lo, n, lane);
  } else {
    const uint32_t num = ctrl->num_mid;
    for (uint32_t k = blockIdx.x; k < num; k += mid_ctas) {
      const uint32_t t = mid_tiles[k];
      const uint2 r = ranges[t];  // possibly clamped already by the tile's warp above: clamping is idempotent
      const uint32_t lo = min(r.x, capacity), hi = min(r.y, capacity), n = hi - lo;
      if (n > TS_SMALL_CAP) sort_segment<TS_MID_CAP, 128>(s_comp, comp, vals, lo, n, threadIdx.x);
      __syncthreads();  // shared memory is reused by the next tile
    }
  }
}

// LARGE: the list is stored from the end of `tiles` backwards (the two lists share one array of T words).
__global__ void __launch_bounds__(512) tile_sort_large_kernel(uint2* ranges, const uint32_t* __restrict__ tiles, int T,
                                                              const GsrControl* __restrict__ ctrl, u64* comp,
                                                              uint32_t* __restrict__ vals) {
  extern __shared__ __align__(16) unsigned char ts_smem[];
  const uint32_t num = ctrl->num_large;
  for (uint32_t k = blockIdx.x; k < num; k += gridDim.x) {
    const uint32_t t = tiles[(uint32_t)T - 1u - k];
    const uint2 r = ranges[t];  // clamped by the launch before
    const uint32_t n = r.y - r.x;
    if (n > TS_SMALL_CAP)
      sort_segment<TS_LARGE_CAP, 512>(reinterpret_cast<u64*>(ts_smem), comp, vals, r.x, n, threadIdx.x);
    __syncthreads();  // shared memory is reused by the next tile
  }
}

cudaError_t launch_tile_sort(int T, uint2* ranges, const uint32_t* large_tiles, const GsrControl* ctrl,
                             unsigned long long* comp, uint32_t* vals, uint32_t capacity, cudaStream_t s) {
  if (T <= 0 || capacity == 0) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(tile_sort_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(TS_LARGE_CAP * 8));
  if (e != cudaSuccess) return e;
  const uint32_t mid_ctas = 148u * 6u;
  tile_sort_kernel<<<mid_ctas + (uint32_t)((T + 3) / 4), 128, 0, s>>>(ranges, large_tiles, T, ctrl, comp, vals, capacity,
                                                                   mid_ctas);
  tile_sort_large_kernel<<<148 * 2, 512, TS_LARGE_CAP * 8, s>>>(ranges, large_tiles, T, ctrl, comp, vals);
  count_launch(2);
  return cudaGetLastError();
}

// ---- export (tests only): the sorted 64-bit keys the reference's binning buffer would hold ----------
__global__ void __launch_bounds__(256) export_segment_keys_kernel(const uint2* __restrict__ ranges,
                                                                  const uint32_t* __restrict__ vals,
                                                                  const GsrRec* __restrict__ recs, uint32_t limit,
                                                                  unsigned long long* __restrict__ keys) {
  const uint32_t t = blockIdx.x;
  const uint2 r = ranges[t];
  for (uint32_t i = r.x + threadIdx.x; i < r.y && i < limit; i += 256)
    keys[i] = ((unsigned long long)t << 32) | (unsigned long long)__float_as_uint(recs[vals[i]].d.w);
}

cudaError_t launch_export_segment_keys(int T, const uint2* ranges, const uint32_t* vals, const GeomState& geom,
                                       uint32_t limit, unsigned long long* keys, cudaStream_t s) {
  if (T <= 0 || limit == 0) return cudaSuccess;
  export_segment_keys_kernel<<<T, 256, 0, s>>>(ranges, vals, geom.recs, limit, keys);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
