// render_backward.cu -- K8: per-tile backward of the alpha compositing for sm_100a.
//
// There is no in-tree restatement of this pass in the reference (its Python rasterizer is
// forward-only); the arithmetic is the analytic derivative of
// gs_lightning/rasterize/rasterize.py:182-261 as stated in oracle/gsr_oracle.cpp
// (gsro_render_backward), which float64 autograd pins in tests/test_oracle_backward.py.
//
// Per pixel, the tile's list is replayed BACK TO FRONT from the last contributor recorded by
// the forward (n_contrib), rebuilding T by division and carrying the colour / inverse-depth
// accumulated behind the current splat.
//
// Design (what differs from a per-pixel-atomics kernel):
//   * same CTA/warp/pixel mapping and batch staging as the forward kernel: 4 warps per tile, a warp owns an
//     8x8 block, a lane the pixels A and B of one column four rows apart, and all per-pixel arithmetic runs
//     on packed pairs {A, B} (FFMA2 / FMUL2 / FADD2, render_common.cuh) -- the second pixel costs no extra
//     issue slots for the FP32 work.  The hit words the forward wrote per 8x4 sub-tile say which list entries
//     touched it; the warp replays the UNION of its two sub-tiles' words and evaluates no culling test
//     (a pixel the entry does not reach recomputes alpha < 1/255 bit for bit and contributes exact zeros);
//   * per (pixel, splat) the kernel produces ten RAW sums instead of the ten finished partial
//     derivatives: with h = opacity * G * dL/dalpha,
//         S0 = h dx, S1 = h dy, S2 = h dx^2, S3 = h dx dy, S4 = h dy^2, S5 = G dL/dalpha,
//         S6..8 = w dL/dC_rgb, S9 = w dL/d(1/z)                       (w = alpha * T)
//     -- everything that multiplies these by per-Gaussian constants (conic, W/2, H/2, -1/2) is
//     applied once per Gaussian in the preprocess backward kernel, not once per pixel here;
//   * the loop carries -T, -alpha and -(accumulated colour) (no packed negation exists): every product and FMA
//     is the exact negation of the textbook one, so the sums are bit-identical to a positive-sign evaluation;
//   * lanes whose pixel the splat does not reach run the same instruction stream with
//     alpha = 0 (an exact no-op of the recurrences), so the hit loop has no divergent branch;
//   * a lane first adds its two pixels' sums and parks the ten sums in a per-warp transposition buffer in shared
//     memory (row = visit slot * 10 + sum, column = lane, row stride 36 floats: conflict-free 128-bit row reads);
//     every third visit the warp adds the 30 parked rows, one row per lane (8 LDS.128 + 31 adds), and 30 lanes issue
//     ONE red.global.add.f32 instruction into the 48-byte-per-Gaussian accumulator rows -- 1/64 of the atomics of
//     the per-pixel scheme.  Measured against a recursive-halving shuffle reduction (12 shuffles + 22 selects + 12
//     adds per visit): 0.873 -> 0.850 ms; a hybrid (one shuffle level, then shared memory) was slower (0.900 ms).
#ifdef GSR_BWD_BATCH
#define GSR_BATCH GSR_BWD_BATCH
#endif
#ifndef GSR_BWD_CTAS
#define GSR_BWD_CTAS 7
#endif
#include "gsr_common.cuh"
#include "render_common.cuh"

// GSR_BWD_PERWARP (default): every warp runs its own double-buffered pipeline over the tile's list, chunk of 32 entries by
// chunk, staging ONLY the entries its hit words name -- no CTA barrier anywhere.  The r2_v9 source-level capture had 10 %
// of all warp samples parked at the per-batch __syncthreads (the four warps of a tile see very different numbers of
// hits).  0: the CTA-wide batches of GSR_BATCH entries.
#ifndef GSR_BWD_PERWARP
#define GSR_BWD_PERWARP 1
#endif
// GSR_BWD_WARP_CTA (with GSR_BWD_PERWARP): one warp per CTA, four CTAs per tile.  Only ~1.5 of a tile's 4 warps have work
// on a given entry; as single-warp CTAs the idle ones retire at once and their slots go to warps that still work.
#ifndef GSR_BWD_WARP_CTA
#define GSR_BWD_WARP_CTA 1
#endif
#if GSR_BWD_PERWARP && GSR_BWD_WARP_CTA
#define GSR_BWD_THREADS 32
#else
#define GSR_BWD_THREADS 128
#endif
#define GSR_BWD_WARPS (GSR_BWD_THREADS / 32)
#if GSR_BWD_PERWARP
#define GSR_BWD_PLANE 512        /* 32 entries x 16 bytes */
#else
#define GSR_BWD_PLANE GSR_PLANE
#endif

namespace gsr {

__global__ void __launch_bounds__(GSR_BWD_THREADS, GSR_BWD_CTAS * (128 / GSR_BWD_THREADS)) render_backward_kernel(
    const GsrView v, const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
    const char* __restrict__ binning,
    const GsrRec* __restrict__ recs, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth, float* __restrict__ grad_acc) {
#if GSR_BWD_PERWARP
  __shared__ float4 sRec[GSR_BWD_WARPS][2][3][32];          // [warp][buffer][plane 0-2 of render_common.cuh][entry of the chunk]
#else
  __shared__ float4 sRec[2][3][GSR_BATCH];      // two staged batches, planes 0-2 of render_common.cuh
  __shared__ uint32_t sHit[2][GSR_BATCH / 32][8];   // [buffer][chunk of the batch][sub-tile column]
  __shared__ uint32_t s_warp_max[4];
#endif
  // reduction over the warp's lanes by transposition through shared memory: RED_STRIDE = 36 floats per row keeps the
  // rows 16-byte aligned and the 128-bit row reads of a quarter warp on 32 distinct banks
  constexpr int RED_STRIDE = 36;
  __shared__ __align__(16) float sRed[GSR_BWD_WARPS][30 * RED_STRIDE];
  __shared__ uint32_t sRedId[GSR_BWD_WARPS][4];

#if GSR_BWD_THREADS == 32
  const int lane = threadIdx.x, warp = blockIdx.x & 3, wslot = 0;      // the tile's warp = the CTA; its shared-memory slot 0
  const int tile_slot = blockIdx.x >> 2;
#else
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, wslot = warp;
  const int tile_slot = blockIdx.x;
#endif
  const int tile = tile_order ? (int)tile_order[tile_slot] : tile_slot;
  const int tx = tile % v.gx, ty = tile / v.gx;
  const int sx0 = tx * GSR_TILE + (warp & 1) * 8, sy0 = ty * GSR_TILE + (warp >> 1) * 8;
  const int pxi = sx0 + (lane & 7), pyA = sy0 + (lane >> 3), pyB = pyA + 4;
  const bool insideA = pxi < v.W && pyA < v.H, insideB = pxi < v.W && pyB < v.H;
  const float pxf = (float)pxi;
  const int colA = (warp >> 1) * 4 + (warp & 1), colB = colA + 2;
  const uint2 range = ranges[tile];
  // shared-memory addresses the hit loop uses, pinned in registers (keep_u32)
  const uint32_t aRed = (uint32_t)__cvta_generic_to_shared(sRed[wslot]);
  const uint32_t aPark0 = keep_u32(aRed + (uint32_t)lane * 4u);                       // column of this lane, slot 0
  const uint32_t aRow = keep_u32(aRed + (uint32_t)lane * (RED_STRIDE * 4u));           // row this lane adds at a flush
  const uint32_t aId = keep_u32((uint32_t)__cvta_generic_to_shared(sRedId[wslot]));
  const int red_visit = lane / 10, red_sum = lane - red_visit * 10;   // the row lane `lane` adds at a flush
  // (pinned: ptxas otherwise recomputes lane / 10, the remainder and the pointer at every flush)
  const unsigned long long acc_col = keep_u64((unsigned long long)__cvta_generic_to_global(grad_acc + red_sum));
  const int lane_pinned = (int)keep_u32((uint32_t)lane);
  const uint32_t aIdMine = keep_u32(aId + (uint32_t)red_visit * 4u);
  uint32_t aPark = aPark0;                                             // parking address of the next visit
  int nslot = 0;                                                       // visits parked in the buffer (warp-uniform)

// add the parked rows: lane l < 10 nslot owns row l = (visit, sum) and issues its red.global.add
#define GSR_BWD_FLUSH()                                                                                   \
  {                                                                                                       \
    __syncwarp();                                                                                         \
    if (lane_pinned < nslot * 10) {                                                                       \
      f32x2 a0, a1, a2, a3, a4, a5, a6, a7, a8, a9, a10, a11, a12, a13, a14, a15;                        \
      lds2x64<0>(aRow, a0, a1); lds2x64<16>(aRow, a2, a3); lds2x64<32>(aRow, a4, a5); lds2x64<48>(aRow, a6, a7);   \
      lds2x64<64>(aRow, a8, a9); lds2x64<80>(aRow, a10, a11); lds2x64<96>(aRow, a12, a13);                \
      lds2x64<112>(aRow, a14, a15);                                                                       \
      /* 32 floats as 16 packed pairs: 15 FADD2 + 1 FADD instead of 31 FADD */                            \
      a0 = add2(a0, a1); a2 = add2(a2, a3); a4 = add2(a4, a5); a6 = add2(a6, a7);                         \
      a8 = add2(a8, a9); a10 = add2(a10, a11); a12 = add2(a12, a13); a14 = add2(a14, a15);                \
      a0 = add2(a0, a2); a4 = add2(a4, a6); a8 = add2(a8, a10); a12 = add2(a12, a14);                     \
      a0 = add2(add2(a0, a4), add2(a8, a12));                                                             \
      float slo, shi;                                                                                     \
      upk2(a0, slo, shi);                                                                                 \
      const float sum = slo + shi;                                                                        \
      asm volatile("red.global.add.f32 [%0], %1;" ::"l"(acc_col + (unsigned long long)lds32u<0>(aIdMine) * 48ull), \
                   "f"(sum) : "memory");                                                                  \
    }                                                                                                     \
    __syncwarp();                                                                                         \
    nslot = 0;                                                                                            \
    aPark = aPark0;                                                                                       \
  }
  const uint32_t aRec = keep_u32((uint32_t)__cvta_generic_to_shared(sRec));

  // the binning buffer describes itself (BinningHeader): sorted point list + hit words
  const BinningHeader* hdr = reinterpret_cast<const BinningHeader*>(binning);
  const size_t cap = hdr->capacity;
  const uint32_t* __restrict__ point_list = reinterpret_cast<const uint32_t*>(binning + gsr_bin_vals_off());
  const uint32_t* __restrict__ hits = reinterpret_cast<const uint32_t*>(binning + gsr_bin_hits_off(cap));
  const uint32_t* __restrict__ hit_rows = hits + hit_slot(range.x, (uint32_t)tile, 0) * 8;

  const size_t HW = (size_t)v.H * v.W;
  float TfA = 0.f, TfB = 0.f, dpA[4] = {0.f, 0.f, 0.f, 0.f}, dpB[4] = {0.f, 0.f, 0.f, 0.f};
  uint32_t lastA = 0, lastB = 0;
  if (insideA) {
    const size_t pid = (size_t)pyA * v.W + pxi;
    TfA = final_T[pid];
    lastA = n_contrib[pid];
    dpA[0] = dL_dpix[pid]; dpA[1] = dL_dpix[HW + pid]; dpA[2] = dL_dpix[2 * HW + pid];
    if (dL_dinvdepth) dpA[3] = dL_dinvdepth[pid];
  }
  if (insideB) {
    const size_t pid = (size_t)pyB * v.W + pxi;
    TfB = final_T[pid];
    lastB = n_contrib[pid];
    dpB[0] = dL_dpix[pid]; dpB[1] = dL_dpix[HW + pid]; dpB[2] = dL_dpix[2 * HW + pid];
    if (dL_dinvdepth) dpB[3] = dL_dinvdepth[pid];
  }
  const float* __restrict__ bg = v.bg;
  // per-pixel state and constants, {A, B} (ptxas keeps each pair in an aligned 64-bit register so that the packed
  // path uses them in place)
  const float dp0A = dpA[0], dp1A = dpA[1], dp2A = dpA[2], dp3A = dpA[3];
  const float dp0B = dpB[0], dp1B = dpB[1], dp2B = dpB[2], dp3B = dpB[3];
  const float TfbgA = TfA * (bg[0] * dp0A + bg[1] * dp1A + bg[2] * dp2A);
  const float TfbgB = TfB * (bg[0] * dp0B + bg[1] * dp1B + bg[2] * dp2B);
  const float npyA = -(float)pyA, npyB = -(float)pyB;

  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, max(lastA, lastB));
#if GSR_BWD_PERWARP
  if (warp_last == 0) return;      // the warps of a tile are independent: no barrier follows
#else
  if (lane == 0) s_warp_max[warp] = warp_last;
  __syncthreads();
  uint32_t max_last = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) max_last = max(max_last, s_warp_max[w]);
  if (max_last == 0) return;
#endif

  float nTA = -TfA, nTB = -TfB;                                  // -T
  float nar0A = 0.f, nar1A = 0.f, nar2A = 0.f, nar3A = 0.f;      // -(colour / inverse depth accumulated behind the splat)
  float nar0B = 0.f, nar1B = 0.f, nar2B = 0.f, nar3B = 0.f;

// one pixel of the lane against the splat just loaded (scalar path: the entry is in one sub-tile's hit word only)
#define GSR_BWD_ONE(S)                                                                                   \
  {                                                                                                      \
    const float dy = my + npy##S;                                                                        \
    float e = __fmaf_rn(__fmul_rn(hC, dy), dy, tt);                                                      \
    e = __fmaf_rn(vv, dy, e);                                                                            \
    const float G = ex2_approx(-e);                                                                      \
    float na = fmaxf(__fmul_rn(no, G), -0.99f);                                                          \
    const bool valid = (idx < last##S) && !(e < 0.f) && (na <= -1.f / 255.f);                            \
    na = valid ? na : 0.f;                                                                               \
    const float rcp = rcp_approx(1.f + na);                                                              \
    nT##S *= rcp;                                                                                        \
    const float w = na * nT##S;                                                                          \
    const float d0 = cr + nar0##S, d1 = cg + nar1##S, d2 = cb + nar2##S, d3 = iz + nar3##S;              \
    float dL = d0 * dp0##S;                                                                              \
    dL = fmaf(d1, dp1##S, dL);                                                                           \
    dL = fmaf(d2, dp2##S, dL);                                                                           \
    dL = fmaf(d3, dp3##S, dL);                                                                           \
    dL = fmaf(dL, nT##S, Tfbg##S * rcp);                                                                 \
    nar0##S = fmaf(na, d0, nar0##S);                                                                     \
    nar1##S = fmaf(na, d1, nar1##S);                                                                     \
    nar2##S = fmaf(na, d2, nar2##S);                                                                     \
    nar3##S = fmaf(na, d3, nar3##S);                                                                     \
    const float n5 = valid ? G * dL : 0.f;                                                               \
    const float h = no * n5;                                                                             \
    s0 = h * dx; s1 = h * dy;                                                                            \
    s2 = s0 * dx; s3 = s0 * dy; s4 = s1 * dy;                                                            \
    s5 = -n5;                                                                                            \
    s6 = w * dp0##S; s7 = w * dp1##S; s8 = w * dp2##S; s9 = w * dp3##S;                                  \
  }

#if GSR_BWD_PERWARP
  // The warp's own pipeline, from the chunk holding its last contributor down to chunk 0.  Per chunk: the two hit
  // words of the warp's sub-tiles and the lane's list entry are loaded two chunks ahead (plain loads), the records of the
  // entries the words name are copied one chunk ahead (three 16-byte cp.async by the lane whose bit is set) into the
  // warp's other buffer, and the current chunk is replayed from shared memory.  __syncwarp is the only synchronisation.
  const uint32_t* __restrict__ plist = point_list + range.x;
  const uint32_t nlist = range.y - range.x;
  const uint32_t aW = keep_u32(aRec + (uint32_t)wslot * (2u * 3u * GSR_BWD_PLANE));
  const uint32_t aMineW = aW + (uint32_t)lane * 16u;
  auto issue_chunk = [&](int buf, uint32_t m, uint32_t gidx) {
    if ((m >> lane) & 1u) {
      const GsrRec* r = recs + gidx;
      const uint32_t dst = aMineW + (uint32_t)buf * (3u * GSR_BWD_PLANE);
      cp_async16(dst, &r->a);
      cp_async16(dst + GSR_BWD_PLANE, &r->b);
      cp_async16(dst + 2 * GSR_BWD_PLANE, &r->c);
    }
    cp_async_commit();
  };
  int c = (int)((warp_last - 1u) >> 5);
  uint32_t maskA = hit_rows[(size_t)c * 8 + colA], maskB = hit_rows[(size_t)c * 8 + colB];
  {
    const uint32_t e = (uint32_t)c * 32u + (uint32_t)lane;
    issue_chunk(0, maskA | maskB, e < nlist ? plist[e] : 0u);
  }
  uint32_t nA = 0, nB = 0, nidx = 0;      // chunk c - 1
  if (c > 0) {
    nA = hit_rows[(size_t)(c - 1) * 8 + colA]; nB = hit_rows[(size_t)(c - 1) * 8 + colB];
    nidx = plist[(uint32_t)(c - 1) * 32u + (uint32_t)lane];
  }
  for (int k = 0; c >= 0; --c, ++k) {
    const int base = c * 32;
    cp_async_wait_all();     // this lane's copies of chunk c have landed ...
    __syncwarp();            // ... every lane's have, and the warp is done with the other buffer
    uint32_t tA = 0, tB = 0, tidx = 0;     // chunk c - 2
    if (c > 0) {
      issue_chunk((k + 1) & 1, nA | nB, nidx);
      if (c > 1) {
        tA = hit_rows[(size_t)(c - 2) * 8 + colA]; tB = hit_rows[(size_t)(c - 2) * 8 + colB];
        tidx = plist[(uint32_t)(c - 2) * 32u + (uint32_t)lane];
      }
    }
    const uint32_t aBuf = keep_u32(aW + (uint32_t)(k & 1) * (3u * GSR_BWD_PLANE));
    {
      const int c0 = 0;
#else
  // Batches of GSR_BATCH entries from the back of the list, double buffered (render_common.cuh): a thread copies ONE
  // entry per batch (three 16-byte cp.async from the entry's record; its Gaussian index was loaded a batch earlier) and
  // the first 32 threads one hit word each, so the copies of the next batch are in flight while the warps replay this
  // one; one barrier per batch.
  const int nb = ((int)max_last + GSR_BATCH - 1) / GSR_BATCH;
  const uint32_t* __restrict__ plist = point_list + range.x;
  const uint32_t aMine = aRec + (uint32_t)tid * 16u;
  const uint32_t aHit = (uint32_t)__cvta_generic_to_shared(sHit);
  const uint32_t aHitA = keep_u32(aHit + (uint32_t)colA * 4u);   // column A of chunk 0, buffer 0 (column B = +8 bytes)
  auto issue = [&](int b, int k, uint32_t gidx) {     // batch b (entries b * GSR_BATCH ...) into buffer k & 1
    if (tid < GSR_BATCH && b * GSR_BATCH + tid < (int)max_last) {
      const GsrRec* r = recs + gidx;
      const uint32_t dst = aMine + (uint32_t)(k & 1) * (3u * GSR_PLANE);
      cp_async16(dst, &r->a);
      cp_async16(dst + GSR_PLANE, &r->b);
      cp_async16(dst + 2 * GSR_PLANE, &r->c);
    }
    // hit words of the batch's chunks: [chunk of the batch][sub-tile column]
    if (tid < (GSR_BATCH / 32) * 8 && b * GSR_BATCH + (tid >> 3) * 32 < (int)max_last)
      cp_async4(aHit + (uint32_t)(k & 1) * (GSR_BATCH / 32) * 32u + (uint32_t)tid * 4u,
                hit_rows + (size_t)(b * (GSR_BATCH / 32) + (tid >> 3)) * 8 + (tid & 7));
    cp_async_commit();
  };
  uint32_t gnext = 0;
  {
    const int e0 = (nb - 1) * GSR_BATCH + tid;
    issue(nb - 1, 0, (tid < GSR_BATCH && e0 < (int)max_last) ? plist[e0] : 0u);
    if (nb > 1 && tid < GSR_BATCH) gnext = plist[(nb - 2) * GSR_BATCH + tid];
  }
  for (int b = nb - 1, k = 0; b >= 0; --b, ++k) {
    const int base = b * GSR_BATCH;
    const int cnt = min(GSR_BATCH, (int)max_last - base);
    cp_async_wait_all();     // this thread's copies of batch b have landed ...
    __syncthreads();         // ... everybody's have, and buffer (k + 1) & 1 (read by the previous batch) is free
    if (b > 0) {
      issue(b - 1, k + 1, gnext);
      if (b > 1 && tid < GSR_BATCH) gnext = plist[(b - 2) * GSR_BATCH + tid];
    }
    if ((int)warp_last <= base) continue;  // nothing in this batch reaches the warp's pixels
    const uint32_t aBuf = keep_u32(aRec + (uint32_t)(k & 1) * (3u * GSR_PLANE));
    const uint32_t aHitBuf = keep_u32(aHitA + (uint32_t)(k & 1) * (GSR_BATCH / 32) * 32u);

    for (int c0 = ((min(cnt, (int)warp_last - base) - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
      const uint32_t maskA = lds32u<0>(aHitBuf + (uint32_t)c0), maskB = lds32u<8>(aHitBuf + (uint32_t)c0);  // (c0 / 32) * 32 B
#endif
      uint32_t mask = maskA | maskB;
      const uint32_t both = maskA & maskB;
      while (mask) {
        const int bit = msb_index(mask);
        const uint32_t bitm = 1u << bit;
        mask ^= bitm;
        const int j = c0 + bit;
        const uint32_t idx = (uint32_t)(base + j);
        const uint32_t addr = aBuf + (uint32_t)j * 16u;
        const float4 ra = lds128<0>(addr);                    // {mean.x, mean.y, hA, hB}
        const float4 rb = lds128<GSR_BWD_PLANE>(addr);            // {hC, -opacity, r, g}
        const float4 rc = lds128<2 * GSR_BWD_PLANE>(addr);        // {b, 1/z, tau2, bits(gaussian id)}
        const float mx = ra.x, my = ra.y, hA = ra.z, hB = ra.w, hC = rb.x, no = rb.y, cr = rb.z, cg = rb.w, cb = rc.x,
                    iz = rc.y;
        const float dx = mx - pxf;
        const float tt = __fmul_rn(__fmul_rn(hA, dx), dx);
        const float vv = __fmul_rn(hB, dx);
        float s0, s1, s2, s3, s4, s5, s6, s7, s8, s9;
        if (both & bitm) {
          // packed path: both pixels in one instruction stream
          const f32x2 dx2 = pk2(dx, dx);
          const f32x2 dy2 = add2(pk2(my, my), pk2(npyA, npyB));
          f32x2 e2 = fma2(mul2(pk2(hC, hC), dy2), dy2, pk2(tt, tt));
          e2 = fma2(pk2(vv, vv), dy2, e2);
          float eA, eB;
          upk2(e2, eA, eB);
          const f32x2 G2 = pk2(ex2_approx(-eA), ex2_approx(-eB));
          float naA, naB;
          const f32x2 no2 = pk2(no, no);
          upk2(mul2(no2, G2), naA, naB);
          naA = fmaxf(naA, -0.99f);
          naB = fmaxf(naB, -0.99f);
          const bool validA = (idx < lastA) && !(eA < 0.f) && (naA <= -1.f / 255.f);
          const bool validB = (idx < lastB) && !(eB < 0.f) && (naB <= -1.f / 255.f);
          const f32x2 na2 = pk2(validA ? naA : 0.f, validB ? naB : 0.f);   // -alpha, 0 where the splat does not reach

          float omA, omB;
          upk2(add2(pk2(1.f, 1.f), na2), omA, omB);                         // 1 - alpha
          const f32x2 rcp2 = pk2(rcp_approx(omA), rcp_approx(omB));
          const f32x2 nT2 = mul2(pk2(nTA, nTB), rcp2);                      // -T in front of this splat
          upk2(nT2, nTA, nTB);
          const f32x2 w2 = mul2(na2, nT2);                                  // alpha T
          // d = colour - (colour accumulated BEHIND this splat)
          const f32x2 d0 = add2(pk2(cr, cr), pk2(nar0A, nar0B)), d1 = add2(pk2(cg, cg), pk2(nar1A, nar1B)),
                      d2 = add2(pk2(cb, cb), pk2(nar2A, nar2B)), d3 = add2(pk2(iz, iz), pk2(nar3A, nar3B));
          const f32x2 dp0 = pk2(dp0A, dp0B), dp1 = pk2(dp1A, dp1B), dp2 = pk2(dp2A, dp2B), dp3 = pk2(dp3A, dp3B);
          f32x2 dL = mul2(d0, dp0);
          dL = fma2(d1, dp1, dL);
          dL = fma2(d2, dp2, dL);
          dL = fma2(d3, dp3, dL);
          dL = fma2(dL, nT2, mul2(pk2(TfbgA, TfbgB), rcp2));                // -(dL/dalpha)
          // fold this splat in for the next (nearer) one: -ar <- -(alpha c + (1 - alpha) ar)
          upk2(fma2(na2, d0, pk2(nar0A, nar0B)), nar0A, nar0B);
          upk2(fma2(na2, d1, pk2(nar1A, nar1B)), nar1A, nar1B);
          upk2(fma2(na2, d2, pk2(nar2A, nar2B)), nar2A, nar2B);
          upk2(fma2(na2, d3, pk2(nar3A, nar3B)), nar3A, nar3B);

          float n5A, n5B;
          upk2(mul2(G2, dL), n5A, n5B);                                     // -(G dL/dalpha)
          const f32x2 n5 = pk2(validA ? n5A : 0.f, validB ? n5B : 0.f);
          const f32x2 h2 = mul2(no2, n5);                                   // opacity G dL/dalpha
          const f32x2 p0 = mul2(h2, dx2), p1 = mul2(h2, dy2);
          const f32x2 p2 = mul2(p0, dx2), p3 = mul2(p0, dy2), p4 = mul2(p1, dy2);
          const f32x2 p6 = mul2(w2, dp0), p7 = mul2(w2, dp1), p8 = mul2(w2, dp2), p9 = mul2(w2, dp3);
          // the lane's two pixels first
          float lo, hi;
          upk2(p0, lo, hi); s0 = lo + hi;
          upk2(p1, lo, hi); s1 = lo + hi;
          upk2(p2, lo, hi); s2 = lo + hi;
          upk2(p3, lo, hi); s3 = lo + hi;
          upk2(p4, lo, hi); s4 = lo + hi;
          upk2(n5, lo, hi); s5 = -lo - hi;
          upk2(p6, lo, hi); s6 = lo + hi;
          upk2(p7, lo, hi); s7 = lo + hi;
          upk2(p8, lo, hi); s8 = lo + hi;
          upk2(p9, lo, hi); s9 = lo + hi;
        } else {
          if (maskA & bitm) GSR_BWD_ONE(A) else GSR_BWD_ONE(B)
        }

        // park the lane's ten sums in the warp's transposition buffer (row = visit slot * 10 + sum, column = lane);
        // every third visit the warp adds the 30 rows, one row per lane (GSR_BWD_FLUSH)
        {
          const uint32_t dst = aPark;
          sts32<0 * RED_STRIDE * 4>(dst, s0); sts32<1 * RED_STRIDE * 4>(dst, s1); sts32<2 * RED_STRIDE * 4>(dst, s2);
          sts32<3 * RED_STRIDE * 4>(dst, s3); sts32<4 * RED_STRIDE * 4>(dst, s4); sts32<5 * RED_STRIDE * 4>(dst, s5);
          sts32<6 * RED_STRIDE * 4>(dst, s6); sts32<7 * RED_STRIDE * 4>(dst, s7); sts32<8 * RED_STRIDE * 4>(dst, s8);
          sts32<9 * RED_STRIDE * 4>(dst, s9);
          asm volatile("st.shared.f32 [%0], %1;" : : "r"(aId + (uint32_t)nslot * 4u), "f"(rc.w) : "memory");  // every lane
          // stores the same word (the Gaussian index): one wavefront, no predicate
          aPark += 10u * RED_STRIDE * 4u;
          if (++nslot == 3) GSR_BWD_FLUSH()
        }
      }
    }
#if GSR_BWD_PERWARP
    maskA = nA; maskB = nB; nA = tA; nB = tB; nidx = tidx;
#endif
  }
  if (nslot) GSR_BWD_FLUSH()
  cp_async_wait_all();
#undef GSR_BWD_FLUSH
#undef GSR_BWD_ONE
}

cudaError_t launch_render_backward(const GsrView& v, const uint2* ranges, const uint32_t* tile_order, const void* binning,
                                   const GsrRec* recs,
                                   const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                                   const float* dL_dinvdepth, float* grad_acc, cudaStream_t s) {
  const int tiles = v.gx * v.gy;
  if (tiles <= 0) return cudaSuccess;
  render_backward_kernel<<<tiles * (128 / GSR_BWD_THREADS), GSR_BWD_THREADS, 0, s>>>(v, ranges, tile_order, (const char*)binning, recs, final_T, n_contrib, dL_dpix,
                                               dL_dinvdepth, grad_acc);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
