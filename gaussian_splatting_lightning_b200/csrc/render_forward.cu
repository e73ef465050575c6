// render_forward.cu -- K7: per-16x16-tile forward alpha compositing for sm_100a.
//
// Semantics (gs_lightning/rasterize/rasterize.py:182-261, render_tools.py:141-144, with the
// CUDA-side rules of SURVEY.md Appendix B3-B5): per pixel, front to back over the tile's
// depth-sorted list,
//     d = mean2D - pixel;  power = -1/2 (A dx^2 + C dy^2) - B dx dy;   skip if power > 0
//     alpha = min(0.99, opacity * exp(power));                          skip if alpha < 1/255
//     stop BEFORE the splat if T (1 - alpha) < 1e-4
//     C += colour * alpha * T;  D += alpha * T / z;  T *= (1 - alpha)
// out = C + T * bg, inverse depth D, final_T, n_contrib (1-based index of the last splat used).
//
// Design:
//   * one CTA (4 warps) per tile; warp w owns the 8x8 pixel block (w&1, w>>1) and every lane TWO pixels of one
//     column, four rows apart: pixel A in the block's upper 8x4 sub-tile, pixel B in the lower one.  All
//     per-pixel arithmetic runs on packed pairs {A, B} (FFMA2 / FMUL2 / FADD2, render_common.cuh): the second
//     pixel costs no extra issue slots for the FP32 work, and these kernels are issue bound;
//   * the tile's list is consumed in batches of 256: each thread gathers two packed 48-byte GsrRec (three
//     128-bit loads each, L2-resident at 1M Gaussians) and stages them in shared memory with the quadratic
//     form pre-scaled into the log2 domain (render_common.cuh; per-splat values enter the packed instructions as
//     broadcast scalar operands), so that the inner loop is 8 FP32 issue slots + two MUFU.EX2 per (pixel pair, splat);
//   * sub-tile culling: before touching a chunk of 32 entries, each lane tests ONE splat against each of
//     the warp's two 8x4 rectangles (exact minimum of the quadratic form over the rectangle vs.
//     tau2 = log2(255 * opacity), conservative margin) and two ballots turn the answers into bit masks; the
//     warp iterates over the UNION of the two masks.  A splat rejected for a sub-tile has alpha < 1/255 on
//     every pixel of it, so evaluating it there anyway (when the other sub-tile needs it) changes nothing;
//   * the hit loop is branch-free per lane (predicated updates); it carries -T and -alpha (see
//     render_common.cuh) -- every product and FMA is the exact negation of the textbook one;
//   * for every chunk of 32 list entries and each 8x4 sub-tile the warp records which entries contributed to
//     at least one of the sub-tile's pixels (one 32-bit "hit word" per sub-tile and chunk): the backward
//     kernel replays exactly those and needs no culling test;
//   * warp ballots retire a warp as soon as its 64 pixels are saturated; the CTA leaves the
//     list when __syncthreads_and says every pixel is done.
#include "gsr_common.cuh"
#include "render_common.cuh"

// GSR_FWD_PERWARP = 1 (measured, NOT the default): one warp per CTA, four CTAs per tile, the warp staging the tile's list
// itself chunk of 32 entries by chunk -- the change that took render_backward from 0.744 to 0.698 ms.  Here every entry
// has to be staged by each of the tile's four warps (the culling test needs all of them, the backward stages only what
// its hit words name): 4 x the L2 -> shared traffic, 0.374 -> 0.410 ms on C3.  Default 0: CTA-wide batches of GSR_BATCH.
#ifndef GSR_FWD_PERWARP
#define GSR_FWD_PERWARP 0
#endif
#if GSR_FWD_PERWARP
#define GSR_FWD_THREADS 32
#define GSR_FWD_PLANE 512          /* 32 entries x 16 bytes */
#define GSR_FWD_MINCTAS 32
#else
#define GSR_FWD_THREADS 128
#define GSR_FWD_PLANE GSR_PLANE
#define GSR_FWD_MINCTAS 9
#endif

namespace gsr {

__global__ void __launch_bounds__(GSR_FWD_THREADS, GSR_FWD_MINCTAS) render_forward_kernel(const GsrView v, const uint2* __restrict__ ranges,
                                                             const uint32_t* __restrict__ tile_order,
                                                             const uint32_t* __restrict__ point_list,
                                                             const GsrRec* __restrict__ recs,
                                                             float* __restrict__ out_color,
                                                             float* __restrict__ out_invdepth,
                                                             float* __restrict__ final_T,
                                                             uint32_t* __restrict__ n_contrib,
                                                             uint32_t* __restrict__ hits) {
#if GSR_FWD_PERWARP
  __shared__ float4 sRec[2][4][32];         // the warp's two staged chunks of four planes; plane 3 = cull plane
  const int lane = threadIdx.x, warp = blockIdx.x & 3;
  const int tile_slot = blockIdx.x >> 2;
#else
  __shared__ float4 sRec[2][4][GSR_BATCH];  // two staged batches of four planes (render_common.cuh); plane 3 = cull plane
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#define tile_slot blockIdx.x
#endif
  const int tile = tile_order ? (int)tile_order[tile_slot] : (int)tile_slot;
#undef tile_slot
  const int tx = tile % v.gx, ty = tile / v.gx;
  // warp -> 8x8 block = sub-tiles A (upper 8x4) and B (lower 8x4); lane -> one pixel in each
  const int sx0 = tx * GSR_TILE + (warp & 1) * 8, sy0 = ty * GSR_TILE + (warp >> 1) * 8;
  const int pxi = sx0 + (lane & 7), pyA = sy0 + (lane >> 3), pyB = pyA + 4;
  const bool insideA = pxi < v.W && pyA < v.H, insideB = pxi < v.W && pyB < v.H;
  const float pxf = (float)pxi;
  const SubTile stA = make_subtile(sx0, sy0, v.W, v.H), stB = make_subtile(sx0, sy0 + 4, v.W, v.H);
  const uint32_t aRec = (uint32_t)__cvta_generic_to_shared(sRec);
  // hit-word columns of the two sub-tiles (8 per chunk: sub-tile (x, y) of the tile at column 2y + x)
  const int colA = (warp >> 1) * 4 + (warp & 1), colB = colA + 2;

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  uint32_t* __restrict__ hit_row = hits + hit_slot(range.x, (uint32_t)tile, 0) * 8;

  // per-pixel state, {A, B}; ptxas keeps each pair in an aligned 64-bit register so that the packed path below
  // uses them in place.  thr = the alpha floor of the pixel, -inf once the pixel is saturated (or outside).
  const float NEG_INF = __int_as_float(0xff800000);
  float nTA = -1.f, nTB = -1.f;                 // -T
  float c0A = 0.f, c0B = 0.f, c1A = 0.f, c1B = 0.f, c2A = 0.f, c2B = 0.f, dA = 0.f, dB = 0.f;
  float thrA = insideA ? -1.f / 255.f : NEG_INF, thrB = insideB ? -1.f / 255.f : NEG_INF;
  uint32_t lastA = 0, lastB = 0;
  const float npyA = -(float)pyA, npyB = -(float)pyB;

// one pixel of the lane against the splat just loaded (scalar path: the splat reaches only one sub-tile)
#define GSR_FWD_ONE(S)                                                                      \
  {                                                                                         \
    const float dy = my + npy##S;                                                           \
    float e = __fmaf_rn(__fmul_rn(hC, dy), dy, tt);                                         \
    e = __fmaf_rn(vv, dy, e);                                                               \
    const float na = fmaxf(__fmul_rn(no, ex2_approx(-e)), -0.99f);                          \
    const float test = __fmaf_rn(nT##S, na, nT##S);                                         \
    bool ok = !(e < 0.f) && (na <= thr##S);                                                 \
    const bool sat = test > -0.0001f;                                                       \
    thr##S = (ok && sat) ? NEG_INF : thr##S;                                                \
    ok = ok && !sat;                                                                        \
    const float w = ok ? __fmul_rn(na, nT##S) : 0.f;                                        \
    c0##S = fmaf(cr, w, c0##S);                                                             \
    c1##S = fmaf(cg, w, c1##S);                                                             \
    c2##S = fmaf(cb, w, c2##S);                                                             \
    d##S = fmaf(iz, w, d##S);                                                               \
    nT##S = ok ? test : nT##S;                                                              \
    ok##S##_bits |= ok ? bitm : 0u;                                                         \
  }

#if GSR_FWD_PERWARP
  const int nchunk = (n + 31) >> 5;
  const uint32_t* __restrict__ plist = point_list + range.x;
  const uint32_t aMine = aRec + (uint32_t)lane * 16u;
  auto issue = [&](int c, uint32_t gidx) {                      // stage entry c * 32 + lane into buffer c & 1
    const uint32_t dst = aMine + (uint32_t)(c & 1) * (4u * GSR_FWD_PLANE);
    if (c * 32 + lane < n) {
      const GsrRec* r = recs + gidx;
      cp_async16(dst, &r->a);
      cp_async16(dst + GSR_FWD_PLANE, &r->b);
      cp_async16(dst + 2 * GSR_FWD_PLANE, &r->c);
      cp_async16(dst + 3 * GSR_FWD_PLANE, &r->d);
    } else {
      // padding: never a hit.  The geometry planes the culling test reads are zeroed too: stale shared memory
      // (another kernel's data) may hold -inf / NaN patterns, and -inf <= -inf would let a padded entry pass.
      sts128(dst, 0.f, 0.f, 0.f, 0.f);
      sts128(dst + GSR_FWD_PLANE, 0.f, 0.f, 0.f, 0.f);
      sts128(dst + 3 * GSR_FWD_PLANE, 0.f, 0.f, NEG_INF, 0.f);
    }
    cp_async_commit();
  };
  uint32_t gnext = 0;
  if (nchunk > 0) {
    issue(0, lane < n ? plist[lane] : 0u);
    if (32 + lane < n) gnext = plist[32 + lane];
  }
  for (int c = 0; c < nchunk; ++c) {
    const int base = c * 32;
    cp_async_wait_all();     // this lane's copies of chunk c have landed ...
    __syncwarp();            // ... every lane's have, and the warp is done with the other buffer
    if (c + 1 < nchunk) {
      issue(c + 1, gnext);
      if (base + 64 + lane < n) gnext = plist[base + 64 + lane];
    }
    const uint32_t aBuf = aRec + (uint32_t)(c & 1) * (4u * GSR_FWD_PLANE);
    const float4* sGeo = &sRec[c & 1][0][0];
    const float4* sHC = &sRec[c & 1][1][0];
    const float4* sCull = &sRec[c & 1][3][0];
    {
      const int c0 = 0;
#else
  // Batches of GSR_BATCH entries, double buffered: a thread copies ONE entry per batch (four 16-byte cp.async from the
  // entry's 64-byte record, whose Gaussian index it loaded a batch earlier), so the copies of batch b + 1 and the index
  // loads of batch b + 2 are in flight while the warps composite batch b; one barrier per batch.
  const int nb = (n + GSR_BATCH - 1) / GSR_BATCH;
  const uint32_t* __restrict__ plist = point_list + range.x;
  const uint32_t aMine = aRec + (uint32_t)tid * 16u;            // this thread's entry slot in buffer 0, plane 0
  auto issue = [&](int b, uint32_t gidx) {                       // stage entry b * GSR_BATCH + tid into buffer b & 1
    const uint32_t dst = aMine + (uint32_t)(b & 1) * (4u * GSR_PLANE);
    if (b * GSR_BATCH + tid < n) {
      const GsrRec* r = recs + gidx;
      cp_async16(dst, &r->a);
      cp_async16(dst + GSR_PLANE, &r->b);
      cp_async16(dst + 2 * GSR_PLANE, &r->c);
      cp_async16(dst + 3 * GSR_PLANE, &r->d);
    } else {
      // padding: never a hit.  The geometry planes the culling test reads are zeroed too: stale shared memory
      // (another kernel's data) may hold -inf / NaN patterns, and -inf <= -inf would let a padded entry pass.
      sts128(dst, 0.f, 0.f, 0.f, 0.f);
      sts128(dst + GSR_PLANE, 0.f, 0.f, 0.f, 0.f);
      sts128(dst + 3 * GSR_PLANE, 0.f, 0.f, NEG_INF, 0.f);
    }
    cp_async_commit();
  };
  uint32_t gnext = 0;
  if (nb > 0) {
    issue(0, tid < n ? plist[tid] : 0u);
    if (GSR_BATCH + tid < n) gnext = plist[GSR_BATCH + tid];
  }
  for (int b = 0; b < nb; ++b) {
    const int base = b * GSR_BATCH;
    cp_async_wait_all();                                         // this thread's copies of batch b have landed ...
    if (__syncthreads_and(thrA == NEG_INF && thrB == NEG_INF)) break;   // ... and after the barrier everybody's have
    if (b + 1 < nb) {                                            // buffer (b + 1) & 1 was read by batch b - 1: free now
      issue(b + 1, gnext);
      if (base + 2 * GSR_BATCH + tid < n) gnext = plist[base + 2 * GSR_BATCH + tid];
    }
    const int cnt = min(GSR_BATCH, n - base);
    if (__all_sync(0xffffffffu, thrA == NEG_INF && thrB == NEG_INF)) continue;  // warp saturated: only the barriers
    const uint32_t aBuf = aRec + (uint32_t)(b & 1) * (4u * GSR_PLANE);
    const float4* sGeo = &sRec[b & 1][0][0];
    const float4* sHC = &sRec[b & 1][1][0];
    const float4* sCull = &sRec[b & 1][3][0];

    for (int c0 = 0; c0 < cnt; c0 += 32) {
#endif
      // one splat per lane against the warp's two sub-tiles
      const float4 geo = sGeo[c0 + lane], cull = sCull[c0 + lane];   // {mean.x, mean.y, hA, hB}, {kx, ky, tauM, -}
      const float hCl = sHC[c0 + lane].x;
      const bool allA = __all_sync(0xffffffffu, thrA == NEG_INF), allB = __all_sync(0xffffffffu, thrB == NEG_INF);
      if (allA && allB) break;
      const uint32_t maskA = __ballot_sync(0xffffffffu, !allA && subtile_may_contribute(geo, hCl, cull, stA));
      const uint32_t maskB = __ballot_sync(0xffffffffu, !allB && subtile_may_contribute(geo, hCl, cull, stB));
      uint32_t mask = maskA | maskB;
      const uint32_t both = maskA & maskB;
      uint32_t okA_bits = 0, okB_bits = 0;  // per lane: entries of the chunk that contributed to pixel A / B
      while (mask) {
        const int bit = __ffs(mask) - 1;
        const uint32_t bitm = mask & (0u - mask);
        mask ^= bitm;
        const uint32_t addr = aBuf + (uint32_t)(c0 + bit) * 16u;
        const float4 ra = lds128<0>(addr);                    // {mean.x, mean.y, hA, hB}
        const float4 rb = lds128<GSR_FWD_PLANE>(addr);            // {hC, -opacity, r, g}
        const float2 rc = lds64<2 * GSR_FWD_PLANE>(addr);         // {b, 1/z}
        const float mx = ra.x, my = ra.y, hA = ra.z, hB = ra.w, hC = rb.x, no = rb.y, cr = rb.z, cg = rb.w, cb = rc.x,
                    iz = rc.y;
        // e = hA dx^2 + hC dy^2 + hB dx dy, per pixel exactly quad_form()'s rounding sequence
        const float dx = mx - pxf;
        const float tt = __fmul_rn(__fmul_rn(hA, dx), dx);
        const float vv = __fmul_rn(hB, dx);
        if (both & bitm) {
          // packed path: both pixels in one instruction stream
          const f32x2 dy2 = add2(pk2(my, my), pk2(npyA, npyB));
          f32x2 e2 = fma2(mul2(pk2(hC, hC), dy2), dy2, pk2(tt, tt));
          e2 = fma2(pk2(vv, vv), dy2, e2);
          float eA, eB;
          upk2(e2, eA, eB);
          float naA, naB;
          upk2(mul2(pk2(no, no), pk2(ex2_approx(-eA), ex2_approx(-eB))), naA, naB);   // (-opacity) G
          naA = fmaxf(naA, -0.99f);
          naB = fmaxf(naB, -0.99f);
          const f32x2 na2 = pk2(naA, naB), nT2 = pk2(nTA, nTB);
          float tA, tB, wA, wB;
          upk2(fma2(nT2, na2, nT2), tA, tB);        // -(T (1 - alpha))
          upk2(mul2(na2, nT2), wA, wB);             // alpha T
          bool okA = !(eA < 0.f) && (naA <= thrA);
          bool okB = !(eB < 0.f) && (naB <= thrB);
          const bool satA = tA > -0.0001f, satB = tB > -0.0001f;   // T (1 - alpha) < 1e-4: stop BEFORE this splat
          thrA = (okA && satA) ? NEG_INF : thrA;
          thrB = (okB && satB) ? NEG_INF : thrB;
          okA = okA && !satA;
          okB = okB && !satB;
          const f32x2 w2 = pk2(okA ? wA : 0.f, okB ? wB : 0.f);
          upk2(fma2(pk2(cr, cr), w2, pk2(c0A, c0B)), c0A, c0B);
          upk2(fma2(pk2(cg, cg), w2, pk2(c1A, c1B)), c1A, c1B);
          upk2(fma2(pk2(cb, cb), w2, pk2(c2A, c2B)), c2A, c2B);
          upk2(fma2(pk2(iz, iz), w2, pk2(dA, dB)), dA, dB);
          nTA = okA ? tA : nTA;
          nTB = okB ? tB : nTB;
          okA_bits |= okA ? bitm : 0u;
          okB_bits |= okB ? bitm : 0u;
        } else {
          if (maskA & bitm) GSR_FWD_ONE(A) else GSR_FWD_ONE(B)
        }
      }
      const uint32_t usedA = __reduce_or_sync(0xffffffffu, okA_bits), usedB = __reduce_or_sync(0xffffffffu, okB_bits);
      if (lane == 0) {
        uint32_t* row = hit_row + (size_t)((base + c0) >> 5) * 8;
        row[colA] = usedA;
        row[colB] = usedB;
      }
      if (okA_bits) lastA = (uint32_t)(base + c0 + msb_index(okA_bits) + 1);
      if (okB_bits) lastB = (uint32_t)(base + c0 + msb_index(okB_bits) + 1);
    }
  }
#undef GSR_FWD_ONE
  cp_async_wait_all();   // a CTA that left the list early must not retire under its own in-flight copies

  const size_t HW = (size_t)v.H * v.W;
  const float* __restrict__ bg = v.bg;
  if (insideA) {
    const size_t pid = (size_t)pyA * v.W + pxi;
    const float T = -nTA;
    final_T[pid] = T;
    n_contrib[pid] = lastA;
    out_color[pid] = fmaf(T, bg[0], c0A);
    out_color[HW + pid] = fmaf(T, bg[1], c1A);
    out_color[2 * HW + pid] = fmaf(T, bg[2], c2A);
    out_invdepth[pid] = dA;
  }
  if (insideB) {
    const size_t pid = (size_t)pyB * v.W + pxi;
    const float T = -nTB;
    final_T[pid] = T;
    n_contrib[pid] = lastB;
    out_color[pid] = fmaf(T, bg[0], c0B);
    out_color[HW + pid] = fmaf(T, bg[1], c1B);
    out_color[2 * HW + pid] = fmaf(T, bg[2], c2B);
    out_invdepth[pid] = dB;
  }
}

cudaError_t launch_render_forward(const GsrView& v, const uint2* ranges, const uint32_t* tile_order,
                                  const uint32_t* point_list, const GsrRec* recs, float* out_color, float* out_invdepth,
                                  float* final_T, uint32_t* n_contrib, uint32_t* hits, cudaStream_t s) {
  const int tiles = v.gx * v.gy;
  if (tiles <= 0) return cudaSuccess;
  render_forward_kernel<<<tiles * (128 / GSR_FWD_THREADS), GSR_FWD_THREADS, 0, s>>>(v, ranges, tile_order, point_list, recs, out_color, out_invdepth, final_T,
                                              n_contrib, hits);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
