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
//   * one CTA per tile, 8 warps; warp w owns the 8x4 pixel sub-tile (w&1, w>>1) so that a warp's
//     pixels are spatially compact;
//   * the tile's list is consumed in batches of 256: each thread gathers one packed 48-byte
//     GsrRec (three 128-bit loads, L2-resident at 1M Gaussians) and stages it in shared memory
//     with the quadratic form pre-scaled into the log2 domain (render_common.cuh), so that the
//     inner loop is 6 FP32 ops + one MUFU.EX2 per (pixel, splat);
//   * sub-tile culling: before touching a batch, each lane tests ONE splat of the batch
//     against the warp's 8x4 pixel rectangle (exact minimum of the quadratic form over the
//     rectangle vs. tau2 = log2(255 * opacity), conservative margin) and a ballot turns the 32
//     answers into a bit mask; the warp then iterates only over set bits.  A splat rejected
//     here has alpha < 1/255 on every pixel of the sub-tile, so results are identical to the
//     unculled loop while most (pixel, splat) pairs are never evaluated;
//   * the hit loop is branch-free per lane (predicated updates);
//   * for every chunk of 32 list entries the warp records which entries contributed to at
//     least one of its pixels (one 32-bit "hit word" per warp and chunk, 1 byte per list entry
//     in total): the backward kernel replays exactly those and needs no culling test;
//   * warp ballots retire a warp as soon as its 32 pixels are saturated; the CTA leaves the
//     list when __syncthreads_and says every pixel is done.
#include "gsr_common.cuh"
#include "render_common.cuh"

namespace gsr {

__global__ void __launch_bounds__(256) render_forward_kernel(const GsrView v, const uint2* __restrict__ ranges,
                                                             const uint32_t* __restrict__ point_list,
                                                             const GsrRec* __restrict__ recs,
                                                             float* __restrict__ out_color,
                                                             float* __restrict__ out_invdepth,
                                                             float* __restrict__ final_T,
                                                             uint32_t* __restrict__ n_contrib,
                                                             uint32_t* __restrict__ hits) {
  __shared__ float4 sRec[4][256];  // staged batch, planes as described in render_common.cuh (3 = cull plane)

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tx = tile % v.gx, ty = tile / v.gx;
  // warp -> 8x4 sub-tile, lane -> pixel inside it
  const int sx0 = tx * GSR_TILE + (warp & 1) * 8, sy0 = ty * GSR_TILE + (warp >> 1) * 4;
  const int pxi = sx0 + (lane & 7), pyi = sy0 + (lane >> 3);
  const bool inside = pxi < v.W && pyi < v.H;
  const float pxf = (float)pxi, pyf = (float)pyi;
  const SubTile st = make_subtile(sx0, sy0, v.W, v.H);
  const uint32_t aRec = (uint32_t)__cvta_generic_to_shared(sRec);

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  uint32_t* __restrict__ hit_row = hits + hit_slot(range.x, (uint32_t)tile, 0) * 8 + warp;

  float T = 1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last = 0;
  bool done = !inside;

  for (int base = 0; base < n; base += 256) {
    if (__syncthreads_and(done)) break;
    const int cnt = min(256, n - base);
    if (tid < cnt) {
      const uint32_t g = point_list[range.x + base + tid];
      const GsrRec* r = recs + g;
      float4 a = r->a, b = r->b, c = r->c;
      a.z *= GSR_HALF_LOG2E; a.w *= GSR_LOG2E; b.x *= GSR_HALF_LOG2E; c.z *= GSR_LOG2E;
      sRec[0][tid] = a; sRec[1][tid] = b; sRec[2][tid] = c;
      sRec[3][tid] = make_cull_plane(a.z, a.w, b.x, c.z);
    } else {
      sRec[3][tid] = make_float4(0.f, 0.f, __int_as_float(0xff800000), 0.f);  // padding: never a hit
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, done)) continue;  // whole warp saturated: only keep the barriers

    for (int c0 = 0; c0 < cnt; c0 += 32) {
      // one splat per lane against the warp's sub-tile
      const bool hit = subtile_may_contribute(sRec[0][c0 + lane], sRec[1][c0 + lane].x, sRec[3][c0 + lane], st);
      uint32_t mask = __ballot_sync(0xffffffffu, hit);
      uint32_t used = 0;
      while (mask) {
        const int bit = __ffs(mask) - 1;
        mask &= mask - 1;
        const uint32_t addr = aRec + (uint32_t)(c0 + bit) * 16u;
        const float4 a = lds128<0>(addr);
        const float4 b = lds128<GSR_PLANE>(addr);
        const float2 c = lds64<2 * GSR_PLANE>(addr);
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float e = quad_form(a.z, a.w, b.x, dx, dy);
        const float alpha = splat_alpha(b.y, ex2_approx(-e));
        const float test_T = __fmaf_rn(-T, alpha, T);
        bool ok = !(e < 0.f) && (alpha >= 1.f / 255.f) && !done;
        done = done || (ok && test_T < 0.0001f);
        ok = ok && !(test_T < 0.0001f);
        const float w = ok ? alpha * T : 0.f;
        C0 = fmaf(b.z, w, C0);
        C1 = fmaf(b.w, w, C1);
        C2 = fmaf(c.x, w, C2);
        D = fmaf(c.y, w, D);
        T = ok ? test_T : T;
        last = ok ? (uint32_t)(base + c0 + bit + 1) : last;
        if (__any_sync(0xffffffffu, ok)) used |= 1u << bit;
      }
      if (lane == 0) hit_row[(size_t)((base + c0) >> 5) * 8] = used;
      if (__all_sync(0xffffffffu, done)) break;
    }
  }

  if (inside) {
    const size_t HW = (size_t)v.H * v.W;
    const size_t pid = (size_t)pyi * v.W + pxi;
    const float* __restrict__ bg = v.bg;
    final_T[pid] = T;
    n_contrib[pid] = last;
    out_color[pid] = fmaf(T, bg[0], C0);
    out_color[HW + pid] = fmaf(T, bg[1], C1);
    out_color[2 * HW + pid] = fmaf(T, bg[2], C2);
    out_invdepth[pid] = D;
  }
}

cudaError_t launch_render_forward(const GsrView& v, const uint2* ranges, const uint32_t* point_list,
                                  const GsrRec* recs, float* out_color, float* out_invdepth, float* final_T,
                                  uint32_t* n_contrib, uint32_t* hits, cudaStream_t s) {
  const int tiles = v.gx * v.gy;
  if (tiles <= 0) return cudaSuccess;
  render_forward_kernel<<<tiles, 256, 0, s>>>(v, ranges, point_list, recs, out_color, out_invdepth, final_T,
                                              n_contrib, hits);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
