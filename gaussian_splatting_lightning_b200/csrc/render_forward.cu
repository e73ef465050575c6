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
//     GsrRec (three 128-bit loads, L2-resident at 1M Gaussians) into shared memory;
//   * sub-tile culling: before touching a batch, each lane tests ONE splat of the batch
//     against the warp's 8x4 pixel rectangle (exact minimum of the quadratic form over the
//     rectangle vs. tau = ln(255 * opacity), conservative margin) and a ballot turns the 32
//     answers into a bit mask; the warp then iterates only over set bits.  A splat rejected
//     here has alpha < 1/255 on every pixel of the sub-tile, so results are identical to the
//     unculled loop while most (pixel, splat) pairs are never evaluated;
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
                                                             uint32_t* __restrict__ n_contrib) {
  __shared__ float4 sA[256];
  __shared__ float4 sB[256];
  __shared__ float4 sC[256];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tx = tile % v.gx, ty = tile / v.gx;
  // warp -> 8x4 sub-tile, lane -> pixel inside it
  const int sx0 = tx * GSR_TILE + (warp & 1) * 8, sy0 = ty * GSR_TILE + (warp >> 1) * 4;
  const int pxi = sx0 + (lane & 7), pyi = sy0 + (lane >> 3);
  const bool inside = pxi < v.W && pyi < v.H;
  const float pxf = (float)pxi, pyf = (float)pyi;
  const SubTile st = make_subtile(sx0, sy0, v.W, v.H);

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);

  float T = 1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last = 0;
  bool done = !inside;

  for (int base = 0; base < n; base += 256) {
    if (__syncthreads_and(done)) break;
    const int cnt = min(256, n - base);
    if (tid < cnt) {
      const uint32_t g = point_list[range.x + base + tid];
      const GsrRec* r = recs + g;
      sA[tid] = r->a; sB[tid] = r->b; sC[tid] = r->c;
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, done)) continue;  // whole warp saturated: only keep the barriers

    for (int c0 = 0; c0 < cnt; c0 += 32) {
      // one splat per lane against the warp's sub-tile
      bool hit = false;
      if (c0 + lane < cnt) hit = subtile_may_contribute(sA[c0 + lane], sB[c0 + lane].x, sC[c0 + lane].z, st);
      uint32_t mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int j = c0 + __ffs(mask) - 1;
        mask &= mask - 1;
        const float4 a = sA[j];
        const float4 b = sB[j];
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float power = -0.5f * (a.z * dx * dx + b.x * dy * dy) - a.w * dx * dy;
        if (power > 0.f) continue;
        const float alpha = fminf(0.99f, b.y * __expf(power));
        if (alpha < 1.f / 255.f || done) continue;
        const float test_T = T * (1.f - alpha);
        if (test_T < 0.0001f) { done = true; continue; }
        const float4 c = sC[j];
        const float w = alpha * T;
        C0 = fmaf(b.z, w, C0);
        C1 = fmaf(b.w, w, C1);
        C2 = fmaf(c.x, w, C2);
        D = fmaf(c.y, w, D);
        T = test_T;
        last = (uint32_t)(base + j + 1);
      }
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
                                  uint32_t* n_contrib, cudaStream_t s) {
  const int tiles = v.gx * v.gy;
  if (tiles <= 0) return cudaSuccess;
  render_forward_kernel<<<tiles, 256, 0, s>>>(v, ranges, point_list, recs, out_color, out_invdepth, final_T,
                                              n_contrib);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
