// render_backward.cu -- K8: per-tile backward of the alpha compositing for sm_100a.
//
// There is no in-tree restatement of this pass in the reference (its Python rasterizer is
// forward-only); the arithmetic is the analytic derivative of
// gs_lightning/rasterize/rasterize.py:182-261 as stated in oracle/gsr_oracle.cpp
// (gsro_render_backward), which float64 autograd pins in tests/test_oracle_backward.py.
//
// Per pixel, the tile's list is replayed BACK TO FRONT from the last contributor recorded by
// the forward (n_contrib), rebuilding T by division and carrying the colour / inverse-depth
// accumulated behind the current splat.  Per (pixel, splat) pair ten partial derivatives are
// produced: d/d{mean2D.xy, conic.ABC, opacity, colour.rgb, 1/z}.
//
// Design (what differs from a per-pixel-atomics kernel):
//   * same CTA/warp/pixel mapping, batch staging and sub-tile culling as the forward kernel
//     (a splat whose alpha is below 1/255 on the whole 8x4 sub-tile produces no gradient);
//   * the ten partials of one splat are reduced over the warp's 32 pixels with a
//     recursive-halving exchange: at each xor-shuffle step a lane keeps half of its values and
//     sends the other half, so the reduction costs 5+3+2+1+1 = 12 shuffles instead of 50, and
//     ends with the ten sums spread over ten lanes;
//   * those ten lanes issue ONE red.global.add.f32 instruction per (warp, splat) into a
//     48-byte-per-Gaussian accumulator row -- 1/32 of the atomics of the per-pixel scheme.
#include "gsr_common.cuh"
#include "render_common.cuh"

namespace gsr {

namespace {

// After warp_reduce10 the sum of value `slot` lives in the lanes for which this returns
// slot (even lanes only); -1 elsewhere.
__device__ __forceinline__ int reduce10_slot(int lane) {
  if (lane & 1) return -1;
  const int tI = (lane & 2) ? 1 : 0;
  const int uI = ((lane & 4) ? 2 : 0) + tI;
  if (lane & 8) {
    if (uI >= 2) return -1;
  } else {
    if (uI >= 3) return -1;
  }
  const int wI = ((lane & 8) ? 3 : 0) + uI;
  return ((lane & 16) ? 5 : 0) + wI;
}

#define GSR_XCHG(dst, lo_val, hi_val, bit)                                  \
  {                                                                         \
    const float keep_ = (lane & (bit)) ? (hi_val) : (lo_val);               \
    const float send_ = (lane & (bit)) ? (lo_val) : (hi_val);               \
    dst = keep_ + __shfl_xor_sync(0xffffffffu, send_, (bit));               \
  }

// Sum each of v[0..9] over the warp; returns the sum selected by reduce10_slot(lane).
__device__ __forceinline__ float warp_reduce10(const float* v, int lane) {
  float w0, w1, w2, w3, w4;
  GSR_XCHG(w0, v[0], v[5], 16);
  GSR_XCHG(w1, v[1], v[6], 16);
  GSR_XCHG(w2, v[2], v[7], 16);
  GSR_XCHG(w3, v[3], v[8], 16);
  GSR_XCHG(w4, v[4], v[9], 16);
  float u0, u1, u2;
  GSR_XCHG(u0, w0, w3, 8);
  GSR_XCHG(u1, w1, w4, 8);
  GSR_XCHG(u2, w2, 0.f, 8);
  float t0, t1;
  GSR_XCHG(t0, u0, u2, 4);
  GSR_XCHG(t1, u1, 0.f, 4);
  float s;
  GSR_XCHG(s, t0, t1, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  return s;
}

}  // namespace

__global__ void __launch_bounds__(256) render_backward_kernel(
    const GsrView v, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
    const GsrRec* __restrict__ recs, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth, float* __restrict__ grad_acc) {
  __shared__ float4 sA[256];
  __shared__ float4 sB[256];
  __shared__ float4 sC[256];
  __shared__ uint32_t sId[256];
  __shared__ uint32_t s_warp_max[8];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tx = tile % v.gx, ty = tile / v.gx;
  const int sx0 = tx * GSR_TILE + (warp & 1) * 8, sy0 = ty * GSR_TILE + (warp >> 1) * 4;
  const int pxi = sx0 + (lane & 7), pyi = sy0 + (lane >> 3);
  const bool inside = pxi < v.W && pyi < v.H;
  const float pxf = (float)pxi, pyf = (float)pyi;
  const SubTile st = make_subtile(sx0, sy0, v.W, v.H);
  const uint2 range = ranges[tile];
  const int slot = reduce10_slot(lane);

  float T_final = 0.f, dp0 = 0.f, dp1 = 0.f, dp2 = 0.f, dinv = 0.f;
  uint32_t last = 0;
  if (inside) {
    const size_t HW = (size_t)v.H * v.W;
    const size_t pid = (size_t)pyi * v.W + pxi;
    T_final = final_T[pid];
    last = n_contrib[pid];
    dp0 = dL_dpix[pid]; dp1 = dL_dpix[HW + pid]; dp2 = dL_dpix[2 * HW + pid];
    if (dL_dinvdepth) dinv = dL_dinvdepth[pid];
  }
  const float* __restrict__ bg = v.bg;
  const float bg_dot = bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2;
  const float ddelx_dx = 0.5f * (float)v.W, ddely_dy = 0.5f * (float)v.H;

  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, last);
  if (lane == 0) s_warp_max[warp] = warp_last;
  __syncthreads();
  uint32_t max_last = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) max_last = max(max_last, s_warp_max[w]);
  if (max_last == 0) return;

  float T = T_final;
  float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
  float accum_invz = 0.f, last_invz = 0.f, last_alpha = 0.f;

  const int nb = ((int)max_last + 255) / 256;
  for (int b = nb - 1; b >= 0; --b) {
    const int base = b * 256;
    const int cnt = min(256, (int)max_last - base);
    __syncthreads();
    if (tid < cnt) {
      const uint32_t g = point_list[range.x + base + tid];
      const GsrRec* r = recs + g;
      sA[tid] = r->a; sB[tid] = r->b; sC[tid] = r->c; sId[tid] = g;
    }
    __syncthreads();
    if ((int)warp_last <= base) continue;  // nothing in this batch reaches the warp's pixels

    for (int c0 = ((cnt - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
      bool hit = false;
      const int jj = c0 + lane;
      if (jj < cnt && base + jj < (int)warp_last)
        hit = subtile_may_contribute(sA[jj], sB[jj].x, sC[jj].z, st);
      uint32_t mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int bit = 31 - __clz(mask);
        mask &= ~(1u << bit);
        const int j = c0 + bit;
        const float4 a = sA[j];
        const float4 bq = sB[j];
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float power = -0.5f * (a.z * dx * dx + bq.x * dy * dy) - a.w * dx * dy;
        const float G = __expf(power);
        const float alpha = fminf(0.99f, bq.y * G);
        const bool valid = ((uint32_t)(base + j) < last) && (power <= 0.f) && (alpha >= 1.f / 255.f);
        if (!__any_sync(0xffffffffu, valid)) continue;

        float o[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) o[k] = 0.f;
        if (valid) {
          const float4 c = sC[j];
          const float one_m_alpha = 1.f - alpha;
          T = __fdividef(T, one_m_alpha);
          const float w = alpha * T;
          float dL_dalpha;
          ar0 = last_alpha * lc0 + (1.f - last_alpha) * ar0; lc0 = bq.z;
          ar1 = last_alpha * lc1 + (1.f - last_alpha) * ar1; lc1 = bq.w;
          ar2 = last_alpha * lc2 + (1.f - last_alpha) * ar2; lc2 = c.x;
          dL_dalpha = (bq.z - ar0) * dp0 + (bq.w - ar1) * dp1 + (c.x - ar2) * dp2;
          o[6] = w * dp0; o[7] = w * dp1; o[8] = w * dp2;
          accum_invz = last_alpha * last_invz + (1.f - last_alpha) * accum_invz;
          last_invz = c.y;
          dL_dalpha += (c.y - accum_invz) * dinv;
          o[9] = w * dinv;
          dL_dalpha *= T;
          last_alpha = alpha;
          dL_dalpha += __fdividef(-T_final, one_m_alpha) * bg_dot;
          const float dL_dG = bq.y * dL_dalpha;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * a.z - gdy * a.w;
          const float dG_ddely = -gdy * bq.x - gdx * a.w;
          o[0] = dL_dG * dG_ddelx * ddelx_dx;
          o[1] = dL_dG * dG_ddely * ddely_dy;
          o[2] = -0.5f * gdx * dx * dL_dG;
          o[3] = -gdx * dy * dL_dG;
          o[4] = -0.5f * gdy * dy * dL_dG;
          o[5] = G * dL_dalpha;
        }
        const float sum = warp_reduce10(o, lane);
        if (slot >= 0) atomicAdd(grad_acc + (size_t)sId[j] * 12 + slot, sum);
      }
    }
  }
}

cudaError_t launch_render_backward(const GsrView& v, const uint2* ranges, const uint32_t* point_list,
                                   const GsrRec* recs, const float* final_T, const uint32_t* n_contrib,
                                   const float* dL_dpix, const float* dL_dinvdepth, float* grad_acc,
                                   cudaStream_t s) {
  const int tiles = v.gx * v.gy;
  if (tiles <= 0) return cudaSuccess;
  render_backward_kernel<<<tiles, 256, 0, s>>>(v, ranges, point_list, recs, final_T, n_contrib, dL_dpix,
                                               dL_dinvdepth, grad_acc);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
