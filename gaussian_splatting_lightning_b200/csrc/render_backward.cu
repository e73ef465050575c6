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
//   * same CTA/warp/pixel mapping and batch staging as the forward kernel; the per-warp hit
//     words the forward wrote say which list entries touched the warp's 8x4 pixels, so only
//     those (warp, splat) pairs are replayed and no culling test is evaluated here;
//   * per (pixel, splat) the kernel produces ten RAW sums instead of the ten finished partial
//     derivatives: with h = opacity * G * dL/dalpha,
//         S0 = h dx, S1 = h dy, S2 = h dx^2, S3 = h dx dy, S4 = h dy^2, S5 = G dL/dalpha,
//         S6..8 = w dL/dC_rgb, S9 = w dL/d(1/z)                       (w = alpha * T)
//     -- everything that multiplies these by per-Gaussian constants (conic, W/2, H/2, -1/2) is
//     applied once per Gaussian in the preprocess backward kernel, not once per pixel here;
//   * lanes whose pixel the splat does not reach run the same instruction stream with
//     alpha = 0 (an exact no-op of the recurrences), so the hit loop has no divergent branch;
//   * the ten sums are reduced over the warp's 32 pixels with a recursive-halving exchange: at
//     each xor-shuffle step a lane keeps half of its values and sends the other half, so the
//     reduction costs 5+3+2+1+1 = 12 shuffles instead of 50, and ends with the ten sums spread
//     over ten lanes;
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

#define GSR_XCHG(dst, lo_val, hi_val, pred, bit)                            \
  {                                                                         \
    const float keep_ = (pred) ? (hi_val) : (lo_val);                       \
    const float send_ = (pred) ? (lo_val) : (hi_val);                       \
    dst = keep_ + __shfl_xor_sync(0xffffffffu, send_, (bit));               \
  }

}  // namespace

__global__ void __launch_bounds__(256) render_backward_kernel(
    const GsrView v, const uint2* __restrict__ ranges, const char* __restrict__ binning,
    const GsrRec* __restrict__ recs, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth, float* __restrict__ grad_acc) {
  __shared__ float4 sRec[3][256];  // staged batch, planes as described in render_common.cuh
  __shared__ uint32_t sHit[8][8];  // [chunk of the batch][warp]
  __shared__ uint32_t s_warp_max[8];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tx = tile % v.gx, ty = tile / v.gx;
  const int sx0 = tx * GSR_TILE + (warp & 1) * 8, sy0 = ty * GSR_TILE + (warp >> 1) * 4;
  const int pxi = sx0 + (lane & 7), pyi = sy0 + (lane >> 3);
  const bool inside = pxi < v.W && pyi < v.H;
  const float pxf = (float)pxi, pyf = (float)pyi;
  const uint2 range = ranges[tile];
  const int slot = reduce10_slot(lane);
  const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2;
  const uint32_t aRec = (uint32_t)__cvta_generic_to_shared(sRec);

  // the binning buffer describes itself (BinningHeader): sorted point list + hit words
  const BinningHeader* hdr = reinterpret_cast<const BinningHeader*>(binning);
  const size_t cap = hdr->capacity;
  const uint32_t* __restrict__ point_list = reinterpret_cast<const uint32_t*>(binning + gsr_bin_vals_off());
  const uint32_t* __restrict__ hits = reinterpret_cast<const uint32_t*>(binning + gsr_bin_hits_off(cap));
  const uint32_t* __restrict__ hit_rows = hits + hit_slot(range.x, (uint32_t)tile, 0) * 8;

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
  const float Tf_bg = T_final * (bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2);

  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, last);
  if (lane == 0) s_warp_max[warp] = warp_last;
  __syncthreads();
  uint32_t max_last = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) max_last = max(max_last, s_warp_max[w]);
  if (max_last == 0) return;

  float T = T_final;
  float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, ar3 = 0.f;  // colour / inverse depth accumulated behind the splat

  const int nb = ((int)max_last + 255) / 256;
  for (int b = nb - 1; b >= 0; --b) {
    const int base = b * 256;
    const int cnt = min(256, (int)max_last - base);
    __syncthreads();
    if (tid < cnt) {
      const uint32_t g = point_list[range.x + base + tid];
      const GsrRec* r = recs + g;
      float4 a = r->a, bq = r->b, c = r->c;
      a.z *= GSR_HALF_LOG2E; a.w *= GSR_LOG2E; bq.x *= GSR_HALF_LOG2E;
      c.w = __uint_as_float(g);
      sRec[0][tid] = a; sRec[1][tid] = bq; sRec[2][tid] = c;
    }
    if (tid < 64 && (tid >> 3) * 32 < cnt) sHit[tid >> 3][tid & 7] = hit_rows[(size_t)((base >> 5) + (tid >> 3)) * 8 + (tid & 7)];
    __syncthreads();
    if ((int)warp_last <= base) continue;  // nothing in this batch reaches the warp's pixels

    for (int c0 = ((min(cnt, (int)warp_last - base) - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
      uint32_t mask = sHit[c0 >> 5][warp];
      while (mask) {
        const int bit = msb_index(mask);
        mask ^= 1u << bit;
        const int j = c0 + bit;
        const uint32_t addr = aRec + (uint32_t)j * 16u;
        const float4 a = lds128<0>(addr);
        const float4 bq = lds128<GSR_PLANE>(addr);
        const float2 c = lds64<2 * GSR_PLANE>(addr);
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float e = quad_form(a.z, a.w, bq.x, dx, dy);
        const float G = ex2_approx(-e);
        const float alpha_raw = splat_alpha(bq.y, G);
        const bool valid = ((uint32_t)(base + j) < last) && !(e < 0.f) && (alpha_raw >= 1.f / 255.f);
        const float alpha = valid ? alpha_raw : 0.f;

        const float rcp = rcp_approx(1.f - alpha);
        T *= rcp;
        const float w = alpha * T;
        // ar = colour / inverse depth accumulated BEHIND this splat (all farther splats folded in)
        const float d0 = bq.z - ar0, d1 = bq.w - ar1, d2 = c.x - ar2, d3 = c.y - ar3;
        float dLda = d0 * dp0;
        dLda = fmaf(d1, dp1, dLda);
        dLda = fmaf(d2, dp2, dLda);
        dLda = fmaf(d3, dinv, dLda);
        dLda = fmaf(dLda, T, -Tf_bg * rcp);
        // fold this splat in for the next (nearer) one: ar <- alpha c + (1 - alpha) ar
        ar0 = fmaf(alpha, d0, ar0);
        ar1 = fmaf(alpha, d1, ar1);
        ar2 = fmaf(alpha, d2, ar2);
        ar3 = fmaf(alpha, d3, ar3);

        const float s5 = valid ? G * dLda : 0.f;
        const float h = bq.y * s5;
        const float s0 = h * dx, s1 = h * dy;
        const float s2 = s0 * dx, s3 = s0 * dy, s4 = s1 * dy;
        const float s6 = w * dp0, s7 = w * dp1, s8 = w * dp2, s9 = w * dinv;

        // recursive-halving reduction of the ten sums over the 32 pixels
        float w0, w1, w2, w3, w4;
        GSR_XCHG(w0, s0, s5, b16, 16);
        GSR_XCHG(w1, s1, s6, b16, 16);
        GSR_XCHG(w2, s2, s7, b16, 16);
        GSR_XCHG(w3, s3, s8, b16, 16);
        GSR_XCHG(w4, s4, s9, b16, 16);
        float u0, u1, u2;
        GSR_XCHG(u0, w0, w3, b8, 8);
        GSR_XCHG(u1, w1, w4, b8, 8);
        GSR_XCHG(u2, w2, 0.f, b8, 8);
        float t0, t1;
        GSR_XCHG(t0, u0, u2, b4, 4);
        GSR_XCHG(t1, u1, 0.f, b4, 4);
        float sum;
        GSR_XCHG(sum, t0, t1, b2, 2);
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        if (slot >= 0) {
          const uint32_t g = lds32u<2 * GSR_PLANE + 12>(addr);
          atomicAdd(grad_acc + (size_t)g * 12 + slot, sum);
        }
      }
    }
  }
}

cudaError_t launch_render_backward(const GsrView& v, const uint2* ranges, const void* binning, const GsrRec* recs,
                                   const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                                   const float* dL_dinvdepth, float* grad_acc, cudaStream_t s) {
  const int tiles = v.gx * v.gy;
  if (tiles <= 0) return cudaSuccess;
  render_backward_kernel<<<tiles, 256, 0, s>>>(v, ranges, (const char*)binning, recs, final_T, n_contrib, dL_dpix,
                                               dL_dinvdepth, grad_acc);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
