// render_common.cuh -- pieces shared by the forward and backward compositing kernels.
#pragma once
#include <cuda_runtime.h>

namespace gsr {

// Pixel-centre extent of a warp's 8x4 sub-tile (inclusive bounds, float pixel coordinates).
struct SubTile {
  float x0, x1, y0, y1;
};

__host__ __device__ __forceinline__ SubTile make_subtile(int sx0, int sy0, int W, int H) {
  SubTile s;
  s.x0 = (float)sx0;
  s.y0 = (float)sy0;
  const int xe = sx0 + 7 < W - 1 ? sx0 + 7 : W - 1;
  const int ye = sy0 + 3 < H - 1 ? sy0 + 3 : H - 1;
  s.x1 = (float)(xe > sx0 ? xe : sx0);
  s.y1 = (float)(ye > sy0 ? ye : sy0);
  return s;
}

__host__ __device__ __forceinline__ float gsr_clampf(float v, float lo, float hi) {
  return fminf(fmaxf(v, lo), hi);
}

// Conservative test "can this splat reach alpha >= 1/255 on ANY pixel of the sub-tile?".
//
// alpha = min(.99, o * exp(power)) >= 1/255  <=>  power >= -tau, tau = ln(255 o), where
// power(d) = -q(d), q(d) = 1/2 (A dx^2 + C dy^2) + B dx dy >= 0 for a positive-definite conic.
// The minimum of the convex q over the rectangle is 0 if the centre lies inside it, else it is
// attained on an edge that faces the centre; on such an edge q is a 1-D parabola whose clamped
// vertex gives the minimum in closed form.  The splat is rejected only if that minimum exceeds
// tau by more than a margin that covers the FP32 evaluation error of `power` in the per-pixel
// loop, of this test, of logf and of the ex2.approx path (bounded by a few 1e-7 * S, S = the
// sum of the absolute terms of q at the farthest corner).  Returns true when in doubt.
//   a = {mean.x, mean.y, conic.A, conic.B}, C = conic.C
__host__ __device__ __forceinline__ bool subtile_may_contribute(const float4 a, const float C, const float tau,
                                                                const SubTile st) {
  const float mx = a.x, my = a.y, A = a.z, B = a.w;
  if (!(tau >= -1e-4f)) return false;  // opacity < 1/255: alpha can never reach the floor
  if (!(A > 0.f && C > 0.f && A * C - B * B > 0.f)) return true;  // not provably convex: keep
  const float dxn = mx - gsr_clampf(mx, st.x0, st.x1);  // dx at the nearest column (0 if inside)
  const float dyn = my - gsr_clampf(my, st.y0, st.y1);
  if (dxn == 0.f && dyn == 0.f) return true;
  const float dx_lo = mx - st.x1, dx_hi = mx - st.x0;  // range of dx over the rectangle
  const float dy_lo = my - st.y1, dy_hi = my - st.y0;
  float minq = 3.0e38f;
  if (dxn != 0.f) {  // facing vertical edge: dx fixed, vertex dy* = -B dx / C
#ifdef __CUDA_ARCH__
    const float dys = gsr_clampf(__fdividef(-B * dxn, C), dy_lo, dy_hi);
#else
    const float dys = gsr_clampf(-B * dxn / C, dy_lo, dy_hi);
#endif
    minq = 0.5f * (A * dxn * dxn + C * dys * dys) + B * dxn * dys;
  }
  if (dyn != 0.f) {  // facing horizontal edge: dy fixed, vertex dx* = -B dy / A
#ifdef __CUDA_ARCH__
    const float dxs = gsr_clampf(__fdividef(-B * dyn, A), dx_lo, dx_hi);
#else
    const float dxs = gsr_clampf(-B * dyn / A, dx_lo, dx_hi);
#endif
    minq = fminf(minq, 0.5f * (A * dxs * dxs + C * dyn * dyn) + B * dxs * dyn);
  }
  const float DX = fmaxf(fabsf(dx_lo), fabsf(dx_hi));
  const float DY = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
  const float S = 0.5f * (A * DX * DX + C * DY * DY) + fabsf(B) * DX * DY;
  return minq <= tau + (4e-6f * S + 2e-5f);
}

}  // namespace gsr
