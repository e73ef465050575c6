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
__host__ __device__ __forceinline__ float __int_as_float_compat(int bits) {
  union { int i; float f; } u;
  u.i = bits;
  return u.f;
}

// ---- staged record: what the compositing kernels keep in shared memory -----------------
// The quadratic form is pre-scaled into the log2 domain when a batch is staged, so that the
// per-(pixel, splat) evaluation is  e = hA dx^2 + hB dx dy + hC dy^2  (= -power * log2 e),
// G = ex2(-e):  hA = A * (log2e / 2), hB = B * log2e, hC = C * (log2e / 2), tau2 = tau * log2e.
#define GSR_LOG2E 1.4426950408889634f
#define GSR_HALF_LOG2E 0.7213475204444817f

#ifdef __CUDACC__
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Shared-memory accesses through explicit 32-bit shared addresses with an immediate byte offset
// (keeps the address arithmetic of the hit loop to one IMAD per splat).
template <int OFF>
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+%5];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ float2 lds64(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2+%3];" : "=f"(v.x), "=f"(v.y) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ uint32_t lds32u(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ void sts32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0+%1], %2;" : : "r"(addr), "n"(OFF), "f"(v) : "memory");
}
// A value the compiler must keep in a register: without this ptxas REMATERIALISES shared-memory addresses inside the hit
// loops (S2R tid, S2R cta id, shifts, LEA, IMAD: 16 instructions per visit in render_backward_kernel's r2 SASS) rather
// than hold them -- cheap by its cost model, but the loops are bound by instruction issue.
__device__ __forceinline__ uint32_t keep_u32(uint32_t v) {
  uint32_t r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}
// the same for a 64-bit value (a per-lane global pointer the hot loop adds into)
__device__ __forceinline__ unsigned long long keep_u64(unsigned long long v) {
  unsigned long long r;
  asm volatile("mov.b64 %0, %1;" : "=l"(r) : "l"(v));
  return r;
}
template <int OFF>
__device__ __forceinline__ void sts32u(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0+%1], %2;" : : "r"(addr), "n"(OFF), "r"(v) : "memory");
}
// index of the most significant set bit (x != 0)
__device__ __forceinline__ int msb_index(uint32_t x) {
  int r;
  asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
  return r;
}

// ---- packed FP32 pairs (sm_100a FFMA2 / FMUL2 / FADD2) ------------------------------------------
// A lane of the compositing kernels owns TWO pixels of the same column, four rows apart (the 8x4 sub-tiles
// A and B of its warp's 8x8 block), and keeps everything per-pixel as a packed pair {A, B} in a 64-bit
// register.  One FFMA2 does the FMA of both pixels in ONE issue slot (same FP32-pipe time as two FFMAs:
// tools/ubench/ffma2.cu measured 73.9 vs 71.5 TFLOP/s), and these kernels are issue bound.  Each half is
// rounded exactly like the scalar instruction, so results are bit-identical to the one-pixel-per-lane code.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// two 64-bit halves of one 16-byte shared-memory word
template <int OFF>
__device__ __forceinline__ void lds2x64(uint32_t addr, f32x2& lo, f32x2& hi) {
  asm volatile("ld.shared.v2.b64 {%0,%1}, [%2+%3];" : "=l"(lo), "=l"(hi) : "r"(addr), "n"(OFF) : "memory");
}
template <int OFF>
__device__ __forceinline__ f32x2 lds1x64(uint32_t addr) {
  f32x2 v;
  asm volatile("ld.shared.b64 %0, [%1+%2];" : "=l"(v) : "r"(addr), "n"(OFF) : "memory");
  return v;
}

// ---- staged batches ----------------------------------------------------------------------------------------------
// A batch of GSR_BATCH list entries lives in shared memory as planes of float4: plane p of entry j at byte offset
// p * GSR_PLANE + j * 16 -- exactly the four words of the Gaussian's GsrRec (gsr_common.cuh), which the preprocess kernel
// already wrote in this form, so staging is four 16-byte cp.async per entry and no arithmetic:
//   plane 0 {mean.x, mean.y, hA, hB}   plane 1 {hC, -opacity, r, g}   plane 2 {b, 1/z, tau2, bits(gaussian id)}
//   plane 3 {kx, ky, tauM, z}  (forward only: the cull plane, make_cull_plane)
// Opacity is stored negated because the loops carry -alpha and -T (T (1 - alpha) = fma(-T, -alpha, -T) negated,
// alpha T = (-alpha)(-T): exact, and no packed negation exists).  The packed instructions take a scalar register as a
// broadcast operand (SASS `R.F32`), so per-splat values need no duplication: pk2(v, v) costs nothing.
// Two buffers: the copies of batch b + 1 are in flight while the warps composite batch b (ONE barrier per batch).
#ifndef GSR_BATCH
#define GSR_BATCH 128
#endif
#define GSR_PLANE (GSR_BATCH * 16)

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" : : "r"(smem_addr), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t smem_addr, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" : : "r"(smem_addr), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" : : : "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" : : : "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, float x, float y, float z, float w) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" : : "r"(addr), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

// e = hA dx^2 + hC dy^2 + hB dx dy with a FIXED rounding sequence: the forward and the backward
// kernel must agree bit for bit on alpha (the backward replays the forward's accept/reject
// decisions), so the chain is spelled out with intrinsics the compiler cannot re-associate.
__device__ __forceinline__ float quad_form(float hA, float hB, float hC, float dx, float dy) {
  const float t = __fmul_rn(hA, dx);
  const float u = __fmul_rn(hC, dy);
  const float v = __fmul_rn(hB, dx);
  float e = __fmul_rn(t, dx);
  e = __fmaf_rn(u, dy, e);
  e = __fmaf_rn(v, dy, e);
  return e;
}
__device__ __forceinline__ float splat_alpha(float opacity, float G) {
  return fminf(0.99f, __fmul_rn(opacity, G));
}
#endif  // __CUDACC__

// Conservative test "can this splat reach alpha >= 1/255 on ANY pixel of the sub-tile?".
//
// alpha = min(.99, o * 2^-e) >= 1/255  <=>  e <= tau2, tau2 = log2(255 o), where
// e(d) = hA dx^2 + hB dx dy + hC dy^2 >= 0 for a positive-definite conic.
// The minimum of the convex e over the rectangle is 0 if the centre lies inside it, else it is
// attained on an edge that faces the centre; on such an edge e is a 1-D parabola whose clamped
// vertex gives the minimum in closed form.  Both candidate points (clamped vertex on the nearest
// vertical and on the nearest horizontal line of the rectangle) are evaluated unconditionally:
// each is a point OF the rectangle, and whenever an edge faces the centre its candidate is that
// edge's minimum, so min(e1, e2) is the exact minimum in every case (and 0 when the centre is
// inside) -- no branches.  The splat is rejected only if that minimum exceeds tau2 by more than
// a margin that covers the FP32 evaluation error of e in the per-pixel loop, of this test, of
// logf and of ex2.approx (a few 1e-7 * S, S >= the sum of the absolute terms of e anywhere in
// the rectangle; by AM-GM |hB| DX DY <= hA DX^2 + hC DY^2 for a convex form).
//
// Everything that does not depend on the rectangle is computed ONCE per staged splat
// (make_cull_plane, by the thread that stages it) instead of once per warp:
//   cull = {kx = -hB / (2 hC), ky = -hB / (2 hA), tauM, -}
//   tauM = tau2 + 3e-5, or +inf when the form is not provably convex (always keep), or -inf when
//   opacity < 1/255 (alpha can never reach the floor) and for padding entries.
__host__ __device__ __forceinline__ float4 make_cull_plane(float hA, float hB, float hC, float tau2) {
  const float inf = __int_as_float_compat(0x7f800000);
  float4 c;
  c.w = 0.f;
  if (!(tau2 >= -1.5e-4f)) { c.x = 0.f; c.y = 0.f; c.z = -inf; return c; }
  if (!(hA > 0.f && hC > 0.f && 4.f * hA * hC - hB * hB > 0.f)) { c.x = 0.f; c.y = 0.f; c.z = inf; return c; }
  c.x = -0.5f * hB / hC;
  c.y = -0.5f * hB / hA;
  c.z = tau2 + 3e-5f;
  return c;
}

//   a = {mean.x, mean.y, hA, hB}
__host__ __device__ __forceinline__ bool subtile_may_contribute(const float4 a, const float hC, const float4 cull,
                                                                const SubTile st) {
  const float mx = a.x, my = a.y, hA = a.z, hB = a.w;
  const float dxn = mx - gsr_clampf(mx, st.x0, st.x1);  // dx at the nearest column (0 if inside)
  const float dyn = my - gsr_clampf(my, st.y0, st.y1);
  const float dx_lo = mx - st.x1, dx_hi = mx - st.x0;  // range of dx over the rectangle
  const float dy_lo = my - st.y1, dy_hi = my - st.y0;
  const float dys = gsr_clampf(cull.x * dxn, dy_lo, dy_hi);  // vertex on the line dx = dxn
  const float dxs = gsr_clampf(cull.y * dyn, dx_lo, dx_hi);  // vertex on the line dy = dyn
  const float e1 = hA * dxn * dxn + hC * dys * dys + hB * dxn * dys;
  const float e2 = hA * dxs * dxs + hC * dyn * dyn + hB * dxs * dyn;
  const float DX = fmaxf(fabsf(dx_lo), fabsf(dx_hi));
  const float DY = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
  const float S = hA * DX * DX + hC * DY * DY;
  return fminf(e1, e2) <= cull.z + 8e-6f * S;
}

// ---- per-(warp, 32-splat chunk) hit words -------------------------------------------------
// The forward records, for every chunk of 32 list entries a warp visited, which of them
// contributed to at least one of the warp's 32 pixels.  The backward replays exactly those.
// Tile t's chunk k lives in slot (range.x >> 5) + t + k (unique and monotonic over tiles);
// a slot holds 8 words, one per warp.
__host__ __device__ __forceinline__ size_t hit_slot(uint32_t range_x, uint32_t tile, uint32_t chunk) {
  return (size_t)(range_x >> 5) + tile + chunk;
}

}  // namespace gsr
