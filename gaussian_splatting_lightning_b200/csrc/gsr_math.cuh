// gsr_math.cuh -- per-Gaussian geometry shared by the preprocess forward and backward kernels.
//
// Canonical FP32 rounding sequence: every expression that feeds an
// integer output (depth bits in the sort key, radii, tile rects) is written as an explicit
// chain of mul / add / fmaf.  The translation units that include this file are compiled with
// -fmad=false so nvcc adds no contraction of its own; the CPU oracle states the same chain
// and is compiled with -ffp-contract=off.  Division and sqrt are the IEEE-rounded defaults
// (no --use_fast_math).
//
// Arithmetic follows the reference's Python restatement:
//   projection   gs_lightning/rasterize/camera_tools.py:5-32
//   cov3D        gs_lightning/rasterize/render_tools.py:56-70
//   cov2D (EWA)  gs_lightning/rasterize/render_tools.py:13-52
//   SH basis     gs_lightning/utils/sh.py:6-81 (degree 3: third_party/gaussian_splatting/utils/sh_utils.py:57-112)
#pragma once
#include "gsr_common.cuh"

namespace gsr {

__device__ __forceinline__ float dot3_chain(float a0, float b0, float a1, float b1, float a2, float b2) {
  return fmaf(a2, b2, fmaf(a1, b1, a0 * b0));
}

// [p,1] @ M, column c (row-vector convention).
__device__ __forceinline__ float xform_col(const float* __restrict__ m, int c, float x, float y, float z) {
  return dot3_chain(m[0 + c], x, m[4 + c], y, m[8 + c], z) + m[12 + c];
}

struct Cov2D {
  float a, b, c;
};

struct JacobianCtx {
  float tx, ty, tz;
  bool gate_x, gate_y;
  float J00, J02, J11, J12;
  float m0[3], m1[3];  // rows of J * R_cw
};

__device__ __forceinline__ void build_jacobian(const float* __restrict__ v, float px, float py, float pz, float fx,
                                               float fy, float tanfovx, float tanfovy, JacobianCtx& c) {
  float tx = xform_col(v, 0, px, py, pz);
  float ty = xform_col(v, 1, px, py, pz);
  const float tz = xform_col(v, 2, px, py, pz);
  const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
  const float txtz = tx / tz, tytz = ty / tz;
  c.gate_x = !(txtz < -limx || txtz > limx);
  c.gate_y = !(tytz < -limy || tytz > limy);
  tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
  ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
  c.tx = tx; c.ty = ty; c.tz = tz;
  const float tz2 = tz * tz;
  c.J00 = fx / tz;
  c.J02 = -(fx * tx) / tz2;
  c.J11 = fy / tz;
  c.J12 = -(fy * ty) / tz2;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    c.m0[j] = fmaf(c.J02, v[4 * j + 2], c.J00 * v[4 * j + 0]);
    c.m1[j] = fmaf(c.J12, v[4 * j + 2], c.J11 * v[4 * j + 1]);
  }
}

__device__ __forceinline__ Cov2D project_cov(const float* s6, const JacobianCtx& c) {
  const float S[3][3] = {{s6[0], s6[1], s6[2]}, {s6[1], s6[3], s6[4]}, {s6[2], s6[4], s6[5]}};
  float u[3], w[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    u[i] = dot3_chain(S[i][0], c.m0[0], S[i][1], c.m0[1], S[i][2], c.m0[2]);
    w[i] = dot3_chain(S[i][0], c.m1[0], S[i][1], c.m1[1], S[i][2], c.m1[2]);
  }
  Cov2D r;
  r.a = dot3_chain(c.m0[0], u[0], c.m0[1], u[1], c.m0[2], u[2]);
  r.b = dot3_chain(c.m1[0], u[0], c.m1[1], u[1], c.m1[2], u[2]);
  r.c = dot3_chain(c.m1[0], w[0], c.m1[1], w[1], c.m1[2], w[2]);
  return r;
}

// q = (w, x, y, z), used as given (the caller normalises, gaussian_model.py:329-330).
__device__ __forceinline__ void quat_to_R(const float4 q, float R[3][3]) {
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  R[0][0] = fmaf(-2.f, fmaf(z, z, y * y), 1.f);
  R[0][1] = 2.f * fmaf(-r, z, x * y);
  R[0][2] = 2.f * fmaf(r, y, x * z);
  R[1][0] = 2.f * fmaf(r, z, x * y);
  R[1][1] = fmaf(-2.f, fmaf(z, z, x * x), 1.f);
  R[1][2] = 2.f * fmaf(-r, x, y * z);
  R[2][0] = 2.f * fmaf(-r, y, x * z);
  R[2][1] = 2.f * fmaf(r, x, y * z);
  R[2][2] = fmaf(-2.f, fmaf(y, y, x * x), 1.f);
}

__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, const float4 q,
                                                     float* out6) {
  float R[3][3];
  quat_to_R(q, R);
  const float s[3] = {mod * sx, mod * sy, mod * sz};
  float L[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) L[i][k] = R[i][k] * s[k];
  out6[0] = dot3_chain(L[0][0], L[0][0], L[0][1], L[0][1], L[0][2], L[0][2]);
  out6[1] = dot3_chain(L[0][0], L[1][0], L[0][1], L[1][1], L[0][2], L[1][2]);
  out6[2] = dot3_chain(L[0][0], L[2][0], L[0][1], L[2][1], L[0][2], L[2][2]);
  out6[3] = dot3_chain(L[1][0], L[1][0], L[1][1], L[1][1], L[1][2], L[1][2]);
  out6[4] = dot3_chain(L[1][0], L[2][0], L[1][1], L[2][1], L[1][2], L[2][2]);
  out6[5] = dot3_chain(L[2][0], L[2][0], L[2][1], L[2][1], L[2][2], L[2][2]);
}

#define GSR_SH_C0 0.28209479177387814f
#define GSR_SH_C1 0.4886025119029199f
#define GSR_SH_C2_0 1.0925484305920792f
#define GSR_SH_C2_1 -1.0925484305920792f
#define GSR_SH_C2_2 0.31539156525252005f
#define GSR_SH_C2_3 -1.0925484305920792f
#define GSR_SH_C2_4 0.5462742152960396f
#define GSR_SH_C3_0 -0.5900435899266435f
#define GSR_SH_C3_1 2.890611442640554f
#define GSR_SH_C3_2 -0.4570457994644658f
#define GSR_SH_C3_3 0.3731763325901154f
#define GSR_SH_C3_4 -0.4570457994644658f
#define GSR_SH_C3_5 1.445305721320277f
#define GSR_SH_C3_6 -0.5900435899266435f

// SH basis values b[0..(deg+1)^2) for a unit direction.
template <int DEG>
__device__ __forceinline__ void sh_basis(float x, float y, float z, float* b) {
  b[0] = GSR_SH_C0;
  if (DEG < 1) return;
  b[1] = -GSR_SH_C1 * y; b[2] = GSR_SH_C1 * z; b[3] = -GSR_SH_C1 * x;
  if (DEG < 2) return;
  const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
  b[4] = GSR_SH_C2_0 * xy;
  b[5] = GSR_SH_C2_1 * yz;
  b[6] = GSR_SH_C2_2 * (2.f * zz - xx - yy);
  b[7] = GSR_SH_C2_3 * xz;
  b[8] = GSR_SH_C2_4 * (xx - yy);
  if (DEG < 3) return;
  b[9] = GSR_SH_C3_0 * y * (3.f * xx - yy);
  b[10] = GSR_SH_C3_1 * xy * z;
  b[11] = GSR_SH_C3_2 * y * (4.f * zz - xx - yy);
  b[12] = GSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
  b[13] = GSR_SH_C3_4 * x * (4.f * zz - xx - yy);
  b[14] = GSR_SH_C3_5 * z * (xx - yy);
  b[15] = GSR_SH_C3_6 * x * (xx - 3.f * yy);
}

// d(basis)/d(x,y,z), x,y,z treated as independent.  Entry 0 is always zero.
template <int DEG>
__device__ __forceinline__ void sh_basis_grad(float x, float y, float z, float* dx, float* dy, float* dz) {
  constexpr int N = (DEG + 1) * (DEG + 1);
#pragma unroll
  for (int i = 0; i < N; ++i) dx[i] = dy[i] = dz[i] = 0.f;
  if (DEG < 1) return;
  dy[1] = -GSR_SH_C1; dz[2] = GSR_SH_C1; dx[3] = -GSR_SH_C1;
  if (DEG < 2) return;
  const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
  dx[4] = GSR_SH_C2_0 * y; dy[4] = GSR_SH_C2_0 * x;
  dy[5] = GSR_SH_C2_1 * z; dz[5] = GSR_SH_C2_1 * y;
  dx[6] = -2.f * GSR_SH_C2_2 * x; dy[6] = -2.f * GSR_SH_C2_2 * y; dz[6] = 4.f * GSR_SH_C2_2 * z;
  dx[7] = GSR_SH_C2_3 * z; dz[7] = GSR_SH_C2_3 * x;
  dx[8] = 2.f * GSR_SH_C2_4 * x; dy[8] = -2.f * GSR_SH_C2_4 * y;
  if (DEG < 3) return;
  dx[9] = GSR_SH_C3_0 * 6.f * xy;             dy[9] = GSR_SH_C3_0 * (3.f * xx - 3.f * yy);
  dx[10] = GSR_SH_C3_1 * yz;                  dy[10] = GSR_SH_C3_1 * xz;  dz[10] = GSR_SH_C3_1 * xy;
  dx[11] = GSR_SH_C3_2 * (-2.f * xy);         dy[11] = GSR_SH_C3_2 * (4.f * zz - xx - 3.f * yy);
  dz[11] = GSR_SH_C3_2 * 8.f * yz;
  dx[12] = GSR_SH_C3_3 * (-6.f * xz);         dy[12] = GSR_SH_C3_3 * (-6.f * yz);
  dz[12] = GSR_SH_C3_3 * (6.f * zz - 3.f * xx - 3.f * yy);
  dx[13] = GSR_SH_C3_4 * (4.f * zz - 3.f * xx - yy); dy[13] = GSR_SH_C3_4 * (-2.f * xy);
  dz[13] = GSR_SH_C3_4 * 8.f * xz;
  dx[14] = GSR_SH_C3_5 * 2.f * xz;            dy[14] = GSR_SH_C3_5 * (-2.f * yz);
  dz[14] = GSR_SH_C3_5 * (xx - yy);
  dx[15] = GSR_SH_C3_6 * (3.f * xx - 3.f * yy); dy[15] = GSR_SH_C3_6 * (-6.f * xy);
}

// Tile rectangle of a splat: C truncation toward zero, clamped to [0, grid]
// (gs_lightning/rasterize/render_tools.py:134-139).
__device__ __forceinline__ void tile_rect(float px, float py, float radius, int gx, int gy, int& x0, int& y0,
                                          int& x1, int& y1) {
  x0 = min(gx, max(0, (int)((px - radius) / (float)GSR_TILE)));
  y0 = min(gy, max(0, (int)((py - radius) / (float)GSR_TILE)));
  x1 = min(gx, max(0, (int)((px + radius + (float)(GSR_TILE - 1)) / (float)GSR_TILE)));
  y1 = min(gy, max(0, (int)((py + radius + (float)(GSR_TILE - 1)) / (float)GSR_TILE)));
}

// ---- mbarrier + 1-D bulk async copy (TMA without a tensor map; SASS: UBLKCP) ----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                              unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared -> global bulk store / bulk reduction (fire and forget, tracked by the bulk async-group)
__device__ __forceinline__ void bulk_copy_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_reduce_add_s2g(float* dst_gmem, const float* src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit_and_wait_read() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
}

}  // namespace gsr
