// preprocess_backward.cu -- K9+K10 fused: per-Gaussian backward through
//   conic -> dilated cov2D -> (cov3D, mean3D via the EWA Jacobian incl. the 1.3*tanfov clamp gate),
//   mean2D -> mean3D via the full projection, 1/z -> mean3D, colour -> (SH, direction -> mean3D)
//   with the clamp mask, cov3D -> (scale, quaternion), and the antialiasing opacity scale.
//
// The forward side of each step is the reference's Python restatement
// (gs_lightning/rasterize/render_tools.py:13-131, camera_tools.py:10-32, utils/sh.py:41-81);
// the reference has no backward for any of it, so the derivative is the one stated in
// oracle/gsr_oracle.cpp (gsro_preprocess_backward) and pinned there by float64 autograd.
//
// Input is the accumulator row the compositing backward left per Gaussian: ten RAW sums over the
// Gaussian's pixels (render_backward.cu); the per-Gaussian factors that turn them into
// dL/d{mean2D, conic, opacity} (conic entries, W/2, H/2, -1/2) are applied here, once per Gaussian.
//
// Fused epilogues: with GSR_GRADS_ACCUMULATE the outputs are += so that the
// views of one optimiser step sum straight into the flat buffer the all-reduce sends, and the
// densification statistics of the view (max radius, || dL_dmean2D.xy ||, count --
// gs_lightning/modules/gaussian_model.py:174-181) are updated from the per-view gradient before
// it is summed.  Outputs whose pointer is NULL are not produced.
//
// HBM-bound: (131 + 12M) B read, (56 + 12M) B written per visible Gaussian.  One thread per
// Gaussian; the thread's dL_dsh row is written with 128-bit stores.
#include "gsr_math.cuh"

namespace gsr {

struct PreBwdArgs {
  GsrView v;
  GsrGaussiansDev g;
  const int32_t* radii;
  const float* cov3D;
  const uint8_t* clamped;
  const float* grad_acc;  // (P,12)
  float* dL_dmeans2D;
  float* dL_dcolors;
  float* dL_dopacity;
  float* dL_dmeans3D;
  float* dL_dcov3D;
  float* dL_dsh;
  float* dL_dscales;
  float* dL_drot;
  int accumulate;
  int use_bulk;  // SH / dL_dsh blocks travel by bulk async copies
  float* max_radii2D;
  float* xyz_grad_accum;
  float* xyz_grad_count;
};

// store or accumulate one float / float4.  Accumulation is an L2 reduction (red.global.add.f32), so
// views rendered concurrently on different streams may accumulate into the same buffers.
template <bool ACC>
__device__ __forceinline__ void put(float* p, float v) {
  if (ACC) atomicAdd(p, v); else *p = v;
}
template <bool ACC>
__device__ __forceinline__ void put4(float4* p, float4 v) {
  if (ACC) {
    float* f = reinterpret_cast<float*>(p);
    atomicAdd(f, v.x); atomicAdd(f + 1, v.y); atomicAdd(f + 2, v.z); atomicAdd(f + 3, v.w);
  } else {
    *p = v;
  }
}
// n consecutive floats (n even, p 8-byte aligned): stores, or vector reductions at the L2 (REDG.E.ADD.F32x2 / F32x4)
template <bool ACC, int N>
__device__ __forceinline__ void put_run(float* p, const float* v) {
  static_assert(N % 2 == 0, "runs of an even number of floats");
  if (!ACC) {
#pragma unroll
    for (int k = 0; k < N; k += 2) *reinterpret_cast<float2*>(p + k) = make_float2(v[k], v[k + 1]);
  } else if ((N % 4) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
#pragma unroll
    for (int k = 0; k < N; k += 4)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p + k), "f"(v[k]), "f"(v[k + 1]), "f"(v[k + 2]),
                   "f"(v[k + 3]) : "memory");
  } else {
#pragma unroll
    for (int k = 0; k < N; k += 2)
      asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p + k), "f"(v[k]), "f"(v[k + 1]) : "memory");
  }
}

template <int DEG, bool ACC>
__device__ __forceinline__ void sh_backward(const float* __restrict__ sh, float* __restrict__ dsh, int M, bool vec4,
                                            float x, float y, float z, const float* g3, float* ddir) {
  constexpr int N = (DEG + 1) * (DEG + 1);
  float basis[N], bx[N], by[N], bz[N];
  sh_basis<DEG>(x, y, z, basis);
  sh_basis_grad<DEG>(x, y, z, bx, by, bz);
  ddir[0] = ddir[1] = ddir[2] = 0.f;
  float out[(3 * N + 3) / 4 * 4];
#pragma unroll
  for (int k = 0; k < N; ++k) {
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      out[3 * k + ch] = basis[k] * g3[ch];
      if (k > 0) {
        const float s = sh[3 * k + ch] * g3[ch];
        ddir[0] += bx[k] * s; ddir[1] += by[k] * s; ddir[2] += bz[k] * s;
      }
    }
  }
  if (!dsh) return;
  // row = 3*M floats; entries past 3*N are zero (left untouched when accumulating)
  const int total = 3 * M;
  if (vec4) {
    float4* d4 = reinterpret_cast<float4*>(dsh);
    constexpr int NV = (3 * N + 3) / 4;
#pragma unroll
    for (int k = 3 * N; k < NV * 4; ++k) out[k] = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) put4<ACC>(d4 + k, make_float4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]));
    if (!ACC)
      for (int k = NV; k < total / 4; ++k) d4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
#pragma unroll
    for (int k = 0; k < 3 * N; ++k) put<ACC>(dsh + k, out[k]);
    if (!ACC)
      for (int k = 3 * N; k < total; ++k) dsh[k] = 0.f;
  }
}

// One Gaussian.  sh_row / dsh_row: this Gaussian's (M,3) coefficient row and gradient row -- in global
// memory, or in the warp's shared-memory block when the rows travel by bulk copy (then DSH_ACC is
// false: the bulk reduction does the += ).
template <bool ACC, bool DSH_ACC>
__device__ __forceinline__ void backward_one(const PreBwdArgs& a, const int i, const int radius,
                                             const float* __restrict__ sh_row, float* __restrict__ dsh_row,
                                             const bool sh_vec4, unsigned long long* sh_bar = nullptr) {
  // sh_bar: the mbarrier behind which the warp's SH block is still arriving in shared memory; it is waited for
  // at the first touch of the block, so that the geometry part and its loads run under the copy.
  const int M = a.g.M;
  if (!(radius > 0)) {
    if (sh_bar) mbar_wait(sh_bar, 0);
    if (!DSH_ACC && dsh_row) {  // a stored (or bulk-reduced) row of a culled Gaussian is all zeros
      if (sh_vec4) for (int k = 0; k < 3 * M / 4; ++k) reinterpret_cast<float4*>(dsh_row)[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      else for (int k = 0; k < 3 * M; ++k) dsh_row[k] = 0.f;
    }
    if (ACC) return;  // accumulating: rows of culled Gaussians keep their value
    if (a.dL_dmeans2D) { a.dL_dmeans2D[3 * (size_t)i] = 0.f; a.dL_dmeans2D[3 * (size_t)i + 1] = 0.f; a.dL_dmeans2D[3 * (size_t)i + 2] = 0.f; }
    if (a.dL_dcolors) { a.dL_dcolors[3 * (size_t)i] = 0.f; a.dL_dcolors[3 * (size_t)i + 1] = 0.f; a.dL_dcolors[3 * (size_t)i + 2] = 0.f; }
    if (a.dL_dopacity) a.dL_dopacity[i] = 0.f;
    for (int k = 0; k < 3; ++k) {
      if (a.dL_dmeans3D) a.dL_dmeans3D[3 * (size_t)i + k] = 0.f;
      if (a.dL_dscales) a.dL_dscales[3 * (size_t)i + k] = 0.f;
    }
    if (a.dL_drot) reinterpret_cast<float4*>(a.dL_drot)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.dL_dcov3D)
      for (int k = 0; k < 6; ++k) a.dL_dcov3D[6 * (size_t)i + k] = 0.f;
    return;
  }

  const float4 g0 = reinterpret_cast<const float4*>(a.grad_acc)[3 * (size_t)i];      // S0 S1 S2 S3
  const float4 g1 = reinterpret_cast<const float4*>(a.grad_acc)[3 * (size_t)i + 1];  // S4 S5 dr dg
  const float4 g2 = reinterpret_cast<const float4*>(a.grad_acc)[3 * (size_t)i + 2];  // db dinvz - -
  // the inputs of the last sections are requested now, with everything else (one round trip, under the SH copy)
  const bool want_sr = !a.g.cov3D_precomp && (a.dL_dscales || a.dL_drot);
  const float4 q_in = want_sr ? reinterpret_cast<const float4*>(a.g.rotations)[i] : make_float4(1.f, 0.f, 0.f, 0.f);
  const float sc0 = want_sr ? a.g.scales[3 * (size_t)i] : 0.f, sc1 = want_sr ? a.g.scales[3 * (size_t)i + 1] : 0.f,
              sc2 = want_sr ? a.g.scales[3 * (size_t)i + 2] : 0.f;
  const unsigned cl_in = sh_row ? a.clamped[i] : 0u;
  const float* __restrict__ view = a.v.view;
  const float* __restrict__ proj = a.v.proj;
  const float fx = a.v.fx, fy = a.v.fy;

  const float px = a.g.means3D[3 * (size_t)i], py = a.g.means3D[3 * (size_t)i + 1], pz = a.g.means3D[3 * (size_t)i + 2];
  float s6[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) s6[k] = a.cov3D[6 * (size_t)i + k];
  JacobianCtx jc;
  build_jacobian(view, px, py, pz, fx, fy, a.v.tanfovx, a.v.tanfovy, jc);
  const Cov2D c0 = project_cov(s6, jc);
  const float ca = c0.a + 0.3f, cc = c0.c + 0.3f, cb = c0.b;
  const float det = fmaf(ca, cc, -(cb * cb));  // same chain as the forward (preprocess.cu)
  const float det_inv = 1.f / det;
  const float cA = cc * det_inv, cB = -cb * det_inv, cC = ca * det_inv;  // the conic the compositing used

  // raw pixel sums -> dL/d{mean2D, conic, opacity}  (render_backward.cu header)
  const float dmx = -(cA * g0.x + cB * g0.y) * (0.5f * (float)a.v.W);
  const float dmy = -(cC * g0.y + cB * g0.x) * (0.5f * (float)a.v.H);
  const float dA = -0.5f * g0.z, dB = -g0.w, dC = -0.5f * g1.x;

  // conic -> dilated cov2D
  const float d2inv = 1.f / (det * det + 0.0000001f);
  float da = d2inv * (-cc * cc * dA + cb * cc * dB - cb * cb * dC);
  float dc = d2inv * (-cb * cb * dA + ca * cb * dB - ca * ca * dC);
  float db = d2inv * (2.f * cb * cc * dA - (ca * cc + cb * cb) * dB + 2.f * ca * cb * dC);

  float dop = g1.y;
  if (a.v.antialiasing) {
    const float det0 = c0.a * c0.c - c0.b * c0.b;
    const float ratio = det0 / det;
    const float h = sqrtf(fmaxf(0.000025f, ratio));
    const float dL_dop_eff = dop;
    dop = dL_dop_eff * h;
    if (ratio > 0.000025f) {
      const float dL_dratio = dL_dop_eff * a.g.opacities[i] / (2.f * h);
      const float inv = 1.f / det, r2 = ratio * inv;
      da += dL_dratio * (c0.c * inv - r2 * cc);
      dc += dL_dratio * (c0.a * inv - r2 * ca);
      db += dL_dratio * (-2.f * cb * inv + r2 * 2.f * cb);
    }
  }
  if (a.dL_dopacity) put<ACC>(a.dL_dopacity + i, dop);

  // cov2D -> cov3D
  const float* m0 = jc.m0;
  const float* m1 = jc.m1;
  float dS[6];
  dS[0] = m0[0] * m0[0] * da + m0[0] * m1[0] * db + m1[0] * m1[0] * dc;
  dS[3] = m0[1] * m0[1] * da + m0[1] * m1[1] * db + m1[1] * m1[1] * dc;
  dS[5] = m0[2] * m0[2] * da + m0[2] * m1[2] * db + m1[2] * m1[2] * dc;
  dS[1] = 2.f * m0[0] * m0[1] * da + (m0[0] * m1[1] + m0[1] * m1[0]) * db + 2.f * m1[0] * m1[1] * dc;
  dS[2] = 2.f * m0[0] * m0[2] * da + (m0[0] * m1[2] + m0[2] * m1[0]) * db + 2.f * m1[0] * m1[2] * dc;
  dS[4] = 2.f * m0[1] * m0[2] * da + (m0[1] * m1[2] + m0[2] * m1[1]) * db + 2.f * m1[1] * m1[2] * dc;
  if (a.dL_dcov3D) {
#pragma unroll
    for (int k = 0; k < 6; ++k) put<ACC>(a.dL_dcov3D + 6 * (size_t)i + k, dS[k]);
  }

  // cov2D -> J*R_cw -> view-space point
  const float S[3][3] = {{s6[0], s6[1], s6[2]}, {s6[1], s6[3], s6[4]}, {s6[2], s6[4], s6[5]}};
  float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float Sm0 = S[r][0] * m0[0] + S[r][1] * m0[1] + S[r][2] * m0[2];
    const float Sm1 = S[r][0] * m1[0] + S[r][1] * m1[1] + S[r][2] * m1[2];
    const float dm0 = 2.f * da * Sm0 + db * Sm1;
    const float dm1 = 2.f * dc * Sm1 + db * Sm0;
    dJ00 += dm0 * view[4 * r + 0]; dJ02 += dm0 * view[4 * r + 2];
    dJ11 += dm1 * view[4 * r + 1]; dJ12 += dm1 * view[4 * r + 2];
  }
  const float tz = jc.tz, tzi = 1.f / tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
  const float dtx = (jc.gate_x ? 1.f : 0.f) * -fx * tz2 * dJ02;
  const float dty = (jc.gate_y ? 1.f : 0.f) * -fy * tz2 * dJ12;
  float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + 2.f * fx * jc.tx * tz3 * dJ02 + 2.f * fy * jc.ty * tz3 * dJ12;
  dtz -= g2.y * tz2;  // inverse depth
  float dmean[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) dmean[j] = dtx * view[4 * j + 0] + dty * view[4 * j + 1] + dtz * view[4 * j + 2];

  // mean2D (NDC-scaled) -> mean3D
  {
    const float hx = xform_col(proj, 0, px, py, pz), hy = xform_col(proj, 1, px, py, pz);
    const float hw = xform_col(proj, 3, px, py, pz);
    const float mw = 1.f / (hw + 0.0000001f);
    const float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
#pragma unroll
    for (int j = 0; j < 3; ++j)
      dmean[j] += (proj[4 * j + 0] * mw - proj[4 * j + 3] * mul1) * dmx +
                  (proj[4 * j + 1] * mw - proj[4 * j + 3] * mul2) * dmy;
  }
  if (a.dL_dmeans2D) {
    put<ACC>(a.dL_dmeans2D + 3 * (size_t)i, dmx);
    put<ACC>(a.dL_dmeans2D + 3 * (size_t)i + 1, dmy);
    if (!ACC) a.dL_dmeans2D[3 * (size_t)i + 2] = 0.f;
  }
  if (a.dL_dcolors) {
    put<ACC>(a.dL_dcolors + 3 * (size_t)i, g1.z);
    put<ACC>(a.dL_dcolors + 3 * (size_t)i + 1, g1.w);
    put<ACC>(a.dL_dcolors + 3 * (size_t)i + 2, g2.x);
  }
  // densification statistics of this view (gaussian_model.py:174-181, visible_mask = radii > 0)
  // (atomics: concurrent views may update the same statistics; non-negative floats order like ints)
  if (a.max_radii2D) atomicMax(reinterpret_cast<int*>(a.max_radii2D + i), __float_as_int((float)radius));
  if (a.xyz_grad_accum) atomicAdd(a.xyz_grad_accum + i, sqrtf(dmx * dmx + dmy * dmy));
  if (a.xyz_grad_count) atomicAdd(a.xyz_grad_count + i, 1.f);

  // SH
  if (sh_bar) mbar_wait(sh_bar, 0);
  if (sh_row) {
    const float* __restrict__ cam = a.v.campos;
    const float dirx = px - cam[0], diry = py - cam[1], dirz = pz - cam[2];
    const float len = sqrtf(dirx * dirx + diry * diry + dirz * dirz);
    const float x = dirx / len, y = diry / len, z = dirz / len;
    const unsigned cl = cl_in;
    const float g3[3] = {(cl & 1u) ? 0.f : g1.z, (cl & 2u) ? 0.f : g1.w, (cl & 4u) ? 0.f : g2.x};
    float ddir[3];
    switch (a.v.deg) {
      case 0: sh_backward<0, DSH_ACC>(sh_row, dsh_row, M, sh_vec4, x, y, z, g3, ddir); break;
      case 1: sh_backward<1, DSH_ACC>(sh_row, dsh_row, M, sh_vec4, x, y, z, g3, ddir); break;
      case 2: sh_backward<2, DSH_ACC>(sh_row, dsh_row, M, sh_vec4, x, y, z, g3, ddir); break;
      default: sh_backward<3, DSH_ACC>(sh_row, dsh_row, M, sh_vec4, x, y, z, g3, ddir); break;
    }
    const float ng = x * ddir[0] + y * ddir[1] + z * ddir[2];
    dmean[0] += (ddir[0] - x * ng) / len;
    dmean[1] += (ddir[1] - y * ng) / len;
    dmean[2] += (ddir[2] - z * ng) / len;
  }
  if (a.dL_dmeans3D) {
#pragma unroll
    for (int j = 0; j < 3; ++j) put<ACC>(a.dL_dmeans3D + 3 * (size_t)i + j, dmean[j]);
  }

  // cov3D -> scale, rotation
  if (!a.g.cov3D_precomp) {
    if (!a.dL_dscales && !a.dL_drot) return;
    const float4 q = q_in;
    float R[3][3];
    quat_to_R(q, R);
    const float mod = a.v.scale_modifier;
    const float s[3] = {mod * sc0, mod * sc1, mod * sc2};
    const float Gm[3][3] = {{dS[0], 0.5f * dS[1], 0.5f * dS[2]},
                            {0.5f * dS[1], dS[3], 0.5f * dS[4]},
                            {0.5f * dS[2], 0.5f * dS[4], dS[5]}};
    float dR[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float dsk = 0.f;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float dLrk = 2.f * (Gm[r][0] * R[0][k] + Gm[r][1] * R[1][k] + Gm[r][2] * R[2][k]) * s[k];
        dsk += dLrk * R[r][k];
        dR[r][k] = dLrk * s[k];
      }
      if (a.dL_dscales) put<ACC>(a.dL_dscales + 3 * (size_t)i + k, mod * dsk);
    }
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    float4 dq;
    dq.x = 2.f * (-z * dR[0][1] + y * dR[0][2] + z * dR[1][0] - x * dR[1][2] - y * dR[2][0] + x * dR[2][1]);
    dq.y = 2.f * (y * dR[0][1] + z * dR[0][2] + y * dR[1][0] - 2.f * x * dR[1][1] - r * dR[1][2] + z * dR[2][0] +
                  r * dR[2][1] - 2.f * x * dR[2][2]);
    dq.z = 2.f * (-2.f * y * dR[0][0] + x * dR[0][1] + r * dR[0][2] + x * dR[1][0] + z * dR[1][2] - r * dR[2][0] +
                  z * dR[2][1] - 2.f * y * dR[2][2]);
    dq.w = 2.f * (-2.f * z * dR[0][0] - r * dR[0][1] + x * dR[0][2] + r * dR[1][0] - 2.f * z * dR[1][1] +
                  y * dR[1][2] + x * dR[2][0] + y * dR[2][1]);
    if (a.dL_drot) put4<ACC>(reinterpret_cast<float4*>(a.dL_drot) + i, dq);
  } else if (!ACC) {
    if (a.dL_dscales)
      for (int k = 0; k < 3; ++k) a.dL_dscales[3 * (size_t)i + k] = 0.f;
    if (a.dL_drot) reinterpret_cast<float4*>(a.dL_drot)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// The kernel proper: one warp per 32 consecutive Gaussians.  The warp's (32, M, 3) block of SH
// coefficients is ONE contiguous range of the (P, M, 3) tensor, and so is its block of dL_dsh:
//   * in:  an elected lane fetches the block with a 1-D bulk async copy (SASS UBLKCP) into the warp's
//          shared-memory block behind the warp's mbarrier -- only if some Gaussian of the warp is
//          visible -- while the lanes run the geometry part in registers;
//   * out: each lane overwrites its own row of the block with its dL_dsh row, and the elected lane
//          sends the block back with one bulk store, or -- accumulating -- one bulk REDUCTION
//          (cp.reduce.async.bulk .add.f32, SASS UBLKRED): the += of 192 B per Gaussian happens at
//          the L2, the SM never reads the old gradient.
// Everything is per warp (no CTA barrier).  Falls back to per-lane global accesses when the block
// is not 16-byte sized / aligned (odd M, unaligned tensors, ragged last warp).
template <bool ACC>
__global__ void __launch_bounds__(256, 3) preprocess_backward_kernel(const PreBwdArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long bars[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int P = a.g.P, M = a.g.M;
  const int w0 = blockIdx.x * 256 + warp * 32;  // first Gaussian of the warp
  if (w0 >= P) return;
  const int i = w0 + lane;
  const int nrows = min(32, P - w0);
  const int radius = i < P ? a.radii[i] : 0;
  const bool any_visible = __any_sync(0xffffffffu, radius > 0);
  if (ACC && !any_visible) return;  // nothing to add anywhere

  const bool need_sh = a.g.shs && (a.dL_dsh || a.dL_dmeans3D);
  const uint32_t row_bytes = (uint32_t)M * 12u, blk_bytes = (uint32_t)nrows * row_bytes;
  const bool bulk = a.use_bulk && need_sh && (blk_bytes % 16u == 0u);
  float* blk = reinterpret_cast<float*>(smem_raw + (size_t)warp * 32 * row_bytes);
  const bool bulk_in = bulk && any_visible;
  if (bulk_in) {
    if (lane == 0) {
      mbar_init(&bars[warp], 1);
      mbar_arrive_expect_tx(&bars[warp], blk_bytes);
      bulk_copy_g2s(blk, a.g.shs + (size_t)w0 * M * 3, blk_bytes, &bars[warp]);
    }
    __syncwarp();
  }

  if (bulk) {
    // rows live in the warp's shared block; 128-bit row accesses need 16-byte rows
    const bool vec4 = (M & 3) == 0;
    if (i < P) backward_one<ACC, false>(a, i, radius, blk + (size_t)lane * M * 3, a.dL_dsh ? blk + (size_t)lane * M * 3 : nullptr, vec4,
                                        bulk_in ? &bars[warp] : nullptr);
    if (bulk_in) mbar_wait(&bars[warp], 0);  // every lane, also those without a row: the block is complete below
    if (a.dL_dsh) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // my generic-proxy row writes -> async proxy
      __syncwarp();
      if (lane == 0) {
        float* dst = a.dL_dsh + (size_t)w0 * M * 3;
        if (ACC) bulk_reduce_add_s2g(dst, blk, blk_bytes);
        else bulk_copy_s2g(dst, blk, blk_bytes);
        bulk_commit_and_wait_read();  // the block must stay intact until the copy engine has read it
      }
    }
  } else if (i < P) {
    const bool vec4 = a.dL_dsh && ((M & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0);
    backward_one<ACC, ACC>(a, i, radius, need_sh ? a.g.shs + (size_t)i * M * 3 : nullptr,
                           a.dL_dsh ? a.dL_dsh + (size_t)i * M * 3 : nullptr, vec4);
  }
}

cudaError_t launch_preprocess_backward(const GsrView& v, const GsrGaussiansDev& g, const int32_t* radii,
                                       const GeomState& geom, const float* grad_acc, const PreBwdOut& o,
                                       cudaStream_t s) {
  if (g.P <= 0) return cudaSuccess;
  PreBwdArgs a;
  a.v = v; a.g = g; a.radii = radii; a.cov3D = geom.cov3D; a.clamped = geom.clamped; a.grad_acc = grad_acc;
  a.dL_dmeans2D = o.dL_dmeans2D; a.dL_dcolors = o.dL_dcolors; a.dL_dopacity = o.dL_dopacity;
  a.dL_dmeans3D = o.dL_dmeans3D; a.dL_dcov3D = o.dL_dcov3D; a.dL_dsh = (g.shs ? o.dL_dsh : nullptr);
  a.dL_dscales = o.dL_dscales; a.dL_drot = o.dL_drot;
  a.accumulate = o.accumulate;
  a.max_radii2D = o.max_radii2D; a.xyz_grad_accum = o.xyz_grad_accum; a.xyz_grad_count = o.xyz_grad_count;
  // bulk copies need 16-byte aligned global addresses; a warp's block offset 32*M*12 is a multiple of 16
  const size_t smem = g.shs ? (size_t)256 * g.M * 12 : 0;
  a.use_bulk = g.shs && smem <= 200 * 1024 && (reinterpret_cast<uintptr_t>(g.shs) & 15) == 0 &&
               (!a.dL_dsh || (reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0);
  auto kern = o.accumulate ? preprocess_backward_kernel<true> : preprocess_backward_kernel<false>;
  if (a.use_bulk && smem > 48 * 1024 - 256) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
  }
  kern<<<(g.P + 255) / 256, 256, a.use_bulk ? smem : 0, s>>>(a);
  count_launch();
  return cudaGetLastError();
}


// ---- all views of a step in ONE pass over the Gaussians -------------------------------------------------------------
// A training step renders V views and sums their gradients.  Run view by view, the kernel above reads the Gaussian's
// inputs (236 B at SH degree 3) and read-modify-writes its 236 B of the step's gradient buffer V times: 8 x 0.143 ms of a
// 11.4 ms C3 step.  Here every Gaussian is visited ONCE by a group of L lanes, one lane per view (views k, k + L, ... for
// V > L): a lane does its view's part of the chain from that view's own 48-byte row of raw pixel sums -- the same
// memory-level parallelism as the per-view kernel -- and the L lanes then combine:
//   * dL/dmean3D, dL/dcov3D, dL/dopacity, dL/dcolor and the statistics by L - 1 shuffle levels;
//   * dL/dsh = sum_v basis_v (x) dL/dcolor_v is a (16 x V)(V x 3) product: the lanes park basis_v and the clamp-masked
//     colour gradient (19 floats) in shared memory and every lane accumulates its own 16 / L coefficients of the row.
// The Gaussian's inputs are loaded once per group (the L lanes read the same addresses: one transaction), its SH row comes
// through L1, the gradient row is written -- or added -- once, cooperatively by the group.  The chain
// cov3D -> (scale, quaternion) is linear in dL/dcov3D with view-independent factors, so it runs once on the sum.  cov3D is
// recomputed from scale and rotation (the forward's exact expression; a Gaussian culled in one view has no stored copy).
// The densification statistics stay PER VIEW quantities (|| dL/dmean2D ||_2 is taken before the views are summed:
// gaussian_model.py:174-181).  dL/dmean2D itself is a per-view output and is not produced here.
// (A first version -- one thread per Gaussian looping over the views, sums in registers, dL/dsh in a second shared-memory
// block -- measured 1.04 ms per 8 views against 8 x 0.143: 120 registers, 16 warps per SM and eight dependent round trips
// per thread.)
constexpr int GSR_MAX_FUSED_VIEWS = 16;
struct ViewSlot {
  GsrView v;
  const int32_t* radii;
  const uint8_t* clamped;
  const float* grad_acc;
};
struct PreBwdViewsArgs {
  int V;
  ViewSlot view[GSR_MAX_FUSED_VIEWS];
  GsrGaussiansDev g;
  float* dL_dcolors; float* dL_dopacity; float* dL_dmeans3D; float* dL_dcov3D; float* dL_dsh; float* dL_dscales;
  float* dL_drot;
  float* max_radii2D; float* xyz_grad_accum; float* xyz_grad_count;
  int use_bulk;   // SH tensor 16-byte aligned: the warps fetch their rows with bulk async copies
};

// The lane's view of the Gaussian's SH row (in shared memory): basis[0..15] (zeros beyond the active degree) goes straight
// to the lane's slot of the dL/dsh exchange buffer; d(colour)/d(direction) is contracted with g3 FIRST
// (t_k = <sh_k, g3>: 3 FMA per coefficient) and then with the basis gradients (3 FMA): 6 instead of 12 per coefficient.
// VEC: the row is 16-byte aligned (M % 4 == 0): twelve 128-bit reads instead of 45 scalar ones.
template <int DEG, bool VEC>
__device__ __forceinline__ void sh_dir_grad(const float* __restrict__ sh, float x, float y, float z, const float* g3,
                                            float4* mine, float* ddir) {
  constexpr int N = (DEG + 1) * (DEG + 1);
  {
    float basis[N];
    sh_basis<DEG>(x, y, z, basis);
    float b16[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) b16[k] = k < N ? basis[k] : 0.f;
    mine[0] = make_float4(b16[0], b16[1], b16[2], b16[3]);
    mine[1] = make_float4(b16[4], b16[5], b16[6], b16[7]);
    mine[2] = make_float4(b16[8], b16[9], b16[10], b16[11]);
    mine[3] = make_float4(b16[12], b16[13], b16[14], b16[15]);
    mine[4] = make_float4(g3[0], g3[1], g3[2], 0.f);
  }
  ddir[0] = ddir[1] = ddir[2] = 0.f;
  if (N == 1) return;
  float c[3 * N];
  if (VEC) {
    static_assert(!VEC || (3 * N) % 4 == 0, "whole 128-bit words");
#pragma unroll
    for (int j = 0; j < (3 * N) / 4; ++j) {
      const float4 w = reinterpret_cast<const float4*>(sh)[j];
      c[4 * j] = w.x; c[4 * j + 1] = w.y; c[4 * j + 2] = w.z; c[4 * j + 3] = w.w;
    }
  } else {
#pragma unroll
    for (int j = 3; j < 3 * N; ++j) c[j] = sh[j];
  }
  float bx[N], by[N], bz[N];
  sh_basis_grad<DEG>(x, y, z, bx, by, bz);
#pragma unroll
  for (int k = 1; k < N; ++k) {
    const float t = fmaf(c[3 * k], g3[0], fmaf(c[3 * k + 1], g3[1], c[3 * k + 2] * g3[2]));
    ddir[0] = fmaf(bx[k], t, ddir[0]); ddir[1] = fmaf(by[k], t, ddir[1]); ddir[2] = fmaf(bz[k], t, ddir[2]);
  }
}

template <bool ACC, int L>
__global__ void __launch_bounds__(256, 3) preprocess_backward_views_kernel(const __grid_constant__ PreBwdViewsArgs a) {
  constexpr int GPW = 32 / L;       // Gaussians per warp
  constexpr int CPL = 16 / L;       // SH coefficients of the row a lane owns
  __shared__ float s_ex[8][32][20]; // per warp and lane: basis[16], masked dL/dcolour[3] of the lane's view
  __shared__ __align__(8) unsigned long long bars[8];
  extern __shared__ __align__(128) unsigned char smem_raw[];   // the warps' SH rows: 8 x GPW x M x 3 floats
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int sub = lane % L, grp = lane / L;
  const int P = a.g.P, M = a.g.M;
  const long long w0 = ((long long)blockIdx.x * 8 + warp) * GPW;   // first Gaussian of the warp
  const long long gi = w0 + grp;
  const bool valid = gi < P;
  const int i = valid ? (int)gi : 0;
  const bool want_sh = a.g.shs != nullptr;
  // The warp's GPW rows of SH coefficients are one contiguous range: in by ONE bulk async copy per warp (waited for at
  // the first use, under the geometry part), read from shared memory by every lane of the group -- through L1 the L
  // lanes' 45 scalar loads of the same row were what bound the kernel (ncu: L1/TEX 84 %, DRAM 14 %).
  float* sh_blk = reinterpret_cast<float*>(smem_raw) + (size_t)warp * GPW * M * 3;
  const int nrows = w0 < P ? (int)min((long long)GPW, (long long)P - w0) : 0;
  const uint32_t blk_bytes = (uint32_t)nrows * (uint32_t)M * 12u;
  const bool stage = want_sh && nrows > 0;
  const bool bulk_in = stage && a.use_bulk && (blk_bytes % 16u == 0u);
  if (bulk_in) {
    if (lane == 0) {
      mbar_init(&bars[warp], 1);
      mbar_arrive_expect_tx(&bars[warp], blk_bytes);
      bulk_copy_g2s(sh_blk, a.g.shs + (size_t)w0 * M * 3, blk_bytes, &bars[warp]);
    }
    __syncwarp();
  } else if (stage) {
    const float* src = a.g.shs + (size_t)w0 * M * 3;
    for (int j = lane; j < nrows * M * 3; j += 32) sh_blk[j] = src[j];
    __syncwarp();
  }
  bool sh_pending = bulk_in;
  const bool vec4 = (M & 3) == 0;
  const uint32_t gmask = (L == 32) ? 0xffffffffu : (((1u << L) - 1u) << (grp * L));   // the group's lanes

  float px = 0.f, py = 0.f, pz = 0.f, opac_in = 0.f, sc0 = 0.f, sc1 = 0.f, sc2 = 0.f;
  float4 q_in = make_float4(1.f, 0.f, 0.f, 0.f);
  float s6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (valid) {
    px = a.g.means3D[3 * (size_t)i]; py = a.g.means3D[3 * (size_t)i + 1]; pz = a.g.means3D[3 * (size_t)i + 2];
    opac_in = a.g.opacities[i];
    if (a.g.cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; ++k) s6[k] = a.g.cov3D_precomp[6 * (size_t)i + k];
    } else {
      q_in = reinterpret_cast<const float4*>(a.g.rotations)[i];
      sc0 = a.g.scales[3 * (size_t)i]; sc1 = a.g.scales[3 * (size_t)i + 1]; sc2 = a.g.scales[3 * (size_t)i + 2];
      cov3d_from_scale_rot(sc0, sc1, sc2, a.view[0].v.scale_modifier, q_in, s6);   // the forward's expression, bit for bit
    }
  }
  const float S[3][3] = {{s6[0], s6[1], s6[2]}, {s6[1], s6[3], s6[4]}, {s6[2], s6[4], s6[5]}};
  const float* __restrict__ sh_row = sh_blk + (size_t)grp * M * 3;

  float dmean[3] = {0.f, 0.f, 0.f}, dS[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, dcol[3] = {0.f, 0.f, 0.f};
  float dop = 0.f, norm_sum = 0.f, count = 0.f;
  int max_radius = 0;
  float dsh_mine[3 * CPL];
#pragma unroll
  for (int k = 0; k < 3 * CPL; ++k) dsh_mine[k] = 0.f;

  for (int k0 = 0; k0 < a.V; k0 += L) {
    const int k = k0 + sub;
    const int radius = (valid && k < a.V) ? a.view[k].radii[i] : 0;
    // the lane's basis_v and masked colour gradient go to shared memory for the dL/dsh product below; zeros when the
    // view does not see the Gaussian
    float4* mine = reinterpret_cast<float4*>(&s_ex[warp][lane][0]);
    if (!(radius > 0)) {
      const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
      mine[0] = z4; mine[1] = z4; mine[2] = z4; mine[3] = z4; mine[4] = z4;
    }
    if (radius > 0) {
      const ViewSlot& vs = a.view[k];
      const float4 g0 = reinterpret_cast<const float4*>(vs.grad_acc)[3 * (size_t)i];      // S0 S1 S2 S3
      const float4 g1 = reinterpret_cast<const float4*>(vs.grad_acc)[3 * (size_t)i + 1];  // S4 S5 dr dg
      const float4 g2 = reinterpret_cast<const float4*>(vs.grad_acc)[3 * (size_t)i + 2];  // db dinvz - -
      const unsigned cl = want_sh ? vs.clamped[i] : 0u;
      const float* __restrict__ view = vs.v.view;
      const float* __restrict__ proj = vs.v.proj;
      const float fx = vs.v.fx, fy = vs.v.fy;
      max_radius = max(max_radius, radius);
      JacobianCtx jc;
      build_jacobian(view, px, py, pz, fx, fy, vs.v.tanfovx, vs.v.tanfovy, jc);
      const Cov2D c0 = project_cov(s6, jc);
      const float ca = c0.a + 0.3f, cc = c0.c + 0.3f, cb = c0.b;
      const float det = fmaf(ca, cc, -(cb * cb));
      const float det_inv = 1.f / det;
      const float cA = cc * det_inv, cB = -cb * det_inv, cC = ca * det_inv;
      const float dmx = -(cA * g0.x + cB * g0.y) * (0.5f * (float)vs.v.W);
      const float dmy = -(cC * g0.y + cB * g0.x) * (0.5f * (float)vs.v.H);
      const float dA = -0.5f * g0.z, dB = -g0.w, dC = -0.5f * g1.x;
      const float d2inv = 1.f / (det * det + 0.0000001f);
      float da = d2inv * (-cc * cc * dA + cb * cc * dB - cb * cb * dC);
      float dc = d2inv * (-cb * cb * dA + ca * cb * dB - ca * ca * dC);
      float db = d2inv * (2.f * cb * cc * dA - (ca * cc + cb * cb) * dB + 2.f * ca * cb * dC);
      float dop_v = g1.y;
      if (vs.v.antialiasing) {
        const float det0 = c0.a * c0.c - c0.b * c0.b;
        const float ratio = det0 / det;
        const float h = sqrtf(fmaxf(0.000025f, ratio));
        const float dL_dop_eff = dop_v;
        dop_v = dL_dop_eff * h;
        if (ratio > 0.000025f) {
          const float dL_dratio = dL_dop_eff * opac_in / (2.f * h);
          const float inv = 1.f / det, r2 = ratio * inv;
          da += dL_dratio * (c0.c * inv - r2 * cc);
          dc += dL_dratio * (c0.a * inv - r2 * ca);
          db += dL_dratio * (-2.f * cb * inv + r2 * 2.f * cb);
        }
      }
      dop += dop_v;
      const float* m0 = jc.m0;
      const float* m1 = jc.m1;
      dS[0] += m0[0] * m0[0] * da + m0[0] * m1[0] * db + m1[0] * m1[0] * dc;
      dS[3] += m0[1] * m0[1] * da + m0[1] * m1[1] * db + m1[1] * m1[1] * dc;
      dS[5] += m0[2] * m0[2] * da + m0[2] * m1[2] * db + m1[2] * m1[2] * dc;
      dS[1] += 2.f * m0[0] * m0[1] * da + (m0[0] * m1[1] + m0[1] * m1[0]) * db + 2.f * m1[0] * m1[1] * dc;
      dS[2] += 2.f * m0[0] * m0[2] * da + (m0[0] * m1[2] + m0[2] * m1[0]) * db + 2.f * m1[0] * m1[2] * dc;
      dS[4] += 2.f * m0[1] * m0[2] * da + (m0[1] * m1[2] + m0[2] * m1[1]) * db + 2.f * m1[1] * m1[2] * dc;
      float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float Sm0 = S[r][0] * m0[0] + S[r][1] * m0[1] + S[r][2] * m0[2];
        const float Sm1 = S[r][0] * m1[0] + S[r][1] * m1[1] + S[r][2] * m1[2];
        const float dm0 = 2.f * da * Sm0 + db * Sm1;
        const float dm1 = 2.f * dc * Sm1 + db * Sm0;
        dJ00 += dm0 * view[4 * r + 0]; dJ02 += dm0 * view[4 * r + 2];
        dJ11 += dm1 * view[4 * r + 1]; dJ12 += dm1 * view[4 * r + 2];
      }
      const float tz = jc.tz, tzi = 1.f / tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
      const float dtx = (jc.gate_x ? 1.f : 0.f) * -fx * tz2 * dJ02;
      const float dty = (jc.gate_y ? 1.f : 0.f) * -fy * tz2 * dJ12;
      float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + 2.f * fx * jc.tx * tz3 * dJ02 + 2.f * fy * jc.ty * tz3 * dJ12;
      dtz -= g2.y * tz2;  // inverse depth
      float dm[3];
#pragma unroll
      for (int j = 0; j < 3; ++j) dm[j] = dtx * view[4 * j + 0] + dty * view[4 * j + 1] + dtz * view[4 * j + 2];
      {
        const float hx = xform_col(proj, 0, px, py, pz), hy = xform_col(proj, 1, px, py, pz);
        const float hw = xform_col(proj, 3, px, py, pz);
        const float mw = 1.f / (hw + 0.0000001f);
        const float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
#pragma unroll
        for (int j = 0; j < 3; ++j)
          dm[j] += (proj[4 * j + 0] * mw - proj[4 * j + 3] * mul1) * dmx +
                   (proj[4 * j + 1] * mw - proj[4 * j + 3] * mul2) * dmy;
      }
      norm_sum += sqrtf(dmx * dmx + dmy * dmy);   // the view's own norm, before the views are summed
      count += 1.f;
      dcol[0] += g1.z; dcol[1] += g1.w; dcol[2] += g2.x;
      if (want_sh) {
        const float* __restrict__ cam = vs.v.campos;
        const float dirx = px - cam[0], diry = py - cam[1], dirz = pz - cam[2];
        const float len = sqrtf(dirx * dirx + diry * diry + dirz * dirz);
        const float x = dirx / len, y = diry / len, z = dirz / len;
        const float g3[3] = {(cl & 1u) ? 0.f : g1.z, (cl & 2u) ? 0.f : g1.w, (cl & 4u) ? 0.f : g2.x};
        float ddir[3];
        if (sh_pending) { mbar_wait(&bars[warp], 0); sh_pending = false; }
        switch (vs.v.deg) {
          case 0: sh_dir_grad<0, false>(sh_row, x, y, z, g3, mine, ddir); break;
          case 1: sh_dir_grad<1, false>(sh_row, x, y, z, g3, mine, ddir); break;
          case 2: sh_dir_grad<2, false>(sh_row, x, y, z, g3, mine, ddir); break;
          default:
            if (vec4) sh_dir_grad<3, true>(sh_row, x, y, z, g3, mine, ddir);
            else sh_dir_grad<3, false>(sh_row, x, y, z, g3, mine, ddir);
            break;
        }
        const float ng = x * ddir[0] + y * ddir[1] + z * ddir[2];
        dm[0] += (ddir[0] - x * ng) / len;
        dm[1] += (ddir[1] - y * ng) / len;
        dm[2] += (ddir[2] - z * ng) / len;
      } else {
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        mine[0] = z4; mine[1] = z4; mine[2] = z4; mine[3] = z4; mine[4] = z4;
      }
      dmean[0] += dm[0]; dmean[1] += dm[1]; dmean[2] += dm[2];
    }
    if (want_sh && a.dL_dsh) {
      // dL/dsh row += sum over the group's views of basis_v (x) g3_v; a lane owns coefficients CPL * sub ...
      __syncwarp();
#pragma unroll
      for (int v = 0; v < L; ++v) {
        const float* src = &s_ex[warp][grp * L + v][0];
        const float4 gv = reinterpret_cast<const float4*>(src)[4];
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
          const float bv = src[CPL * sub + c];
          dsh_mine[3 * c + 0] = fmaf(bv, gv.x, dsh_mine[3 * c + 0]);
          dsh_mine[3 * c + 1] = fmaf(bv, gv.y, dsh_mine[3 * c + 1]);
          dsh_mine[3 * c + 2] = fmaf(bv, gv.z, dsh_mine[3 * c + 2]);
        }
      }
      __syncwarp();
    }
  }
  // ---- combine the group's lanes -----------------------------------------------------------------------------------
#pragma unroll
  for (int off = 1; off < L; off <<= 1) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      dmean[j] += __shfl_xor_sync(0xffffffffu, dmean[j], off);
      dcol[j] += __shfl_xor_sync(0xffffffffu, dcol[j], off);
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) dS[j] += __shfl_xor_sync(0xffffffffu, dS[j], off);
    dop += __shfl_xor_sync(0xffffffffu, dop, off);
    norm_sum += __shfl_xor_sync(0xffffffffu, norm_sum, off);
    count += __shfl_xor_sync(0xffffffffu, count, off);
    max_radius = max(max_radius, __shfl_xor_sync(0xffffffffu, max_radius, off));
  }
  (void)gmask;
  if (sh_pending) mbar_wait(&bars[warp], 0);   // the warp must not retire under an in-flight copy
  const bool seen = count > 0.f;
  if (!valid || (ACC && !seen)) return;   // accumulating: rows of Gaussians no view saw keep their value
  // ---- the Gaussian's row of every output, once: scalars by the group's first lane, dL/dsh by all of them ---------
  if (want_sh && a.dL_dsh) {
    float* row = a.dL_dsh + (size_t)i * M * 3;
    if (M >= 16 && (reinterpret_cast<uintptr_t>(row) & 7) == 0) {
      put_run<ACC, 3 * CPL>(row + 3 * CPL * sub, dsh_mine);   // the lane's 3 * CPL consecutive floats of the row
    } else {
#pragma unroll
      for (int c = 0; c < CPL; ++c) {
        const int coef = CPL * sub + c;
        if (coef < M) {
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) put<ACC>(row + 3 * coef + ch, dsh_mine[3 * c + ch]);
        }
      }
    }
    if (!ACC)   // stored coefficients beyond 16 (never used by a degree <= 3 evaluation) read as zeros
      for (int coef = 16 + sub; coef < M; coef += L)
        for (int ch = 0; ch < 3; ++ch) row[3 * coef + ch] = 0.f;
  }
  if (sub != 0) return;
  if (a.dL_dopacity) put<ACC>(a.dL_dopacity + i, dop);
  if (a.dL_dmeans3D)
    for (int j = 0; j < 3; ++j) put<ACC>(a.dL_dmeans3D + 3 * (size_t)i + j, dmean[j]);
  if (a.dL_dcolors)
    for (int j = 0; j < 3; ++j) put<ACC>(a.dL_dcolors + 3 * (size_t)i + j, dcol[j]);
  if (a.dL_dcov3D)
    for (int k = 0; k < 6; ++k) put<ACC>(a.dL_dcov3D + 6 * (size_t)i + k, dS[k]);
  if (seen) {
    if (a.max_radii2D) atomicMax(reinterpret_cast<int*>(a.max_radii2D + i), __float_as_int((float)max_radius));
    if (a.xyz_grad_accum) atomicAdd(a.xyz_grad_accum + i, norm_sum);
    if (a.xyz_grad_count) atomicAdd(a.xyz_grad_count + i, count);
  }
  float4 dq = make_float4(0.f, 0.f, 0.f, 0.f);
  float dsc[3] = {0.f, 0.f, 0.f};
  if (seen && !a.g.cov3D_precomp && (a.dL_dscales || a.dL_drot)) {
    const float4 q = q_in;
    float R[3][3];
    quat_to_R(q, R);
    const float mod = a.view[0].v.scale_modifier;
    const float s[3] = {mod * sc0, mod * sc1, mod * sc2};
    const float Gm[3][3] = {{dS[0], 0.5f * dS[1], 0.5f * dS[2]},
                            {0.5f * dS[1], dS[3], 0.5f * dS[4]},
                            {0.5f * dS[2], 0.5f * dS[4], dS[5]}};
    float dR[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float dsk = 0.f;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float dLrk = 2.f * (Gm[r][0] * R[0][k] + Gm[r][1] * R[1][k] + Gm[r][2] * R[2][k]) * s[k];
        dsk += dLrk * R[r][k];
        dR[r][k] = dLrk * s[k];
      }
      dsc[k] = mod * dsk;
    }
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    dq.x = 2.f * (-z * dR[0][1] + y * dR[0][2] + z * dR[1][0] - x * dR[1][2] - y * dR[2][0] + x * dR[2][1]);
    dq.y = 2.f * (y * dR[0][1] + z * dR[0][2] + y * dR[1][0] - 2.f * x * dR[1][1] - r * dR[1][2] + z * dR[2][0] +
                  r * dR[2][1] - 2.f * x * dR[2][2]);
    dq.z = 2.f * (-2.f * y * dR[0][0] + x * dR[0][1] + r * dR[0][2] + x * dR[1][0] + z * dR[1][2] - r * dR[2][0] +
                  z * dR[2][1] - 2.f * y * dR[2][2]);
    dq.w = 2.f * (-2.f * z * dR[0][0] - r * dR[0][1] + x * dR[0][2] + r * dR[1][0] - 2.f * z * dR[1][1] +
                  y * dR[1][2] + x * dR[2][0] + y * dR[2][1]);
  }
  if (a.dL_dscales)
    for (int k = 0; k < 3; ++k) put<ACC>(a.dL_dscales + 3 * (size_t)i + k, dsc[k]);
  if (a.dL_drot) {
    const float dqv[4] = {dq.x, dq.y, dq.z, dq.w};
    put_run<ACC, 4>(a.dL_drot + 4 * (size_t)i, dqv);
  }
}

// All views must share sh_degree, scale_modifier and the Gaussian tensors.  The caller falls back to the per-view
// kernel for what this one does not take.
bool preprocess_backward_views_supported(const GsrGaussiansDev& g, const PreBwdOut& o) {
  return !o.dL_dmeans2D;
}

cudaError_t launch_preprocess_backward_views(int V, const GsrView* views, const int32_t* const* radii,
                                             const GeomState* geoms, const float* const* grad_accs,
                                             const GsrGaussiansDev& g, const PreBwdOut& o, cudaStream_t s) {
  if (g.P <= 0 || V <= 0) return cudaSuccess;
  PreBwdOut out = o;
  for (int v0 = 0; v0 < V; v0 += GSR_MAX_FUSED_VIEWS) {
    PreBwdViewsArgs a;
    a.V = min(GSR_MAX_FUSED_VIEWS, V - v0);
    for (int k = 0; k < a.V; ++k) {
      a.view[k].v = views[v0 + k];
      a.view[k].radii = radii[v0 + k];
      a.view[k].clamped = geoms[v0 + k].clamped;
      a.view[k].grad_acc = grad_accs[v0 + k];
    }
    a.g = g;
    a.dL_dcolors = out.dL_dcolors; a.dL_dopacity = out.dL_dopacity; a.dL_dmeans3D = out.dL_dmeans3D;
    a.dL_dcov3D = out.dL_dcov3D; a.dL_dsh = g.shs ? out.dL_dsh : nullptr; a.dL_dscales = out.dL_dscales;
    a.dL_drot = out.dL_drot;
    a.max_radii2D = out.max_radii2D; a.xyz_grad_accum = out.xyz_grad_accum; a.xyz_grad_count = out.xyz_grad_count;
    // L lanes per Gaussian, one per view: 8 for a step of many views, 4 / 2 for few
    const int L = a.V > 4 ? 8 : (a.V > 2 ? 4 : 2);
    const int per_cta = 256 / L;
    const unsigned grid = (unsigned)((g.P + per_cta - 1) / per_cta);
    const size_t smem = g.shs ? (size_t)per_cta * g.M * 12 : 0;   // the CTA's SH rows
    a.use_bulk = g.shs && (reinterpret_cast<uintptr_t>(g.shs) & 15) == 0;
    void (*kern)(const PreBwdViewsArgs);
    if (out.accumulate)
      kern = L == 8 ? preprocess_backward_views_kernel<true, 8>
                    : (L == 4 ? preprocess_backward_views_kernel<true, 4> : preprocess_backward_views_kernel<true, 2>);
    else
      kern = L == 8 ? preprocess_backward_views_kernel<false, 8>
                    : (L == 4 ? preprocess_backward_views_kernel<false, 4> : preprocess_backward_views_kernel<false, 2>);
    if (smem + 21 * 1024 > 48 * 1024) {
      if (smem > 200 * 1024) return cudaErrorInvalidValue;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    kern<<<grid, 256, smem, s>>>(a);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    out.accumulate = 1;   // further groups of views add to the first
  }
  return cudaSuccess;
}

}  // namespace gsr
