// optimizer.cu -- what runs right after the preprocess backward in every training step
// (SURVEY.md 8(f) row 1): the densification-statistics fold and the Adam update, each ONE launch.
//
//   adam_kernel            multi-tensor Adam over the six parameter groups of
//                          gs_lightning/lightning/gs_lightning_module.py:120-127 (torch.optim.Adam,
//                          configs: eps 1e-15, betas (0.9, 0.999), no weight decay / amsgrad), reading the
//                          gradients where the backward kernel left them (the flat all-reduce buffer) and
//                          optionally zeroing them for the next step (optimizer.zero_grad(), :168-170).
//                          HBM bound: 16 B read + 12 B written (+4 with zero_grad) per parameter.
//   densify_fold_kernel    folds the step's (all-reduced) statistics deltas into the model's running
//                          statistics, gs_lightning/modules/gaussian_model.py:174-181.  20 B read, 12 B written
//                          per Gaussian.
//
// The arithmetic is one explicit chain of IEEE mul / add / div / sqrt (this unit is compiled with
// -fmad=false), so oracle/training.py reproduces it bit for bit in numpy float32.
#include "../../include/gsr_b200.h"
#include "gsr_common.cuh"

namespace {

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_VEC_PER_THREAD = 4;                                      // float4s per thread per chunk
constexpr long long ADAM_CHUNK = (long long)ADAM_THREADS * ADAM_VEC_PER_THREAD * 4;  // 4096 floats

struct AdamArgs {
  float* param[GSR_ADAM_MAX_GROUPS];
  float* grad[GSR_ADAM_MAX_GROUPS];
  float* m[GSR_ADAM_MAX_GROUPS];
  float* v[GSR_ADAM_MAX_GROUPS];
  long long n[GSR_ADAM_MAX_GROUPS];
  long long chunk_end[GSR_ADAM_MAX_GROUPS];  // exclusive prefix end of each group's chunks
  float neg_step_size[GSR_ADAM_MAX_GROUPS];  // -(lr / (1 - beta1^t))
  int vec_ok[GSR_ADAM_MAX_GROUPS];           // all four pointers 16-byte aligned
  int groups;
  float w1;        // 1 - beta1
  float beta2;
  float w2;        // 1 - beta2
  float bc2_sqrt;  // sqrt(1 - beta2^t)
  float eps;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float nss, float w1, float beta2,
                                         float w2, float bc2_sqrt, float eps) {
  m = __fadd_rn(m, __fmul_rn(w1, __fsub_rn(g, m)));                       // exp_avg.lerp_(grad, 1 - beta1)
  v = __fadd_rn(__fmul_rn(v, beta2), __fmul_rn(__fmul_rn(w2, g), g));     // mul_(beta2).addcmul_(g, g, 1 - beta2)
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps); // sqrt / bias_correction2_sqrt + eps
  p = __fadd_rn(p, __fmul_rn(nss, __fdiv_rn(m, denom)));                  // addcdiv_(exp_avg, denom, -step_size)
}

template <bool ZERO_GRAD>
__global__ void __launch_bounds__(ADAM_THREADS) adam_kernel(const AdamArgs a) {
  const long long total = a.chunk_end[a.groups - 1];
  for (long long c = blockIdx.x; c < total; c += gridDim.x) {
    int gi = 0;
    while (c >= a.chunk_end[gi]) ++gi;
    const long long first = gi ? a.chunk_end[gi - 1] : 0;
    const long long base = (c - first) * ADAM_CHUNK;
    const long long n = a.n[gi];
    float* __restrict__ P = a.param[gi];
    float* __restrict__ G = a.grad[gi];
    float* __restrict__ M = a.m[gi];
    float* __restrict__ V = a.v[gi];
    const float nss = a.neg_step_size[gi];
    if (a.vec_ok[gi] && base + ADAM_CHUNK <= n) {
      float4 p[ADAM_VEC_PER_THREAD], g[ADAM_VEC_PER_THREAD], m[ADAM_VEC_PER_THREAD], v[ADAM_VEC_PER_THREAD];
#pragma unroll
      for (int k = 0; k < ADAM_VEC_PER_THREAD; ++k) {
        const long long i = (base >> 2) + k * ADAM_THREADS + threadIdx.x;
        p[k] = reinterpret_cast<const float4*>(P)[i];
        g[k] = reinterpret_cast<const float4*>(G)[i];
        m[k] = reinterpret_cast<const float4*>(M)[i];
        v[k] = reinterpret_cast<const float4*>(V)[i];
      }
#pragma unroll
      for (int k = 0; k < ADAM_VEC_PER_THREAD; ++k) {
        adam_one(p[k].x, g[k].x, m[k].x, v[k].x, nss, a.w1, a.beta2, a.w2, a.bc2_sqrt, a.eps);
        adam_one(p[k].y, g[k].y, m[k].y, v[k].y, nss, a.w1, a.beta2, a.w2, a.bc2_sqrt, a.eps);
        adam_one(p[k].z, g[k].z, m[k].z, v[k].z, nss, a.w1, a.beta2, a.w2, a.bc2_sqrt, a.eps);
        adam_one(p[k].w, g[k].w, m[k].w, v[k].w, nss, a.w1, a.beta2, a.w2, a.bc2_sqrt, a.eps);
      }
#pragma unroll
      for (int k = 0; k < ADAM_VEC_PER_THREAD; ++k) {
        const long long i = (base >> 2) + k * ADAM_THREADS + threadIdx.x;
        reinterpret_cast<float4*>(P)[i] = p[k];
        reinterpret_cast<float4*>(M)[i] = m[k];
        reinterpret_cast<float4*>(V)[i] = v[k];
        if (ZERO_GRAD) reinterpret_cast<float4*>(G)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    } else {
      const long long end = (base + ADAM_CHUNK < n) ? base + ADAM_CHUNK : n;
      for (long long i = base + threadIdx.x; i < end; i += ADAM_THREADS) {
        float p = P[i], m = M[i], v = V[i];
        adam_one(p, G[i], m, v, nss, a.w1, a.beta2, a.w2, a.bc2_sqrt, a.eps);
        P[i] = p; M[i] = m; V[i] = v;
        if (ZERO_GRAD) G[i] = 0.f;
      }
    }
  }
}

__global__ void __launch_bounds__(256) densify_fold_kernel(int P, const float* __restrict__ d_max_radii,
                                                           const float* __restrict__ d_accum,
                                                           const float* __restrict__ d_count,
                                                           float* __restrict__ max_radii2D,
                                                           float* __restrict__ xyz_grad_accum,
                                                           float* __restrict__ xyz_grad_count) {
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
    const float cnt = d_count[i];
    if (cnt > 0.f) {  // seen by at least one view of the step: radii > 0 there
      if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], d_max_radii[i]);
      if (xyz_grad_accum) xyz_grad_accum[i] = __fadd_rn(xyz_grad_accum[i], d_accum[i]);
      if (xyz_grad_count) xyz_grad_count[i] = __fadd_rn(xyz_grad_count[i], cnt);
    }
  }
}

}  // namespace

namespace gsr {

cudaError_t launch_adam(const GsrAdamGroup* groups, int num_groups, int step, double beta1, double beta2, double eps,
                        int zero_grad, cudaStream_t s) {
  AdamArgs a;
  long long chunks = 0;
  int k = 0;
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  for (int i = 0; i < num_groups; ++i) {
    if (groups[i].n <= 0) continue;
    a.param[k] = groups[i].param; a.grad[k] = groups[i].grad; a.m[k] = groups[i].exp_avg; a.v[k] = groups[i].exp_avg_sq;
    a.n[k] = groups[i].n;
    chunks += (groups[i].n + ADAM_CHUNK - 1) / ADAM_CHUNK;
    a.chunk_end[k] = chunks;
    a.neg_step_size[k] = (float)(-(groups[i].lr / bc1));
    a.vec_ok[k] = ((((uintptr_t)a.param[k]) | ((uintptr_t)a.grad[k]) | ((uintptr_t)a.m[k]) | ((uintptr_t)a.v[k])) & 15) == 0;
    ++k;
  }
  if (k == 0) return cudaSuccess;
  for (int i = k; i < GSR_ADAM_MAX_GROUPS; ++i) { a.chunk_end[i] = chunks; a.n[i] = 0; }
  a.groups = k;
  a.w1 = (float)(1.0 - beta1);
  a.beta2 = (float)beta2;
  a.w2 = (float)(1.0 - beta2);
  a.bc2_sqrt = (float)sqrt(bc2);
  a.eps = (float)eps;
  const long long want = chunks < 148LL * 8 ? chunks : 148LL * 8;   // 8 CTAs of 256 threads per SM, grid-stride
  if (zero_grad) adam_kernel<true><<<(int)want, ADAM_THREADS, 0, s>>>(a);
  else           adam_kernel<false><<<(int)want, ADAM_THREADS, 0, s>>>(a);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_densify_fold(int P, const float* d_max_radii, const float* d_accum, const float* d_count,
                                float* max_radii2D, float* xyz_grad_accum, float* xyz_grad_count, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  int ctas = (P + 255) / 256;
  if (ctas > 148 * 8) ctas = 148 * 8;
  densify_fold_kernel<<<ctas, 256, 0, s>>>(P, d_max_radii, d_accum, d_count, max_radii2D, xyz_grad_accum,
                                           xyz_grad_count);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
