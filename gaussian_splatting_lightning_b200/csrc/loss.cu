// loss.cu -- the reconstruction loss on the other side of the rasterizer in training_step
// (SURVEY.md 8(f) row 2): nn.L1Loss() of gs_lightning/lightning/gs_lightning_module.py:99,278,
// forward value and unit gradient in ONE pass over the two images.
//
//   loss           = mean(|image - target|)
//   dL_dimage[i]   = sign(image[i] - target[i]) / n          (sign(0) = 0, as torch's L1 backward)
//
// HBM bound: 8 B read (+ 4 B written when the gradient is wanted) per element, one launch.  The sum is
// deterministic: every CTA writes one double partial, the last CTA to finish (ticket) adds the partials
// in index order and resets the ticket, so the scratch can be reused by the next call on the same stream.
#include "../../include/gsr_b200.h"
#include "gsr_common.cuh"

namespace {

constexpr int L1_THREADS = 256;
constexpr int L1_MAX_CTAS = 148 * 8;

__device__ __forceinline__ float sgn(float d) { return (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f); }

template <bool GRAD>
__global__ void __launch_bounds__(L1_THREADS) l1_loss_kernel(const float* __restrict__ image,
                                                             const float* __restrict__ target, long long n,
                                                             float inv_n, float* __restrict__ dL_dimage,
                                                             float* __restrict__ out_loss, double* partials,
                                                             unsigned int* ticket) {
  const long long n4 = n >> 2;
  const float4* a4 = reinterpret_cast<const float4*>(image);
  const float4* b4 = reinterpret_cast<const float4*>(target);
  float4* g4 = reinterpret_cast<float4*>(dL_dimage);
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * L1_THREADS;
  for (long long i = (long long)blockIdx.x * L1_THREADS + threadIdx.x; i < n4; i += stride) {
    const float4 a = __ldg(a4 + i), b = __ldg(b4 + i);
    const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, dw = a.w - b.w;
    acc += (fabsf(dx) + fabsf(dy)) + (fabsf(dz) + fabsf(dw));
    if (GRAD) g4[i] = make_float4(sgn(dx) * inv_n, sgn(dy) * inv_n, sgn(dz) * inv_n, sgn(dw) * inv_n);
  }
  // ragged tail (n not a multiple of 4): the first CTA's first threads
  if (blockIdx.x == 0) {
    const long long i = (n4 << 2) + threadIdx.x;
    if (i < n) {
      const float d = image[i] - target[i];
      acc += fabsf(d);
      if (GRAD) dL_dimage[i] = sgn(d) * inv_n;
    }
  }
  // CTA reduction in double (fixed order: lane tree, then warp order)
  double v = (double)acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  __shared__ double warp_sum[L1_THREADS / 32];
  __shared__ bool last;
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < L1_THREADS / 32; ++w) s += warp_sum[w];
    partials[blockIdx.x] = s;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  // the last CTA: partials in index order, strided over the threads, then the same fixed tree
  double t = 0.0;
  for (unsigned int i = threadIdx.x; i < gridDim.x; i += L1_THREADS) t += ((volatile double*)partials)[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < L1_THREADS / 32; ++w) s += warp_sum[w];
    *out_loss = (float)(s / (double)n);
    *ticket = 0u;
  }
}

}  // namespace

namespace gsr {
size_t l1_scratch_bytes() { return (size_t)L1_MAX_CTAS * sizeof(double) + 128; }

cudaError_t launch_l1_loss(const float* image, const float* target, long long n, float* dL_dimage, float* out_loss,
                           void* scratch, cudaStream_t s) {
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 128);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  long long want = ((n >> 2) + L1_THREADS - 1) / L1_THREADS;
  int ctas = (int)(want < 1 ? 1 : (want > L1_MAX_CTAS ? L1_MAX_CTAS : want));
  const float inv_n = 1.0f / (float)n;
  if (dL_dimage)
    l1_loss_kernel<true><<<ctas, L1_THREADS, 0, s>>>(image, target, n, inv_n, dL_dimage, out_loss, partials, ticket);
  else
    l1_loss_kernel<false><<<ctas, L1_THREADS, 0, s>>>(image, target, n, inv_n, nullptr, out_loss, partials, ticket);
  count_launch();
  return cudaGetLastError();
}
}  // namespace gsr
