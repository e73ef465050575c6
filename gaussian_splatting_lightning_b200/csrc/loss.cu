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
#include "render_common.cuh"

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
  long long i = (long long)blockIdx.x * L1_THREADS + threadIdx.x;
  // two grid strides per trip: four 16-byte loads in flight per thread (the sum keeps the order of the one-stride loop)
  for (; i + stride < n4; i += 2 * stride) {
    const float4 a = __ldg(a4 + i), b = __ldg(b4 + i), c = __ldg(a4 + i + stride), d = __ldg(b4 + i + stride);
    const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, dw = a.w - b.w;
    const float ex = c.x - d.x, ey = c.y - d.y, ez = c.z - d.z, ew = c.w - d.w;
    acc += (fabsf(dx) + fabsf(dy)) + (fabsf(dz) + fabsf(dw));
    acc += (fabsf(ex) + fabsf(ey)) + (fabsf(ez) + fabsf(ew));
    if (GRAD) {
      g4[i] = make_float4(sgn(dx) * inv_n, sgn(dy) * inv_n, sgn(dz) * inv_n, sgn(dw) * inv_n);
      g4[i + stride] = make_float4(sgn(ex) * inv_n, sgn(ey) * inv_n, sgn(ez) * inv_n, sgn(ew) * inv_n);
    }
  }
  for (; i < n4; i += stride) {
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

// ---- gsr_l1_loss_noalloc: the same kernel on scratch that lives in the library image -----------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The
// scratch (ticket + one double per CTA: GSR_L1_NOALLOC_SCRATCH_BYTES = 9600 bytes) is a file-scope __device__
// array: every GPU that loads the library holds it until exit.  In HOST code the array's name is only a stand-in
// for the symbol, not a device address: its address is asked for once per device (that device current) with
// cudaGetSymbolAddress and kept.  All calls share the array: calls that overlap on the GPU (two streams, two
// threads) race on the ticket and the partial sums.
static __device__ __align__(256) unsigned char g_l1_noalloc_scratch[GSR_L1_NOALLOC_SCRATCH_BYTES];

namespace gsr {
cudaError_t launch_l1_loss_noalloc(const float* image, const float* target, long long n, float* dL_dimage,
                                   float* out_loss, cudaStream_t s) {
  static_assert(GSR_L1_NOALLOC_SCRATCH_BYTES >= (size_t)L1_MAX_CTAS * sizeof(double) + 128, "scratch too small");
  constexpr int MAX_DEVICES = 64;
  static void* volatile addr_of_device[MAX_DEVICES] = {};   // set once per device; a race sets the same value twice
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
  void* scratch = addr_of_device[dev];
  if (!scratch) {
    e = cudaGetSymbolAddress(&scratch, g_l1_noalloc_scratch);   // the CURRENT device's copy of the array
    if (e != cudaSuccess) return e;
    if (!scratch) return cudaErrorInvalidSymbol;
    addr_of_device[dev] = scratch;
  }
  // the array keeps the last call's contents: the ticket must start at zero (the partials are written before read)
  e = cudaMemsetAsync(scratch, 0, 128, s);
  if (e != cudaSuccess) return e;
  return launch_l1_loss(image, target, n, dL_dimage, out_loss, scratch, s);
}
}  // namespace gsr

// ---- fused SSIM: value + gradient in ONE kernel ----------------------------------------------------------
// Reference: submodules/fused-ssim (ssim.cu:187-366 fusedssimCUDA / fusedssim_backwardCUDA, fused_ssim/__init__.py:8-41),
// used as `1 - fused_ssim(rendered, gt)` at gs_lightning/lightning/gs_lightning_module.py:100,279.
//
//   mu1 = g*x, mu2 = g*y, s11 = g*(x x) - mu1^2, s22 = g*(y y) - mu2^2, s12 = g*(x y) - mu1 mu2     (g: 11x11 Gaussian,
//   A = mu1^2 + mu2^2 + C1, B = s11 + s22 + C2, C = 2 mu1 mu2 + C1, D = 2 s12 + C2, m = C D / (A B)   sigma 1.5, zero pad)
//   ssim = mean(m)  over the image ("same") or over the image minus a 5-pixel border ("valid")
//   d ssim / d x = g*(w dm/dmu1) + 2 x g*(w dm/ds11) + y g*(w dm/ds12),  w = 1/N inside the averaged region, else 0
//
// The reference runs two kernels with three full-size derivative maps in between (28 B/element written, 32 read, plus
// four memsets).  Here a CTA recomputes the derivative maps on its tile + 5-pixel halo from the images on a 10-pixel
// halo, all in shared memory: HBM traffic is the two images in and the gradient out, 12 B/element, one launch.
namespace {

constexpr int SS_TX = 32, SS_TY = 32, SS_R = 5;
constexpr int SS_W2 = SS_TX + 2 * SS_R, SS_H2 = SS_TY + 2 * SS_R;      // derivative-map region 42 x 42
constexpr int SS_W1 = SS_TX + 4 * SS_R, SS_H1 = SS_TY + 4 * SS_R;      // image region 52 x 52
constexpr int SS_THREADS = 256;
// the derivative maps (3 x 42 x 42 floats) live in the image region (2 x 52 x 52), which is dead once the horizontal sums
// are taken: 65.3 KB per CTA, THREE CTAs of 256 threads per SM (86.5 KB and 2 x 384 threads before: the six barrier-
// separated phases of a CTA leave the SM to the other CTAs while it waits)
constexpr size_t SS_SMEM = (size_t)(2 * SS_H1 * SS_W1 + 5 * SS_H1 * SS_W2) * sizeof(float);
static_assert(3 * SS_H2 * SS_W2 <= 2 * SS_H1 * SS_W1, "the derivative maps must fit into the image region");

__constant__ float kGauss[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f,
                                 0.10936068743467331f,  0.21300552785396576f,   0.26601171493530273f,
                                 0.21300552785396576f,  0.10936068743467331f,   0.036000773310661316f,
                                 0.0075987582094967365f, 0.001028380123898387f};

// Shared-memory layout (floats): the two images interleaved as (x, y) pairs on the 52 x 52 region; the horizontal
// sums as (sum x, sum y) pairs, (sum xx, sum yy) pairs and sum xy on 52 x 42; the weighted derivative maps as
// (dm/dmu1, dm/ds11) pairs and dm/ds12 on 42 x 42.  Pairs go through the packed FP32 instructions (FFMA2) with the
// Gaussian weight broadcast; every thread owns a strip of 3-4 consecutive outputs along the filter direction so that
// a loaded value feeds up to 11 x 3 multiply-adds from registers.
constexpr int SS_N1 = SS_H1 * SS_W1, SS_NH = SS_H1 * SS_W2, SS_N2 = SS_H2 * SS_W2;

// loss = w_l1 * mean|x - y| + w_ssim * ssim + w_const   (fused_ssim alone: 0, 1, 0; the reference's training loss
// (1 - w) L1 + w (1 - ssim), gs_lightning_module.py:281-284: 1 - w, -w, w).  inv_n_l1 = 1 / (planes H W).
struct SsimLossWeights { float w_l1, w_ssim, w_const, inv_n_l1; };

template <bool GRAD>
__global__ void __launch_bounds__(SS_THREADS, 3) fused_ssim_kernel(const float* __restrict__ img1,
                                                                const float* __restrict__ img2, int H, int W, float C1,
                                                                float C2, int valid, float inv_n,
                                                                const SsimLossWeights lw, float* __restrict__ dL_dimg1,
                                                                float* out_loss, float* out_ssim, float* out_l1, double* partials,
                                                                unsigned int* ticket) {
  using gsr::f32x2; using gsr::pk2; using gsr::upk2; using gsr::fma2; using gsr::mul2;
  extern __shared__ __align__(16) float smem[];
  float2* sxy = reinterpret_cast<float2*>(smem);                    // [52][52] (x, y)
  float2* hA = reinterpret_cast<float2*>(smem + 2 * SS_N1);         // [52][42] (sum x, sum y)
  float2* hB = hA + SS_NH;                                          // [52][42] (sum xx, sum yy)
  float* hC = reinterpret_cast<float*>(hB + SS_NH);                 // [52][42] sum xy
  float2* pA = reinterpret_cast<float2*>(smem);                     // [42][42] w (dm/dmu1, dm/ds11): over sxy, after step 2
  float* pC = reinterpret_cast<float*>(pA + SS_N2);                 // [42][42] w dm/ds12
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * SS_TX, y0 = blockIdx.y * SS_TY;
  const size_t plane = (size_t)blockIdx.z * H * W;
  const float* __restrict__ a = img1 + plane;
  const float* __restrict__ b = img2 + plane;
  float g[11];
#pragma unroll
  for (int k = 0; k < 11; ++k) g[k] = kGauss[k];

  // 1. the two images on the tile + 10-pixel halo (zero outside the image); all of a thread's loads are issued
  //    before the first store so that their latencies overlap.  The L1 term is taken here, from the tile's own pixels.
  float local = 0.f, local_l1 = 0.f;
  {
    constexpr int IT = (SS_N1 + SS_THREADS - 1) / SS_THREADS;
    float2 t[IT];
#pragma unroll
    for (int k = 0; k < IT; ++k) {
      const int i = tid + k * SS_THREADS;
      const int r = i / SS_W1, c = i - r * SS_W1;
      const int y = y0 - 2 * SS_R + r, x = x0 - 2 * SS_R + c;
      const bool in = i < SS_N1 && (unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W;
      const size_t o = in ? (size_t)y * W + x : 0;
      t[k] = in ? make_float2(__ldg(a + o), __ldg(b + o)) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int k = 0; k < IT; ++k) {
      const int i = tid + k * SS_THREADS;
      if (i < SS_N1) sxy[i] = t[k];
      const int r = i / SS_W1, c = i - r * SS_W1;
      const bool centre = i < SS_N1 && r >= 2 * SS_R && r < 2 * SS_R + SS_TY && c >= 2 * SS_R && c < 2 * SS_R + SS_TX;
      if (centre) local_l1 += fabsf(t[k].x - t[k].y);     // zero outside the image
    }
  }
  __syncthreads();
  // 2. horizontal 11-tap sums: a thread owns 3 consecutive columns of one row
  for (int it = tid; it < SS_H1 * (SS_W2 / 3); it += SS_THREADS) {
    const int r = it / (SS_W2 / 3), c = (it - r * (SS_W2 / 3)) * 3;
    f32x2 v[13], q[13];
    float xy[13];
#pragma unroll
    for (int k = 0; k < 13; ++k) {
      const float2 t = sxy[r * SS_W1 + c + k];
      v[k] = pk2(t.x, t.y);
      q[k] = mul2(v[k], v[k]);
      xy[k] = t.x * t.y;
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      f32x2 s01 = 0ull, s23 = 0ull;
      float s4 = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) {
        const f32x2 w2 = pk2(g[k], g[k]);
        s01 = fma2(w2, v[j + k], s01);
        s23 = fma2(w2, q[j + k], s23);
        s4 = fmaf(g[k], xy[j + k], s4);
      }
      const int o = r * SS_W2 + c + j;
      float lo, hi;
      upk2(s01, lo, hi); hA[o] = make_float2(lo, hi);
      upk2(s23, lo, hi); hB[o] = make_float2(lo, hi);
      hC[o] = s4;
    }
  }
  __syncthreads();
  // 3. vertical sums -> statistics -> m and its three partial derivatives: a thread owns 3 consecutive rows of one column
  for (int it = tid; it < (SS_H2 / 3) * SS_W2; it += SS_THREADS) {
    const int rs = it / SS_W2, c = it - rs * SS_W2, r0 = rs * 3;
    f32x2 vA[13], vB[13];
    float vC[13];
#pragma unroll
    for (int k = 0; k < 13; ++k) {
      const int o = (r0 + k) * SS_W2 + c;
      const float2 ta = hA[o], tb = hB[o];
      vA[k] = pk2(ta.x, ta.y);
      vB[k] = pk2(tb.x, tb.y);
      vC[k] = hC[o];
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      f32x2 m12 = 0ull, e12 = 0ull;
      float exy = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) {
        const f32x2 w2 = pk2(g[k], g[k]);
        m12 = fma2(w2, vA[j + k], m12);
        e12 = fma2(w2, vB[j + k], e12);
        exy = fmaf(g[k], vC[j + k], exy);
      }
      float mu1, mu2, exx, eyy;
      upk2(m12, mu1, mu2);
      upk2(e12, exx, eyy);
      const int r = r0 + j;
      const int y = y0 - SS_R + r, x = x0 - SS_R + c;
      const bool in_img = (unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W;
      const bool counted = in_img && (!valid || (y >= SS_R && y < H - SS_R && x >= SS_R && x < W - SS_R));
      const float s11 = exx - mu1 * mu1, s22 = eyy - mu2 * mu2, s12 = exy - mu1 * mu2;
      const float A = mu1 * mu1 + mu2 * mu2 + C1, B = s11 + s22 + C2;
      const float C = 2.f * mu1 * mu2 + C1, D = 2.f * s12 + C2;
      const float rA = __frcp_rn(A), rB = __frcp_rn(B);
      const float rAB = rA * rB;
      const float CD = C * D;
      const float m = CD * rAB;
      const bool centre = r >= SS_R && r < SS_R + SS_TY && c >= SS_R && c < SS_R + SS_TX;
      if (centre && counted) local += m;
      if (GRAD) {
        const float wgt = counted ? inv_n : 0.f;
        const float wm = wgt * m;                                     // w C D / (A B)
        const float d_mu1 = 2.f * (wgt * rAB) * mu2 * (D - C) + 2.f * mu1 * wm * (rB - rA);
        const int o = r * SS_W2 + c;
        pA[o] = make_float2(d_mu1, -wm * rB);
        pC[o] = 2.f * C * (wgt * rAB);
      }
    }
  }
  if (GRAD) {
    __syncthreads();
    // 4. horizontal sums of the weighted derivative maps: a thread owns 4 consecutive columns of one row;
    //    results reuse the storage of step 2
    for (int it = tid; it < SS_H2 * (SS_TX / 4); it += SS_THREADS) {
      const int r = it / (SS_TX / 4), c = (it - r * (SS_TX / 4)) * 4;
      f32x2 v[14];
      float u[14];
#pragma unroll
      for (int k = 0; k < 14; ++k) {
        const float2 t = pA[r * SS_W2 + c + k];
        v[k] = pk2(t.x, t.y);
        u[k] = pC[r * SS_W2 + c + k];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        f32x2 s01 = 0ull;
        float s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
          s01 = fma2(pk2(g[k], g[k]), v[j + k], s01);
          s2 = fmaf(g[k], u[j + k], s2);
        }
        float lo, hi;
        upk2(s01, lo, hi);
        hA[r * SS_TX + c + j] = make_float2(lo, hi);
        hC[r * SS_TX + c + j] = s2;
      }
    }
    __syncthreads();
    // 5. vertical sums and the gradient: a thread owns 4 consecutive rows of one column
    for (int it = tid; it < (SS_TY / 4) * SS_TX; it += SS_THREADS) {
      const int rs = it / SS_TX, c = it - rs * SS_TX, r0 = rs * 4;
      f32x2 v[14];
      float u[14];
#pragma unroll
      for (int k = 0; k < 14; ++k) {
        const float2 t = hA[(r0 + k) * SS_TX + c];
        v[k] = pk2(t.x, t.y);
        u[k] = hC[(r0 + k) * SS_TX + c];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        f32x2 t01 = 0ull;
        float t2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
          t01 = fma2(pk2(g[k], g[k]), v[j + k], t01);
          t2 = fmaf(g[k], u[j + k], t2);
        }
        const int y = y0 + r0 + j, x = x0 + c;
        if (y < H && x < W) {
          float t0, t1;
          upk2(t01, t0, t1);
          // the pixel itself: from global memory again (an L2 hit; its shared-memory copy holds the derivative maps now)
          const float2 px = make_float2(__ldg(a + (size_t)y * W + x), __ldg(b + (size_t)y * W + x));
          const float dss = t0 + 2.f * px.x * t1 + px.y * t2;          // d ssim / d x
          const float d = px.x - px.y;
          const float sg = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
          dL_dimg1[plane + (size_t)y * W + x] = lw.w_ssim * dss + (lw.w_l1 * lw.inv_n_l1) * sg;
        }
      }
    }
  }

  // deterministic sums of m and |x - y|: CTA partials in double, the last CTA adds the partials in index order
  double v = (double)local, v1 = (double)local_l1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v += __shfl_down_sync(0xffffffffu, v, o);
    v1 += __shfl_down_sync(0xffffffffu, v1, o);
  }
  __shared__ double warp_sum[2][SS_THREADS / 32];
  __shared__ bool last;
  if ((tid & 31) == 0) { warp_sum[0][tid >> 5] = v; warp_sum[1][tid >> 5] = v1; }
  __syncthreads();
  const unsigned int nblocks = gridDim.x * gridDim.y * gridDim.z;
  const unsigned int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  if (tid == 0) {
    double s = 0.0, s1 = 0.0;
#pragma unroll
    for (int w = 0; w < SS_THREADS / 32; ++w) { s += warp_sum[0][w]; s1 += warp_sum[1][w]; }
    partials[2 * bid] = s;
    partials[2 * bid + 1] = s1;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == nblocks - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double t = 0.0, t1 = 0.0;
  for (unsigned int i = tid; i < nblocks; i += SS_THREADS) {
    t += ((volatile double*)partials)[2 * i];
    t1 += ((volatile double*)partials)[2 * i + 1];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    t += __shfl_down_sync(0xffffffffu, t, o);
    t1 += __shfl_down_sync(0xffffffffu, t1, o);
  }
  __syncthreads();
  if ((tid & 31) == 0) { warp_sum[0][tid >> 5] = t; warp_sum[1][tid >> 5] = t1; }
  __syncthreads();
  if (tid == 0) {
    double s = 0.0, s1 = 0.0;
#pragma unroll
    for (int w = 0; w < SS_THREADS / 32; ++w) { s += warp_sum[0][w]; s1 += warp_sum[1][w]; }
    const float ssim = (float)(s * (double)inv_n), l1 = (float)(s1 * (double)lw.inv_n_l1);
    if (out_loss) *out_loss = lw.w_l1 * l1 + lw.w_ssim * ssim + lw.w_const;
    if (out_ssim) *out_ssim = ssim;
    if (out_l1) *out_l1 = l1;
    *ticket = 0u;
  }
}

}  // namespace

namespace gsr {
size_t ssim_scratch_bytes(int planes, int H, int W) {
  const size_t blocks = (size_t)((W + SS_TX - 1) / SS_TX) * ((H + SS_TY - 1) / SS_TY) * (size_t)(planes > 0 ? planes : 1);
  return 128 + blocks * 2 * sizeof(double);
}

// *out_loss = w_l1 * L1 + w_ssim * ssim + w_const, *out_ssim = ssim, *out_l1 = L1 (each may be NULL)
cudaError_t launch_fused_ssim(const float* img1, const float* img2, int planes, int H, int W, float C1, float C2,
                              int valid, float w_l1, float w_ssim, float w_const, float* dL_dimg1, float* out_loss,
                              float* out_ssim, float* out_l1, void* scratch, cudaStream_t s) {
  // per call: the attribute belongs to the (function, device) pair and a process may drive several devices
  cudaError_t e = cudaFuncSetAttribute(fused_ssim_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SS_SMEM);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(fused_ssim_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SS_SMEM);
  if (e != cudaSuccess) return e;
  const long long hh = valid ? H - 2 * SS_R : H, ww = valid ? W - 2 * SS_R : W;
  const double n = (double)planes * (double)(hh > 0 ? hh : 0) * (double)(ww > 0 ? ww : 0);
  const float inv_n = n > 0 ? (float)(1.0 / n) : 0.f;
  SsimLossWeights lw;
  lw.w_l1 = w_l1; lw.w_ssim = w_ssim; lw.w_const = w_const;
  lw.inv_n_l1 = (float)(1.0 / ((double)planes * H * W));
  dim3 grid((W + SS_TX - 1) / SS_TX, (H + SS_TY - 1) / SS_TY, planes);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 128);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  if (dL_dimg1)
    fused_ssim_kernel<true><<<grid, SS_THREADS, SS_SMEM, s>>>(img1, img2, H, W, C1, C2, valid, inv_n, lw, dL_dimg1, out_loss,
                                                             out_ssim, out_l1, partials, ticket);
  else
    fused_ssim_kernel<false><<<grid, SS_THREADS, SS_SMEM, s>>>(img1, img2, H, W, C1, C2, valid, inv_n, lw, nullptr, out_loss,
                                                              out_ssim, out_l1, partials, ticket);
  count_launch();
  return cudaGetLastError();
}
}  // namespace gsr
