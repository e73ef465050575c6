// preprocess.cu -- K1 (markVisible) and K2 (per-Gaussian preprocess) for sm_100a.
//
// Compiled with -fmad=false: see gsr_math.cuh for the canonical rounding sequence.
//
// K2 design (HBM-bound: 52 B read per Gaussian + 12*M B of SH for visible ones, 83 B written):
//   * one thread per Gaussian, 256 per CTA;
//   * the CTA's 256 x M x 3 SH coefficients are ONE contiguous range of the (P,M,3) tensor, so
//     a single elected thread fetches it with a 1-D bulk async copy (cp.async.bulk, SASS
//     UBLKCP) into shared memory behind an mbarrier while all threads do projection, EWA
//     covariance and culling in registers; SH evaluation then reads the thread's row with
//     128-bit shared loads.  This replaces upstream's 48 stride-192-B scalar loads per thread.
//   * outputs: radii, tiles_touched, a packed 64-byte record per Gaussian (GsrRec), cov3D,
//     clamp bits.
#include "gsr_math.cuh"
#include "render_common.cuh"

namespace gsr {

// a4 / K1: camera_tools.py:5-8, rasterize.py:23-26.
__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                           const float* __restrict__ view,
                                                           uint8_t* __restrict__ present) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float x = means3D[3 * i], y = means3D[3 * i + 1], z = means3D[3 * i + 2];
  present[i] = xform_col(view, 2, x, y, z) > GSR_NEAR ? 1 : 0;
}

cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
  count_launch();
  return cudaGetLastError();
}

struct PreArgs {
  GsrView v;
  GsrGaussiansDev g;
  int32_t* radii;
  GsrRec* recs;
  float* cov3D;
  uint8_t* clamped;
  uint32_t* tiles_touched;
  uint32_t* tile_count;  // per-tile instance counters (tile-segmented binning) or NULL
  int use_bulk;  // SH rows staged by cp.async.bulk
};

// Evaluates colour channel sums for DEG from a row of SH coefficients (3 floats per
// coefficient, coefficient-major).  Chain order is fixed: acc = b0*sh0; acc = fma(bk, shk, acc).
template <int DEG>
__device__ __forceinline__ void eval_sh_row(const float* __restrict__ sh, float dx, float dy, float dz,
                                            float* rgb, unsigned& clamp_bits) {
  constexpr int N = (DEG + 1) * (DEG + 1);
  float basis[N];
  sh_basis<DEG>(dx, dy, dz, basis);
  clamp_bits = 0;
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    float acc = basis[0] * sh[ch];
#pragma unroll
    for (int k = 1; k < N; ++k) acc = fmaf(basis[k], sh[3 * k + ch], acc);
    acc += 0.5f;
    if (acc < 0.f) clamp_bits |= 1u << ch;
    rgb[ch] = fmaxf(acc, 0.f);
  }
}

template <int DEG>
__device__ __forceinline__ void load_and_eval_sh(const float* __restrict__ row, bool vec4, float dx, float dy,
                                                 float dz, float* rgb, unsigned& clamp_bits) {
  constexpr int N = (DEG + 1) * (DEG + 1);
  constexpr int NF = 3 * N;
  float sh[(NF + 3) / 4 * 4];
  if (vec4) {
    const float4* r4 = reinterpret_cast<const float4*>(row);
#pragma unroll
    for (int k = 0; k < (NF + 3) / 4; ++k) {
      const float4 t = r4[k];
      sh[4 * k] = t.x; sh[4 * k + 1] = t.y; sh[4 * k + 2] = t.z; sh[4 * k + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < NF; ++k) sh[k] = row[k];
  }
  eval_sh_row<DEG>(sh, dx, dy, dz, rgb, clamp_bits);
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS) preprocess_kernel(const PreArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long sh_bar;
  const int tid = threadIdx.x;
  const int base = blockIdx.x * THREADS;
  const int i = base + tid;
  const int P = a.g.P, M = a.g.M;
  const int nvalid = min(THREADS, P - base);
  const bool want_sh = a.g.colors_precomp == nullptr;
  const uint32_t chunk_bytes = (uint32_t)nvalid * (uint32_t)M * 12u;
  const bool bulk = want_sh && a.use_bulk && (chunk_bytes % 16u == 0u);
  float* sh_smem = reinterpret_cast<float*>(smem_raw);

  if (bulk) {
    if (tid == 0) mbar_init(&sh_bar, 1);
    __syncthreads();
    if (tid == 0) {
      mbar_arrive_expect_tx(&sh_bar, chunk_bytes);
      bulk_copy_g2s(sh_smem, a.g.shs + (size_t)base * M * 3, chunk_bytes, &sh_bar);
    }
  }

  bool alive = i < P;
  float px = 0.f, py = 0.f, pz = 0.f, zview = 0.f;
  const float* __restrict__ view = a.v.view;
  const float* __restrict__ proj = a.v.proj;
  // every input of the Gaussian is requested before the first one is used (one round trip instead of a chain
  // mean -> cull -> scale / rotation -> opacity); a culled Gaussian's loads are wasted, culled Gaussians are few
  float4 q_in = make_float4(1.f, 0.f, 0.f, 0.f);
  float sc_in[3] = {0.f, 0.f, 0.f}, op_in = 0.f;
  if (alive) {
    if (!a.g.cov3D_precomp) {
      q_in = reinterpret_cast<const float4*>(a.g.rotations)[i];
      sc_in[0] = a.g.scales[3 * i]; sc_in[1] = a.g.scales[3 * i + 1]; sc_in[2] = a.g.scales[3 * i + 2];
    }
    op_in = a.g.opacities[i];
  }
  if (alive) {
    px = a.g.means3D[3 * i]; py = a.g.means3D[3 * i + 1]; pz = a.g.means3D[3 * i + 2];
    zview = xform_col(view, 2, px, py, pz);
    alive = zview > GSR_NEAR;  // camera_tools.py:7 (x/y frustum test is disabled there)
  }

  float s6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float cA = 0.f, cB = 0.f, cC = 0.f, opac = 0.f, radius = 0.f, pixx = 0.f, pixy = 0.f;
  int area = 0;
  if (alive) {
    const float hx = xform_col(proj, 0, px, py, pz);
    const float hy = xform_col(proj, 1, px, py, pz);
    const float hw = xform_col(proj, 3, px, py, pz);
    const float pw = 1.0f / (hw + 0.0000001f);  // camera_tools.py:17
    const float ndcx = hx * pw, ndcy = hy * pw;

    if (a.g.cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; ++k) s6[k] = a.g.cov3D_precomp[6 * (size_t)i + k];
    } else {
      cov3d_from_scale_rot(sc_in[0], sc_in[1], sc_in[2], a.v.scale_modifier, q_in, s6);
    }
    JacobianCtx jc;
    build_jacobian(view, px, py, pz, a.v.fx, a.v.fy, a.v.tanfovx, a.v.tanfovy, jc);
    const Cov2D c2 = project_cov(s6, jc);

    const float det0 = fmaf(c2.a, c2.c, -(c2.b * c2.b));
    const float ca = c2.a + 0.3f, cc = c2.c + 0.3f, cb = c2.b;  // render_tools.py:79-80
    const float det = fmaf(ca, cc, -(cb * cb));
    float hscale = 1.f;
    if (a.v.antialiasing) hscale = sqrtf(fmaxf(0.000025f, det0 / det));  // render_tools.py:84
    if (det == 0.f) {
      alive = false;  // render_tools.py:89
    } else {
      const float det_inv = 1.f / det;
      cA = cc * det_inv; cB = -cb * det_inv; cC = ca * det_inv;
      const float mid = 0.5f * (ca + cc);  // render_tools.py:106-115
      const float disc = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
      const float lambda1 = mid + disc, lambda2 = mid - disc;
      radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
      pixx = fmaf(ndcx + 1.0f, (float)a.v.W, -1.0f) * 0.5f;  // camera_tools.py:29-32
      pixy = fmaf(ndcy + 1.0f, (float)a.v.H, -1.0f) * 0.5f;
      int x0, y0, x1, y1;
      tile_rect(pixx, pixy, radius, a.v.gx, a.v.gy, x0, y0, x1, y1);
      area = (x1 - x0) * (y1 - y0);
      if (area == 0) alive = false;
      if (a.tile_count && area) {  // one 64-bit RED per pair of tiles (gsr_common.cuh); they fly under the SH copy
        unsigned long long* cnt = reinterpret_cast<unsigned long long*>(a.tile_count);
        for (int y = y0; y < y1; ++y)
          for (int xp = x0 >> 1; xp <= (x1 - 1) >> 1; ++xp) {
            const unsigned long long inc = (2 * xp >= x0 ? 1ull : 0ull) | (2 * xp + 1 < x1 ? 1ull << 32 : 0ull);
            atomicAdd(cnt + gsr_tile_pair_word(2 * xp, y, a.v.gx), inc);
          }
      }
      opac = op_in * hscale;  // rasterize.py:72
    }
  }

  float rgb[3] = {0.f, 0.f, 0.f};
  unsigned clamp_bits = 0;
  if (bulk) mbar_wait(&sh_bar, 0);  // every thread waits: the CTA must not retire under an in-flight copy
  if (alive) {
    if (!want_sh) {
      rgb[0] = a.g.colors_precomp[3 * (size_t)i];
      rgb[1] = a.g.colors_precomp[3 * (size_t)i + 1];
      rgb[2] = a.g.colors_precomp[3 * (size_t)i + 2];
    } else {
      const float* __restrict__ cam = a.v.campos;
      float dx = px - cam[0], dy = py - cam[1], dz = pz - cam[2];
      const float len = sqrtf(dot3_chain(dx, dx, dy, dy, dz, dz));  // render_tools.py:124-125
      dx = dx / len; dy = dy / len; dz = dz / len;
      const float* row = bulk ? (sh_smem + (size_t)tid * M * 3) : (a.g.shs + (size_t)i * M * 3);
      // 128-bit row reads need 16-byte aligned rows: M*12 % 16 == 0 and an aligned base.
      const bool vec4 = ((M & 3) == 0) && (bulk || ((reinterpret_cast<uintptr_t>(a.g.shs) & 15) == 0));
      // (a row read rounded up to whole float4s must stay inside the row: 4*ceil(3n/4) <= 3M)
      switch (a.v.deg) {
        case 0: load_and_eval_sh<0>(row, vec4 && M >= 4, dx, dy, dz, rgb, clamp_bits); break;
        case 1: load_and_eval_sh<1>(row, vec4, dx, dy, dz, rgb, clamp_bits); break;
        case 2: load_and_eval_sh<2>(row, vec4 && M >= 12, dx, dy, dz, rgb, clamp_bits); break;
        default: load_and_eval_sh<3>(row, vec4, dx, dy, dz, rgb, clamp_bits); break;
      }
    }
  }

  if (i < P) {
    a.radii[i] = alive ? (int32_t)radius : 0;
    a.tiles_touched[i] = alive ? (uint32_t)area : 0u;
    if (alive) {
      // the record is written in the form the compositing kernels stage (gsr_common.cuh): log2-domain quadratic
      // form, negated opacity, the Gaussian's own index, the forward's cull plane
      GsrRec r;
      const float hA = cA * GSR_HALF_LOG2E, hB = cB * GSR_LOG2E, hC = cC * GSR_HALF_LOG2E;
      // tau: a pixel can only receive alpha >= 1/255 from this splat where power >= -tau.
      const float tau = opac > 0.f ? logf(255.f * opac) : -1.f;
      const float tau2 = tau * GSR_LOG2E;
      r.a = make_float4(pixx, pixy, hA, hB);
      r.b = make_float4(hC, -opac, rgb[0], rgb[1]);
      r.c = make_float4(rgb[2], 1.f / zview, tau2, __uint_as_float((uint32_t)i));
      r.d = make_cull_plane(hA, hB, hC, tau2);
      r.d.w = zview;
      a.recs[i] = r;
#pragma unroll
      for (int k = 0; k < 6; ++k) a.cov3D[6 * (size_t)i + k] = s6[k];
      a.clamped[i] = (uint8_t)clamp_bits;
    }
  }
}

cudaError_t launch_preprocess(const GsrView& v, const GsrGaussiansDev& g, int32_t* radii, const GeomState& geom,
                              uint32_t* tile_count, cudaStream_t s) {
  if (g.P <= 0) return cudaSuccess;
  PreArgs a;
  a.v = v; a.g = g; a.radii = radii;
  a.recs = geom.recs; a.cov3D = geom.cov3D; a.clamped = geom.clamped; a.tiles_touched = geom.tiles_touched;
  a.tile_count = tile_count;
  static const int threads = env_threads("GSR_PRE_THREADS", 64);
  const size_t smem = g.shs ? (size_t)threads * g.M * 12 : 0;
  // bulk copies need a 16-byte aligned source; the per-CTA offset threads*M*12 is always a multiple of 16
  a.use_bulk = g.shs && ((reinterpret_cast<uintptr_t>(g.shs) & 15) == 0) && smem <= 200 * 1024;
  void (*kern)(const PreArgs) = threads == 32 ? preprocess_kernel<32> : threads == 64 ? preprocess_kernel<64>
                              : threads == 128 ? preprocess_kernel<128> : preprocess_kernel<256>;
  if (a.use_bulk && smem > 48 * 1024 - 64) {  // per-device attribute: set on every launch that needs it
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
  }
  kern<<<(g.P + threads - 1) / threads, threads, a.use_bulk ? smem : 0, s>>>(a);
  count_launch();
  return cudaGetLastError();
}

// Debug accessor: unpack the geometry records into the canonical arrays of SURVEY.md Appendix C.
__global__ void __launch_bounds__(256) export_geom_kernel(int P, const int32_t* __restrict__ radii, GeomState g,
                                                          float* depths, float* means2D, float* cov3D,
                                                          float* conic_opacity, float* rgb, uint8_t* clamped,
                                                          uint32_t* tiles_touched, uint32_t* point_offsets) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const bool vis = radii[i] > 0;
  GsrRec r;
  r.a = r.b = r.c = r.d = make_float4(0.f, 0.f, 0.f, 0.f);
  if (vis) r = g.recs[i];
  if (depths) depths[i] = r.d.w;
  if (means2D) { means2D[2 * i] = r.a.x; means2D[2 * i + 1] = r.a.y; }
  if (conic_opacity) {  // back from the log2 domain (two roundings: the parity tests compare these at 1e-6 relative)
    conic_opacity[4 * i] = r.a.z / GSR_HALF_LOG2E; conic_opacity[4 * i + 1] = r.a.w / GSR_LOG2E;
    conic_opacity[4 * i + 2] = r.b.x / GSR_HALF_LOG2E; conic_opacity[4 * i + 3] = vis ? -r.b.y : 0.f;
  }
  if (rgb) { rgb[3 * i] = r.b.z; rgb[3 * i + 1] = r.b.w; rgb[3 * i + 2] = r.c.x; }
  if (cov3D)
    for (int k = 0; k < 6; ++k) cov3D[6 * (size_t)i + k] = vis ? g.cov3D[6 * (size_t)i + k] : 0.f;
  if (clamped) {
    const unsigned b = vis ? g.clamped[i] : 0u;
    for (int k = 0; k < 3; ++k) clamped[3 * (size_t)i + k] = (b >> k) & 1u;
  }
  if (tiles_touched) tiles_touched[i] = g.tiles_touched[i];
  if (point_offsets) point_offsets[i] = g.point_offsets[i];
}

cudaError_t launch_export_geom(int P, const int32_t* radii, const GeomState& geom, float* depths, float* means2D,
                               float* cov3D, float* conic_opacity, float* rgb, uint8_t* clamped,
                               uint32_t* tiles_touched, uint32_t* point_offsets, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  export_geom_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, geom, depths, means2D, cov3D, conic_opacity, rgb,
                                                     clamped, tiles_touched, point_offsets);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gsr
