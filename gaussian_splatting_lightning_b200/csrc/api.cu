// api.cu -- the C-ABI of libgsr_b200.so (declared in include/gsr_b200.h).
//
// Host-side orchestration only: validates arguments, carves the three scratch buffers,
// enqueues the stage kernels on the caller's stream and reads back num_rendered.  No torch
// types, no CPU fallback: every stage is a CUDA kernel of this library.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/gsr_b200.h"
#include "gsr_common.cuh"

#ifndef GSR_PRIORITY_LANES_DEFAULT
#define GSR_PRIORITY_LANES_DEFAULT 0
#endif

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};

int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
  g_err = what;
  if (e != cudaSuccess) {
    g_err += ": ";
    g_err += cudaGetErrorString(e);
  }
  return code;
}

#define GSR_CUDA(expr, what)                                   \
  do {                                                         \
    cudaError_t e_ = (expr);                                   \
    if (e_ != cudaSuccess) return fail(GSR_ECUDA, what, e_);   \
  } while (0)

// Optional per-stage timing (bench.py's roofline): CUDA events on the launching stream around
// every stage, folded into per-stage totals by gsr_profile_collect().
enum Stage { ST_PREPROCESS, ST_SCAN, ST_DUPLICATE, ST_SORT, ST_RANGES, ST_RENDER_FWD, ST_RENDER_BWD, ST_PREPROCESS_BWD,
             ST_COUNT };
const char* const kStageNames[ST_COUNT] = {"preprocess", "scan", "duplicate", "sort", "tile_ranges", "render_forward",
                                           "render_backward", "preprocess_backward"};
struct ProfRec { int stage; cudaEvent_t a, b; };
std::mutex g_prof_mu;
bool g_prof_on = false;
std::vector<ProfRec> g_prof_pending;
std::vector<cudaEvent_t> g_prof_free;
double g_prof_ms[ST_COUNT] = {0};
long long g_prof_n[ST_COUNT] = {0};

cudaEvent_t prof_event() {
  if (!g_prof_free.empty()) { cudaEvent_t e = g_prof_free.back(); g_prof_free.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
struct ProfScope {
  bool on; int stage; cudaStream_t s; cudaEvent_t a;
  ProfScope(int stage_, cudaStream_t s_) : on(g_prof_on), stage(stage_), s(s_), a(nullptr) {
    if (on) { std::lock_guard<std::mutex> l(g_prof_mu); a = prof_event(); cudaEventRecord(a, s); }
  }
  ~ProfScope() {
    if (on) {
      std::lock_guard<std::mutex> l(g_prof_mu);
      cudaEvent_t b = prof_event();
      cudaEventRecord(b, s);
      g_prof_pending.push_back({stage, a, b});
    }
  }
};

// With settings->debug every stage is followed by a synchronise + error check that names
// the stage (the CHECK_CUDA behaviour of notes/rasterizer_note.h:44-53).
#define GSR_STAGE(id, expr, what, debug, stream)                              \
  do {                                                                        \
    cudaError_t e_;                                                           \
    {                                                                         \
      ProfScope prof_(id, stream);                                            \
      e_ = (expr);                                                            \
    }                                                                         \
    if (e_ != cudaSuccess) return fail(GSR_ECUDA, what, e_);                  \
    if (debug) {                                                              \
      e_ = cudaStreamSynchronize(stream);                                     \
      if (e_ != cudaSuccess) return fail(GSR_ECUDA, what " (debug sync)", e_); \
    }                                                                         \
  } while (0)

// Pinned mailbox + event for the asynchronous read of num_rendered, per thread and device.
struct Mailbox {
  unsigned int* host = nullptr;
  cudaEvent_t ev = nullptr;
};
constexpr int MAX_DEVICES = 64;
thread_local Mailbox g_mail[MAX_DEVICES];

// Which binning the forward runs: tile-segmented (tile_binning.cu, default) or the global onesweep
// radix sort (GSR_BINNING=onesweep; kept for A/B measurements and as the gsr_sort_pairs primitive).
// Which binning the forward runs: tile-segmented (tile_binning.cu, default) or the global onesweep radix sort (kept for
// A/B measurements and as the gsr_sort_pairs primitive).  Initial value from GSR_BINNING=onesweep; gsr_set_binning() changes
// it at run time (process-wide; forward, backward and export of a frame must see the same mode).
std::atomic<int> g_binning{-1};
bool binning_segmented() {
  int m = g_binning.load(std::memory_order_relaxed);
  if (m < 0) {
    const char* e = getenv("GSR_BINNING");
    m = (e && strcmp(e, "onesweep") == 0) ? GSR_BINNING_ONESWEEP : GSR_BINNING_TILE;
    g_binning.store(m, std::memory_order_relaxed);
  }
  return m == GSR_BINNING_TILE;
}

int get_mailbox(Mailbox** out) {
  int dev = 0;
  GSR_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
  if (dev < 0 || dev >= MAX_DEVICES) return fail(GSR_EINVAL, "device index out of range");
  Mailbox& m = g_mail[dev];
  if (!m.host) {
    GSR_CUDA(cudaMallocHost((void**)&m.host, 64), "cudaMallocHost(mailbox)");
    GSR_CUDA(cudaEventCreateWithFlags(&m.ev, cudaEventDisableTiming), "cudaEventCreate(mailbox)");
  }
  *out = &m;
  return GSR_OK;
}

// Priority lanes.  The stages before the compositing kernel (preprocess, tile scan, scatter, sort) are bound by memory
// latency and L2 atomics, the compositing kernels by instruction issue: side by side on an SM they hardly slow each other
// down.  But kernels of equal priority are dispatched grid after grid -- a binning kernel launched behind another
// view's compositing grid gets its first CTA when that grid has no pending CTA left, so the views' streams overlap at
// the grids' tails only.  With lanes on, gsr_forward runs the stages before compositing on a HIGH-PRIORITY stream of
// its own (one per caller stream and device; fork / join by events), whose CTAs the block scheduler places ahead of
// pending compositing CTAs as SM resources free up.  GSR_PRIORITY_LANES=0/1 sets the initial value, gsr_set_priority_lanes()
// changes it at run time.
struct Lane {
  cudaStream_t hp = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
std::mutex g_lane_mu;
std::map<std::pair<int, cudaStream_t>, Lane> g_lanes;
std::atomic<int> g_lanes_on{-1};
bool lanes_on() {
  int m = g_lanes_on.load(std::memory_order_relaxed);
  if (m < 0) {
    const char* e = getenv("GSR_PRIORITY_LANES");
    m = e ? (atoi(e) != 0) : GSR_PRIORITY_LANES_DEFAULT;
    g_lanes_on.store(m, std::memory_order_relaxed);
  }
  return m != 0;
}
int get_lane(cudaStream_t caller, Lane* out) {
  int dev = 0;
  GSR_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
  std::lock_guard<std::mutex> l(g_lane_mu);
  Lane& ln = g_lanes[std::make_pair(dev, caller)];
  if (!ln.hp) {
    int lo = 0, hi = 0;   // hi = the numerically lowest value = the greatest priority
    GSR_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi), "cudaDeviceGetStreamPriorityRange");
    GSR_CUDA(cudaStreamCreateWithPriority(&ln.hp, cudaStreamNonBlocking, hi), "cudaStreamCreateWithPriority");
    GSR_CUDA(cudaEventCreateWithFlags(&ln.fork, cudaEventDisableTiming), "cudaEventCreate(lane)");
    GSR_CUDA(cudaEventCreateWithFlags(&ln.join, cudaEventDisableTiming), "cudaEventCreate(lane)");
  }
  *out = ln;
  return GSR_OK;
}

int make_view(const GsrSettings* s, int M, GsrView* v) {
  if (!s) return fail(GSR_EINVAL, "settings is NULL");
  if (s->image_width <= 0 || s->image_height <= 0) return fail(GSR_EINVAL, "image size must be positive");
  if (!s->bg || !s->viewmatrix || !s->projmatrix || !s->campos)
    return fail(GSR_EINVAL, "bg / viewmatrix / projmatrix / campos must be device pointers");
  if (s->sh_degree < 0 || s->sh_degree > 3) return fail(GSR_EINVAL, "sh_degree must be in [0,3]");
  v->W = s->image_width; v->H = s->image_height;
  v->gx = (v->W + GSR_TILE - 1) / GSR_TILE; v->gy = (v->H + GSR_TILE - 1) / GSR_TILE;
  v->tanfovx = s->tanfovx; v->tanfovy = s->tanfovy;
  v->fx = (float)v->W / (2.f * s->tanfovx);  // gs_lightning/rasterize/rasterize.py:64-65
  v->fy = (float)v->H / (2.f * s->tanfovy);
  v->scale_modifier = s->scale_modifier;
  v->deg = s->sh_degree; v->M = M;
  v->antialiasing = s->antialiasing ? 1 : 0;
  v->view = s->viewmatrix; v->proj = s->projmatrix; v->campos = s->campos; v->bg = s->bg;
  return GSR_OK;
}

int make_gaussians(const GsrGaussians* g, int deg, GsrGaussiansDev* d) {
  if (!g) return fail(GSR_EINVAL, "gaussians is NULL");
  if (g->P < 0) return fail(GSR_EINVAL, "P must be >= 0");
  if (g->P > 0) {
    if (!g->means3D) return fail(GSR_EINVAL, "means3D must have dimensions (num_points, 3)");
    if (!g->opacities) return fail(GSR_EINVAL, "opacities is NULL");
    if ((g->shs == nullptr) == (g->colors_precomp == nullptr))
      return fail(GSR_EINVAL, "Please provide excatly one of either SHs or precomputed colors!");
    const bool sr = g->scales != nullptr && g->rotations != nullptr;
    if (sr == (g->cov3D_precomp != nullptr) || (g->scales != nullptr) != (g->rotations != nullptr))
      return fail(GSR_EINVAL, "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    if (g->shs && g->M < (deg + 1) * (deg + 1)) return fail(GSR_EINVAL, "shs holds fewer coefficients than sh_degree needs");
    if (g->rotations && (reinterpret_cast<uintptr_t>(g->rotations) & 15))
      return fail(GSR_EINVAL, "rotations must be 16-byte aligned");
  }
  d->P = g->P; d->M = g->shs ? g->M : 0;
  d->means3D = g->means3D; d->opacities = g->opacities; d->shs = g->shs; d->colors_precomp = g->colors_precomp;
  d->scales = g->scales; d->rotations = g->rotations; d->cov3D_precomp = g->cov3D_precomp;
  return GSR_OK;
}

__global__ void set_u32_kernel(unsigned int* p, unsigned int v) { *p = v; }
__global__ void iota_u32_kernel(unsigned int* p, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = (unsigned int)i;
}

}  // namespace

namespace gsr {
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace gsr

extern "C" {

const char* gsr_version(void) { return "gsr_b200 0.1 sm_100a"; }
const char* gsr_last_error(void) { return g_err.c_str(); }
int64_t gsr_launch_count(void) { return (int64_t)g_launches.load(); }
int gsr_set_binning(int32_t mode) {
  if (mode != GSR_BINNING_TILE && mode != GSR_BINNING_ONESWEEP) return fail(GSR_EINVAL, "set_binning: unknown mode");
  g_binning.store(mode, std::memory_order_relaxed);
  return GSR_OK;
}
int gsr_set_priority_lanes(int32_t on) {
  g_lanes_on.store(on ? 1 : 0, std::memory_order_relaxed);
  return GSR_OK;
}
int32_t gsr_get_priority_lanes(void) { return lanes_on() ? 1 : 0; }
int32_t gsr_get_binning(void) { return binning_segmented() ? GSR_BINNING_TILE : GSR_BINNING_ONESWEEP; }

size_t gsr_geom_bytes(int32_t P) { return GeomState::carve(nullptr, P).bytes; }
size_t gsr_img_bytes(int32_t W, int32_t H) { return ImgState::carve(nullptr, W, H).bytes; }
size_t gsr_binning_bytes(int64_t capacity, int32_t W, int32_t H) {
  const int T = ((W + GSR_TILE - 1) / GSR_TILE) * ((H + GSR_TILE - 1) / GSR_TILE);
  return BinningState::carve(nullptr, capacity, T).bytes;
}
size_t gsr_backward_scratch_bytes(int32_t P) { return gsr_align_up((size_t)(P > 0 ? P : 1) * 12 * sizeof(float)); }

int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, uint8_t* present, void* stream) {
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return fail(GSR_EINVAL, "mark_visible: bad arguments");
  GSR_CUDA(gsr::launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream), "mark_visible kernel");
  return GSR_OK;
}

int gsr_forward(const GsrSettings* settings, const GsrGaussians* gaussians, float* out_color, int32_t* radii,
                float* out_invdepth, gsr_alloc_fn alloc, void* alloc_user, int64_t binning_hint,
                int64_t* num_rendered, int64_t* binning_capacity, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GsrView v;
  GsrGaussiansDev g;
  int rc = make_gaussians(gaussians, settings ? settings->sh_degree : 0, &g);
  if (rc) return rc;
  rc = make_view(settings, g.M, &v);
  if (rc) return rc;
  if (!out_color || !out_invdepth || !alloc || (g.P > 0 && !radii)) return fail(GSR_EINVAL, "forward: NULL output or allocator");
  const bool debug = settings->debug != 0;
  const int P = g.P;
  const int num_tiles = v.gx * v.gy;
  const int end_bit = gsr_key_bits(num_tiles);
  const int npass = gsr_sort_passes(end_bit);

  char* geom_ptr = (char*)alloc(alloc_user, GSR_BUF_GEOM, GeomState::carve(nullptr, P).bytes);
  char* img_ptr = (char*)alloc(alloc_user, GSR_BUF_IMG, ImgState::carve(nullptr, v.W, v.H).bytes);
  if (!geom_ptr || !img_ptr) return fail(GSR_EALLOC, "allocator returned NULL");
  GeomState geom = GeomState::carve(geom_ptr, P);
  ImgState img = ImgState::carve(img_ptr, v.W, v.H);

  // GSR_BINNING=onesweep selects the global radix sort (K3-K6); the default is the tile-segmented binning.
  const bool segmented = binning_segmented();
  // the stages before compositing: on the caller's stream, or on its high-priority lane (fork here, join before the
  // compositing kernel; see Lane above)
  Lane lane;
  const bool lanes = lanes_on() && !debug && P > 0;
  cudaStream_t work = stream;
  if (lanes) {
    rc = get_lane(stream, &lane);
    if (rc) return rc;
    work = lane.hp;
    GSR_CUDA(cudaEventRecord(lane.fork, stream), "event record (lane fork)");
    GSR_CUDA(cudaStreamWaitEvent(work, lane.fork, 0), "stream wait (lane fork)");
  }
  // one memset clears ranges + tile counters + control block, one the scan status words
  GSR_CUDA(cudaMemsetAsync(img_ptr + img.ranges_off, 0, img.bytes - img.ranges_off, work), "memset(ranges)");
  if (!segmented) GSR_CUDA(cudaMemsetAsync(geom.scan_state, 0, geom.scan_state_bytes(P), work), "memset(scan state)");

  GSR_STAGE(ST_PREPROCESS, gsr::launch_preprocess(v, g, radii, geom, segmented ? img.tile_count : nullptr, work),
            "preprocess kernel", debug, work);
  if (segmented)
    GSR_STAGE(ST_SCAN, gsr::launch_tile_scan(P > 0 ? v.gx : 0, v.gy, img.tile_count, img.ranges, img.tile_cursor, img.large_tiles,
                                             img.tile_scan_state, img.ctrl, img.tile_order, work),
              "tile scan kernel", debug, work);
  else
    GSR_STAGE(ST_SCAN, gsr::launch_scan(P, geom, img.ctrl, work), "scan kernel", debug, work);
  if (!segmented || P == 0) {  // the tile scan writes the compositing kernels' tile order; without it: raster order
    iota_u32_kernel<<<(num_tiles + 255) / 256, 256, 0, work>>>(img.tile_order, num_tiles);
    gsr::count_launch();
  }

  Mailbox* mail = nullptr;
  rc = get_mailbox(&mail);
  if (rc) return rc;
  GSR_CUDA(cudaMemcpyAsync(mail->host, &img.ctrl->num_rendered, sizeof(unsigned int), cudaMemcpyDeviceToHost, work),
           "memcpy(num_rendered)");
  GSR_CUDA(cudaEventRecord(mail->ev, work), "event record");

  int64_t R = -1;
  int64_t cap = binning_hint;
  if (cap <= 0) {
    GSR_CUDA(cudaEventSynchronize(mail->ev), "read num_rendered");
    R = (int64_t)*mail->host;
    cap = R;
  }
  for (int attempt = 0; attempt < 2; ++attempt) {
    if (cap >= (1ll << 30)) return fail(GSR_ELIMIT, "more than 2^30 tile instances");
    char* bin_ptr = (char*)alloc(alloc_user, GSR_BUF_BINNING, BinningState::carve(nullptr, cap, num_tiles).bytes);
    if (!bin_ptr) return fail(GSR_EALLOC, "allocator returned NULL");
    BinningState bin = BinningState::carve(bin_ptr, cap, num_tiles);
    if (lanes && attempt > 0) {  // the truncated frame's compositing kernel still reads what the stages rewrite
      GSR_CUDA(cudaEventRecord(lane.fork, stream), "event record (lane fork)");
      GSR_CUDA(cudaStreamWaitEvent(work, lane.fork, 0), "stream wait (lane fork)");
    }
    if (cap > 0 && P > 0 && segmented) {
      if (attempt > 0) {  // rebuild the un-truncated ranges, rewind the cursors (scan state and list counters zeroed)
        GSR_CUDA(cudaMemsetAsync(img.tile_scan_state, 0, img.tile_scan_state_bytes, work), "memset(tile scan state)");
        GSR_CUDA(cudaMemsetAsync(img.ctrl, 0, sizeof(GsrControl), work), "memset(control)");
        GSR_STAGE(ST_SCAN, gsr::launch_tile_scan(v.gx, v.gy, img.tile_count, img.ranges, img.tile_cursor, img.large_tiles,
                                                 img.tile_scan_state, img.ctrl, img.tile_order, work),
                  "tile scan kernel", debug, work);
      }
      GSR_STAGE(ST_DUPLICATE, gsr::launch_tile_scatter(v, P, radii, geom, img.tile_cursor, bin.keys_a,
                                                       (uint32_t)cap, img.ctrl, bin.header, work),
                "tile scatter kernel", debug, work);
      GSR_STAGE(ST_SORT, gsr::launch_tile_sort(num_tiles, img.ranges, img.large_tiles, img.ctrl, bin.keys_a, bin.vals_a, (uint32_t)cap,
                                                    work),
                "tile sort kernel", debug, work);
    } else if (cap > 0 && P > 0) {
      GSR_CUDA(cudaMemsetAsync(bin.sort_ws_base, 0, bin.sort_ws.bytes, work), "memset(sort workspace)");
      // start in the buffer that makes the last pass land in (keys_a, vals_a)
      unsigned long long* k0 = (npass % 2 == 0) ? bin.keys_a : bin.keys_b;
      uint32_t* v0 = (npass % 2 == 0) ? bin.vals_a : bin.vals_b;
      unsigned long long* k1 = (npass % 2 == 0) ? bin.keys_b : bin.keys_a;
      uint32_t* v1 = (npass % 2 == 0) ? bin.vals_b : bin.vals_a;
      GSR_STAGE(ST_DUPLICATE, gsr::launch_duplicate(v, P, radii, geom, k0, v0, (uint32_t)cap, img.ctrl, bin.header, bin.sort_ws.hist, npass, work), "duplicate kernel",
                debug, work);
      GSR_STAGE(ST_SORT, gsr::launch_onesweep(k0, v0, k1, v1, &img.ctrl->num_rendered, (uint32_t)cap, end_bit, bin.sort_ws,
                                     true, work),
                "onesweep radix sort", debug, work);
      GSR_STAGE(ST_RANGES, gsr::launch_tile_ranges(bin.keys_a, &img.ctrl->num_rendered, (uint32_t)cap, img.ranges, work),
                "tile ranges kernel", debug, work);
    }
    if (lanes) {
      GSR_CUDA(cudaEventRecord(lane.join, work), "event record (lane join)");
      GSR_CUDA(cudaStreamWaitEvent(stream, lane.join, 0), "stream wait (lane join)");
    }
    GSR_STAGE(ST_RENDER_FWD, gsr::launch_render_forward(v, img.ranges, img.tile_order, bin.vals_a, geom.recs, out_color,
                                                        out_invdepth, img.final_T, img.n_contrib, bin.hits, stream),
              "render forward kernel", debug, stream);
    if (R < 0) {
      GSR_CUDA(cudaEventSynchronize(mail->ev), "read num_rendered");
      R = (int64_t)*mail->host;
    }
    if (R <= cap) break;
    // the hint was too small: the frame above rendered a truncated list. Redo binning + compositing.
    cap = R + R / 8 + 1024;
    if (!segmented)
      GSR_CUDA(cudaMemsetAsync(img.ranges, 0, (size_t)num_tiles * sizeof(uint2), stream), "memset(ranges)");  // before the next fork
  }
  if (R >= 0xFFFFFFFFll) return fail(GSR_ELIMIT, "more than 2^32 tile instances");
  if (num_rendered) *num_rendered = R;
  if (binning_capacity) *binning_capacity = cap;
  return GSR_OK;
}

int gsr_backward(const GsrSettings* settings, const GsrGaussians* gaussians, const int32_t* radii,
                 const float* dL_dcolor, const float* dL_dinvdepth, const void* geom_, const void* binning_,
                 const void* img_, int64_t num_rendered, void* scratch, const GsrGrads* grads, int32_t flags,
                 const GsrDensifyStats* stats, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GsrView v;
  GsrGaussiansDev g;
  int rc = make_gaussians(gaussians, settings ? settings->sh_degree : 0, &g);
  if (rc) return rc;
  rc = make_view(settings, g.M, &v);
  if (rc) return rc;
  if (flags & ~(GSR_GRADS_ACCUMULATE | GSR_BWD_RENDER_ONLY)) return fail(GSR_EINVAL, "backward: unknown flag");
  const bool render_only = (flags & GSR_BWD_RENDER_ONLY) != 0;
  static const GsrGrads no_grads = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  if (render_only && !grads) grads = &no_grads;
  if (!grads) return fail(GSR_EINVAL, "backward: grads is NULL");
  const int P = g.P;
  if (P == 0) return GSR_OK;
  if (!radii || !dL_dcolor || !geom_ || !binning_ || !img_ || !scratch)
    return fail(GSR_EINVAL, "backward: NULL buffer");
  if (grads->dL_drotations && (reinterpret_cast<uintptr_t>(grads->dL_drotations) & 15))
    return fail(GSR_EINVAL, "dL_drotations must be 16-byte aligned");
  const bool debug = settings->debug != 0;
  GeomState geom = GeomState::carve((char*)geom_, P);
  ImgState img = ImgState::carve((char*)img_, v.W, v.H);
  float* grad_acc = (float*)scratch;
  gsr::PreBwdOut out;
  out.dL_dmeans2D = grads->dL_dmeans2D; out.dL_dcolors = grads->dL_dcolors; out.dL_dopacity = grads->dL_dopacity;
  out.dL_dmeans3D = grads->dL_dmeans3D; out.dL_dcov3D = grads->dL_dcov3D; out.dL_dsh = grads->dL_dsh;
  out.dL_dscales = grads->dL_dscales; out.dL_drot = grads->dL_drotations;
  out.accumulate = (flags & GSR_GRADS_ACCUMULATE) ? 1 : 0;
  out.max_radii2D = stats ? stats->max_radii2D : nullptr;
  out.xyz_grad_accum = stats ? stats->xyz_grad_accum : nullptr;
  out.xyz_grad_count = stats ? stats->xyz_grad_count : nullptr;

  GSR_CUDA(cudaMemsetAsync(grad_acc, 0, (size_t)P * 12 * sizeof(float), stream), "memset(grad accumulators)");
  if (num_rendered > 0)
    GSR_STAGE(ST_RENDER_BWD, gsr::launch_render_backward(v, img.ranges, img.tile_order, binning_, geom.recs, img.final_T, img.n_contrib, dL_dcolor,
                                          dL_dinvdepth, grad_acc, stream),
              "render backward kernel", debug, stream);
  if (render_only) return GSR_OK;   // the raw sums stay in `scratch` for gsr_preprocess_backward_views
  GSR_STAGE(ST_PREPROCESS_BWD, gsr::launch_preprocess_backward(v, g, radii, geom, grad_acc, out, stream),
            "preprocess backward kernel", debug, stream);
  return GSR_OK;
}

int gsr_preprocess_backward_views(const GsrViewState* views, int32_t num_views, const GsrGaussians* gaussians,
                                  const GsrGrads* grads, int32_t flags, const GsrDensifyStats* stats, void* stream_) {
  return gsr_preprocess_backward_views_range(views, num_views, gaussians, grads, flags, stats, 0,
                                             gaussians ? gaussians->P : 0, stream_);
}

int gsr_preprocess_backward_views_range(const GsrViewState* views, int32_t num_views, const GsrGaussians* gaussians,
                                        const GsrGrads* grads, int32_t flags, const GsrDensifyStats* stats,
                                        int32_t first, int32_t count, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!views || num_views < 0 || !grads) return fail(GSR_EINVAL, "preprocess_backward_views: bad arguments");
  if (flags & ~GSR_GRADS_ACCUMULATE) return fail(GSR_EINVAL, "preprocess_backward_views: unknown flag");
  if (grads->dL_dmeans2D) return fail(GSR_EINVAL, "preprocess_backward_views: dL_dmeans2D is a per-view output (pass NULL)");
  if (num_views == 0) return GSR_OK;
  for (int k = 0; k < num_views; ++k)
    if (!views[k].settings) return fail(GSR_EINVAL, "preprocess_backward_views: settings is NULL");
  GsrGaussiansDev g;
  int rc = make_gaussians(gaussians, views[0].settings->sh_degree, &g);
  if (rc) return rc;
  const int P = g.P;
  if (first < 0 || count < 0 || (long long)first + count > P) return fail(GSR_EINVAL, "preprocess_backward_views: range outside [0, P)");
  if (P == 0 || count == 0) return GSR_OK;
  if (grads->dL_drotations && (reinterpret_cast<uintptr_t>(grads->dL_drotations) & 15))
    return fail(GSR_EINVAL, "dL_drotations must be 16-byte aligned");
  std::vector<GsrView> vv(num_views);
  std::vector<const int32_t*> radii(num_views);
  std::vector<GeomState> geoms(num_views);
  std::vector<const float*> accs(num_views);
  for (int k = 0; k < num_views; ++k) {
    const GsrViewState& s = views[k];
    rc = make_view(s.settings, g.M, &vv[k]);
    if (rc) return rc;
    if (!s.radii || !s.geom || !s.scratch) return fail(GSR_EINVAL, "preprocess_backward_views: NULL buffer in a view");
    if (s.settings->sh_degree != views[0].settings->sh_degree || s.settings->scale_modifier != views[0].settings->scale_modifier ||
        (s.settings->antialiasing != 0) != (views[0].settings->antialiasing != 0))
      return fail(GSR_EINVAL, "preprocess_backward_views: the views must share sh_degree, scale_modifier and antialiasing");
    radii[k] = s.radii + first;
    geoms[k] = GeomState::carve((char*)s.geom, P);
    geoms[k].clamped += first;     // the geometry arrays the backward reads per Gaussian
    geoms[k].cov3D += (size_t)first * 6;
    accs[k] = (const float*)s.scratch + (size_t)first * 12;
  }
  // the Gaussians of the range: every input pointer moves to the range's first row
  g.P = count;
  g.means3D += (size_t)first * 3; g.opacities += first;
  if (g.shs) g.shs += (size_t)first * g.M * 3;
  if (g.colors_precomp) g.colors_precomp += (size_t)first * 3;
  if (g.scales) g.scales += (size_t)first * 3;
  if (g.rotations) g.rotations += (size_t)first * 4;
  if (g.cov3D_precomp) g.cov3D_precomp += (size_t)first * 6;
  gsr::PreBwdOut out;
  out.dL_dmeans2D = nullptr; out.dL_dcolors = grads->dL_dcolors; out.dL_dopacity = grads->dL_dopacity;
  out.dL_dmeans3D = grads->dL_dmeans3D; out.dL_dcov3D = grads->dL_dcov3D; out.dL_dsh = grads->dL_dsh;
  out.dL_dscales = grads->dL_dscales; out.dL_drot = grads->dL_drotations;
  out.accumulate = (flags & GSR_GRADS_ACCUMULATE) ? 1 : 0;
  out.max_radii2D = stats ? stats->max_radii2D : nullptr;
  out.xyz_grad_accum = stats ? stats->xyz_grad_accum : nullptr;
  out.xyz_grad_count = stats ? stats->xyz_grad_count : nullptr;
  const bool debug = views[0].settings->debug != 0;
  // (a single view: the per-view kernel -- one thread per Gaussian, SH block by bulk copy -- is the better shape)
  if (num_views > 1 && gsr::preprocess_backward_views_supported(g, out)) {
    GSR_STAGE(ST_PREPROCESS_BWD, gsr::launch_preprocess_backward_views(num_views, vv.data(), radii.data(), geoms.data(),
                                                                       accs.data(), g, out, stream),
              "preprocess backward kernel (all views)", debug, stream);
  } else {  // layouts the fused kernel does not take (SH rows that are not 16-byte multiples, unaligned tensors)
    for (int k = 0; k < num_views; ++k) {
      GSR_STAGE(ST_PREPROCESS_BWD, gsr::launch_preprocess_backward(vv[k], g, radii[k], geoms[k], accs[k], out, stream),
                "preprocess backward kernel", debug, stream);
      out.accumulate = 1;
    }
  }
  return GSR_OK;
}

int gsr_export_state(const GsrSettings* settings, int32_t P, const int32_t* radii, const void* geom_,
                     const void* binning_, const void* img_, int64_t num_rendered, int64_t binning_capacity,
                     const GsrStateExport* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GsrView v;
  int rc = make_view(settings, 0, &v);
  if (rc) return rc;
  if (!out || !geom_ || !img_ || (P > 0 && !radii)) return fail(GSR_EINVAL, "export: NULL buffer");
  GeomState geom = GeomState::carve((char*)geom_, P);
  ImgState img = ImgState::carve((char*)img_, v.W, v.H);
  const int64_t cap = binning_capacity > 0 ? binning_capacity : num_rendered;
  const size_t R = (size_t)num_rendered;
  if (binning_segmented() && P > 0) {  // the segmented forward has no per-Gaussian offsets: make them for the export
    GSR_CUDA(cudaMemsetAsync(geom.scan_state, 0, geom.scan_state_bytes(P), stream), "memset(scan state)");
    GSR_CUDA(gsr::launch_scan(P, geom, img.ctrl, stream), "export scan kernel");
  }
  GSR_CUDA(gsr::launch_export_geom(P, radii, geom, out->depths, out->means2D, out->cov3D, out->conic_opacity, out->rgb,
                                   out->clamped, out->tiles_touched, out->point_offsets, stream),
           "export kernel");
  if (R > 0 && binning_) {
    BinningState bin = BinningState::carve((char*)binning_, cap, v.gx * v.gy);
    if (out->keys_sorted) {
      if (binning_segmented())  // the segments hold composites: rebuild (tile << 32 | depth) from the sorted list
        GSR_CUDA(gsr::launch_export_segment_keys(v.gx * v.gy, img.ranges, bin.vals_a, geom, (uint32_t)R,
                                                 (unsigned long long*)out->keys_sorted, stream), "export keys");
      else
        GSR_CUDA(cudaMemcpyAsync(out->keys_sorted, bin.keys_a, R * 8, cudaMemcpyDeviceToDevice, stream), "export keys");
    }
    if (out->vals_sorted)
      GSR_CUDA(cudaMemcpyAsync(out->vals_sorted, bin.vals_a, R * 4, cudaMemcpyDeviceToDevice, stream), "export vals");
    if (out->keys_unsorted && out->vals_unsorted)
      GSR_CUDA(gsr::launch_duplicate(v, P, radii, geom, (unsigned long long*)out->keys_unsorted, out->vals_unsorted,
                                     (uint32_t)R, img.ctrl, nullptr, nullptr, 0, stream),
               "export duplicate kernel");
  }
  const size_t HW = (size_t)v.W * v.H;
  if (out->ranges)
    GSR_CUDA(cudaMemcpyAsync(out->ranges, img.ranges, (size_t)v.gx * v.gy * 8, cudaMemcpyDeviceToDevice, stream),
             "export ranges");
  if (out->final_T)
    GSR_CUDA(cudaMemcpyAsync(out->final_T, img.final_T, HW * 4, cudaMemcpyDeviceToDevice, stream), "export final_T");
  if (out->n_contrib)
    GSR_CUDA(cudaMemcpyAsync(out->n_contrib, img.n_contrib, HW * 4, cudaMemcpyDeviceToDevice, stream),
             "export n_contrib");
  return GSR_OK;
}

size_t gsr_l1_loss_scratch_bytes(void) { return gsr::l1_scratch_bytes(); }

int gsr_l1_loss(const float* image, const float* target, int64_t n, float* dL_dimage, float* out_loss, void* scratch,
                void* stream) {
  if (n <= 0 || !image || !target || !out_loss || !scratch) return fail(GSR_EINVAL, "l1_loss: bad arguments");
  if ((((uintptr_t)image) | ((uintptr_t)target) | ((uintptr_t)dL_dimage) | ((uintptr_t)scratch)) & 15)
    return fail(GSR_EINVAL, "l1_loss: image, target, dL_dimage and scratch must be 16-byte aligned");
  GSR_CUDA(gsr::launch_l1_loss(image, target, (long long)n, dL_dimage, out_loss, scratch, (cudaStream_t)stream),
           "l1_loss kernel");
  return GSR_OK;
}

int gsr_l1_loss_noalloc(const float* image, const float* target, int64_t n, float* dL_dimage, float* out_loss,
                        void* stream) {
  if (n <= 0 || !image || !target || !out_loss) return fail(GSR_EINVAL, "l1_loss_noalloc: bad arguments");
  if (n > (int64_t)GSR_L1_NOALLOC_MAX_ELEMENTS)
    return fail(GSR_ELIMIT, "l1_loss_noalloc: n exceeds GSR_L1_NOALLOC_MAX_ELEMENTS (2^28); use gsr_l1_loss with caller scratch");
  if ((((uintptr_t)image) | ((uintptr_t)target) | ((uintptr_t)dL_dimage)) & 15)
    return fail(GSR_EINVAL, "l1_loss_noalloc: image, target and dL_dimage must be 16-byte aligned");
  GSR_CUDA(gsr::launch_l1_loss_noalloc(image, target, (long long)n, dL_dimage, out_loss, (cudaStream_t)stream),
           "l1_loss_noalloc (cudaGetSymbolAddress / memset / kernel)");
  return GSR_OK;
}

size_t gsr_fused_ssim_scratch_bytes(int32_t planes, int32_t H, int32_t W) { return gsr::ssim_scratch_bytes(planes, H, W); }

int gsr_fused_ssim(const float* img1, const float* img2, int32_t planes, int32_t H, int32_t W, float C1, float C2,
                   int32_t padding_valid, float* dL_dimg1, float* out_ssim, void* scratch, void* stream) {
  if (planes <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !out_ssim || !scratch)
    return fail(GSR_EINVAL, "fused_ssim: bad arguments");
  if (planes > 65535) return fail(GSR_ELIMIT, "fused_ssim: more than 65535 image planes");
  if (padding_valid && (H <= 10 || W <= 10)) return fail(GSR_EINVAL, "fused_ssim: padding 'valid' needs H, W > 10");
  if (((uintptr_t)scratch) & 15) return fail(GSR_EINVAL, "fused_ssim: scratch must be 16-byte aligned");
  GSR_CUDA(gsr::launch_fused_ssim(img1, img2, planes, H, W, C1, C2, padding_valid ? 1 : 0, 0.f, 1.f, 0.f, dL_dimg1, nullptr,
                                  out_ssim, nullptr, scratch, (cudaStream_t)stream), "fused_ssim kernel");
  return GSR_OK;
}

int gsr_photometric_loss(const float* image, const float* target, int32_t planes, int32_t H, int32_t W, float w_ssim,
                         float* dL_dimage, float* out_loss, float* out_l1, float* out_ssim, void* scratch,
                         void* stream) {
  if (planes <= 0 || H <= 0 || W <= 0 || !image || !target || !out_loss || !scratch)
    return fail(GSR_EINVAL, "photometric_loss: bad arguments");
  if (planes > 65535) return fail(GSR_ELIMIT, "photometric_loss: more than 65535 image planes");
  if (((uintptr_t)scratch) & 15) return fail(GSR_EINVAL, "photometric_loss: scratch must be 16-byte aligned");
  // (1 - w) L1 + w (1 - ssim) = (1 - w) L1 - w ssim + w
  GSR_CUDA(gsr::launch_fused_ssim(image, target, planes, H, W, 0.01f * 0.01f, 0.03f * 0.03f, 0, 1.f - w_ssim, -w_ssim,
                                  w_ssim, dL_dimage, out_loss, out_ssim, out_l1, scratch, (cudaStream_t)stream),
           "photometric_loss kernel");
  return GSR_OK;
}

int gsr_adam_step(const GsrAdamGroup* groups, int32_t num_groups, int32_t step, double beta1, double beta2, double eps,
                  int32_t flags, void* stream) {
  if (!groups || num_groups < 0 || num_groups > GSR_ADAM_MAX_GROUPS || step < 1)
    return fail(GSR_EINVAL, "adam_step: bad arguments (1..8 groups, step >= 1)");
  if (!(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0)) return fail(GSR_EINVAL, "adam_step: betas outside [0,1)");
  for (int i = 0; i < num_groups; ++i) {
    const GsrAdamGroup& g = groups[i];
    if (g.n < 0 || (g.n > 0 && (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq)))
      return fail(GSR_EINVAL, "adam_step: NULL pointer in a non-empty group");
  }
  GSR_CUDA(gsr::launch_adam(groups, num_groups, step, beta1, beta2, eps, (flags & GSR_ADAM_ZERO_GRAD) ? 1 : 0,
                            (cudaStream_t)stream), "adam kernel");
  return GSR_OK;
}

int gsr_densify_fold(int32_t P, const float* d_max_radii, const float* d_accum, const float* d_count,
                     float* max_radii2D, float* xyz_grad_accum, float* xyz_grad_count, void* stream) {
  if (P < 0 || (P > 0 && (!d_count || (max_radii2D && !d_max_radii) || (xyz_grad_accum && !d_accum))))
    return fail(GSR_EINVAL, "densify_fold: bad arguments");
  GSR_CUDA(gsr::launch_densify_fold(P, d_max_radii, d_accum, d_count, max_radii2D, xyz_grad_accum, xyz_grad_count,
                                    (cudaStream_t)stream), "densify_fold kernel");
  return GSR_OK;
}

size_t gsr_sort_workspace_bytes(int64_t n) {
  return gsr_sort::Workspace::carve(nullptr, n).bytes + 128;  // + the device-side element count
}

int gsr_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                   uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n, int32_t end_bit, void* workspace,
                   void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n < 0 || n >= (1ll << 30)) return fail(GSR_ELIMIT, "sort: n must be in [0, 2^30)");
  if (end_bit < 1 || end_bit > 64) return fail(GSR_EINVAL, "sort: end_bit must be in [1,64]");
  if (n == 0) return GSR_OK;
  if (!keys_in || !vals_in || !keys_out || !vals_out || !keys_tmp || !vals_tmp || !workspace)
    return fail(GSR_EINVAL, "sort: NULL buffer");
  const size_t ws_bytes = gsr_sort::Workspace::carve(nullptr, n).bytes;
  GSR_CUDA(cudaMemsetAsync(workspace, 0, ws_bytes + 128, stream), "memset(sort workspace)");
  gsr_sort::Workspace ws = gsr_sort::Workspace::carve((char*)workspace, n);
  unsigned int* d_n = (unsigned int*)((char*)workspace + ws_bytes);
  set_u32_kernel<<<1, 1, 0, stream>>>(d_n, (unsigned int)n);
  gsr::count_launch();
  const int npass = gsr_sort_passes(end_bit);
  unsigned long long* ka = (unsigned long long*)((npass % 2 == 0) ? keys_out : keys_tmp);
  uint32_t* va = (npass % 2 == 0) ? vals_out : vals_tmp;
  unsigned long long* kb = (unsigned long long*)((npass % 2 == 0) ? keys_tmp : keys_out);
  uint32_t* vb = (npass % 2 == 0) ? vals_tmp : vals_out;
  GSR_CUDA(cudaMemcpyAsync(ka, keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, stream), "sort: copy keys");
  GSR_CUDA(cudaMemcpyAsync(va, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream), "sort: copy vals");
  GSR_CUDA(gsr::launch_onesweep(ka, va, kb, vb, d_n, (uint32_t)n, end_bit, ws, false, stream), "onesweep radix sort");
  return GSR_OK;
}

int gsr_profile_enable(int on) {
  std::lock_guard<std::mutex> l(g_prof_mu);
  g_prof_on = on != 0;
  return GSR_OK;
}

void gsr_profile_reset(void) {
  std::lock_guard<std::mutex> l(g_prof_mu);
  for (int i = 0; i < ST_COUNT; ++i) { g_prof_ms[i] = 0.0; g_prof_n[i] = 0; }
}

int gsr_profile_collect(void) {
  std::lock_guard<std::mutex> l(g_prof_mu);
  for (const ProfRec& r : g_prof_pending) {
    GSR_CUDA(cudaEventSynchronize(r.b), "profile: event synchronize");
    float ms = 0.f;
    GSR_CUDA(cudaEventElapsedTime(&ms, r.a, r.b), "profile: elapsed time");
    g_prof_ms[r.stage] += ms;
    g_prof_n[r.stage] += 1;
    g_prof_free.push_back(r.a);
    g_prof_free.push_back(r.b);
  }
  g_prof_pending.clear();
  return GSR_OK;
}

int gsr_profile_stage_count(void) { return ST_COUNT; }
const char* gsr_profile_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }
double gsr_profile_stage_ms(int i) { return (i >= 0 && i < ST_COUNT) ? g_prof_ms[i] : 0.0; }
int64_t gsr_profile_stage_calls(int i) { return (i >= 0 && i < ST_COUNT) ? (int64_t)g_prof_n[i] : 0; }

}  // extern "C"
