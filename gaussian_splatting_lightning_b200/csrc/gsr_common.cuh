// gsr_common.cuh -- shared declarations of the B200 rasterizer kernels.
//
// HBM layout of the three scratch buffers (opaque uint8 tensors to the caller; the
// forward writes them, the backward only reads them):
//
//   geom (per Gaussian, P)     GsrRec recs[P]           64 B  packed record the tile kernels gather
//                              float  cov3D[6P]         24 B  kept for the backward
//                              u8     clamped[P]         1 B  bit c set: SH colour channel c was clamped at 0
//                              u32    tiles_touched[P]   4 B
//                              u32    point_offsets[P]   4 B  inclusive scan of tiles_touched
//                              u64    scan_state[...]         decoupled look-back words of the scan
//   binning (per instance, R)  128-byte header {u32 capacity, u32 num_tiles} (written on the device; the backward
//                              locates everything else from it), u32 vals_a[cap] (sorted point list ends here),
//                              u32 hits[(cap/32 + T + 2) * 8] (per-warp hit words of the forward compositing),
//                              vals_b[cap], u64 keys_a[cap], keys_b[cap],
//                              sort control (digit histograms, tile counters, look-back words)
//   img                        float final_T[HW], u32 n_contrib[HW], uint2 ranges[T], GsrControl
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define GSR_TILE 16
#define GSR_TILE_PIX 256
#define GSR_NEAR 0.2f

// One packed per-Gaussian record (AoS, 64 B = two 32-byte sectors, 64-byte aligned): everything the forward and
// backward compositing kernels need, ALREADY in the form they keep in shared memory, so that staging a batch is a pure
// copy (four 16-byte cp.async per entry, issued a whole batch ahead of its use) with no arithmetic.  The quadratic form
// is pre-scaled into the log2 domain (render_common.cuh):
//   hA = conic.A * log2(e)/2, hB = conic.B * log2(e), hC = conic.C * log2(e)/2, tau2 = ln(255 * opacity) * log2(e)
struct __align__(64) GsrRec {
  float4 a;  // {mean2D.x, mean2D.y, hA, hB}
  float4 b;  // {hC, -opacity (AA-scaled), r, g}
  float4 c;  // {b, 1/z_view, tau2, bits(Gaussian index)}
  float4 d;  // {kx, ky, tauM, z_view}: the cull plane of the forward (make_cull_plane) + the depth the sort keys use
};
static_assert(sizeof(GsrRec) == 64, "GsrRec is two sectors");

// Per-frame device-side control block (zeroed by one memset at the start of a frame).
struct GsrControl {
  unsigned int num_rendered;   // R = sum(tiles_touched), written by the scan
  unsigned int num_mid;        // tiles with 513..2048 instances (tile_binning.cu)
  unsigned int overflow;       // set when R exceeded the binning capacity
  unsigned int num_large;      // tiles with more than 2048 instances
};

// Kernel-side view of one camera + flags.
struct GsrView {
  int W, H, gx, gy;
  float tanfovx, tanfovy, fx, fy;
  float scale_modifier;
  int deg, M;
  int antialiasing;
  const float* view;
  const float* proj;
  const float* campos;
  const float* bg;
};

static inline size_t gsr_align_up(size_t x, size_t a = 128) { return (x + a - 1) / a * a; }

// ---- geometry buffer carving -------------------------------------------------------
struct GeomState {
  GsrRec* recs;
  float* cov3D;
  uint8_t* clamped;
  uint32_t* tiles_touched;
  uint32_t* point_offsets;
  unsigned long long* scan_state;
  size_t bytes;
  static constexpr int SCAN_ITEMS = 2048;  // items per scan tile (256 threads x 8)
  static GeomState carve(char* base, int P) {
    GeomState g;
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off = gsr_align_up(off + n); return base ? base + o : (char*)nullptr; };
    const size_t Pp = (size_t)(P > 0 ? P : 1);
    g.recs = (GsrRec*)take(Pp * sizeof(GsrRec));
    g.cov3D = (float*)take(Pp * 6 * sizeof(float));
    g.clamped = (uint8_t*)take(Pp);
    g.tiles_touched = (uint32_t*)take(Pp * 4);
    g.point_offsets = (uint32_t*)take(Pp * 4);
    const size_t ntiles = (Pp + SCAN_ITEMS - 1) / SCAN_ITEMS;
    g.scan_state = (unsigned long long*)take(ntiles * 8);
    g.bytes = off;
    return g;
  }
  size_t scan_state_bytes(int P) const { return ((size_t)(P > 0 ? P : 1) + SCAN_ITEMS - 1) / SCAN_ITEMS * 8; }
};

// ---- image buffer carving ----------------------------------------------------------
// Per-tile instance counters.  The tiles (2k, y) and (2k + 1, y) share ONE 64-bit word (low / high half): a splat
// that covers both is counted -- and later given its two slots -- by a single 64-bit atomic.  One word per 32-byte
// sector: the L2 atomic unit serialises the updates of a sector, and neighbouring tiles are updated together.
constexpr int GSR_TILE_COUNT_STRIDE = 8;  // 32-bit words per TILE of storage (a pair uses one sector of its two)
__host__ __device__ inline size_t gsr_tile_pair_word(int x, int y, int gx) {  // index of the pair's 64-bit word
  return ((size_t)y * (size_t)((gx + 1) >> 1) + (size_t)(x >> 1)) * 4;
}

struct ImgState {
  float* final_T;
  uint32_t* n_contrib;
  uint2* ranges;
  uint32_t* tile_count;   // instances per tile, counted by the preprocess kernel (tile_binning.cu)
  uint32_t* tile_cursor;  // the scatter's per-tile cursors (same layout), set to the segment starts by the tile scan
  uint32_t* large_tiles;  // lists of the tiles the MID (from the front) and LARGE (from the back) sort launches take
  uint32_t* tile_order;   // the tiles by descending list length: blockIdx -> tile of the compositing kernels
  unsigned long long* tile_scan_state;  // status words of the chained tile scan, one per 1024 tiles
  size_t tile_scan_state_bytes;
  GsrControl* ctrl;
  size_t ranges_off, bytes;
  static ImgState carve(char* base, int W, int H) {
    ImgState s;
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off = gsr_align_up(off + n); return base ? base + o : (char*)nullptr; };
    const size_t HW = (size_t)W * H;
    const size_t T = (size_t)((W + GSR_TILE - 1) / GSR_TILE) * ((H + GSR_TILE - 1) / GSR_TILE);
    s.final_T = (float*)take(HW * 4);
    s.n_contrib = (uint32_t*)take(HW * 4);
    s.ranges_off = off;
    s.ranges = (uint2*)take(T * 8);
    s.tile_count = (uint32_t*)take(T * 4 * GSR_TILE_COUNT_STRIDE);
    s.tile_cursor = (uint32_t*)take(T * 4 * GSR_TILE_COUNT_STRIDE);
    s.large_tiles = (uint32_t*)take(T * 4);
    s.tile_order = (uint32_t*)take(T * 4);
    s.tile_scan_state_bytes = ((T + 1023) / 1024 + 1) * 8;
    s.tile_scan_state = (unsigned long long*)take(s.tile_scan_state_bytes);
    s.ctrl = (GsrControl*)take(sizeof(GsrControl));
    s.bytes = off;
    return s;
  }
};

// ---- onesweep sort configuration ------------------------------------------------------
namespace gsr_sort {
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int THREADS = 512;
constexpr int WARPS = THREADS / 32;
constexpr int IPT = 8;                        // items per thread
constexpr int TILE_ITEMS = THREADS * IPT;     // 4096 pairs per tile
constexpr int MAX_PASSES = 8;
constexpr int LB_GROUP = 32;                  // tiles per look-back group (radix_sort.cu)

struct Workspace {
  uint32_t* hist;          // [MAX_PASSES][RADIX] digit histograms of all passes
  uint32_t* tile_counter;  // [MAX_PASSES] dynamic tile ids
  uint32_t* lookback;      // [MAX_PASSES][ntiles][RADIX]
  uint32_t* lookback_group;  // [MAX_PASSES][ngroups][RADIX]
  size_t ntiles, ngroups;
  size_t bytes;
  static Workspace carve(char* base, int64_t capacity) {
    Workspace w;
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off = gsr_align_up(off + n); return base ? base + o : (char*)nullptr; };
    w.ntiles = (size_t)((capacity + TILE_ITEMS - 1) / TILE_ITEMS);
    if (w.ntiles == 0) w.ntiles = 1;
    w.hist = (uint32_t*)take((size_t)MAX_PASSES * RADIX * 4);
    w.tile_counter = (uint32_t*)take((size_t)MAX_PASSES * 4);
    w.lookback = (uint32_t*)take((size_t)MAX_PASSES * w.ntiles * RADIX * 4);
    w.ngroups = (w.ntiles + LB_GROUP - 1) / LB_GROUP;
    w.lookback_group = (uint32_t*)take((size_t)MAX_PASSES * w.ngroups * RADIX * 4);
    w.bytes = off;
    return w;
  }
};
}  // namespace gsr_sort

// ---- binning buffer carving ----------------------------------------------------------
struct BinningHeader {
  uint32_t capacity;   // instance capacity the buffer was laid out for
  uint32_t num_tiles;
  uint32_t segmented;  // 1: laid out by the tile-segmented binning (tile_binning.cu): keys_a holds composites
  uint32_t pad[29];
};
static_assert(sizeof(BinningHeader) == 128, "binning header is one 128-byte line");

// Offsets of the two arrays the backward needs; a pure function of (capacity, num_tiles) so that
// device code can locate them from the header alone.
__host__ __device__ inline size_t gsr_bin_align(size_t x) { return (x + 127) / 128 * 128; }
__host__ __device__ inline size_t gsr_bin_vals_off() { return sizeof(BinningHeader); }
__host__ __device__ inline size_t gsr_bin_hit_words(size_t cap, size_t num_tiles) { return (cap / 32 + num_tiles + 2) * 8; }
__host__ __device__ inline size_t gsr_bin_hits_off(size_t cap) { return gsr_bin_vals_off() + gsr_bin_align(cap * 4); }

struct BinningState {
  BinningHeader* header;
  unsigned long long* keys_a;
  unsigned long long* keys_b;
  uint32_t* vals_a;
  uint32_t* vals_b;
  uint32_t* hits;
  gsr_sort::Workspace sort_ws;
  char* sort_ws_base;
  size_t bytes;
  static BinningState carve(char* base, int64_t capacity, int num_tiles) {
    BinningState b;
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off = gsr_align_up(off + n); return base ? base + o : (char*)nullptr; };
    const size_t cap = (size_t)(capacity > 0 ? capacity : 1);
    const size_t T = (size_t)(num_tiles > 0 ? num_tiles : 1);
    b.header = (BinningHeader*)take(sizeof(BinningHeader));
    // vals_a right behind the header: the sorted point list always ends in vals_a.
    b.vals_a = (uint32_t*)take(cap * 4);
    b.hits = (uint32_t*)take(gsr_bin_hit_words(cap, T) * 4);
    b.vals_b = (uint32_t*)take(cap * 4);
    b.keys_a = (unsigned long long*)take(cap * 8);
    b.keys_b = (unsigned long long*)take(cap * 8);
    const gsr_sort::Workspace probe = gsr_sort::Workspace::carve(nullptr, (int64_t)cap);
    b.sort_ws_base = take(probe.bytes);
    b.sort_ws = gsr_sort::Workspace::carve(b.sort_ws_base, (int64_t)cap);
    b.bytes = off;
    return b;
  }
};

// Number of key bits the sort must cover: 32 depth bits + bits of the largest tile id.
static inline int gsr_key_bits(int num_tiles) {
  int bits = 0;
  unsigned int t = num_tiles > 1 ? (unsigned int)(num_tiles - 1) : 1u;
  while (t) { ++bits; t >>= 1; }
  return 32 + bits;
}
static inline int gsr_sort_passes(int end_bit) {
  return (end_bit + gsr_sort::RADIX_BITS - 1) / gsr_sort::RADIX_BITS;
}

// ---- launch declarations (each stage lives in its own translation unit) ---------------
struct GsrGaussiansDev {
  int P, M;
  const float* means3D;
  const float* opacities;
  const float* shs;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
};

namespace gsr {
void count_launch(int n = 1);

cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);
// tile_count (may be NULL): per-tile instance counters the kernel adds every Gaussian's tiles to.
cudaError_t launch_preprocess(const GsrView& v, const GsrGaussiansDev& g, int32_t* radii, const GeomState& geom,
                              uint32_t* tile_count, cudaStream_t s);
// tile-segmented binning (tile_binning.cu)
// order (may be NULL): receives the tiles by (approximately) descending list length, blockIdx -> tile of the
// compositing kernels
cudaError_t launch_tile_scan(int gx, int gy, const uint32_t* tile_count, uint2* ranges, uint32_t* tile_cursor,
                             uint32_t* large_tiles, unsigned long long* state, GsrControl* ctrl, uint32_t* order,
                             cudaStream_t s);
cudaError_t launch_tile_scatter(const GsrView& v, int P, const int32_t* radii, const GeomState& geom, uint32_t* cursor,
                                unsigned long long* comp, uint32_t capacity, GsrControl* ctrl, BinningHeader* header,
                                cudaStream_t s);
cudaError_t launch_tile_sort(int T, uint2* ranges, const uint32_t* large_tiles, const GsrControl* ctrl,
                             unsigned long long* comp, uint32_t* vals, uint32_t capacity, cudaStream_t s);
cudaError_t launch_export_segment_keys(int T, const uint2* ranges, const uint32_t* vals, const GeomState& geom,
                                       uint32_t limit, unsigned long long* keys, cudaStream_t s);
cudaError_t launch_scan(int P, const GeomState& geom, GsrControl* ctrl, cudaStream_t s);
cudaError_t launch_duplicate(const GsrView& v, int P, const int32_t* radii, const GeomState& geom,
                             unsigned long long* keys, uint32_t* vals, uint32_t capacity, GsrControl* ctrl,
                             BinningHeader* header, uint32_t* hist, int npass, cudaStream_t s);
// Sorts min(*d_n, capacity) pairs; the result lands in (keys_a, vals_a) when the pass
// count is even, else in (keys_b, vals_b).  The workspace must be zeroed before the histograms
// are built; hist_ready = the caller (key duplication) already filled ws.hist.
cudaError_t launch_onesweep(unsigned long long* keys_a, uint32_t* vals_a, unsigned long long* keys_b,
                            uint32_t* vals_b, const unsigned int* d_n, uint32_t capacity, int end_bit,
                            const gsr_sort::Workspace& ws, bool hist_ready, cudaStream_t s);
cudaError_t launch_tile_ranges(const unsigned long long* keys_sorted, const unsigned int* d_n, uint32_t capacity,
                               uint2* ranges, cudaStream_t s);
// tile_order (may be NULL = raster order): blockIdx -> tile
cudaError_t launch_render_forward(const GsrView& v, const uint2* ranges, const uint32_t* tile_order,
                                  const uint32_t* point_list, const GsrRec* recs, float* out_color, float* out_invdepth,
                                  float* final_T, uint32_t* n_contrib, uint32_t* hits, cudaStream_t s);
// binning: the forward's binning buffer (header + sorted point list + hit words).
// grad_acc: (P,12) floats, ten raw pixel sums per Gaussian (render_backward.cu) + 2 pad.
cudaError_t launch_render_backward(const GsrView& v, const uint2* ranges, const uint32_t* tile_order, const void* binning,
                                   const GsrRec* recs, const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                                   const float* dL_dinvdepth, float* grad_acc, cudaStream_t s);
struct PreBwdOut {
  float* dL_dmeans2D; float* dL_dcolors; float* dL_dopacity; float* dL_dmeans3D; float* dL_dcov3D; float* dL_dsh;
  float* dL_dscales; float* dL_drot;
  int accumulate;
  float* max_radii2D; float* xyz_grad_accum; float* xyz_grad_count;
};
cudaError_t launch_preprocess_backward(const GsrView& v, const GsrGaussiansDev& g, const int32_t* radii,
                                       const GeomState& geom, const float* grad_acc, const PreBwdOut& out,
                                       cudaStream_t s);
// all views of a step in one pass over the Gaussians (preprocess_backward.cu); supported() says whether the fused
// kernel can take these tensors (else the caller runs the per-view kernel V times, accumulating)
bool preprocess_backward_views_supported(const GsrGaussiansDev& g, const PreBwdOut& out);
cudaError_t launch_preprocess_backward_views(int V, const GsrView* views, const int32_t* const* radii,
                                             const GeomState* geoms, const float* const* grad_accs,
                                             const GsrGaussiansDev& g, const PreBwdOut& out, cudaStream_t s);
cudaError_t launch_export_geom(int P, const int32_t* radii, const GeomState& geom, float* depths, float* means2D,
                               float* cov3D, float* conic_opacity, float* rgb, uint8_t* clamped,
                               uint32_t* tiles_touched, uint32_t* point_offsets, cudaStream_t s);
// loss.cu / optimizer.cu
// CTA size of a kernel from the environment (32 / 64 / 128 / 256; anything else: the default), read once
inline int env_threads(const char* name, int dflt) {
  const char* e = getenv(name);
  if (!e) return dflt;
  const int v = atoi(e);
  return (v == 32 || v == 64 || v == 128 || v == 256) ? v : dflt;
}
size_t l1_scratch_bytes();
cudaError_t launch_l1_loss(const float* image, const float* target, long long n, float* dL_dimage, float* out_loss,
                           void* scratch, cudaStream_t s);
cudaError_t launch_l1_loss_noalloc(const float* image, const float* target, long long n, float* dL_dimage,
                                   float* out_loss, cudaStream_t s);
size_t ssim_scratch_bytes(int planes, int H, int W);
cudaError_t launch_fused_ssim(const float* img1, const float* img2, int planes, int H, int W, float C1, float C2,
                              int valid, float w_l1, float w_ssim, float w_const, float* dL_dimg1, float* out_loss,
                              float* out_ssim, float* out_l1, void* scratch, cudaStream_t s);
cudaError_t launch_adam(const struct GsrAdamGroup* groups, int num_groups, int step, double beta1, double beta2,
                        double eps, int zero_grad, cudaStream_t s);
cudaError_t launch_densify_fold(int P, const float* d_max_radii, const float* d_accum, const float* d_count,
                                float* max_radii2D, float* xyz_grad_accum, float* xyz_grad_count, cudaStream_t s);
}  // namespace gsr
