/* gsr_b200.h -- C-ABI of the B200-native 3D Gaussian Splatting rasterizer (libgsr_b200.so).
 *
 * This is the drop-in boundary for the rasterizer hot path of
 * pomelyu/gaussian_splatting_lightning.  Each entry point replaces one function of the
 * `diff_gaussian_rasterization._C` extension the reference calls (the extension's source,
 * graphdeco-inria/diff-gaussian-rasterization @ 9c5c2028, is an un-vendored submodule:
 * /root/reference/.gitmodules:1-4, .SUBMODULES.json:9-15), so the citations below are the
 * reference's CALL SITES, which fix the contract:
 *
 *   gsr_forward       <- _C.rasterize_gaussians            gs_lightning/lightning/gs_lightning_module.py:337-348
 *                                                            scripts/render_trained_image.py:113-124
 *   gsr_backward      <- _C.rasterize_gaussians_backward    autograd of the call above; the dummy
 *                                                            means2D leaf at gs_lightning_module.py:316-317,177
 *   gsr_mark_visible  <- _C.mark_visible                    tests/rasterizer_python/test_mark_visible.py:13-14
 *                                                            (python twin gs_lightning/rasterize/rasterize.py:23-26)
 *   gsr_alloc_fn      <- the resizeFunctional buffer lambda notes/rasterizer_note.h:27-40
 *
 * Conventions
 *   - every pointer marked "device" is a CUDA device pointer on the CURRENT device; the
 *     caller selects the device (cudaSetDevice / torch.cuda.device) before the call.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - all entry points return GSR_OK (0) or a negative GSR_E* code; gsr_last_error() gives
 *     the message of the last failure on the calling thread.
 *   - matrices use the reference's row-vector convention: p_view = [p,1] @ viewmatrix,
 *     element (r,c) at flat index 4r+c (tests/rasterizer_python/test_cases.py:23-28).
 *   - nothing here takes or returns a torch type.
 */
#ifndef GSR_B200_H
#define GSR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_OK 0
#define GSR_EINVAL (-1)  /* bad argument (shape, NULL, alignment)        */
#define GSR_ECUDA (-2)   /* a CUDA runtime call or kernel failed         */
#define GSR_EALLOC (-3)  /* the allocation callback returned NULL        */
#define GSR_ELIMIT (-4)  /* a size exceeds what the kernels index (2^30) */

#define GSR_BUF_GEOM 0
#define GSR_BUF_BINNING 1
#define GSR_BUF_IMG 2

/* Grows buffer `which` (GSR_BUF_*) to at least `bytes` and returns its device pointer
 * (128-byte aligned).  May be called more than once per buffer in one gsr_forward; the
 * last returned pointer is the live one.  notes/rasterizer_note.h:27-40. */
typedef void* (*gsr_alloc_fn)(void* user, int which, size_t bytes);

/* The 13 fields of GaussianRasterizationSettings (tests/rasterizer_python/test_cases.py:7-18,
 * gs_lightning_module.py:322-336).  Tensor-valued fields stay on the device so that a call
 * never forces a device-to-host copy. */
typedef struct GsrSettings {
  int32_t image_height;
  int32_t image_width;
  float tanfovx;
  float tanfovy;
  float scale_modifier;
  int32_t sh_degree;
  int32_t prefiltered;       /* accepted, no effect (reference always passes False) */
  int32_t debug;             /* synchronise + check after every kernel             */
  int32_t antialiasing;
  const float* bg;           /* device (3)  */
  const float* viewmatrix;   /* device (16) */
  const float* projmatrix;   /* device (16) */
  const float* campos;       /* device (3)  */
} GsrSettings;

/* Borrowed, read-only Gaussian parameters exactly as GaussianModel.get_* returns them
 * (gs_lightning/modules/gaussian_model.py:317-330): contiguous fp32, AoS. */
typedef struct GsrGaussians {
  int32_t P;                    /* number of Gaussians                                  */
  int32_t M;                    /* SH coefficients stored per Gaussian (shs is P x M x 3) */
  const float* means3D;         /* device (P,3)                                          */
  const float* opacities;       /* device (P,1)                                          */
  const float* shs;             /* device (P,M,3) or NULL when colors_precomp is given   */
  const float* colors_precomp;  /* device (P,3) or NULL                                  */
  const float* scales;          /* device (P,3) or NULL when cov3D_precomp is given      */
  const float* rotations;       /* device (P,4) (w,x,y,z) or NULL                        */
  const float* cov3D_precomp;   /* device (P,6) or NULL                                  */
} GsrGaussians;

/* Gradients produced by gsr_backward; same order and shapes as the tuple
 * _C.rasterize_gaussians_backward returns.  All device.  ANY pointer may be NULL: that
 * gradient is then not wanted and is neither computed nor written (a training step that
 * passes shs + scales/rotations never reads dL_dcolors / dL_dcov3D). */
typedef struct GsrGrads {
  float* dL_dmeans2D;   /* (P,3)  z = 0; consumer takes [:, :2] (gaussian_model.py:180) */
  float* dL_dcolors;    /* (P,3)                                                         */
  float* dL_dopacity;   /* (P,1)                                                         */
  float* dL_dmeans3D;   /* (P,3)                                                         */
  float* dL_dcov3D;     /* (P,6)                                                         */
  float* dL_dsh;        /* (P,M,3) or NULL                                               */
  float* dL_dscales;    /* (P,3)                                                         */
  float* dL_drotations; /* (P,4)                                                         */
} GsrGrads;

/* gsr_backward flags */
#define GSR_GRADS_ACCUMULATE 1 /* outputs are += (several views into one gradient buffer, the layout the
                                  view-parallel all-reduce sends); rows of culled Gaussians are not touched */

#define GSR_BWD_RENDER_ONLY 2   /* stop after the compositing backward: `scratch` then holds the view's raw pixel sums
                                  (P,12) for gsr_preprocess_backward_views; grads may be NULL, nothing else is written */

/* Densification statistics of one view, fused into the backward so that the per-view norm is
 * taken BEFORE gradients of several views are summed.  Mirrors
 * GaussianModel.update_max_radii2D / update_xyz_gradient (gs_lightning/modules/gaussian_model.py:174-181)
 * with visible_mask = radii > 0 (gs_lightning_module.py:192-195).  Any pointer may be NULL. */
typedef struct GsrDensifyStats {
  float* max_radii2D;     /* (P)  = max(max_radii2D, radii)              where radii > 0 */
  float* xyz_grad_accum;  /* (P) += || dL_dmeans2D[:, :2] ||_2           where radii > 0 */
  float* xyz_grad_count;  /* (P) += 1                                     where radii > 0 */
} GsrDensifyStats;

/* Canonical-layout export of the opaque scratch buffers for the
 * bit-exact parity checks.  Any pointer may be NULL (skipped). All device. */
typedef struct GsrStateExport {
  float* depths;            /* (P)    0 where culled */
  float* means2D;           /* (P,2)                 */
  float* cov3D;             /* (P,6)                 */
  float* conic_opacity;     /* (P,4)                 */
  float* rgb;               /* (P,3)                 */
  uint8_t* clamped;         /* (P,3)                 */
  uint32_t* tiles_touched;  /* (P)                   */
  uint32_t* point_offsets;  /* (P) inclusive scan    */
  uint64_t* keys_unsorted;  /* (R)                   */
  uint32_t* vals_unsorted;  /* (R)                   */
  uint64_t* keys_sorted;    /* (R)                   */
  uint32_t* vals_sorted;    /* (R) = point_list      */
  uint32_t* ranges;         /* (T,2)                 */
  float* final_T;           /* (H,W)                 */
  uint32_t* n_contrib;      /* (H,W)                 */
} GsrStateExport;

/* Library / build identification: "gsr_b200 <version> sm_100a". */
const char* gsr_version(void);
const char* gsr_last_error(void);

/* Scratch sizes (bytes).  binning is a function of the instance capacity, not of P. */
size_t gsr_geom_bytes(int32_t P);
size_t gsr_img_bytes(int32_t image_width, int32_t image_height);
size_t gsr_binning_bytes(int64_t capacity, int32_t image_width, int32_t image_height);
size_t gsr_backward_scratch_bytes(int32_t P);

/* Forward: preprocess (+ per-tile instance counts) -> tile scan (ranges) -> scatter of the
 * (depth, index) composites into the tile segments -> per-tile sort -> compositing; the lists
 * equal those of a global stable sort of (tile << 32 | depth, index) pairs.  With the
 * environment variable GSR_BINNING=onesweep or gsr_set_binning(GSR_BINNING_ONESWEEP): preprocess -> scan ->
 * duplicate keys -> onesweep radix sort -> tile ranges -> compositing.  Writes out_color (3,H,W), radii (P) int32, out_invdepth (1,H,W).
 *
 * binning_hint = 0 reproduces upstream's behaviour: one blocking read of num_rendered
 * after the scan, then the binning buffer is sized exactly.  binning_hint > 0 sizes the
 * binning buffer to that many instances up front and enqueues the whole frame without a
 * host round trip; num_rendered is read afterwards and, if it exceeded the hint, the
 * binning + compositing stages are re-run with a large enough buffer (results are always
 * exact).  *num_rendered receives R = sum(tiles_touched); *binning_capacity (may be NULL)
 * the instance capacity the live binning buffer was laid out for. */
int gsr_forward(const GsrSettings* settings, const GsrGaussians* gaussians, float* out_color,
                int32_t* radii, float* out_invdepth, gsr_alloc_fn alloc, void* alloc_user,
                int64_t binning_hint, int64_t* num_rendered, int64_t* binning_capacity, void* stream);

/* Backward from the three scratch buffers the forward left behind (read-only here).
 * dL_dinvdepth may be NULL.  `scratch` is gsr_backward_scratch_bytes(P) of device memory.
 * flags: 0 or GSR_GRADS_ACCUMULATE.  stats: NULL or the densification statistics to update. */
int gsr_backward(const GsrSettings* settings, const GsrGaussians* gaussians, const int32_t* radii,
                 const float* dL_dcolor, const float* dL_dinvdepth, const void* geom,
                 const void* binning, const void* img, int64_t num_rendered, void* scratch,
                 const GsrGrads* grads, int32_t flags, const GsrDensifyStats* stats, void* stream);

/* The preprocess backward of ALL views of a training step in one pass over the Gaussians (the reference trains one view
 * per step, gs_lightning_module.py:139-140; the view-parallel step of SURVEY.md 8(e) sums V views per rank).  Each view
 * was rendered by gsr_forward and taken through gsr_backward(..., GSR_BWD_RENDER_ONLY): its settings, radii, geometry
 * buffer and scratch are handed over here.  The Gaussian's inputs are read once, its gradient row written (or, with
 * GSR_GRADS_ACCUMULATE, added) once; the densification statistics stay per-view quantities (the norm of dL_dmeans2D is
 * taken per view before the sum).  grads->dL_dmeans2D must be NULL (a per-view output).  All views must share
 * sh_degree, scale_modifier and antialiasing.  Results equal the sum of per-view gsr_backward calls up to float
 * summation order. */
typedef struct GsrViewState {
  const GsrSettings* settings;
  const int32_t* radii;   /* device (P)                                              */
  const void* geom;       /* the view's geometry buffer (gsr_forward)                */
  const void* scratch;    /* the view's (P,12) raw sums (gsr_backward RENDER_ONLY)    */
} GsrViewState;
int gsr_preprocess_backward_views(const GsrViewState* views, int32_t num_views, const GsrGaussians* gaussians,
                                  const GsrGrads* grads, int32_t flags, const GsrDensifyStats* stats, void* stream);
/* The same pass over the Gaussians [first, first + count) only.  `gaussians` and the views describe the whole set; the
 * pointers of `grads` and `stats` address the RANGE's own rows (row 0 = Gaussian `first`), so that a step buffer laid out
 * range by range can be exchanged range by range: the all-reduce of range c runs under the pass over range c + 1
 * (view_parallel.StepBuffers(chunks=...)). */
int gsr_preprocess_backward_views_range(const GsrViewState* views, int32_t num_views, const GsrGaussians* gaussians,
                                        const GsrGrads* grads, int32_t flags, const GsrDensifyStats* stats,
                                        int32_t first, int32_t count, void* stream);

/* present[i] = ([p_i,1] @ viewmatrix).z > 0.2   (camera_tools.py:5-8). */
int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, uint8_t* present,
                     void* stream);

/* Debug accessor over the buffers of one gsr_forward call (binning_capacity as returned by
 * it).  keys_unsorted / vals_unsorted are regenerated by re-running the duplication kernel
 * because the sort ping-pongs over its input. */
int gsr_export_state(const GsrSettings* settings, int32_t P, const int32_t* radii, const void* geom,
                     const void* binning, const void* img, int64_t num_rendered, int64_t binning_capacity,
                     const GsrStateExport* out, void* stream);

/* Stand-alone entry to the hand-written onesweep radix sort (stable, LSD, 8-bit digits)
 * used between key duplication and tile ranges: sorts n (key,value) pairs on key bits
 * [0, end_bit).  workspace: gsr_sort_workspace_bytes(n) bytes of device memory.
 * keys_out/vals_out receive the result; keys_tmp/vals_tmp are ping-pong storage of n
 * elements each. */
size_t gsr_sort_workspace_bytes(int64_t n);
int gsr_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                   uint32_t* vals_out, uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n,
                   int32_t end_bit, void* workspace, void* stream);

/* ---- the steps on either side of the rasterizer in training_step ---- */

/* nn.L1Loss() of gs_lightning/lightning/gs_lightning_module.py:99,278: *out_loss = mean(|image - target|)
 * (device scalar) and, when dL_dimage is not NULL, the unit gradient sign(image - target) / n in the same
 * pass.  image / target / dL_dimage: n contiguous floats, 16-byte aligned.  scratch:
 * gsr_l1_loss_scratch_bytes() of device memory whose first 128 bytes are zero on first use (the kernel
 * leaves them zero).  The sum is deterministic. */
size_t gsr_l1_loss_scratch_bytes(void);
int gsr_l1_loss(const float* image, const float* target, int64_t n, float* dL_dimage, float* out_loss,
                void* scratch, void* stream);

/* gsr_l1_loss without any allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Same kernel, same result as gsr_l1_loss; the scratch (GSR_L1_NOALLOC_SCRATCH_BYTES) is a static
 * __device__ array inside the library, held by every GPU that loads the library until the process exits.  n above
 * GSR_L1_NOALLOC_MAX_ELEMENTS returns GSR_ELIMIT.  Calls no cudaMalloc / cudaFree / allocating routine (one
 * cudaGetSymbolAddress per device on first use, a 128-byte cudaMemsetAsync and the kernel per call, on `stream`
 * of the CURRENT device).  All calls share the one array: calls that may overlap on the GPU (different streams
 * or host threads) race -- serialise them, or use gsr_l1_loss with a scratch buffer per stream. */
#define GSR_L1_NOALLOC_MAX_ELEMENTS (1 << 28) /* 2^28 floats per image (1 GiB): 3 x 8192 x 8192 fits */
#define GSR_L1_NOALLOC_SCRATCH_BYTES 9600     /* 128 (ticket) + 148 * 8 CTAs x 8 bytes               */
int gsr_l1_loss_noalloc(const float* image, const float* target, int64_t n, float* dL_dimage, float* out_loss,
                        void* stream);

/* fused_ssim(img1, img2, padding, train) of submodules/fused-ssim (fused_ssim/__init__.py:33-41, ssim.cu:187-444),
 * used as `1 - fused_ssim(rendered, gt)` at gs_lightning_module.py:100,279: *out_ssim = mean of the SSIM map
 * (11x11 Gaussian window, sigma 1.5, zero padding; padding_valid != 0 averages the map without its 5-pixel border)
 * and, when dL_dimg1 is not NULL, d(out_ssim)/d(img1) -- both from ONE launch (the reference's fusedssim +
 * fusedssim_backward pair with dL_dmap = 1/N).  img1 / img2 / dL_dimg1: planes x H x W contiguous floats
 * (planes = batch x channels).  scratch: gsr_fused_ssim_scratch_bytes() of device memory whose first 128 bytes
 * are zero on first use (the kernel leaves them zero).  Deterministic. */
size_t gsr_fused_ssim_scratch_bytes(int32_t planes, int32_t H, int32_t W);
int gsr_fused_ssim(const float* img1, const float* img2, int32_t planes, int32_t H, int32_t W, float C1, float C2,
                   int32_t padding_valid, float* dL_dimg1, float* out_ssim, void* scratch, void* stream);

/* The training loss of GSLightningModule.calculate_loss (gs_lightning_module.py:278-284) in ONE launch:
 *   *out_loss = (1 - w_ssim) * L1(image, target) + w_ssim * (1 - fused_ssim(image, target))       (padding "same")
 * and, when dL_dimage is not NULL, d(out_loss)/d(image).  *out_l1 (may be NULL) receives the L1 term and
 * *out_ssim (may be NULL) the mean SSIM (the reference logs recon = L1 and ssim = 1 - mean SSIM).  Same kernel,
 * scratch and determinism as gsr_fused_ssim: the images are read once for both terms. */
int gsr_photometric_loss(const float* image, const float* target, int32_t planes, int32_t H, int32_t W, float w_ssim,
                         float* dL_dimage, float* out_loss, float* out_l1, float* out_ssim, void* scratch, void* stream);

/* torch.optim.Adam over the parameter groups of gs_lightning_module.py:120-127 in ONE launch
 * (configs/*.yaml: eps 1e-15, default betas, no weight decay, no amsgrad).  `step` is the 1-based step
 * count after the increment, as torch uses it for the bias corrections; lr, betas and eps are doubles and are
 * folded into the float32 kernel constants the way torch folds its Python scalars.  All pointers device. */
#define GSR_ADAM_MAX_GROUPS 8
#define GSR_ADAM_ZERO_GRAD 1 /* also zero the gradients (optimizer.zero_grad(), gs_lightning_module.py:168-170) */
typedef struct GsrAdamGroup {
  float* param;       /* (n) updated in place                                   */
  float* grad;        /* (n) read; zeroed with GSR_ADAM_ZERO_GRAD               */
  float* exp_avg;     /* (n) first moment, updated in place                     */
  float* exp_avg_sq;  /* (n) second moment, updated in place                    */
  int64_t n;
  double lr;          /* the group's learning rate of this step                 */
} GsrAdamGroup;
int gsr_adam_step(const GsrAdamGroup* groups, int32_t num_groups, int32_t step, double beta1, double beta2,
                  double eps, int32_t flags, void* stream);

/* Folds the statistics deltas of one step (GsrDensifyStats buffers the backward filled, summed / maxed over
 * the ranks) into the model's running statistics, gs_lightning/modules/gaussian_model.py:174-181: where
 * d_count > 0:  max_radii2D = max(., d_max_radii); xyz_grad_accum += d_accum; xyz_grad_count += d_count.
 * The three outputs may be NULL. */
int gsr_densify_fold(int32_t P, const float* d_max_radii, const float* d_accum, const float* d_count,
                     float* max_radii2D, float* xyz_grad_accum, float* xyz_grad_count, void* stream);

/* The binning the forward runs: GSR_BINNING_TILE (default: tile-segmented, csrc/tile_binning.cu) or GSR_BINNING_ONESWEEP
 * (scan + key duplication + onesweep radix sort + ranges).  Same sorted lists and ranges, bit for bit.  Process-wide;
 * initial value from the environment variable GSR_BINNING=onesweep.  The forward, backward and export of one frame must
 * run under the same mode. */
#define GSR_BINNING_TILE 0
#define GSR_BINNING_ONESWEEP 1
int gsr_set_binning(int32_t mode);
int32_t gsr_get_binning(void);

/* Priority lanes (no counterpart in the reference: a scheduling choice of this library).  With lanes on, gsr_forward
 * runs the stages before the compositing kernel (preprocess, tile scan, scatter, sort: bound by latency and L2 atomics)
 * on a high-priority stream of its own -- one per (device, caller stream), forked from and joined back into the
 * caller's stream by events -- so that, when several views are in flight on several streams, those stages run BESIDE
 * another view's issue-bound compositing kernel instead of queueing behind its grid.  Results are identical.
 * Process-wide; initial value from the environment variable GSR_PRIORITY_LANES (0 / 1). */
int gsr_set_priority_lanes(int32_t on);
int32_t gsr_get_priority_lanes(void);

/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
int64_t gsr_launch_count(void);

/* Optional per-stage device timing (CUDA events on the launching stream around every stage of
 * gsr_forward / gsr_backward); used by bench.py for the per-kernel roofline.  Stages:
 * preprocess, scan, duplicate, sort, tile_ranges, render_forward, render_backward,
 * preprocess_backward.  gsr_profile_collect() waits for the recorded events and folds them into
 * the per-stage totals read by gsr_profile_stage_ms / _calls. */
int gsr_profile_enable(int on);
void gsr_profile_reset(void);
int gsr_profile_collect(void);
int gsr_profile_stage_count(void);
const char* gsr_profile_stage_name(int i);
double gsr_profile_stage_ms(int i);
int64_t gsr_profile_stage_calls(int i);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H */
