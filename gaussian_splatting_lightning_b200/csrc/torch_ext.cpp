// torch_ext.cpp -- `_C`: the thin PyTorch C++ extension between Python and libgsr_b200.so.
//
// Same entry points, argument order and returns as the `diff_gaussian_rasterization._C` module the reference's
// package calls (un-vendored submodule graphdeco-inria/diff-gaussian-rasterization @ 9c5c2028; contract by the call
// sites gs_lightning/lightning/gs_lightning_module.py:337-348, scripts/render_trained_image.py:113-124,
// tests/rasterizer_python/test_mark_visible.py:13-14 and SURVEY.md 8(b)):
//
//   rasterize_gaussians(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp,
//                       viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, degree, campos,
//                       prefiltered, antialiasing, debug)
//       -> (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepth)
//   rasterize_gaussians_backward(bg, means3D, radii, colors_precomp, opacities, scales, rotations, scale_modifier,
//                       cov3D_precomp, viewmatrix, projmatrix, tanfovx, tanfovy, dL_dout_color, dL_dout_invdepth, sh,
//                       degree, campos, geomBuffer, R, binningBuffer, imageBuffer, antialiasing, debug)
//       -> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
//   mark_visible(means3D, viewmatrix, projmatrix) -> bool (P)
//
// plus the `_ex` variants the Python layer of this repository uses (explicit binning hint, gradient selection,
// fused accumulation, densification statistics).  Nothing here computes: every function checks its tensors, takes a
// device guard and the current stream, and calls the C-ABI of include/gsr_b200.h.  The three scratch buffers are
// uint8 tensors grown through tensor.resize_ by the C-ABI's allocation callback (the resizeFunctional idiom of
// notes/rasterizer_note.h:27-40); torch owns all memory.
#include <torch/extension.h>

#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <array>
#include <mutex>
#include <optional>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "../../include/gsr_b200.h"

namespace {

using torch::Tensor;
using OptTensor = std::optional<Tensor>;

bool present(const OptTensor& t) { return t.has_value() && t->defined() && t->numel() > 0; }

// contiguous float32 on `dev`, 16-byte aligned; absent / empty -> undefined tensor
Tensor f32c(const OptTensor& t, const c10::Device& dev) {
  if (!present(t)) return Tensor();
  Tensor x = *t;
  if (x.device() != dev) x = x.to(dev);
  if (x.scalar_type() != torch::kFloat32) x = x.to(torch::kFloat32);
  x = x.contiguous();
  if (reinterpret_cast<uintptr_t>(x.data_ptr()) & 15) x = x.clone(at::MemoryFormat::Contiguous);
  return x;
}
const float* fp(const Tensor& t) { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; }
float* fpm(const Tensor& t) { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; }

void check(int rc, const char* what) {
  TORCH_CHECK(rc == GSR_OK, what, " failed (code ", rc, "): ", gsr_last_error());
}

struct Scratch {
  std::array<Tensor, 3> buf;
};
void* resize_cb(void* user, int which, size_t bytes) {
  auto* s = static_cast<Scratch*>(user);
  if (which < 0 || which > 2) return nullptr;
  try {
    s->buf[which].resize_({(int64_t)(bytes ? bytes : 1)});
    return s->buf[which].data_ptr();
  } catch (...) {
    return nullptr;  // gsr_forward reports GSR_EALLOC
  }
}

struct Frame {
  c10::Device dev;
  Tensor means3D, sh, colors, opacities, scales, rotations, cov3D, bg, view, proj, campos;
  GsrSettings st;
  GsrGaussians g;
  Frame(const Tensor& means3D_, const OptTensor& sh_, const OptTensor& colors_, const Tensor& opacities_,
        const OptTensor& scales_, const OptTensor& rotations_, const OptTensor& cov3D_, const Tensor& bg_,
        const Tensor& view_, const Tensor& proj_, const Tensor& campos_, double tanfovx, double tanfovy, int64_t H,
        int64_t W, double scale_modifier, int64_t degree, bool prefiltered, bool antialiasing, bool debug)
      : dev(means3D_.device()) {
    TORCH_CHECK(means3D_.dim() == 2 && means3D_.size(1) == 3, "means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D_.is_cuda(), "gaussian_splatting_lightning_b200 rasterizes on CUDA devices only (no CPU fallback)");
    const int64_t P = means3D_.size(0);
    means3D = P ? f32c(means3D_, dev) : means3D_;
    sh = f32c(sh_, dev); colors = f32c(colors_, dev); opacities = f32c(opacities_, dev);
    scales = f32c(scales_, dev); rotations = f32c(rotations_, dev); cov3D = f32c(cov3D_, dev);
    bg = f32c(bg_, dev); view = f32c(view_, dev); proj = f32c(proj_, dev); campos = f32c(campos_, dev);
    st.image_height = (int32_t)H; st.image_width = (int32_t)W;
    st.tanfovx = (float)tanfovx; st.tanfovy = (float)tanfovy; st.scale_modifier = (float)scale_modifier;
    st.sh_degree = (int32_t)degree; st.prefiltered = prefiltered; st.debug = debug; st.antialiasing = antialiasing;
    st.bg = fp(bg); st.viewmatrix = fp(view); st.projmatrix = fp(proj); st.campos = fp(campos);
    g.P = (int32_t)P;
    g.M = sh.defined() ? (int32_t)sh.size(1) : 0;
    g.means3D = fp(means3D); g.opacities = fp(opacities); g.shs = fp(sh); g.colors_precomp = fp(colors);
    g.scales = fp(scales); g.rotations = fp(rotations); g.cov3D_precomp = fp(cov3D);
  }
};

// Last binning capacity that was enough, per (device, W, H): the upstream-signature entry point sizes the next
// frame's binning buffer from it so that the frame is enqueued without a device-to-host round trip.
std::mutex g_hint_mu;
std::unordered_map<uint64_t, int64_t> g_hint;
uint64_t hint_key(int dev, int64_t W, int64_t H) { return ((uint64_t)(dev & 0xff) << 56) | ((uint64_t)W << 28) | (uint64_t)H; }
int64_t bucket(int64_t n) {  // 4-8 steps per octave: sizes repeat, the caching allocator reuses blocks as they are
  if (n <= 65536) return 65536;
  int b = 0;
  for (int64_t t = n; t; t >>= 1) ++b;
  const int64_t step = (int64_t)1 << (b - 3);
  return (n + step - 1) / step * step;
}

std::tuple<int64_t, Tensor, Tensor, Tensor, Tensor, Tensor, Tensor, int64_t> rasterize_gaussians_ex(
    const Tensor& bg, const Tensor& means3D, const OptTensor& colors, const Tensor& opacities, const OptTensor& scales,
    const OptTensor& rotations, double scale_modifier, const OptTensor& cov3D, const Tensor& viewmatrix,
    const Tensor& projmatrix, double tanfovx, double tanfovy, int64_t image_height, int64_t image_width,
    const OptTensor& sh, int64_t degree, const Tensor& campos, bool prefiltered, bool antialiasing, bool debug,
    int64_t binning_hint) {
  Frame fr(means3D, sh, colors, opacities, scales, rotations, cov3D, bg, viewmatrix, projmatrix, campos, tanfovx,
           tanfovy, image_height, image_width, scale_modifier, degree, prefiltered, antialiasing, debug);
  const c10::cuda::CUDAGuard guard(fr.dev);
  auto fopt = torch::TensorOptions().dtype(torch::kFloat32).device(fr.dev);
  Tensor color = torch::empty({3, image_height, image_width}, fopt);
  Tensor invdepth = torch::empty({1, image_height, image_width}, fopt);
  Tensor radii = torch::empty({(int64_t)fr.g.P}, fopt.dtype(torch::kInt32));
  Scratch sc;
  for (auto& b : sc.buf) b = torch::empty({0}, fopt.dtype(torch::kUInt8));
  int64_t R = 0, cap = 0;
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(fr.dev.index()).stream();
  int rc;
  {
    pybind11::gil_scoped_release nogil;  // the call may block on the num_rendered mailbox; no Python object is touched
    rc = gsr_forward(&fr.st, &fr.g, color.data_ptr<float>(), fr.g.P ? radii.data_ptr<int32_t>() : nullptr,
                     invdepth.data_ptr<float>(), resize_cb, &sc, binning_hint, &R, &cap, (void*)stream);
  }
  check(rc, "gsr_forward");
  return {R, color, radii, sc.buf[GSR_BUF_GEOM], sc.buf[GSR_BUF_BINNING], sc.buf[GSR_BUF_IMG], invdepth, cap};
}

std::tuple<int64_t, Tensor, Tensor, Tensor, Tensor, Tensor, Tensor> rasterize_gaussians(
    const Tensor& bg, const Tensor& means3D, const OptTensor& colors, const Tensor& opacities, const OptTensor& scales,
    const OptTensor& rotations, double scale_modifier, const OptTensor& cov3D, const Tensor& viewmatrix,
    const Tensor& projmatrix, double tanfovx, double tanfovy, int64_t image_height, int64_t image_width,
    const OptTensor& sh, int64_t degree, const Tensor& campos, bool prefiltered, bool antialiasing, bool debug) {
  const uint64_t key = hint_key(means3D.is_cuda() ? means3D.get_device() : 0, image_width, image_height);
  int64_t hint = 0;
  {
    std::lock_guard<std::mutex> l(g_hint_mu);
    auto it = g_hint.find(key);
    if (it != g_hint.end()) hint = it->second;
  }
  auto out = rasterize_gaussians_ex(bg, means3D, colors, opacities, scales, rotations, scale_modifier, cov3D, viewmatrix,
                                    projmatrix, tanfovx, tanfovy, image_height, image_width, sh, degree, campos,
                                    prefiltered, antialiasing, debug, hint);
  {
    std::lock_guard<std::mutex> l(g_hint_mu);
    const int64_t want = bucket(std::get<0>(out) + std::get<0>(out) / 8 + 4096);
    int64_t& h = g_hint[key];
    h = std::max(want, h - h / 16);  // grows at once, shrinks slowly
  }
  return {std::get<0>(out), std::get<1>(out), std::get<2>(out), std::get<3>(out), std::get<4>(out), std::get<5>(out),
          std::get<6>(out)};
}

// order of the gradient tuple (= GsrGrads): dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
// dL_dscales, dL_drotations
constexpr int NUM_GRADS = 8;

// want_mask: bit i set -> gradient i is produced (others come back as None, not computed).
// accumulate_into: empty, or 8 optional tensors: the given ones receive `+=` (GSR_GRADS_ACCUMULATE), nothing else is made.
// stats: empty, or 3 optional tensors (max_radii2D, xyz_grad_accum, xyz_grad_count), (P) float32 each.
std::vector<OptTensor> rasterize_gaussians_backward_ex(
    const Tensor& bg, const Tensor& means3D, const Tensor& radii, const OptTensor& colors, const Tensor& opacities,
    const OptTensor& scales, const OptTensor& rotations, double scale_modifier, const OptTensor& cov3D,
    const Tensor& viewmatrix, const Tensor& projmatrix, double tanfovx, double tanfovy, const OptTensor& dL_dout_color,
    const OptTensor& dL_dout_invdepth, const OptTensor& sh, int64_t degree, const Tensor& campos, const Tensor& geomBuffer,
    int64_t R, const Tensor& binningBuffer, const Tensor& imageBuffer, bool antialiasing, bool debug, int64_t want_mask,
    const std::vector<OptTensor>& accumulate_into, const std::vector<OptTensor>& stats) {
  TORCH_CHECK(imageBuffer.defined() && geomBuffer.defined(), "backward: scratch buffers missing");
  int64_t H = 0, W = 0;
  if (present(dL_dout_color)) { H = dL_dout_color->size(-2); W = dL_dout_color->size(-1); }
  else if (present(dL_dout_invdepth)) { H = dL_dout_invdepth->size(-2); W = dL_dout_invdepth->size(-1); }
  TORCH_CHECK(H > 0 && W > 0, "backward: one of dL_dout_color / dL_dout_invdepth must be given");
  Frame fr(means3D, sh, colors, opacities, scales, rotations, cov3D, bg, viewmatrix, projmatrix, campos, tanfovx,
           tanfovy, H, W, scale_modifier, degree, false, antialiasing, debug);
  const c10::cuda::CUDAGuard guard(fr.dev);
  const int64_t P = fr.g.P, M = fr.g.M;
  auto fopt = torch::TensorOptions().dtype(torch::kFloat32).device(fr.dev);
  const std::array<std::vector<int64_t>, NUM_GRADS> shapes = {{{P, 3}, {P, 3}, {P, 1}, {P, 3}, {P, 6}, {P, M, 3}, {P, 3}, {P, 4}}};
  std::vector<OptTensor> out(NUM_GRADS);
  const bool acc = !accumulate_into.empty();
  if (acc) {
    TORCH_CHECK((int)accumulate_into.size() == NUM_GRADS, "accumulate_into must hold 8 entries");
    for (int i = 0; i < NUM_GRADS; ++i) {
      if (!accumulate_into[i].has_value() || !accumulate_into[i]->defined()) continue;
      const Tensor& t = *accumulate_into[i];
      int64_t n = 1;
      for (auto d : shapes[i]) n *= d;
      TORCH_CHECK(t.is_cuda() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() && t.numel() == n,
                  "accumulate_into[", i, "] must be a contiguous float32 CUDA tensor of ", n, " elements");
      out[i] = t;
    }
  } else {
    for (int i = 0; i < NUM_GRADS; ++i)
      if (want_mask & (1ll << i)) out[i] = torch::empty(shapes[i], fopt);
  }
  if (P == 0) return out;
  Tensor dC = f32c(dL_dout_color, fr.dev);
  if (!dC.defined()) dC = torch::zeros({3, H, W}, fopt);
  Tensor dD = f32c(dL_dout_invdepth, fr.dev);
  GsrGrads gr;
  float** slots[NUM_GRADS] = {&gr.dL_dmeans2D, &gr.dL_dcolors, &gr.dL_dopacity, &gr.dL_dmeans3D, &gr.dL_dcov3D, &gr.dL_dsh,
                              &gr.dL_dscales, &gr.dL_drotations};
  for (int i = 0; i < NUM_GRADS; ++i) *slots[i] = out[i].has_value() ? fpm(*out[i]) : nullptr;
  GsrDensifyStats ds{nullptr, nullptr, nullptr};
  bool have_stats = false;
  if (!stats.empty()) {
    TORCH_CHECK(stats.size() == 3, "stats must hold 3 entries");
    float** sp[3] = {&ds.max_radii2D, &ds.xyz_grad_accum, &ds.xyz_grad_count};
    for (int i = 0; i < 3; ++i) {
      if (!stats[i].has_value() || !stats[i]->defined()) continue;
      const Tensor& t = *stats[i];
      TORCH_CHECK(t.is_cuda() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() && t.numel() == P,
                  "stats[", i, "] must be a contiguous float32 CUDA tensor of ", P, " elements");
      *sp[i] = t.data_ptr<float>();
      have_stats = true;
    }
  }
  Tensor scratch = torch::empty({(int64_t)gsr_backward_scratch_bytes((int32_t)P)}, fopt.dtype(torch::kUInt8));
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(fr.dev.index()).stream();
  Tensor radii_c = radii.contiguous();
  check(gsr_backward(&fr.st, &fr.g, radii_c.data_ptr<int32_t>(), dC.data_ptr<float>(), fp(dD), geomBuffer.data_ptr(),
                     binningBuffer.defined() && binningBuffer.numel() ? binningBuffer.data_ptr() : nullptr,
                     imageBuffer.data_ptr(), R, scratch.data_ptr(), &gr, acc ? GSR_GRADS_ACCUMULATE : 0,
                     have_stats ? &ds : nullptr, (void*)stream),
        "gsr_backward");
  return out;
}

std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor, Tensor, Tensor, Tensor> rasterize_gaussians_backward(
    const Tensor& bg, const Tensor& means3D, const Tensor& radii, const OptTensor& colors, const Tensor& opacities,
    const OptTensor& scales, const OptTensor& rotations, double scale_modifier, const OptTensor& cov3D,
    const Tensor& viewmatrix, const Tensor& projmatrix, double tanfovx, double tanfovy, const Tensor& dL_dout_color,
    const OptTensor& dL_dout_invdepth, const OptTensor& sh, int64_t degree, const Tensor& campos, const Tensor& geomBuffer,
    int64_t R, const Tensor& binningBuffer, const Tensor& imageBuffer, bool antialiasing, bool debug) {
  auto g = rasterize_gaussians_backward_ex(bg, means3D, radii, colors, opacities, scales, rotations, scale_modifier, cov3D,
                                           viewmatrix, projmatrix, tanfovx, tanfovy, dL_dout_color, dL_dout_invdepth, sh,
                                           degree, campos, geomBuffer, R, binningBuffer, imageBuffer, antialiasing, debug,
                                           0xff, {}, {});
  return {*g[0], *g[1], *g[2], *g[3], *g[4], *g[5], *g[6], *g[7]};
}

// The compositing backward alone (gsr_backward with GSR_BWD_RENDER_ONLY): returns the view's (P,12) raw pixel sums for
// preprocess_backward_views.
Tensor render_backward_only(
    const Tensor& bg, const Tensor& means3D, const Tensor& radii, const OptTensor& colors, const Tensor& opacities,
    const OptTensor& scales, const OptTensor& rotations, double scale_modifier, const OptTensor& cov3D,
    const Tensor& viewmatrix, const Tensor& projmatrix, double tanfovx, double tanfovy, const OptTensor& dL_dout_color,
    const OptTensor& dL_dout_invdepth, const OptTensor& sh, int64_t degree, const Tensor& campos, const Tensor& geomBuffer,
    int64_t R, const Tensor& binningBuffer, const Tensor& imageBuffer, bool antialiasing, bool debug) {
  int64_t H = 0, W = 0;
  if (present(dL_dout_color)) { H = dL_dout_color->size(-2); W = dL_dout_color->size(-1); }
  else if (present(dL_dout_invdepth)) { H = dL_dout_invdepth->size(-2); W = dL_dout_invdepth->size(-1); }
  TORCH_CHECK(H > 0 && W > 0, "backward: one of dL_dout_color / dL_dout_invdepth must be given");
  Frame fr(means3D, sh, colors, opacities, scales, rotations, cov3D, bg, viewmatrix, projmatrix, campos, tanfovx,
           tanfovy, H, W, scale_modifier, degree, false, antialiasing, debug);
  const c10::cuda::CUDAGuard guard(fr.dev);
  auto fopt = torch::TensorOptions().dtype(torch::kFloat32).device(fr.dev);
  const int64_t P = fr.g.P;
  Tensor scratch = torch::empty({(int64_t)gsr_backward_scratch_bytes((int32_t)P)}, fopt.dtype(torch::kUInt8));
  if (P == 0) return scratch;
  Tensor dC = f32c(dL_dout_color, fr.dev);
  if (!dC.defined()) dC = torch::zeros({3, H, W}, fopt);
  Tensor dD = f32c(dL_dout_invdepth, fr.dev);
  Tensor radii_c = radii.contiguous();
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(fr.dev.index()).stream();
  check(gsr_backward(&fr.st, &fr.g, radii_c.data_ptr<int32_t>(), dC.data_ptr<float>(), fp(dD), geomBuffer.data_ptr(),
                     binningBuffer.defined() && binningBuffer.numel() ? binningBuffer.data_ptr() : nullptr,
                     imageBuffer.data_ptr(), R, scratch.data_ptr(), nullptr, GSR_BWD_RENDER_ONLY, nullptr, (void*)stream),
        "gsr_backward");
  return scratch;
}

// One view of a step for preprocess_backward_views:
// (bg, viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, campos, radii, geomBuffer, scratch)
using ViewTuple = std::tuple<Tensor, Tensor, Tensor, double, double, int64_t, int64_t, Tensor, Tensor, Tensor, Tensor>;

// gsr_preprocess_backward_views: the preprocess backward of all views of a step in one pass over the Gaussians.
// Returns the 8-tuple layout of the backward (dL_dmeans2D is a per-view output and always None here).
std::vector<OptTensor> preprocess_backward_views(
    const std::vector<ViewTuple>& views, const Tensor& means3D, const OptTensor& colors, const Tensor& opacities,
    const OptTensor& scales, const OptTensor& rotations, double scale_modifier, const OptTensor& cov3D, const OptTensor& sh,
    int64_t degree, bool antialiasing, bool debug, int64_t want_mask, const std::vector<OptTensor>& accumulate_into,
    const std::vector<OptTensor>& stats, bool overwrite, int64_t first, int64_t count) {
  // [first, first + count): the Gaussians this call covers (count < 0: all); outputs and statistics hold `count` rows
  // overwrite: the tensors of accumulate_into are WRITTEN (every row, zeros where no view saw the Gaussian) instead of
  // added to -- the step's buffer needs no clearing and the kernel no reductions at the L2
  TORCH_CHECK(!views.empty(), "preprocess_backward_views: no views");
  const c10::Device dev = means3D.device();
  std::vector<Frame> frames;   // keeps every view's device-side settings alive over the call
  frames.reserve(views.size());
  std::vector<Tensor> keep;
  std::vector<GsrViewState> vs(views.size());
  for (size_t k = 0; k < views.size(); ++k) {
    const auto& [bg, view, proj, tfx, tfy, H, W, campos, radii, geom, scratch] = views[k];
    frames.emplace_back(means3D, sh, colors, opacities, scales, rotations, cov3D, bg, view, proj, campos, tfx, tfy, H, W,
                        scale_modifier, degree, false, antialiasing, debug);
    keep.push_back(radii.contiguous());
  }
  for (size_t k = 0; k < views.size(); ++k) {
    vs[k].settings = &frames[k].st;
    vs[k].radii = keep[k].data_ptr<int32_t>();
    vs[k].geom = std::get<9>(views[k]).data_ptr();
    vs[k].scratch = std::get<10>(views[k]).data_ptr();
  }
  const Frame& fr = frames[0];
  const c10::cuda::CUDAGuard guard(dev);
  const int64_t Pall = fr.g.P, M = fr.g.M;
  if (count < 0) { first = 0; count = Pall; }
  TORCH_CHECK(first >= 0 && first + count <= Pall, "preprocess_backward_views: range outside [0, P)");
  const int64_t P = count;
  auto fopt = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  const std::array<std::vector<int64_t>, NUM_GRADS> shapes = {{{P, 3}, {P, 3}, {P, 1}, {P, 3}, {P, 6}, {P, M, 3}, {P, 3}, {P, 4}}};
  std::vector<OptTensor> out(NUM_GRADS);
  const bool acc = !accumulate_into.empty();
  if (acc) {
    TORCH_CHECK((int)accumulate_into.size() == NUM_GRADS, "accumulate_into must hold 8 entries");
    for (int i = 1; i < NUM_GRADS; ++i) {
      if (!accumulate_into[i].has_value() || !accumulate_into[i]->defined()) continue;
      const Tensor& t = *accumulate_into[i];
      int64_t n = 1;
      for (auto d : shapes[i]) n *= d;
      TORCH_CHECK(t.is_cuda() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() && t.numel() == n,
                  "accumulate_into[", i, "] must be a contiguous float32 CUDA tensor of ", n, " elements");
      out[i] = t;
    }
  } else {
    for (int i = 1; i < NUM_GRADS; ++i)
      if (want_mask & (1ll << i)) out[i] = torch::empty(shapes[i], fopt);
  }
  if (P == 0) return out;
  GsrGrads gr;
  float** slots[NUM_GRADS] = {&gr.dL_dmeans2D, &gr.dL_dcolors, &gr.dL_dopacity, &gr.dL_dmeans3D, &gr.dL_dcov3D, &gr.dL_dsh,
                              &gr.dL_dscales, &gr.dL_drotations};
  for (int i = 0; i < NUM_GRADS; ++i) *slots[i] = out[i].has_value() ? fpm(*out[i]) : nullptr;
  GsrDensifyStats ds{nullptr, nullptr, nullptr};
  bool have_stats = false;
  if (!stats.empty()) {
    TORCH_CHECK(stats.size() == 3, "stats must hold 3 entries");
    float** sp[3] = {&ds.max_radii2D, &ds.xyz_grad_accum, &ds.xyz_grad_count};
    for (int i = 0; i < 3; ++i) {
      if (!stats[i].has_value() || !stats[i]->defined()) continue;
      const Tensor& t = *stats[i];
      TORCH_CHECK(t.is_cuda() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() && t.numel() == P,
                  "stats[", i, "] must be a contiguous float32 CUDA tensor of ", P, " elements");
      *sp[i] = t.data_ptr<float>();
      have_stats = true;
    }
  }
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(dev.index()).stream();
  check(gsr_preprocess_backward_views_range(vs.data(), (int32_t)vs.size(), &fr.g, &gr,
                                            (acc && !overwrite) ? GSR_GRADS_ACCUMULATE : 0, have_stats ? &ds : nullptr,
                                            (int32_t)first, (int32_t)count, (void*)stream),
        "gsr_preprocess_backward_views");
  return out;
}

Tensor mark_visible(const Tensor& means3D, const Tensor& viewmatrix, const OptTensor& /*projmatrix*/) {
  TORCH_CHECK(means3D.is_cuda(), "mark_visible runs on CUDA devices only (no CPU fallback)");
  const c10::Device dev = means3D.device();
  const c10::cuda::CUDAGuard guard(dev);
  const int64_t P = means3D.size(0);
  Tensor present_ = torch::zeros({P}, torch::TensorOptions().dtype(torch::kUInt8).device(dev));
  if (P) {
    Tensor p = f32c(means3D, dev), v = f32c(viewmatrix, dev);
    check(gsr_mark_visible((int32_t)P, fp(p), fp(v), present_.data_ptr<uint8_t>(),
                           (void*)at::cuda::getCurrentCUDAStream(dev.index()).stream()),
          "gsr_mark_visible");
  }
  return present_.to(torch::kBool);
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "thin torch binding of libgsr_b200.so (include/gsr_b200.h); upstream diff_gaussian_rasterization._C signatures";
  m.def("rasterize_gaussians", &rasterize_gaussians);
  m.def("rasterize_gaussians_backward", &rasterize_gaussians_backward);
  m.def("mark_visible", &mark_visible, pybind11::arg("means3D"), pybind11::arg("viewmatrix"),
        pybind11::arg("projmatrix") = pybind11::none());
  m.def("rasterize_gaussians_ex", &rasterize_gaussians_ex);
  m.def("rasterize_gaussians_backward_ex", &rasterize_gaussians_backward_ex);
  m.def("render_backward_only", &render_backward_only);
  m.def("preprocess_backward_views", &preprocess_backward_views);
  m.def("version", [] { return std::string(gsr_version()); });
}
