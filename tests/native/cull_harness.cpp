// Brute-force check of the conservative sub-tile culling test of the compositing kernels
// (gaussian_splatting_lightning_b200/csrc/render_common.cuh): for random splats and 8x4 pixel
// rectangles, a splat the test rejects must have alpha < 1/255 on every pixel of the rectangle.
// Built with g++ by tests/test_cull_host.py; prints "pairs false_rejects false_accepts accepts".
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <random>

#include "render_common.cuh"

int main(int argc, char** argv) {
  const long n = argc > 1 ? atol(argv[1]) : 1000000;
  std::mt19937_64 rng(12345);
  std::uniform_real_distribution<float> U(0.f, 1.f);
  std::normal_distribution<float> N(0.f, 1.f);
  long false_reject = 0, false_accept = 0, accepts = 0;
  for (long it = 0; it < n; ++it) {
    const int W = 64 + (int)(U(rng) * 64), H = 48 + (int)(U(rng) * 64);
    const int sx0 = 8 * (int)(U(rng) * ((W + 7) / 8)), sy0 = 4 * (int)(U(rng) * ((H + 3) / 4));
    const gsr::SubTile st = gsr::make_subtile(sx0, sy0, W, H);
    // covariance = R diag(s1^2, s2^2) R^T + 0.3 I, like the preprocess kernel produces
    const float s1 = std::exp(U(rng) * 4.5f - 1.5f), s2 = s1 * std::exp(-U(rng) * 3.f);
    const float th = U(rng) * 6.2831853f, c = std::cos(th), s = std::sin(th);
    const float ca = c * c * s1 * s1 + s * s * s2 * s2 + 0.3f, cc = s * s * s1 * s1 + c * c * s2 * s2 + 0.3f;
    const float cb = c * s * (s1 * s1 - s2 * s2);
    const float det = ca * cc - cb * cb;
    const float A = cc / det, B = -cb / det, C = ca / det;
    const float opacity = (it % 7 == 0) ? U(rng) * 0.01f : U(rng);
    const float reach = 3.5f * s1 + 6.f;
    const float mx = sx0 + 3.5f + N(rng) * reach, my = sy0 + 1.5f + N(rng) * reach;
    const float hA = A * GSR_HALF_LOG2E, hB = B * GSR_LOG2E, hC = C * GSR_HALF_LOG2E;
    const float tau = opacity > 0.f ? std::log(255.f * opacity) : -1.f;  // as preprocess.cu stores it
    const float tau2 = tau * GSR_LOG2E;
    const float4 a = make_float4(mx, my, hA, hB);
    const bool keep = gsr::subtile_may_contribute(a, hC, gsr::make_cull_plane(hA, hB, hC, tau2), st);
    bool any = false;
    for (int y = (int)st.y0; y <= (int)st.y1 && !any; ++y)
      for (int x = (int)st.x0; x <= (int)st.x1; ++x) {
        const float dx = mx - (float)x, dy = my - (float)y;
        // the kernels' rounding sequence (quad_form), then alpha in double with slack for ex2.approx
        const float t = hA * dx, u = hC * dy, w = hB * dx;
        float e = t * dx;
        e = std::fmaf(u, dy, e);
        e = std::fmaf(w, dy, e);
        if (e < 0.f) continue;
        const double alpha = std::fmin(0.99, (double)opacity * std::exp2(-(double)e));
        if (alpha >= (1.0 / 255.0) * (1.0 - 2e-6)) { any = true; break; }
      }
    if (any && !keep) ++false_reject;
    if (!any && keep) ++false_accept;
    if (keep) ++accepts;
  }
  std::printf("%ld %ld %ld %ld\n", n, false_reject, false_accept, accepts);
  return false_reject == 0 ? 0 : 1;
}
