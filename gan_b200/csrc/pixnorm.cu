// pixnorm.cu — pixelwise feature-vector normalisation (ProGAN), the `Pixnorm` layer of the reference's generator_11
// (generators/generator_11.py:9-16): y[n, c, p] = x[n, c, p] / sqrt(eps + mean_c x[n, c, p]^2), fp32 NCHW.
// HBM-bound: one thread per pixel walks the channels (consecutive threads = consecutive pixels, so every channel step is
// a coalesced row); the forward reads x twice (the second pass hits L2 for the plane sizes of the model) and writes y
// and the per-pixel 1 / norm, which the backward reuses:  gx = r * (gy - y * mean_c(gy * y)).
#include "common.cuh"
#include <algorithm>

namespace ganb200 {

__global__ void __launch_bounds__(256) pixnorm_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                          float* __restrict__ rnorm, int64_t pixels, int C, int64_t HW,
                                                          float eps) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < pixels; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / HW, p = i - n * HW;
    const float* xp = x + n * C * HW + p;
    float s = 0.f;
    for (int c = 0; c < C; ++c) { const float v = __ldg(xp + c * HW); s = fmaf(v, v, s); }
    const float r = 1.f / sqrtf(eps + s / (float)C);
    rnorm[i] = r;
    float* yp = y + n * C * HW + p;
    for (int c = 0; c < C; ++c) yp[c * HW] = __ldg(xp + c * HW) * r;
  }
}

__global__ void __launch_bounds__(256) pixnorm_bwd_kernel(const float* __restrict__ gy, const float* __restrict__ y,
                                                          const float* __restrict__ rnorm, float* __restrict__ gx,
                                                          int64_t pixels, int C, int64_t HW) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < pixels; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / HW, p = i - n * HW;
    const int64_t base = n * C * HW + p;
    float d = 0.f;
    for (int c = 0; c < C; ++c) d = fmaf(__ldg(gy + base + c * HW), __ldg(y + base + c * HW), d);
    d /= (float)C;
    const float r = rnorm[i];
    for (int c = 0; c < C; ++c) gx[base + c * HW] = r * (__ldg(gy + base + c * HW) - __ldg(y + base + c * HW) * d);
  }
}

static int pix_grid(int64_t pixels) {
  return (int)std::min<int64_t>((pixels + 255) / 256, (int64_t)kNumSMs * 8);
}

}  // namespace ganb200
using namespace ganb200;

extern "C" int ganb200_pixnorm_fwd(const float* x, float* y, float* rnorm, int N, int C, int64_t HW, float eps,
                                   void* stream) {
  if ((int64_t)N * C * HW == 0) return 0;
  GB_REQUIRE(x && y && rnorm && N > 0 && C > 0 && HW > 0, GANB200_BAD_ARG, "pixnorm_fwd: bad argument");
  const int64_t pixels = (int64_t)N * HW;
  pixnorm_fwd_kernel<<<pix_grid(pixels), 256, 0, (cudaStream_t)stream>>>(x, y, rnorm, pixels, C, HW, eps);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_pixnorm_bwd(const float* gy, const float* y, const float* rnorm, float* gx, int N, int C,
                                   int64_t HW, void* stream) {
  if ((int64_t)N * C * HW == 0) return 0;
  GB_REQUIRE(gy && y && rnorm && gx && N > 0 && C > 0 && HW > 0, GANB200_BAD_ARG, "pixnorm_bwd: bad argument");
  const int64_t pixels = (int64_t)N * HW;
  pixnorm_bwd_kernel<<<pix_grid(pixels), 256, 0, (cudaStream_t)stream>>>(gy, y, rnorm, gx, pixels, C, HW);
  GB_CHECK_LAUNCH();
  return 0;
}
