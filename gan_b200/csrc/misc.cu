// misc.cu — channel-slice copies (cat / slice / zero-pad along C without ATen), device counters, and the
// spectral-norm power iteration.  All HBM-bound; algorithmic bytes = 4 B read + 4 B written per element.
#include "common.cuh"
#include "det_reduce.cuh"

namespace ganb200 {

// dst[n, dst_c0 + c, hw] (+)= src[n, src_c0 + c, hw] for c < ncopy
template <bool ACC>
__global__ void __launch_bounds__(256) chan_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, int N,
                                                        int src_C, int src_c0, int dst_C, int dst_c0, int ncopy,
                                                        int HW, bool vec) {
  // the (n, c) pair of a source row: consecutive copied rows are contiguous in neither tensor, only within a row
  if (vec) {
    const int hw4 = HW >> 2;
    grid_stride((int64_t)N * ncopy * hw4, [&](auto i) {
      const int j = (int)(i % hw4);
      const auto r = i / hw4;
      const int c = (int)(r % ncopy), n = (int)(r / ncopy);
      float4 v = __ldg(reinterpret_cast<const float4*>(src + ((int64_t)n * src_C + src_c0 + c) * HW) + j);
      float4* d = reinterpret_cast<float4*>(dst + ((int64_t)n * dst_C + dst_c0 + c) * HW) + j;
      if (ACC) { float4 o = *d; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
      *d = v;
    });
  } else {
    grid_stride((int64_t)N * ncopy * HW, [&](auto i) {
      const int j = (int)(i % HW);
      const auto r = i / HW;
      const int c = (int)(r % ncopy), n = (int)(r / ncopy);
      float v = src[((int64_t)n * src_C + src_c0 + c) * HW + j];
      float* d = dst + ((int64_t)n * dst_C + dst_c0 + c) * HW + j;
      *d = ACC ? *d + v : v;
    });
  }
}

__global__ void counter_add_kernel(int64_t* ctr, int64_t inc) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *ctr += inc;
}

// Spectral-norm power iteration, torch.nn.utils.parametrizations._SpectralNorm._power_method order:
//   u = normalize(W v);  v = normalize(W^T u);  (n_iter times)   sigma = u . (W v)
// One CTA per weight matrix (blockIdx.x indexes a table of matrices so that all the spectrally-normalised convs of a
// model advance in ONE launch); u and v live in shared memory for the whole iteration.
struct SpectralJob {
  const float* W;
  float* u;
  float* v;
  float* sigma;
  int rows, cols;
};

__device__ __forceinline__ void spectral_Wv(const float* __restrict__ W, const float* sv, float* su, int rows, int cols,
                                            int tid, int nt) {
  int lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
  for (int r = wid; r < rows; r += nw) {  // one warp per row, coalesced along the row
    float acc = 0.f;
    for (int c = lane; c < cols; c += 32) acc = fmaf(W[(int64_t)r * cols + c], sv[c], acc);
    acc = warp_sum(acc);
    if (lane == 0) su[r] = acc;
  }
}

__device__ __forceinline__ void spectral_normalize(float* x, int n, float eps, int tid, int nt) {
  float n2 = 0.f;
  for (int i = tid; i < n; i += nt) n2 = fmaf(x[i], x[i], n2);
  float tot = block_sum<256>(n2);
  float inv = 1.f / fmaxf(sqrtf(tot), eps);
  for (int i = tid; i < n; i += nt) x[i] *= inv;
  __syncthreads();
}

__device__ void spectral_power_iter_body(const SpectralJob& job, int n_iter, float eps, float* sm) {
  const float* __restrict__ W = job.W;
  const int rows = job.rows, cols = job.cols;
  float* sv = sm;          // [cols]
  float* su = sm + cols;   // [rows]
  float* st = su + rows;   // [rows] scratch for W v of the final sigma
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < rows; i += nt) su[i] = job.u[i];
  for (int i = tid; i < cols; i += nt) sv[i] = job.v[i];
  __syncthreads();
  for (int it = 0; it < n_iter; ++it) {
    spectral_Wv(W, sv, su, rows, cols, tid, nt);
    __syncthreads();
    spectral_normalize(su, rows, eps, tid, nt);
    for (int c = tid; c < cols; c += nt) {  // v = W^T u: thread per column, coalesced across the warp
      float acc = 0.f;
      for (int r = 0; r < rows; ++r) acc = fmaf(W[(int64_t)r * cols + c], su[r], acc);
      sv[c] = acc;
    }
    __syncthreads();
    spectral_normalize(sv, cols, eps, tid, nt);
  }
  spectral_Wv(W, sv, st, rows, cols, tid, nt);
  __syncthreads();
  float part = 0.f;
  for (int r = tid; r < rows; r += nt) part = fmaf(st[r], su[r], part);
  float s = block_sum<256>(part);
  if (tid == 0) job.sigma[0] = s;
  for (int i = tid; i < rows; i += nt) job.u[i] = su[i];
  for (int i = tid; i < cols; i += nt) job.v[i] = sv[i];
}

__global__ void __launch_bounds__(256) spectral_power_iter_kernel(SpectralJob job, int n_iter, float eps) {
  extern __shared__ float sm[];
  spectral_power_iter_body(job, n_iter, eps, sm);
}

__global__ void __launch_bounds__(256) spectral_power_iter_batched_kernel(const SpectralJob* __restrict__ jobs,
                                                                          int n_iter, float eps) {
  extern __shared__ float sm[];
  SpectralJob job = jobs[blockIdx.x];
  spectral_power_iter_body(job, n_iter, eps, sm);
}

// out = W / sigma[0]   (the parametrised weight)
__global__ void __launch_bounds__(256) spectral_scale_kernel(const float* __restrict__ W, const float* __restrict__ sigma,
                                                             float* __restrict__ out, int64_t n) {
  float inv = 1.f / sigma[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = W[i] * inv;
}

// dot[0] = <G, W>   (det_reduce.cuh: ordered sum of the CTA partials)
__global__ void __launch_bounds__(256) spectral_dot_kernel(const float* __restrict__ G, const float* __restrict__ W,
                                                           float* __restrict__ dot, int64_t n, int region) {
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    acc = fmaf(G[i], W[i], acc);
  float t = block_sum<256>(acc);
  det_group_store<256, 1>(region, 0, (int)blockIdx.x, (int)gridDim.x, 1.f, t, dot);
}

// gW[r,c] = G[r,c]/sigma - dot/sigma^2 * u[r] v[c]     (backward of W/sigma with sigma = u^T W v, u and v constants)
__global__ void __launch_bounds__(256) spectral_scale_bwd_kernel(const float* __restrict__ G, const float* __restrict__ u,
                                                                 const float* __restrict__ v,
                                                                 const float* __restrict__ sigma,
                                                                 const float* __restrict__ dot, float* __restrict__ gW,
                                                                 int rows, int cols) {
  float inv = 1.f / sigma[0];
  float k = dot[0] * inv * inv;
  int64_t n = (int64_t)rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int r = (int)(i / cols), c = (int)(i % cols);
    gW[i] = G[i] * inv - k * u[r] * v[c];
  }
}

// ------------------------------------------------------------------ preview grid (train/trainunits.py:52-56)
// out[0] = min(x), out[1] = max(x).  One CTA: the preview batch is at most a few MB, and one CTA keeps the result
// independent of any launch geometry.
__global__ void __launch_bounds__(1024) minmax_kernel(const float* __restrict__ x, int64_t n, float* __restrict__ out) {
  float lo = INFINITY, hi = -INFINITY;
  for (int64_t i = threadIdx.x; i < n; i += 1024) {
    const float v = x[i];
    lo = fminf(lo, v);
    hi = fmaxf(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  __shared__ float slo[32], shi[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { slo[warp] = lo; shi[warp] = hi; }
  __syncthreads();
  if (warp == 0) {
    lo = slo[lane]; hi = shi[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) { out[0] = lo; out[1] = hi; }
  }
}

// torchvision.utils.make_grid(x, nrow, padding, normalize=True) followed by save_image's quantisation, as ONE pass that
// writes the uint8 HWC image a PNG encoder takes:
//   v = (clamp(x, lo, hi) - lo) / max(hi - lo, 1e-5);  pixel = (uint8) clamp(v * 255 + 0.5, 0, 255);  padding pixels = 0.
// The arithmetic is torch's, operation by operation (fp32 subtract, IEEE divide, multiply, add — no FMA contraction, the
// divisor formed in double like the Python expression), so the bytes equal torchvision's.
__global__ void __launch_bounds__(256) make_grid_u8_kernel(const float* __restrict__ x, const float* __restrict__ mm,
                                                           unsigned char* __restrict__ out, int N, int C, int H, int W,
                                                           int xmaps, int ymaps, int pad) {
  const int cell_h = H + pad, cell_w = W + pad;
  const int GH = (N == 1) ? H : cell_h * ymaps + pad, GW = (N == 1) ? W : cell_w * xmaps + pad;
  const float lo = mm[0], hi = mm[1];
  const float div = (float)fmax((double)hi - (double)lo, 1e-5);
  const int64_t total = (int64_t)GH * GW * 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ch = (int)(i % 3);
    const int gx = (int)((i / 3) % GW), gy = (int)(i / (3 * (int64_t)GW));
    float v = 0.f;
    int n = -1, h = 0, w = 0;
    if (N == 1) { n = 0; h = gy; w = gx; }
    else {
      const int cy = gy / cell_h, cx = gx / cell_w;
      h = gy - cy * cell_h - pad; w = gx - cx * cell_w - pad;
      if (cy < ymaps && cx < xmaps && h >= 0 && w >= 0 && cy * xmaps + cx < N) n = cy * xmaps + cx;
    }
    if (n >= 0) {
      const float xv = x[(((int64_t)n * C + (C == 1 ? 0 : ch)) * H + h) * W + w];
      v = __fdiv_rn(__fsub_rn(fminf(fmaxf(xv, lo), hi), lo), div);
    }
    float q = __fadd_rn(__fmul_rn(v, 255.f), 0.5f);
    q = fminf(fmaxf(q, 0.f), 255.f);
    out[i] = (unsigned char)q;
  }
}

// ------------------------------------------------------------------ input pipeline (units/dataloader.py:5-32)
// batch[b] = cache[idx[b]] with the columns reversed when flip[b] != 0 (RandomHorizontalFlip on the cached, already
// resized + normalised image: the flip commutes with the symmetric bicubic resize and the per-channel normalisation).
// cache: [n][CHW] fp32 resident in HBM; one launch per batch, no host work.
__global__ void __launch_bounds__(256) gather_flip_kernel(const float* __restrict__ cache, const int64_t* __restrict__ idx,
                                                          const unsigned char* __restrict__ flip, float* __restrict__ out,
                                                          int B, int CH, int W) {
  const int64_t per = (int64_t)CH * W;
  grid_stride((int64_t)B * per, [&](auto i) {
    const int w = (int)(i % W);
    const auto t = i / W;
    const int ch = (int)(t % CH);
    const int b = (int)(t / CH);
    const int ws = flip[b] ? W - 1 - w : w;
    out[i] = __ldg(cache + idx[b] * per + (int64_t)ch * W + ws);
  });
}

}  // namespace ganb200
using namespace ganb200;

extern "C" int ganb200_gather_flip(const float* cache, const int64_t* idx, const unsigned char* flip, float* out, int B,
                                   int C, int H, int W, void* stream) {
  if ((int64_t)B * C * H * W <= 0) return 0;
  GB_REQUIRE(cache && idx && flip && out, GANB200_BAD_ARG, "gather_flip: null pointer");
  gather_flip_kernel<<<ew_grid((int64_t)B * C * H * W, 256, 4), 256, 0, (cudaStream_t)stream>>>(cache, idx, flip, out, B,
                                                                                             C * H, W);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_minmax(const float* x, int64_t n, float* out2, void* stream) {
  GB_REQUIRE(n > 0 && x && out2, GANB200_BAD_ARG, "minmax: empty input");
  minmax_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(x, n, out2);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_make_grid_u8(const float* x, const float* minmax, unsigned char* out, int N, int C, int H, int W,
                                    int nrow, int padding, void* stream) {
  GB_REQUIRE(N > 0 && H > 0 && W > 0 && (C == 1 || C == 3), GANB200_BAD_SHAPE, "make_grid_u8: need [N,1|3,H,W]");
  GB_REQUIRE(nrow > 0 && padding >= 0, GANB200_BAD_ARG, "make_grid_u8: nrow > 0, padding >= 0");
  const int xmaps = nrow < N ? nrow : N, ymaps = (N + xmaps - 1) / xmaps;
  const int64_t GH = N == 1 ? H : (int64_t)(H + padding) * ymaps + padding;
  const int64_t GW = N == 1 ? W : (int64_t)(W + padding) * xmaps + padding;
  make_grid_u8_kernel<<<ew_grid(GH * GW * 3, 256, 4), 256, 0, (cudaStream_t)stream>>>(x, minmax, out, N, C, H, W, xmaps,
                                                                                    ymaps, padding);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_chan_copy(const float* src, float* dst, int N, int src_C, int src_c0, int dst_C, int dst_c0,
                                 int ncopy, int HW, int accumulate, void* stream) {
  if ((int64_t)N * ncopy * HW <= 0) return 0;
  GB_REQUIRE(src_c0 >= 0 && dst_c0 >= 0 && src_c0 + ncopy <= src_C && dst_c0 + ncopy <= dst_C, GANB200_BAD_SHAPE,
             "chan_copy: slice out of range");
  bool vec = ((HW & 3) == 0) && ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  int64_t total = (int64_t)N * ncopy * (vec ? HW / 4 : HW);
  int grid = ew_grid(total, 256, 2);
  if (accumulate)
    chan_copy_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, N, src_C, src_c0, dst_C, dst_c0, ncopy, HW, vec);
  else
    chan_copy_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, N, src_C, src_c0, dst_C, dst_c0, ncopy, HW, vec);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_counter_add(int64_t* counter, int64_t inc, void* stream) {
  counter_add_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(counter, inc);
  GB_CHECK_LAUNCH();
  return 0;
}

static int spectral_smem(int rows, int cols, size_t* out) {
  size_t smem = sizeof(float) * ((size_t)2 * rows + cols);
  GB_REQUIRE(smem <= 200 * 1024, GANB200_BAD_SHAPE, "spectral_power_iter: 2*rows+cols too large for one CTA");
  *out = smem;
  return 0;
}

extern "C" int ganb200_spectral_power_iter(const float* W, float* u, float* v, float* sigma, int rows, int cols,
                                           int n_iter, float eps, void* stream) {
  GB_REQUIRE(rows > 0 && cols > 0 && n_iter >= 0, GANB200_BAD_SHAPE, "spectral_power_iter: bad shape");
  size_t smem;
  if (int rc = spectral_smem(rows, cols, &smem)) return rc;
  static bool attr = false;
  if (!attr) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(spectral_power_iter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr = true;
  }
  SpectralJob job{W, u, v, sigma, rows, cols};
  spectral_power_iter_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(job, n_iter, eps);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_spectral_power_iter_batched(const void* jobs, int n_jobs, int max_rows2_plus_cols, int n_iter,
                                                   float eps, void* stream) {
  if (n_jobs <= 0) return 0;
  size_t smem = sizeof(float) * (size_t)max_rows2_plus_cols;
  GB_REQUIRE(smem <= 200 * 1024, GANB200_BAD_SHAPE, "spectral_power_iter_batched: matrix too large for one CTA");
  static bool attr = false;
  if (!attr) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(spectral_power_iter_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       200 * 1024));
    attr = true;
  }
  spectral_power_iter_batched_kernel<<<n_jobs, 256, smem, (cudaStream_t)stream>>>((const SpectralJob*)jobs, n_iter, eps);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_spectral_scale(const float* W, const float* sigma, float* out, int64_t n, void* stream) {
  if (n <= 0) return 0;
  spectral_scale_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(W, sigma, out, n);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_spectral_scale_bwd(const float* G, const float* W, const float* u, const float* v,
                                          const float* sigma, float* dot_scratch, float* gW, int rows, int cols,
                                          void* stream) {
  int64_t n = (int64_t)rows * cols;
  if (n <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = ew_grid(n, 256, 4), region = det_region_for(s, 1, grid);
  if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(dot_scratch, 0, sizeof(float), s));
  spectral_dot_kernel<<<grid, 256, 0, s>>>(G, W, dot_scratch, n, region);
  spectral_scale_bwd_kernel<<<ew_grid(n, 256, 4), 256, 0, s>>>(G, u, v, sigma, dot_scratch, gW, rows, cols);
  GB_CHECK_LAUNCH();
  return 0;
}
