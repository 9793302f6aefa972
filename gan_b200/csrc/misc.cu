// misc.cu — channel-slice copies (cat / slice / zero-pad along C without ATen), device counters, and the
// spectral-norm power iteration.  All HBM-bound; algorithmic bytes = 4 B read + 4 B written per element.
#include "common.cuh"

namespace ganb200 {

// dst[n, dst_c0 + c, hw] (+)= src[n, src_c0 + c, hw] for c < ncopy
template <bool ACC>
__global__ void __launch_bounds__(256) chan_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, int N,
                                                        int src_C, int src_c0, int dst_C, int dst_c0, int ncopy,
                                                        int HW, bool vec) {
  if (vec) {
    int hw4 = HW >> 2;
    int64_t total = (int64_t)N * ncopy * hw4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
      int j = (int)(i % hw4);
      int64_t r = i / hw4;
      int c = (int)(r % ncopy), n = (int)(r / ncopy);
      float4 v = reinterpret_cast<const float4*>(src + ((int64_t)n * src_C + src_c0 + c) * HW)[j];
      float4* d = reinterpret_cast<float4*>(dst + ((int64_t)n * dst_C + dst_c0 + c) * HW) + j;
      if (ACC) { float4 o = *d; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
      *d = v;
    }
  } else {
    int64_t total = (int64_t)N * ncopy * HW;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
      int j = (int)(i % HW);
      int64_t r = i / HW;
      int c = (int)(r % ncopy), n = (int)(r / ncopy);
      float v = src[((int64_t)n * src_C + src_c0 + c) * HW + j];
      float* d = dst + ((int64_t)n * dst_C + dst_c0 + c) * HW + j;
      *d = ACC ? *d + v : v;
    }
  }
}

__global__ void counter_add_kernel(int64_t* ctr, int64_t inc) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *ctr += inc;
}

// One CTA per weight matrix would be enough for D7's small convs; here one CTA runs the whole iteration for one
// matrix: v = normalize(W^T u); u = normalize(W v); sigma = u . (W v).  rows x cols is at most a few 10^5 elements.
__global__ void __launch_bounds__(256) spectral_power_iter_kernel(const float* __restrict__ W, float* __restrict__ u,
                                                                  float* __restrict__ v, float* __restrict__ sigma,
                                                                  int rows, int cols, int n_iter, float eps) {
  extern __shared__ float sm[];  // [cols] for v-candidate, [rows] for u-candidate
  float* sv = sm;
  float* su = sm + cols;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < rows; i += nt) su[i] = u[i];
  for (int i = tid; i < cols; i += nt) sv[i] = v[i];
  __syncthreads();
  for (int it = 0; it < n_iter; ++it) {
    // v = W^T u : column sums, coalesced over columns
    float nrm = 0.f;
    for (int c = tid; c < cols; c += nt) {
      float acc = 0.f;
      for (int r = 0; r < rows; ++r) acc = fmaf(W[(int64_t)r * cols + c], su[r], acc);
      sv[c] = acc;
      nrm = fmaf(acc, acc, nrm);
    }
    float tot = block_sum<256>(nrm);
    float inv = 1.f / fmaxf(sqrtf(tot), eps);
    for (int c = tid; c < cols; c += nt) sv[c] *= inv;
    __syncthreads();
    // u = W v : one warp per row
    int lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    for (int r = wid; r < rows; r += nw) {
      float acc = 0.f;
      for (int c = lane; c < cols; c += 32) acc = fmaf(W[(int64_t)r * cols + c], sv[c], acc);
      acc = warp_sum(acc);
      if (lane == 0) su[r] = acc;
    }
    __syncthreads();
    float n2 = 0.f;
    for (int r = tid; r < rows; r += nt) n2 = fmaf(su[r], su[r], n2);
    float tot2 = block_sum<256>(n2);
    float inv2 = 1.f / fmaxf(sqrtf(tot2), eps);
    for (int r = tid; r < rows; r += nt) su[r] *= inv2;
    __syncthreads();
  }
  // sigma = u . (W v)
  float part = 0.f;
  {
    int lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    for (int r = wid; r < rows; r += nw) {
      float acc = 0.f;
      for (int c = lane; c < cols; c += 32) acc = fmaf(W[(int64_t)r * cols + c], sv[c], acc);
      acc = warp_sum(acc);
      if (lane == 0) part = fmaf(acc, su[r], part);
    }
  }
  float s = block_sum<256>(part);
  if (tid == 0) sigma[0] = s;
  for (int i = tid; i < rows; i += nt) u[i] = su[i];
  for (int i = tid; i < cols; i += nt) v[i] = sv[i];
}

}  // namespace ganb200
using namespace ganb200;

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

extern "C" int ganb200_spectral_power_iter(const float* W, float* u, float* v, float* sigma, int rows, int cols,
                                           int n_iter, float eps, void* stream) {
  GB_REQUIRE(rows > 0 && cols > 0 && n_iter >= 0, GANB200_BAD_SHAPE, "spectral_power_iter: bad shape");
  size_t smem = sizeof(float) * ((size_t)rows + cols);
  GB_REQUIRE(smem <= 200 * 1024, GANB200_BAD_SHAPE, "spectral_power_iter: rows+cols too large for one CTA");
  if (smem > 48 * 1024)
    GB_CHECK_CUDA(cudaFuncSetAttribute(spectral_power_iter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  spectral_power_iter_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(W, u, v, sigma, rows, cols, n_iter, eps);
  GB_CHECK_LAUNCH();
  return 0;
}
