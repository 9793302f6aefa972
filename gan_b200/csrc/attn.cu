// attn.cu — the two operators of the SAGAN-style self-attention of the reference's discriminator_9 family
// (discriminators/discriminator_9_3.py:58-80: s = f^T g, beta = softmax(s, dim=1), v = h beta) that the rest of the
// library does not have: a batched fp32 matrix product with arbitrary element strides (so that every transpose of the
// forward, backward and double-backward formulas is a stride pattern, not a copy) and a softmax over the rows of a
// matrix.  CUDA-core fp32: the critic's penalties differentiate these twice and the parity bar for that path is 1e-5.
#include "common.cuh"
#include <algorithm>

namespace ganb200 {

constexpr int BMM_T = 64, BMM_K = 16;

struct BmmParams {
  const float* a; const float* b; float* c;
  int M, N, K;
  int64_t sab, sam, sak;   // A[b][m][k] = a[b * sab + m * sam + k * sak]
  int64_t sbb, sbk, sbn;   // B[b][k][n] = b[b * sbb + k * sbk + n * sbn]
  float alpha;
};

// C[b][m][n] = alpha * sum_k A[b][m][k] * B[b][k][n]; one 64 x 64 tile per block, 4 x 4 outputs per thread, k ascending
__global__ void __launch_bounds__(256) bmm_kernel(BmmParams p) {
  __shared__ float As[BMM_K][BMM_T + 1];
  __shared__ float Bs[BMM_K][BMM_T + 1];
  const int bz = blockIdx.z, m0 = blockIdx.y * BMM_T, n0 = blockIdx.x * BMM_T;
  const float* A = p.a + bz * p.sab;
  const float* B = p.b + bz * p.sbb;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // loaders: the fast index of a thread follows the unit-stride dimension of the operand where there is one
  const bool a_mfast = p.sam == 1, b_nfast = p.sbn == 1;
  for (int k0 = 0; k0 < p.K; k0 += BMM_K) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int idx = threadIdx.x + e * 256;
      const int m = a_mfast ? idx % BMM_T : idx / BMM_K, k = a_mfast ? idx / BMM_T : idx % BMM_K;
      const int gm = m0 + m, gk = k0 + k;
      As[k][m] = (gm < p.M && gk < p.K) ? A[gm * p.sam + gk * p.sak] : 0.f;
      const int n = b_nfast ? idx % BMM_T : idx / BMM_K, kb = b_nfast ? idx / BMM_T : idx % BMM_K;
      const int gn = n0 + n, gkb = k0 + kb;
      Bs[kb][n] = (gn < p.N && gkb < p.K) ? B[gkb * p.sbk + gn * p.sbn] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BMM_K; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Bs[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* C = p.c + (int64_t)bz * p.M * p.N;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn < p.N) C[(int64_t)gm * p.N + gn] = p.alpha * acc[i][j];
    }
  }
}

// y[r][:] = softmax(x[r][:]) over rows of n values; BWD: gx = y * (gy - sum(gy * y)).  One block per row.
template <bool BWD>
__global__ void __launch_bounds__(256) softmax_rows_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                           float* __restrict__ out, int64_t rows, int n) {
  __shared__ float red[32];
  __shared__ float bcast;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    const float* xr = x + r * n;
    const float* yr = BWD ? y + r * n : nullptr;
    float* o = out + r * n;
    if (!BWD) {
      float m = -INFINITY;
      for (int i = threadIdx.x; i < n; i += 256) m = fmaxf(m, xr[i]);
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
      __syncthreads();
      if (lane == 0) red[wid] = m;
      __syncthreads();
      if (threadIdx.x == 0) {
        float t = red[0];
        for (int w = 1; w < 8; ++w) t = fmaxf(t, red[w]);
        bcast = t;
      }
      __syncthreads();
      m = bcast;
      float s = 0.f;
      for (int i = threadIdx.x; i < n; i += 256) s += expf(xr[i] - m);
      s = block_sum<256>(s);
      const float inv = 1.f / s;
      for (int i = threadIdx.x; i < n; i += 256) o[i] = expf(xr[i] - m) * inv;
    } else {
      float d = 0.f;
      for (int i = threadIdx.x; i < n; i += 256) d = fmaf(xr[i], yr[i], d);   // x = gy here
      d = block_sum<256>(d);
      for (int i = threadIdx.x; i < n; i += 256) o[i] = yr[i] * (xr[i] - d);
    }
  }
}

}  // namespace ganb200
using namespace ganb200;

extern "C" int ganb200_bmm(const float* a, const float* b, float* c, int batch, int M, int N, int K, int64_t sab,
                           int64_t sam, int64_t sak, int64_t sbb, int64_t sbk, int64_t sbn, float alpha, void* stream) {
  if ((int64_t)batch * M * N == 0) return 0;
  GB_REQUIRE(a && b && c, GANB200_BAD_ARG, "bmm: null argument");
  GB_REQUIRE(batch > 0 && M > 0 && N > 0 && K > 0 && batch <= 65535, GANB200_BAD_SHAPE, "bmm: bad shape");
  BmmParams p;
  p.a = a; p.b = b; p.c = c; p.M = M; p.N = N; p.K = K;
  p.sab = sab; p.sam = sam; p.sak = sak; p.sbb = sbb; p.sbk = sbk; p.sbn = sbn; p.alpha = alpha;
  dim3 grid((N + BMM_T - 1) / BMM_T, (M + BMM_T - 1) / BMM_T, batch);
  GB_REQUIRE(grid.y <= 65535, GANB200_BAD_SHAPE, "bmm: M too large");
  bmm_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_softmax_rows_fwd(const float* x, float* y, int64_t rows, int n, void* stream) {
  if (rows * n == 0) return 0;
  GB_REQUIRE(x && y && n > 0, GANB200_BAD_ARG, "softmax_rows_fwd: bad argument");
  const int grid = (int)std::min<int64_t>(rows, (int64_t)kNumSMs * 32);
  softmax_rows_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(x, nullptr, y, rows, n);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_softmax_rows_bwd(const float* gy, const float* y, float* gx, int64_t rows, int n, void* stream) {
  if (rows * n == 0) return 0;
  GB_REQUIRE(gy && y && gx && n > 0, GANB200_BAD_ARG, "softmax_rows_bwd: bad argument");
  const int grid = (int)std::min<int64_t>(rows, (int64_t)kNumSMs * 32);
  softmax_rows_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(gy, y, gx, rows, n);
  GB_CHECK_LAUNCH();
  return 0;
}
