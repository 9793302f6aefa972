// conv_linear.cu — fp32 CUDA-core kernels for convolutions on 1x1 maps with 1x1 filters, i.e. the nn.Linear layers of
// the models (EqualizedLinear: generators/generator_13_5.py:19-26, discriminators/discriminator_9_4.py:20-27; the
// 519 per-layer style MLPs and the 12-layer mapping network are B x 256 x 256 products).
//
// One training iteration of the final pair issues ~7 000 of these with M = batch rows (256).  On the tcgen05 engine
// each costs an activation pack, a weight pack, the MMA kernel (two CTAs) and, for wgrad, a memset and a split
// reduction: ~140 ms of 4-5 us launches per iteration (CUPTI, r01).  Here every operation is ONE launch that reads the
// fp32 operands as they are (scales, alpha and bias folded in), computes in fp32 and writes the result: no workspace.
//
//   fprop  y [n,k] = alpha * osc[n,k] * sum_c w[k,c] * (isc[n,c] * x[n,c]) + bias[k]
//   dgrad  dx[n,c] = alpha * isc[n,c] * sum_k w[k,c] * (osc[n,k] * dy[n,k])
//   wgrad  dw[k,c] = alpha * sum_n (osc[n,k] * dy[n,k]) * (isc[n,c] * x[n,c])        (one CTA sums all n: deterministic)
//
// All three are out[i,j] = sum_r A(i,r) * B(j,r) on 32 x 32 output tiles, 256 threads (2 x 2 outputs each), the
// reduction staged through shared memory in chunks of 128 as As[r][i], Bs[r][j] (rows padded to 34 words).
#include "common.cuh"
#include "conv_common.cuh"

namespace ganb200 {

constexpr int LT = 32;        // output tile edge
constexpr int LR = 128;       // reduction chunk staged per round trip
constexpr int LP = LT + 2;    // padded shared-memory row (even: the 2 x 2 register tile reads 8-byte pairs)

// element (i, r) of an operand whose storage is [rows][cols] row-major with an optional same-shape scale
struct LinOperand {
  const float* p;       // data
  const float* scale;   // nullptr or same indexing as p
  int ld;               // row pitch
  int red_is_row;       // 1: reduction index r is the storage row, output index i the column; 0: the other way round
};

// 32 x 128 elements of one operand -> S[r][i]: 16 loads per thread, ALL issued before the first shared-memory store
// (unconditional loads from a clamped index, masked afterwards — with an early-out per element the compiler emitted
// load / wait / store sixteen times in a row, one global round trip each), coalesced along the storage's contiguous axis.
__device__ __forceinline__ void lin_stage(const LinOperand& o, float (*S)[LP], int i0, int r0, int ni, int nr) {
  const int la = threadIdx.x & 31, lb = threadIdx.x >> 5;
  constexpr int NV = LR * LT / 256;
  float v[NV];
  int64_t idx[NV];
  bool ok[NV];
#pragma unroll
  for (int e = 0; e < NV; ++e) {
    // red_is_row: e -> r = lb + 8e, i = la.   otherwise: e = 4q + t -> i = lb + 8q, r = la + 32t
    const int i = o.red_is_row ? i0 + la : i0 + lb + 8 * (e >> 2);
    const int r = o.red_is_row ? r0 + lb + 8 * e : r0 + la + 32 * (e & 3);
    ok[e] = i < ni && r < nr;
    idx[e] = ok[e] ? (o.red_is_row ? (int64_t)r * o.ld + i : (int64_t)i * o.ld + r) : 0;
    v[e] = __ldg(o.p + idx[e]);
  }
  if (o.scale) {
    float sc[NV];
#pragma unroll
    for (int e = 0; e < NV; ++e) sc[e] = __ldg(o.scale + idx[e]);
#pragma unroll
    for (int e = 0; e < NV; ++e) v[e] *= sc[e];
  }
#pragma unroll
  for (int e = 0; e < NV; ++e) {
    const float val = ok[e] ? v[e] : 0.f;
    if (o.red_is_row) S[lb + 8 * e][la] = val;
    else S[la + 32 * (e & 3)][lb + 8 * (e >> 2)] = val;
  }
}

// out[i][j] (row pitch ldo) = alpha * oscale(i,j) * sum_r A(i,r) B(j,r) + bias[j]
// These calls are latency-bound (64 CTAs, a few microseconds): the reduction is staged 128 deep per barrier pair so that
// a 256-long product is two global round trips, not eight.
__global__ void __launch_bounds__(256) linear_small_kernel(LinOperand A, LinOperand B, float* __restrict__ out, int ldo,
                                                           int ni, int nj, int nr, float alpha,
                                                           const float* __restrict__ oscale,
                                                           const float* __restrict__ bias) {
  __shared__ __align__(16) float As[LR][LP];
  __shared__ __align__(16) float Bs[LR][LP];
  const int i0 = blockIdx.y * LT, j0 = blockIdx.x * LT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc00 = 0.f, acc01 = 0.f, acc10 = 0.f, acc11 = 0.f;
  for (int r0 = 0; r0 < nr; r0 += LR) {
    lin_stage(A, As, i0, r0, ni, nr);
    lin_stage(B, Bs, j0, r0, nj, nr);
    __syncthreads();
    const int rn = min(LR, nr - r0);
#pragma unroll 8
    for (int r = 0; r < rn; ++r) {
      const float2 a = *reinterpret_cast<const float2*>(&As[r][2 * ty]);
      const float2 b = *reinterpret_cast<const float2*>(&Bs[r][2 * tx]);
      acc00 = fmaf(a.x, b.x, acc00); acc01 = fmaf(a.x, b.y, acc01);
      acc10 = fmaf(a.y, b.x, acc10); acc11 = fmaf(a.y, b.y, acc11);
    }
    __syncthreads();
  }
  const float accs[2][2] = {{acc00, acc01}, {acc10, acc11}};
#pragma unroll
  for (int di = 0; di < 2; ++di)
#pragma unroll
    for (int dj = 0; dj < 2; ++dj) {
      const int i = i0 + 2 * ty + di, j = j0 + 2 * tx + dj;
      if (i < ni && j < nj) {
        float v = accs[di][dj] * alpha;
        if (oscale) v *= __ldg(oscale + (int64_t)i * ldo + j);
        if (bias) v += __ldg(bias + j);
        out[(int64_t)i * ldo + j] = v;
      }
    }
}

int linear_eligible(const ganb200_conv_desc* d) {
  // up to 2^26 multiply-adds (256 x 512 x 512): beyond that (the critic's 4100 x 4100 head) the tensor cores win
  return d->H == 1 && d->W == 1 && d->R == 1 && d->S == 1 && d->P == 1 && d->Q == 1 && d->N <= 1024 &&
         (int64_t)d->N * d->C * d->K <= ((int64_t)1 << 26);
}

static int launch(const LinOperand& A, const LinOperand& B, float* out, int ldo, int ni, int nj, int nr, float alpha,
                  const float* oscale, const float* bias, cudaStream_t s) {
  dim3 grid(ceil_div(nj, LT), ceil_div(ni, LT));
  linear_small_kernel<<<grid, 256, 0, s>>>(A, B, out, ldo, ni, nj, nr, alpha, oscale, bias);
  GB_CHECK_LAUNCH();
  return 0;
}

int linear_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias, const float* iscale,
                 const float* oscale, float* y, cudaStream_t s) {
  LinOperand A{x, iscale, d->C, 0}, B{w, nullptr, d->C, 0};
  return launch(A, B, y, d->K, d->N, d->K, d->C, d->alpha, oscale, bias, s);
}

int linear_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale, const float* oscale,
                 float* dx, cudaStream_t s) {
  LinOperand A{dy, oscale, d->K, 0}, B{w, nullptr, d->C, 1};   // B(j = c, r = k) = w[k][c]
  return launch(A, B, dx, d->C, d->N, d->C, d->K, d->alpha, iscale, nullptr, s);
}

int linear_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale, const float* oscale,
                 float* dw, cudaStream_t s) {
  LinOperand A{dy, oscale, d->K, 1}, B{x, iscale, d->C, 1};    // A(i = k, r = n) = dy[n][k], B(j = c, r = n) = x[n][c]
  return launch(A, B, dw, d->C, d->K, d->C, d->N, d->alpha, nullptr, nullptr, s);
}

}  // namespace ganb200
