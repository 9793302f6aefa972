// group.cu — grouped launches for families of small independent layers: ONE launch per layer KIND instead of one per layer.
//
// generator_13_5 owns 519 modulated convolutions and each has its own style MLP
//   to_style = [EqualizedLinear(256,256), BatchNorm1d(256), PReLU(256)], EqualizedLinear(256,Cin), BatchNorm1d(Cin)
// (generators/generator_13_5.py:205-216, 223-227) evaluated on the SAME input w (B x 256).  Layer by layer that is
// 519 x (2 linear + 2 batch-norm) launches forward, twice that backward — ~12 000 launches of 8-9 us per iteration
// (profiles/r02_kernel_time_cupti_b256_start.txt: linear_small 71 ms + bn_small 36 ms of 1 412 ms), each far too small
// to fill 148 SMs.  Here a device table of jobs describes the whole family and
//   linear_group_kernel            : out_g = alpha_g * A_g B_g^T (+ bias_g), any mix of sizes (ragged N), fp32
//   bn1d_group_{stats,apply}       : training-mode BatchNorm1d (+ the following per-channel PReLU), running statistics
//   bn1d_group_bwd_{reduce,apply}  : its backward (+ PReLU slope gradient, + the bias gradient of the linear before it)
// run every member in one launch.  The statistics and the gradient sums are separate kernels from the apply passes so that
// the data-parallel exchange (sync-BN, gan_b200/dist.py) sits between them as ONE all-reduce for the whole family.
// All sums run in a fixed order: results do not depend on the launch (deterministic).
#include "common.cuh"

namespace ganb200 {

constexpr int GT = 32;        // output tile edge
constexpr int GR = 128;       // reduction chunk per shared-memory round trip
constexpr int GP = GT + 2;    // padded row

template <class Job>
__device__ __forceinline__ int find_job(const Job* __restrict__ jobs, int n_jobs, int blk) {
  int lo = 0, hi = n_jobs - 1;   // last job whose first_block <= blk
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].first_block <= blk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// 32 x 128 elements of one operand -> S[r][i]; element (i, r) lives at p[r*ld + i] (red_is_row) or p[i*ld + r].
// All loads are issued before the first shared-memory store (see conv_linear.cu: lin_stage).
__device__ __forceinline__ void g_stage(const float* __restrict__ p, int ld, int red_is_row, float (*S)[GP], int i0,
                                        int r0, int ni, int nr) {
  const int la = threadIdx.x & 31, lb = threadIdx.x >> 5;
  constexpr int NV = GR * GT / 256;
  float v[NV];
#pragma unroll
  for (int e = 0; e < NV; ++e) {
    const int i = red_is_row ? i0 + la : i0 + lb + 8 * (e >> 2);
    const int r = red_is_row ? r0 + lb + 8 * e : r0 + la + 32 * (e & 3);
    const bool ok = i < ni && r < nr;
    const int64_t idx = ok ? (red_is_row ? (int64_t)r * ld + i : (int64_t)i * ld + r) : 0;
    v[e] = __ldg(p + idx);
    if (!ok) v[e] = 0.f;
  }
#pragma unroll
  for (int e = 0; e < NV; ++e) {
    if (red_is_row) S[lb + 8 * e][la] = v[e];
    else S[la + 32 * (e & 3)][lb + 8 * (e >> 2)] = v[e];
  }
}

__global__ void __launch_bounds__(256) linear_group_kernel(const ganb200_glin_job* __restrict__ jobs, int n_jobs,
                                                           const float* __restrict__ a_base,
                                                           const float* __restrict__ b_base, float* __restrict__ o_base) {
  __shared__ __align__(16) float As[GR][GP];
  __shared__ __align__(16) float Bs[GR][GP];
  __shared__ ganb200_glin_job J;
  if (threadIdx.x == 0) J = jobs[find_job(jobs, n_jobs, (int)blockIdx.x)];
  __syncthreads();
  const int tiles_j = (J.nj + GT - 1) / GT;
  const int t = (int)blockIdx.x - J.first_block;
  const int i0 = (t / tiles_j) * GT, j0 = (t % tiles_j) * GT;
  const float* A = J.a ? J.a : a_base + J.a_off;
  const float* B = J.b ? J.b : b_base + J.b_off;
  float* out = J.out ? J.out : o_base + J.out_off;
  const int ni = J.ni, nj = J.nj, nr = J.nr;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc00 = 0.f, acc01 = 0.f, acc10 = 0.f, acc11 = 0.f;
  for (int r0 = 0; r0 < nr; r0 += GR) {
    g_stage(A, J.lda, J.a_red_is_row, As, i0, r0, ni, nr);
    g_stage(B, J.ldb, J.b_red_is_row, Bs, j0, r0, nj, nr);
    __syncthreads();
    const int rn = min(GR, nr - r0);
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
  const float* bias = J.bias;
  const float alpha = J.alpha;
  const int ldo = J.ldo;
#pragma unroll
  for (int di = 0; di < 2; ++di)
#pragma unroll
    for (int dj = 0; dj < 2; ++dj) {
      const int i = i0 + 2 * ty + di, j = j0 + 2 * tx + dj;
      if (i < ni && j < nj) {
        float v = accs[di][dj] * alpha;
        if (bias) v += __ldg(bias + j);
        out[(int64_t)i * ldo + j] = v;
      }
    }
}

// ------------------------------------------------------------------------------------------------ BatchNorm1d family
// One CTA = one job x 32 channels; 8 warps walk the N rows (warp w: rows w, w+8, ...), lane = channel, so every global
// access is a 128-byte row segment.  Column sums go through shared memory in a fixed order.
struct BnTile {
  const float* gamma; const float* beta; const float* slope;
  float* rm; float* rv; int64_t* nbt;
  int64_t x_off, stat_off;
  int C, c, tile;
  bool ok;
};

__device__ __forceinline__ BnTile bn_tile(const ganb200_gbn_job* __restrict__ jobs, int n_jobs, ganb200_gbn_job* Js) {
  if (threadIdx.x == 0) *Js = jobs[find_job(jobs, n_jobs, (int)blockIdx.x)];
  __syncthreads();
  BnTile t;
  t.gamma = Js->gamma; t.beta = Js->beta; t.slope = Js->slope;
  t.rm = Js->running_mean; t.rv = Js->running_var; t.nbt = Js->num_batches_tracked;
  t.x_off = Js->x_off; t.stat_off = Js->stat_off; t.C = Js->C;
  t.tile = (int)blockIdx.x - Js->first_block;
  t.c = t.tile * 32 + (threadIdx.x & 31);
  t.ok = t.c < t.C;
  return t;
}

// sum over the 8 warps, result valid for every thread of lane l (broadcast through shared memory)
__device__ __forceinline__ float col_sum(float v, float (*red)[33]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  red[warp][lane] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) t += red[k][lane];
  return t;
}

__global__ void __launch_bounds__(256) bn1d_group_stats_kernel(const ganb200_gbn_job* __restrict__ jobs, int n_jobs,
                                                               const float* __restrict__ x_base, int N,
                                                               float* __restrict__ mean_out, float* __restrict__ var_out) {
  __shared__ ganb200_gbn_job Js;
  __shared__ float red[8][33];
  const BnTile t = bn_tile(jobs, n_jobs, &Js);
  const int warp = threadIdx.x >> 5;
  const float* x = x_base + t.x_off + t.c;
  float s = 0.f;
  if (t.ok)
    for (int n = warp; n < N; n += 8) s += __ldg(x + (int64_t)n * t.C);
  const float mean = col_sum(s, red) / (float)N;
  float q = 0.f;
  if (t.ok)
    for (int n = warp; n < N; n += 8) {
      const float d = __ldg(x + (int64_t)n * t.C) - mean;
      q = fmaf(d, d, q);
    }
  const float var = col_sum(q, red) / (float)N;
  if (warp == 0 && t.ok) {
    mean_out[t.stat_off + t.c] = mean;
    var_out[t.stat_off + t.c] = var;
  }
}

// y = PReLU((x - mean) * invstd * gamma + beta); running statistics with the GLOBAL count (data parallel: N * world)
__global__ void __launch_bounds__(256) bn1d_group_apply_kernel(const ganb200_gbn_job* __restrict__ jobs, int n_jobs,
                                                               const float* __restrict__ x_base, float* __restrict__ y_base,
                                                               int N, const float* __restrict__ mean,
                                                               const float* __restrict__ var, float* __restrict__ invstd_out,
                                                               float eps, float momentum, int64_t count_total) {
  __shared__ ganb200_gbn_job Js;
  const BnTile t = bn_tile(jobs, n_jobs, &Js);
  const int warp = threadIdx.x >> 5;
  if (t.tile == 0 && threadIdx.x == 0 && t.nbt) *t.nbt += 1;
  if (!t.ok) return;
  const float m = mean[t.stat_off + t.c], v = var[t.stat_off + t.c];
  const float is = 1.f / sqrtf(v + eps);
  const float gm = t.gamma ? t.gamma[t.c] : 1.f, bt = t.beta ? t.beta[t.c] : 0.f;
  const float sl = t.slope ? t.slope[t.c] : 1.f;
  const float* x = x_base + t.x_off + t.c;
  float* y = y_base + t.x_off + t.c;
  for (int n = warp; n < N; n += 8) {
    const float u = (__ldg(x + (int64_t)n * t.C) - m) * is * gm + bt;
    y[(int64_t)n * t.C] = u > 0.f ? u : sl * u;
  }
  if (warp == 0) {
    invstd_out[t.stat_off + t.c] = is;
    const float cnt = (float)count_total;
    if (t.rm) t.rm[t.c] = (1.f - momentum) * t.rm[t.c] + momentum * m;
    if (t.rv) t.rv[t.c] = (1.f - momentum) * t.rv[t.c] + momentum * (count_total > 1 ? v * cnt / (cnt - 1.f) : v);
  }
}

// gu = gy * (u > 0 ? 1 : slope);  s1 = sum_n gu;  s2 = sum_n gu * xhat;  gslope = sum_n gy * min(u, 0)
__global__ void __launch_bounds__(256) bn1d_group_bwd_reduce_kernel(const ganb200_gbn_job* __restrict__ jobs, int n_jobs,
                                                                    const float* __restrict__ gy_base,
                                                                    const float* __restrict__ x_base, int N,
                                                                    const float* __restrict__ mean,
                                                                    const float* __restrict__ invstd,
                                                                    float* __restrict__ s1_out, float* __restrict__ s2_out,
                                                                    float* __restrict__ gslope_out) {
  __shared__ ganb200_gbn_job Js;
  __shared__ float red[8][33];
  const BnTile t = bn_tile(jobs, n_jobs, &Js);
  const int warp = threadIdx.x >> 5;
  float s1 = 0.f, s2 = 0.f, ps = 0.f;
  const bool has_slope = t.slope != nullptr;
  if (t.ok) {
    const float m = mean[t.stat_off + t.c], is = invstd[t.stat_off + t.c];
    const float gm = t.gamma ? t.gamma[t.c] : 1.f, bt = t.beta ? t.beta[t.c] : 0.f;
    const float sl = has_slope ? t.slope[t.c] : 1.f;
    const float* x = x_base + t.x_off + t.c;
    const float* gy = gy_base + t.x_off + t.c;
    for (int n = warp; n < N; n += 8) {
      const float xh = (__ldg(x + (int64_t)n * t.C) - m) * is;
      float g = __ldg(gy + (int64_t)n * t.C);
      if (has_slope) {
        const float u = xh * gm + bt;
        if (!(u > 0.f)) { ps = fmaf(g, u, ps); g *= sl; }
      }
      s1 += g;
      s2 = fmaf(g, xh, s2);
    }
  }
  const float a = col_sum(s1, red), b = col_sum(s2, red);
  const float c = has_slope ? col_sum(ps, red) : 0.f;
  if (warp == 0 && t.ok) {
    s1_out[t.stat_off + t.c] = a;
    s2_out[t.stat_off + t.c] = b;
    if (has_slope && gslope_out) gslope_out[t.stat_off + t.c] = c;
  }
}

// gx = gamma * invstd * (gu - s1/cnt - xhat * s2/cnt) with the GLOBAL sums;  gbias = sum_n gx (the gradient of the bias of
// the linear layer that feeds this norm: analytically zero, numerically the rounding noise the reference also computes)
__global__ void __launch_bounds__(256) bn1d_group_bwd_apply_kernel(const ganb200_gbn_job* __restrict__ jobs, int n_jobs,
                                                                   const float* __restrict__ gy_base,
                                                                   const float* __restrict__ x_base,
                                                                   float* __restrict__ gx_base, int N,
                                                                   const float* __restrict__ mean,
                                                                   const float* __restrict__ invstd,
                                                                   const float* __restrict__ s1, const float* __restrict__ s2,
                                                                   int64_t count_total, float* __restrict__ gbias_out) {
  __shared__ ganb200_gbn_job Js;
  __shared__ float red[8][33];
  const BnTile t = bn_tile(jobs, n_jobs, &Js);
  const int warp = threadIdx.x >> 5;
  float sb = 0.f;
  if (t.ok) {
    const bool has_slope = t.slope != nullptr;
    const float m = mean[t.stat_off + t.c], is = invstd[t.stat_off + t.c];
    const float gm = t.gamma ? t.gamma[t.c] : 1.f, bt = t.beta ? t.beta[t.c] : 0.f;
    const float sl = has_slope ? t.slope[t.c] : 1.f;
    const float inv_cnt = 1.f / (float)count_total;
    const float a = s1[t.stat_off + t.c] * inv_cnt, b = s2[t.stat_off + t.c] * inv_cnt;
    const float k = gm * is;
    const float* x = x_base + t.x_off + t.c;
    const float* gy = gy_base + t.x_off + t.c;
    float* gx = gx_base + t.x_off + t.c;
    for (int n = warp; n < N; n += 8) {
      const float xh = (__ldg(x + (int64_t)n * t.C) - m) * is;
      float g = __ldg(gy + (int64_t)n * t.C);
      if (has_slope) {
        const float u = xh * gm + bt;
        if (!(u > 0.f)) g *= sl;
      }
      const float o = k * (g - a - xh * b);
      gx[(int64_t)n * t.C] = o;
      sb += o;
    }
  }
  if (gbias_out) {
    const float tot = col_sum(sb, red);
    if (warp == 0 && t.ok) gbias_out[t.stat_off + t.c] = tot;
  }
}

}  // namespace ganb200
using namespace ganb200;

extern "C" int ganb200_linear_group(const ganb200_glin_job* jobs, int n_jobs, int n_blocks, const float* a_base,
                                    const float* b_base, float* out_base, void* stream) {
  if (n_jobs <= 0 || n_blocks <= 0) return 0;
  GB_REQUIRE(jobs != nullptr, GANB200_BAD_ARG, "linear_group: null job table");
  linear_group_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(jobs, n_jobs, a_base, b_base, out_base);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn1d_group_stats(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* x_base, int N,
                                        float* mean, float* var, void* stream) {
  if (n_jobs <= 0 || n_blocks <= 0) return 0;
  GB_REQUIRE(jobs && N >= 1, GANB200_BAD_ARG, "bn1d_group_stats: bad arguments");
  bn1d_group_stats_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(jobs, n_jobs, x_base, N, mean, var);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn1d_group_apply(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* x_base,
                                        float* y_base, int N, const float* mean, const float* var, float* invstd, float eps,
                                        float momentum, int64_t count_total, void* stream) {
  if (n_jobs <= 0 || n_blocks <= 0) return 0;
  GB_REQUIRE(jobs && N >= 1 && count_total >= 1, GANB200_BAD_ARG, "bn1d_group_apply: bad arguments");
  bn1d_group_apply_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(jobs, n_jobs, x_base, y_base, N, mean, var, invstd,
                                                                      eps, momentum, count_total);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn1d_group_bwd_reduce(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* gy_base,
                                             const float* x_base, int N, const float* mean, const float* invstd, float* s1,
                                             float* s2, float* gslope, void* stream) {
  if (n_jobs <= 0 || n_blocks <= 0) return 0;
  GB_REQUIRE(jobs && N >= 1, GANB200_BAD_ARG, "bn1d_group_bwd_reduce: bad arguments");
  bn1d_group_bwd_reduce_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(jobs, n_jobs, gy_base, x_base, N, mean, invstd,
                                                                           s1, s2, gslope);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn1d_group_bwd_apply(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* gy_base,
                                            const float* x_base, float* gx_base, int N, const float* mean,
                                            const float* invstd, const float* s1, const float* s2, int64_t count_total,
                                            float* gbias, void* stream) {
  if (n_jobs <= 0 || n_blocks <= 0) return 0;
  GB_REQUIRE(jobs && N >= 1 && count_total >= 1, GANB200_BAD_ARG, "bn1d_group_bwd_apply: bad arguments");
  bn1d_group_bwd_apply_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(jobs, n_jobs, gy_base, x_base, gx_base, N, mean,
                                                                          invstd, s1, s2, count_total, gbias);
  GB_CHECK_LAUNCH();
  return 0;
}
