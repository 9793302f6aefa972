// layers.cu — PReLU, BatchNorm, softmax-over-branches, MiniBatchStdDev, BCE, separable resampling.
// All HBM-bound; see DESIGN.md for the per-kernel algorithmic bytes.
#include "common.cuh"
#include "det_reduce.cuh"
#include "peer.cuh"
#include <algorithm>

namespace ganb200 {

// ------------------------------------------------------------------ PReLU family on [N,C,HW], a_n in {1, C}
// MODE 0: y = x>0 ? x : a*x          (g unused)
// MODE 1: y = g * (x>0 ? 1 : a)
template <int MODE>
__global__ void __launch_bounds__(256) prelu_kernel(const float* __restrict__ g, const float* __restrict__ x,
                                                    const float* __restrict__ a, float* __restrict__ y,
                                                    int64_t rows, int C, int HW, int a_n) {
  if (HW >= 64) {
    bool vec = ((HW & 3) == 0) && aligned16(x) && aligned16(y) && (MODE == 0 || aligned16(g));
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
      float slope = a[a_n == 1 ? 0 : (int)(row % C)];
      const float* xr = x + row * HW;
      const float* gr = MODE == 1 ? g + row * HW : nullptr;
      float* yr = y + row * HW;
      if (vec) {
        for (int i = threadIdx.x; i < (HW >> 2); i += blockDim.x) {
          float4 v = reinterpret_cast<const float4*>(xr)[i], r;
          if (MODE == 0) {
            r.x = v.x > 0.f ? v.x : slope * v.x; r.y = v.y > 0.f ? v.y : slope * v.y;
            r.z = v.z > 0.f ? v.z : slope * v.z; r.w = v.w > 0.f ? v.w : slope * v.w;
          } else {
            float4 q = reinterpret_cast<const float4*>(gr)[i];
            r.x = v.x > 0.f ? q.x : slope * q.x; r.y = v.y > 0.f ? q.y : slope * q.y;
            r.z = v.z > 0.f ? q.z : slope * q.z; r.w = v.w > 0.f ? q.w : slope * q.w;
          }
          reinterpret_cast<float4*>(yr)[i] = r;
        }
      } else {
        for (int i = threadIdx.x; i < HW; i += blockDim.x) {
          float v = xr[i];
          float src = MODE == 0 ? v : gr[i];
          yr[i] = v > 0.f ? src : slope * src;
        }
      }
    }
  } else {
    int64_t total = rows * HW;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
      float slope = a[a_n == 1 ? 0 : (int)((i / HW) % C)];
      float v = x[i];
      float src = MODE == 0 ? v : g[i];
      y[i] = v > 0.f ? src : slope * src;
    }
  }
}

// out[c or 0] += sum over (n,hw) of  MODE 0: gy*min(x,0)   MODE 1: u*v*[x<=0]
// grid (groups, nsplit): groups = a_n==1 ? 1 : C
template <int MODE>
__global__ void __launch_bounds__(256) prelu_reduce_kernel(const float* __restrict__ u, const float* __restrict__ v,
                                                           const float* __restrict__ x, float* __restrict__ out,
                                                           int N, int C, int HW, int a_n, int region) {
  float acc = 0.f;
  int split = blockIdx.y, nsplit = gridDim.y;
  if (a_n == 1) {
    int64_t len = (int64_t)N * C * HW;
    int64_t chunk = (len + nsplit - 1) / nsplit;
    int64_t lo = split * chunk, hi = lo + chunk < len ? lo + chunk : len;
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      float xv = x[i];
      if (!(xv > 0.f)) acc += MODE == 0 ? u[i] * xv : u[i] * v[i];
    }
  } else {
    int c = blockIdx.x;
    if (HW >= 1024) {
      for (int n = split; n < N; n += nsplit) {
        int64_t base = ((int64_t)n * C + c) * HW;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) {
          float xv = x[base + i];
          if (!(xv > 0.f)) acc += MODE == 0 ? u[base + i] * xv : u[base + i] * v[base + i];
        }
      }
    } else {
      // small planes (4x4 ... 7x7): threads over (n, hw) pairs so that all 256 lanes have work
      const int total = N * HW;
      for (int j = split * 256 + threadIdx.x; j < total; j += nsplit * 256) {
        const int n = j / HW, i = j - n * HW;
        const int64_t idx = ((int64_t)n * C + c) * HW + i;
        float xv = x[idx];
        if (!(xv > 0.f)) acc += MODE == 0 ? u[idx] * xv : u[idx] * v[idx];
      }
    }
  }
  float t = block_sum<256>(acc);
  const int g = a_n == 1 ? 0 : (int)blockIdx.x;
  det_group_store<256, 1>(region, g, split, nsplit, 1.f, t, out + g);
}

// Fused PReLU backward: gx = gy * (x > 0 ? 1 : a) and ga[c] += sum gy * min(x, 0) in ONE pass over (gy, x)
// (12 B/element instead of 20).  ROWS: one warp per (n,c) row, float4 (HW >= 64, HW % 4 == 0);
// otherwise one block per channel with threads over (n, hw) (small planes).  ga must be zeroed by the caller.
template <bool ROWS>
__global__ void __launch_bounds__(256) prelu_bwd_fused_kernel(const float* __restrict__ gy, const float* __restrict__ x,
                                                              const float* __restrict__ a, float* __restrict__ gx,
                                                              float* __restrict__ ga, int N, int C, int HW, int a_n,
                                                              int region) {
  if (ROWS) {
    const int lane = threadIdx.x & 31;
    const int64_t rows = (int64_t)N * C;
    const int hw4 = HW >> 2;
    // rows_fit: every (n, c) row sum goes to g_det_part[row]; the last CTA then adds them per channel in n order
    const bool rows_fit = region >= 0;
    float* part = g_det_part + (size_t)(rows_fit ? region : 0) * kDetPartials;
    unsigned int* count = g_det_count + (rows_fit ? region : 0) * kDetGroups;
    float keep = 0.f;  // !rows_fit && a_n == 1: one atomic per warp at the end
    for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < rows; row += (int64_t)gridDim.x * 8) {
      const int c = a_n == 1 ? 0 : (int)(row % C);
      const float slope = __ldg(a + c);
      const float4* g4 = reinterpret_cast<const float4*>(gy + row * HW);
      const float4* x4 = reinterpret_cast<const float4*>(x + row * HW);
      float4* o4 = reinterpret_cast<float4*>(gx + row * HW);
      float acc = 0.f;
      for (int i = lane; i < hw4; i += 32) {
        const float4 g = __ldg(g4 + i), v = __ldg(x4 + i);
        float4 r;
        r.x = v.x > 0.f ? g.x : slope * g.x; r.y = v.y > 0.f ? g.y : slope * g.y;
        r.z = v.z > 0.f ? g.z : slope * g.z; r.w = v.w > 0.f ? g.w : slope * g.w;
        o4[i] = r;
        acc += (v.x > 0.f ? 0.f : g.x * v.x) + (v.y > 0.f ? 0.f : g.y * v.y) + (v.z > 0.f ? 0.f : g.z * v.z) +
               (v.w > 0.f ? 0.f : g.w * v.w);
      }
      if (rows_fit) {
        acc = warp_sum(acc);
        if (lane == 0) part[row] = acc;
      } else if (a_n == 1) {
        keep += acc;
      } else {
        acc = warp_sum(acc);
        if (lane == 0) atomicAdd(ga + c, acc);
      }
    }
    if (rows_fit) {
      __shared__ bool last;
      __threadfence();
      __syncthreads();
      if (threadIdx.x == 0) {
        last = atomicAdd(count, 1u) == gridDim.x - 1;
        if (last) *count = 0u;
      }
      __syncthreads();
      if (last) {
        __threadfence();
        if (a_n == 1) {
          float acc = 0.f;
          for (int64_t r = threadIdx.x; r < rows; r += 256) acc += __ldcg(part + r);
          acc = block_sum<256>(acc);
          if (threadIdx.x == 0) ga[0] = acc;
        } else {
          // one warp per channel: lanes over the samples, then the warp tree — a fixed order
          for (int c = threadIdx.x >> 5; c < C; c += 8) {
            float acc = 0.f;
            for (int n = lane; n < N; n += 32) acc += __ldcg(part + (int64_t)n * C + c);
            acc = warp_sum(acc);
            if (lane == 0) ga[c] = acc;
          }
        }
      }
    } else if (a_n == 1) {
      keep = warp_sum(keep);
      if (lane == 0) atomicAdd(ga, keep);
    }
  } else {
    const int c = blockIdx.x, split = blockIdx.y, nsplit = gridDim.y;
    const float slope = __ldg(a + c);
    const int total = N * HW;
    float acc = 0.f;
    for (int j = split * 256 + threadIdx.x; j < total; j += nsplit * 256) {
      const int n = j / HW, i = j - n * HW;
      const int64_t idx = ((int64_t)n * C + c) * HW + i;
      const float g = gy[idx], v = x[idx];
      gx[idx] = v > 0.f ? g : slope * g;
      if (!(v > 0.f)) acc += g * v;
    }
    float t = block_sum<256>(acc);
    det_group_store<256, 1>(region, c, split, nsplit, 1.f, t, ga + c);
  }
}

}  // namespace ganb200
using namespace ganb200;

static int rows_grid(int64_t rows, int HW) {
  if (HW >= 64) return (int)(rows < (int64_t)kNumSMs * 16 ? rows : (int64_t)kNumSMs * 16);
  return ew_grid(rows * HW, 256, 4);
}

static void prelu_reduce_dims(int N, int C, int HW, int a_n, dim3* grid) {
  int groups = a_n == 1 ? 1 : C;
  int64_t per_group = (int64_t)N * C * HW / groups;
  int max_split = a_n == 1 ? (int)((per_group + 2047) / 2048) : (HW >= 1024 ? N : (int)(((int64_t)N * HW + 255) / 256));
  int want = (kNumSMs * 4 + groups - 1) / groups;
  int nsplit = want < max_split ? want : max_split;
  if (nsplit < 1) nsplit = 1;
  *grid = dim3(groups, nsplit);
}

extern "C" int ganb200_prelu_fwd(const float* x, const float* a, float* y, int N, int C, int HW, int a_n, void* stream) {
  GB_REQUIRE(a_n == 1 || a_n == C, GANB200_BAD_SHAPE, "prelu: a_n must be 1 or C");
  int64_t rows = (int64_t)N * C;
  if (rows * HW == 0) return 0;
  prelu_kernel<0><<<rows_grid(rows, HW), 256, 0, (cudaStream_t)stream>>>(nullptr, x, a, y, rows, C, HW, a_n);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_maskscale(const float* g, const float* x, const float* a, float* y, int N, int C, int HW, int a_n,
                                 void* stream) {
  GB_REQUIRE(a_n == 1 || a_n == C, GANB200_BAD_SHAPE, "maskscale: a_n must be 1 or C");
  int64_t rows = (int64_t)N * C;
  if (rows * HW == 0) return 0;
  prelu_kernel<1><<<rows_grid(rows, HW), 256, 0, (cudaStream_t)stream>>>(g, x, a, y, rows, C, HW, a_n);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_prelu_bwd(const float* gy, const float* x, const float* a, float* gx, float* ga, int N, int C,
                                 int HW, int a_n, void* stream) {
  GB_REQUIRE(a_n == 1 || a_n == C, GANB200_BAD_SHAPE, "prelu_bwd: a_n must be 1 or C");
  cudaStream_t s = (cudaStream_t)stream;
  int64_t rows = (int64_t)N * C;
  if (rows * HW == 0) return 0;
  if (gx && ga) {
    const bool al = ((((uintptr_t)gy) | ((uintptr_t)x) | ((uintptr_t)gx)) & 15) == 0;
    if (HW >= 64 && (HW & 3) == 0 && al) {
      const int region = rows <= kDetPartials ? det_region(s) : -1;   // per-row partials
      if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(ga, 0, sizeof(float) * a_n, s));
      int grid = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)kNumSMs * 16);
      prelu_bwd_fused_kernel<true><<<grid, 256, 0, s>>>(gy, x, a, gx, ga, N, C, HW, a_n, region);
      GB_CHECK_LAUNCH();
      return 0;
    }
    if (HW < 64 && a_n == C && (int64_t)N * HW < (int64_t)1 << 30) {
      int nsplit = std::max(1, std::min((kNumSMs * 4 + C - 1) / C, (N * HW + 255) / 256));
      const int region = det_region_for(s, C, nsplit);
      if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(ga, 0, sizeof(float) * a_n, s));
      prelu_bwd_fused_kernel<false><<<dim3(C, nsplit), 256, 0, s>>>(gy, x, a, gx, ga, N, C, HW, a_n, region);
      GB_CHECK_LAUNCH();
      return 0;
    }
  }
  if (gx) {
    prelu_kernel<1><<<rows_grid(rows, HW), 256, 0, s>>>(gy, x, a, gx, rows, C, HW, a_n);
    GB_CHECK_LAUNCH();
  }
  if (ga) {
    dim3 grid;
    prelu_reduce_dims(N, C, HW, a_n, &grid);
    const int region = det_region_for(s, grid.x, grid.y);
    if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(ga, 0, sizeof(float) * a_n, s));
    prelu_reduce_kernel<0><<<grid, 256, 0, s>>>(gy, nullptr, x, ga, N, C, HW, a_n, region);
    GB_CHECK_LAUNCH();
  }
  return 0;
}

extern "C" int ganb200_masked_dot(const float* u, const float* v, const float* x, float* out, int N, int C, int HW,
                                  int a_n, void* stream) {
  GB_REQUIRE(a_n == 1 || a_n == C, GANB200_BAD_SHAPE, "masked_dot: a_n must be 1 or C");
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid;
  prelu_reduce_dims(N, C, HW, a_n, &grid);
  const int region = det_region_for(s, grid.x, grid.y);
  if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * a_n, s));
  prelu_reduce_kernel<1><<<grid, 256, 0, s>>>(u, v, x, out, N, C, HW, a_n, region);
  GB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ BatchNorm (training mode)
namespace ganb200 {

// PASS 0: out_a[c] += sum x * inv_count                     (mean)
// PASS 1: out_a[c] += sum (x-mean[c])^2 * inv_count          (biased variance)
// PASS 2: out_a[c] += sum gy ; out_b[c] += sum gy*(x-mean)*invstd
template <int PASS>
__global__ void __launch_bounds__(256) bn_reduce_kernel(const float* __restrict__ x, const float* __restrict__ gy,
                                                        const float* __restrict__ mean, const float* __restrict__ invstd,
                                                        float* __restrict__ out_a, float* __restrict__ out_b,
                                                        int N, int C, int HW, float inv_count, int region) {
  int c = blockIdx.x, split = blockIdx.y, nsplit = gridDim.y;
  float m = PASS >= 1 ? mean[c] : 0.f;
  float is = PASS == 2 ? invstd[c] : 0.f;
  float a = 0.f, b = 0.f;
  if (HW >= 1024) {
    for (int n = split; n < N; n += nsplit) {
      int64_t base = ((int64_t)n * C + c) * HW;
      for (int i = threadIdx.x; i < HW; i += blockDim.x) {
        float xv = x[base + i];
        if (PASS == 0) a += xv;
        else if (PASS == 1) { float d = xv - m; a += d * d; }
        else { float g = gy[base + i]; a += g; b += g * (xv - m) * is; }
      }
    }
  } else {
    // small HW: threads over (n, hw) pairs
    const int total = N * HW;  // < 2^31: N * HW < N * 1024
    for (int j = split * 256 + threadIdx.x; j < total; j += nsplit * 256) {
      const int n = j / HW, i = j - n * HW;
      int64_t idx = ((int64_t)n * C + c) * HW + i;
      float xv = x[idx];
      if (PASS == 0) a += xv;
      else if (PASS == 1) { float d = xv - m; a += d * d; }
      else { float g = gy[idx]; a += g; b += g * (xv - m) * is; }
    }
  }
  float ta = block_sum<256>(a);
  if (PASS == 2) {
    float tb = block_sum<256>(b);
    det_group_store<256, 2>(region, c, split, nsplit, 1.f, ta, out_a + c, tb, out_b + c);
  } else {
    det_group_store<256, 1>(region, c, split, nsplit, inv_count, ta, out_a + c);
  }
}

// MODE 0 (forward): y = (x-mean)*invstd*gamma + beta ; row n==0 also writes invstd + running stats
// MODE 1 (backward): gx = gamma*invstd*(gy - sgy/cnt - xhat*sgx/cnt) ; row n==0 writes ggamma/gbeta
template <int MODE>
__global__ void __launch_bounds__(256) bn_apply_kernel(const float* __restrict__ x, const float* __restrict__ gy,
                                                       const float* __restrict__ mean, const float* __restrict__ var_or_invstd,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       const float* __restrict__ sgy, const float* __restrict__ sgx,
                                                       float* __restrict__ y, float* __restrict__ o0, float* __restrict__ o1,
                                                       float* __restrict__ o2, int N, int C, int HW, float eps,
                                                       float momentum, float count, int64_t* nbt) {
  if (MODE == 0 && nbt && blockIdx.x == 0 && threadIdx.x == 0) *nbt += 1;  // num_batches_tracked
  grid_stride((int64_t)N * C * HW, [&](auto i) {
    const auto row = i / HW;
    int c = (int)(row % C);
    float m = mean[c];
    float gm = gamma ? gamma[c] : 1.f;
    if (MODE == 0) {
      float is = 1.f / sqrtf(var_or_invstd[c] + eps);
      float bt = beta ? beta[c] : 0.f;
      y[i] = (x[i] - m) * is * gm + bt;
      if (row < C && (i % HW) == 0) {  // n == 0, first element of the row
        float var = var_or_invstd[c];
        if (o0) o0[c] = is;
        if (o1) o1[c] = (1.f - momentum) * o1[c] + momentum * m;
        if (o2) o2[c] = (1.f - momentum) * o2[c] + momentum * (count > 1.f ? var * count / (count - 1.f) : var);
      }
    } else {
      float is = var_or_invstd[c];
      float xhat = (x[i] - m) * is;
      float a = sgy[c] / count, b = sgx[c] / count;
      y[i] = gm * is * (gy[i] - a - xhat * b);
      if (row < C && (i % HW) == 0) {
        if (o0) o0[c] = sgx[c];
        if (o1) o1[c] = sgy[c];
      }
    }
  });
}

}  // namespace ganb200

// ---- fused training-mode batch norm for small per-channel populations (N*HW <= BN_SMALL_MAX = 8192) -------------------
// The generator runs 1 888 batch norms per forward, nearly all BatchNorm1d over (256, C) or BatchNorm2d on 5x5 / 3x3 /
// 1x1 attention maps.  The generic path is memset, memset, mean pass, variance pass, apply: five graph nodes of 1-7 us
// each (and atomics).  Here one block owns one channel, keeps its N*HW values in registers, and does mean, variance
// (two-pass, as the reference's native_batch_norm), normalisation, running statistics in ONE launch; the backward
// (sums, dx, dgamma, dbeta) likewise.  Deterministic: fixed-order block reductions, no atomics.
// Data parallel (sync-BN, `peer.bufs` set): between its statistics and its apply phase every block swaps two floats with
// the block of the same channel on every other rank over NVLink (peer_ll_sum2, peer.cuh) — (mean, var + mean^2) forward,
// (sum dy, sum dy * xhat) backward — so the layer stays ONE launch under data parallelism.  All C blocks must be
// co-resident (they wait for their partners; the host wrapper checks C against the occupancy).
namespace ganb200 {
constexpr int BN_SMALL_PER = 32;                                   // values per thread
constexpr int BN_SMALL_MAX = 256 * BN_SMALL_PER;                   // 8192: 128 threads up to 4096 values, 256 above

template <int BN_SMALL_THREADS>
__global__ void __launch_bounds__(BN_SMALL_THREADS) bn_small_fwd_kernel(
    const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ y,
    float* __restrict__ mean_out, float* __restrict__ invstd_out, float* __restrict__ running_mean,
    float* __restrict__ running_var, int64_t* nbt, int N, int C, int HW, float eps, float momentum,
    const float* __restrict__ slope, PeerRef peer) {
  const int c = blockIdx.x, cnt = N * HW;
  const unsigned int seq = peer.bufs ? peer_seq_begin(peer) : 0u;
  float v[BN_SMALL_PER];
  float s = 0.f;
#pragma unroll
  for (int e = 0; e < BN_SMALL_PER; ++e) {
    const int j = threadIdx.x + e * BN_SMALL_THREADS;
    const bool ok = j < cnt;
    const int n = ok ? j / HW : 0, i = ok ? j - n * HW : 0;
    v[e] = __ldg(x + ((int64_t)n * C + c) * HW + i);
    if (!ok) v[e] = 0.f;
    s += v[e];
  }
  float mean = block_sum<BN_SMALL_THREADS>(s) / (float)cnt;
  float q = 0.f;
#pragma unroll
  for (int e = 0; e < BN_SMALL_PER; ++e) {
    const float d = (threadIdx.x + e * BN_SMALL_THREADS < cnt) ? v[e] - mean : 0.f;
    q = fmaf(d, d, q);
  }
  float var = block_sum<BN_SMALL_THREADS>(q) / (float)cnt;
  float cnt_g = (float)cnt;
  if (peer.bufs) {   // equal shards: global mean = mean of means, global biased variance = E[var + mean^2] - mean_g^2
    const float2 t = peer_ll_sum2(peer, seq, c, mean, fmaf(mean, mean, var));
    const float inv = 1.f / (float)peer.world;
    mean = t.x * inv;
    var = fmaf(-mean, mean, t.y * inv);
    cnt_g *= (float)peer.world;
  }
  const float is = 1.f / sqrtf(var + eps);
  const float gm = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
  const float sl = slope ? slope[c] : 1.f;   // following nn.PReLU(C), fused: y = u > 0 ? u : slope * u
#pragma unroll
  for (int e = 0; e < BN_SMALL_PER; ++e) {
    const int j = threadIdx.x + e * BN_SMALL_THREADS;
    if (j < cnt) {
      const int n = j / HW, i = j - n * HW;
      const float u = (v[e] - mean) * is * gm + bt;
      y[((int64_t)n * C + c) * HW + i] = u > 0.f ? u : sl * u;
    }
  }
  if (threadIdx.x == 0) {
    mean_out[c] = mean;
    invstd_out[c] = is;
    if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    if (running_var)
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (cnt_g > 1.f ? var * cnt_g / (cnt_g - 1.f) : var);
    if (nbt && c == 0) *nbt += 1;
  }
  if (peer.bufs) peer_seq_commit(peer, seq);
}

template <int BN_SMALL_THREADS>
__global__ void __launch_bounds__(BN_SMALL_THREADS) bn_small_bwd_kernel(
    const float* __restrict__ gy, const float* __restrict__ x, const float* __restrict__ mean,
    const float* __restrict__ invstd, const float* __restrict__ gamma, float* __restrict__ gx,
    float* __restrict__ ggamma, float* __restrict__ gbeta, int N, int C, int HW, const float* __restrict__ beta,
    const float* __restrict__ slope, float* __restrict__ gslope, PeerRef peer) {
  const int c = blockIdx.x, cnt = N * HW;
  const unsigned int seq = peer.bufs ? peer_seq_begin(peer) : 0u;
  const float m = mean[c], is = invstd[c], gm = gamma ? gamma[c] : 1.f;
  const float bt = beta ? beta[c] : 0.f, sl = slope ? slope[c] : 1.f;
  float g[BN_SMALL_PER], xh[BN_SMALL_PER];
  float a = 0.f, b = 0.f, ps = 0.f;
#pragma unroll
  for (int e = 0; e < BN_SMALL_PER; ++e) {
    const int j = threadIdx.x + e * BN_SMALL_THREADS;
    const bool ok = j < cnt;
    const int n = ok ? j / HW : 0, i = ok ? j - n * HW : 0;
    const int64_t idx = ((int64_t)n * C + c) * HW + i;
    g[e] = __ldg(gy + idx);
    xh[e] = (__ldg(x + idx) - m) * is;
    if (!ok) { g[e] = 0.f; xh[e] = 0.f; }
    if (slope) {   // fused PReLU: the batch-norm output u is recomputed; d/dslope = sum gy * min(u, 0)
      const float u = xh[e] * gm + bt;
      ps = fmaf(g[e], fminf(u, 0.f), ps);
      g[e] *= u > 0.f ? 1.f : sl;
    }
    a += g[e];
    b = fmaf(g[e], xh[e], b);
  }
  const float s1 = block_sum<BN_SMALL_THREADS>(a);
  const float s2 = block_sum<BN_SMALL_THREADS>(b);
  const float s3 = slope ? block_sum<BN_SMALL_THREADS>(ps) : 0.f;
  float ia = s1 / (float)cnt, ib = s2 / (float)cnt;
  if (peer.bufs) {   // dx needs the sums over the global batch; dgamma / dbeta / dslope stay rank-local partial sums
    const float2 t = peer_ll_sum2(peer, seq, c, s1, s2);
    const float cnt_g = (float)cnt * (float)peer.world;
    ia = t.x / cnt_g;
    ib = t.y / cnt_g;
  }
#pragma unroll
  for (int e = 0; e < BN_SMALL_PER; ++e) {
    const int j = threadIdx.x + e * BN_SMALL_THREADS;
    if (j < cnt) {
      const int n = j / HW, i = j - n * HW;
      gx[((int64_t)n * C + c) * HW + i] = gm * is * (g[e] - ia - xh[e] * ib);
    }
  }
  if (threadIdx.x == 0) {
    if (ggamma) ggamma[c] = s2;
    if (gbeta) gbeta[c] = s1;
    if (gslope) gslope[c] = s3;
  }
  if (peer.bufs) peer_seq_commit(peer, seq);
}
}  // namespace ganb200

extern "C" int ganb200_bn_small_eligible(int N, int C, int HW) {
  return N > 0 && C > 0 && HW > 0 && (int64_t)N * HW <= BN_SMALL_MAX;
}

static int bn_small_fwd_launch(const float* x, const float* gamma, const float* beta, float* y, float* mean, float* invstd,
                               float* running_mean, float* running_var, int64_t* nbt, int N, int C, int HW, float eps,
                               float momentum, const float* slope, PeerRef peer, cudaStream_t s) {
  if ((int64_t)N * HW <= 128 * BN_SMALL_PER)
    bn_small_fwd_kernel<128><<<C, 128, 0, s>>>(x, gamma, beta, y, mean, invstd, running_mean, running_var, nbt, N, C, HW, eps,
                                               momentum, slope, peer);
  else
    bn_small_fwd_kernel<256><<<C, 256, 0, s>>>(x, gamma, beta, y, mean, invstd, running_mean, running_var, nbt, N, C, HW, eps,
                                               momentum, slope, peer);
  GB_CHECK_LAUNCH();
  return 0;
}

static int bn_small_bwd_launch(const float* gy, const float* x, const float* mean, const float* invstd, const float* gamma,
                               float* gx, float* ggamma, float* gbeta, int N, int C, int HW, const float* beta,
                               const float* slope, float* gslope, PeerRef peer, cudaStream_t s) {
  if ((int64_t)N * HW <= 128 * BN_SMALL_PER)
    bn_small_bwd_kernel<128><<<C, 128, 0, s>>>(gy, x, mean, invstd, gamma, gx, ggamma, gbeta, N, C, HW, beta, slope, gslope,
                                               peer);
  else
    bn_small_bwd_kernel<256><<<C, 256, 0, s>>>(gy, x, mean, invstd, gamma, gx, ggamma, gbeta, N, C, HW, beta, slope, gslope,
                                               peer);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn_small_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean,
                                    float* invstd, float* running_mean, float* running_var,
                                    int64_t* num_batches_tracked, int N, int C, int HW, float eps, float momentum,
                                    const float* prelu_slope, void* stream) {
  GB_REQUIRE(ganb200_bn_small_eligible(N, C, HW), GANB200_BAD_SHAPE, "bn_small_fwd: N*HW must be in [1, 8192]");
  return bn_small_fwd_launch(x, gamma, beta, y, mean, invstd, running_mean, running_var, num_batches_tracked, N, C, HW, eps,
                             momentum, prelu_slope, PeerRef{nullptr, 0, 1}, (cudaStream_t)stream);
}

// channels (= blocks that wait for their partners on the other ranks) the fused sync kernels can keep co-resident
extern "C" int ganb200_bn_small_sync_max_channels(void) {
  static int cap = -1;
  if (cap < 0) {
    int dev = 0, sms = 0, a = 0, b = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, bn_small_fwd_kernel<256>, 256, 0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, bn_small_bwd_kernel<256>, 256, 0) != cudaSuccess)
      return 0;
    // half of what fits: other kernels (NCCL on the bucket stream) may hold SM slots
    cap = std::min(std::min(a, b) * sms / 2, kPeerLLSlots);
  }
  return cap;
}

extern "C" int ganb200_bn_small_fwd_sync(const float* x, const float* gamma, const float* beta, float* y, float* mean,
                                         float* invstd, float* running_mean, float* running_var,
                                         int64_t* num_batches_tracked, int N, int C, int HW, float eps, float momentum,
                                         const float* prelu_slope, const void* peer_bufs_dev, int rank, int world,
                                         void* stream) {
  GB_REQUIRE(ganb200_bn_small_eligible(N, C, HW), GANB200_BAD_SHAPE, "bn_small_fwd_sync: N*HW must be in [1, 8192]");
  GB_REQUIRE(peer_bufs_dev && world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, GANB200_BAD_ARG,
             "bn_small_fwd_sync: bad peer table / rank / world");
  GB_REQUIRE(C <= ganb200_bn_small_sync_max_channels(), GANB200_BAD_SHAPE, "bn_small_fwd_sync: too many channels to be co-resident");
  return bn_small_fwd_launch(x, gamma, beta, y, mean, invstd, running_mean, running_var, num_batches_tracked, N, C, HW, eps,
                             momentum, prelu_slope, PeerRef{(char* const*)peer_bufs_dev, rank, world}, (cudaStream_t)stream);
}

extern "C" int ganb200_bn_small_bwd(const float* gy, const float* x, const float* mean, const float* invstd,
                                    const float* gamma, float* gx, float* ggamma, float* gbeta, int N, int C, int HW,
                                    const float* beta, const float* prelu_slope, float* gslope, void* stream) {
  GB_REQUIRE(ganb200_bn_small_eligible(N, C, HW), GANB200_BAD_SHAPE, "bn_small_bwd: N*HW must be in [1, 8192]");
  GB_REQUIRE(!prelu_slope || gslope, GANB200_BAD_ARG, "bn_small_bwd: prelu_slope without gslope");
  return bn_small_bwd_launch(gy, x, mean, invstd, gamma, gx, ggamma, gbeta, N, C, HW, beta, prelu_slope, gslope,
                             PeerRef{nullptr, 0, 1}, (cudaStream_t)stream);
}

extern "C" int ganb200_bn_small_bwd_sync(const float* gy, const float* x, const float* mean, const float* invstd,
                                         const float* gamma, float* gx, float* ggamma, float* gbeta, int N, int C, int HW,
                                         const float* beta, const float* prelu_slope, float* gslope,
                                         const void* peer_bufs_dev, int rank, int world, void* stream) {
  GB_REQUIRE(ganb200_bn_small_eligible(N, C, HW), GANB200_BAD_SHAPE, "bn_small_bwd_sync: N*HW must be in [1, 8192]");
  GB_REQUIRE(!prelu_slope || gslope, GANB200_BAD_ARG, "bn_small_bwd_sync: prelu_slope without gslope");
  GB_REQUIRE(peer_bufs_dev && world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, GANB200_BAD_ARG,
             "bn_small_bwd_sync: bad peer table / rank / world");
  GB_REQUIRE(C <= ganb200_bn_small_sync_max_channels(), GANB200_BAD_SHAPE, "bn_small_bwd_sync: too many channels to be co-resident");
  return bn_small_bwd_launch(gy, x, mean, invstd, gamma, gx, ggamma, gbeta, N, C, HW, beta, prelu_slope, gslope,
                             PeerRef{(char* const*)peer_bufs_dev, rank, world}, (cudaStream_t)stream);
}

static dim3 bn_reduce_grid(int N, int C, int HW) {
  int want = (kNumSMs * 4 + C - 1) / C;
  int max_split = HW >= 1024 ? N : (int)(((int64_t)N * HW + 255) / 256);
  int nsplit = want < max_split ? want : max_split;
  if (nsplit < 1) nsplit = 1;
  return dim3(C, nsplit);
}

extern "C" int ganb200_bn_stats(const float* x, float* mean, float* var_biased, int N, int C, int HW, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  GB_REQUIRE(N > 0 && C > 0 && HW > 0, GANB200_BAD_SHAPE, "bn_stats: empty");
  float inv = 1.f / ((float)N * (float)HW);
  dim3 grid = bn_reduce_grid(N, C, HW);
  const int region = det_region_for(s, grid.x, grid.y);
  if (region < 0) {
    GB_CHECK_CUDA(cudaMemsetAsync(mean, 0, sizeof(float) * C, s));
    GB_CHECK_CUDA(cudaMemsetAsync(var_biased, 0, sizeof(float) * C, s));
  }
  bn_reduce_kernel<0><<<grid, 256, 0, s>>>(x, nullptr, nullptr, nullptr, mean, nullptr, N, C, HW, inv, region);
  bn_reduce_kernel<1><<<grid, 256, 0, s>>>(x, nullptr, mean, nullptr, var_biased, nullptr, N, C, HW, inv, region);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn_apply(const float* x, const float* mean, const float* var_biased, const float* gamma,
                                const float* beta, float* y, float* save_invstd, float* running_mean,
                                float* running_var, int64_t* num_batches_tracked, int N, int C, int HW, float eps,
                                float momentum, int64_t count_total, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int64_t total = (int64_t)N * C * HW;
  bn_apply_kernel<0><<<ew_grid(total, 256, 4), 256, 0, s>>>(x, nullptr, mean, var_biased, gamma, beta, nullptr, nullptr, y,
                                                           save_invstd, running_mean, running_var, N, C, HW, eps,
                                                           momentum, (float)count_total, num_batches_tracked);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn_bwd_reduce(const float* gy, const float* x, const float* mean, const float* invstd,
                                     float* sum_gy, float* sum_gy_xhat, int N, int C, int HW, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const dim3 grid = bn_reduce_grid(N, C, HW);
  const int region = det_region_for(s, grid.x, grid.y, 2);
  if (region < 0) {
    GB_CHECK_CUDA(cudaMemsetAsync(sum_gy, 0, sizeof(float) * C, s));
    GB_CHECK_CUDA(cudaMemsetAsync(sum_gy_xhat, 0, sizeof(float) * C, s));
  }
  bn_reduce_kernel<2><<<grid, 256, 0, s>>>(x, gy, mean, invstd, sum_gy, sum_gy_xhat, N, C, HW, 1.f, region);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bn_bwd_apply(const float* gy, const float* x, const float* mean, const float* invstd,
                                    const float* gamma, const float* sum_gy, const float* sum_gy_xhat, float* gx,
                                    float* ggamma, float* gbeta, int N, int C, int HW, int64_t count_total,
                                    void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int64_t total = (int64_t)N * C * HW;
  bn_apply_kernel<1><<<ew_grid(total, 256, 4), 256, 0, s>>>(x, gy, mean, invstd, gamma, nullptr, sum_gy, sum_gy_xhat, gx,
                                                           ggamma, gbeta, nullptr, N, C, HW, 0.f, 0.f,
                                                           (float)count_total, nullptr);
  GB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ softmax over M of [B,M,C]
namespace ganb200 {

__global__ void __launch_bounds__(256) softmax_m_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int B,
                                                            int M, int C) {
  int64_t total = (int64_t)B * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int b = (int)(i / C), c = (int)(i % C);
    const float* xp = x + (int64_t)b * M * C + c;
    float* yp = y + (int64_t)b * M * C + c;
    float mx = -INFINITY;
    for (int m = 0; m < M; ++m) mx = fmaxf(mx, xp[(int64_t)m * C]);
    float sum = 0.f;
    for (int m = 0; m < M; ++m) sum += expf(xp[(int64_t)m * C] - mx);
    float inv = 1.f / sum;
    for (int m = 0; m < M; ++m) yp[(int64_t)m * C] = expf(xp[(int64_t)m * C] - mx) * inv;
  }
}

__global__ void __launch_bounds__(256) softmax_m_bwd_kernel(const float* __restrict__ gy, const float* __restrict__ y,
                                                            float* __restrict__ gx, int B, int M, int C) {
  int64_t total = (int64_t)B * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int b = (int)(i / C), c = (int)(i % C);
    int64_t base = (int64_t)b * M * C + c;
    float dot = 0.f;
    for (int m = 0; m < M; ++m) dot += gy[base + (int64_t)m * C] * y[base + (int64_t)m * C];
    for (int m = 0; m < M; ++m) {
      int64_t k = base + (int64_t)m * C;
      gx[k] = y[k] * (gy[k] - dot);
    }
  }
}

// ------------------------------------------------------------------ MiniBatchStdDev scalar, x as [G,L]
// MODE 0 fwd: out[0] += sum_j sigma_j / L
// MODE 1 bwd: gx[q,j] = g * (x_qj - mu_j) / ((G-1) L sigma_j)
// MODE 2 bwd_bwd: out[0] += sum ggx_qj (x_qj-mu_j)/((G-1) L sigma_j);
//                 gx[r,j] = g/((G-1)L) * ((ggx_rj - mean_q ggx_qj)/sigma_j - (x_rj-mu_j)/((G-1) sigma_j^3) * sum_q ggx_qj (x_qj-mu_j))
// REG: the G values of a column live in registers (G <= 16); otherwise they are re-read (L2 hits; the columns of a
// warp are contiguous so every pass is coalesced).
template <int MODE, bool REG>
__global__ void __launch_bounds__(256) mbstd_kernel(const float* __restrict__ g, const float* __restrict__ x,
                                                    const float* __restrict__ ggx, float* __restrict__ out,
                                                    float* __restrict__ gx, int G, int64_t L, float eps, int region) {
  float acc = 0.f;
  float gv = MODE >= 1 ? g[0] : 0.f;
  float invL = 1.f / (float)L, invG1 = 1.f / (float)(G - 1);
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < L; j += (int64_t)gridDim.x * blockDim.x) {
    float xv[REG ? 16 : 1];
    float mu = 0.f;
    for (int q = 0; q < G; ++q) {
      float v = x[(int64_t)q * L + j];
      if (REG) xv[q] = v;
      mu += v;
    }
    mu /= (float)G;
    float ss = 0.f;
    for (int q = 0; q < G; ++q) {
      float d = (REG ? xv[q] : x[(int64_t)q * L + j]) - mu;
      if (REG) xv[q] = d;
      ss += d * d;
    }
    float sigma = sqrtf(ss * invG1 + eps);
    if (MODE == 0) {
      acc += sigma * invL;
    } else if (MODE == 1) {
      float k = gv * invG1 * invL / sigma;
      for (int q = 0; q < G; ++q) gx[(int64_t)q * L + j] = k * (REG ? xv[q] : x[(int64_t)q * L + j] - mu);
    } else {
      float gg[REG ? 16 : 1];
      float gmean = 0.f, dot = 0.f;
      for (int q = 0; q < G; ++q) {
        float v = ggx[(int64_t)q * L + j];
        if (REG) gg[q] = v;
        gmean += v;
        dot += v * (REG ? xv[q] : x[(int64_t)q * L + j] - mu);
      }
      gmean /= (float)G;
      acc += dot * invG1 * invL / sigma;
      float k = gv * invG1 * invL;
      float c3 = dot * invG1 / (sigma * sigma * sigma);
      for (int q = 0; q < G; ++q) {
        float d = REG ? xv[q] : x[(int64_t)q * L + j] - mu;
        float gq = REG ? gg[q] : ggx[(int64_t)q * L + j];
        gx[(int64_t)q * L + j] = k * ((gq - gmean) / sigma - d * c3);
      }
    }
  }
  if (MODE != 1) {
    float t = block_sum<256>(acc);
    det_group_store<256, 1>(region, 0, (int)blockIdx.x, (int)gridDim.x, 1.f, t, out);
  }
}

// ------------------------------------------------------------------ BCE (mean), torch semantics
__global__ void __launch_bounds__(256) bce_fwd_kernel(const float* __restrict__ p, const float* __restrict__ t,
                                                      float* __restrict__ out, int* __restrict__ flag, int64_t n,
                                                      int region) {
  float acc = 0.f;
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float pv = p[i], tv = t[i];
    if (!(pv >= 0.f && pv <= 1.f)) bad = true;
    float lp = fmaxf(logf(pv), -100.f), lq = fmaxf(logf(1.f - pv), -100.f);
    acc += -(tv * lp + (1.f - tv) * lq);
  }
  if (bad && flag) atomicExch(flag, 1);
  float s = block_sum<256>(acc);
  det_group_store<256, 1>(region, 0, (int)blockIdx.x, (int)gridDim.x, 1.f / (float)n, s, out);
}

__global__ void __launch_bounds__(256) bce_bwd_kernel(const float* __restrict__ g, const float* __restrict__ p,
                                                      const float* __restrict__ t, float* __restrict__ gp, int64_t n) {
  float gv = g[0] / (float)n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float pv = p[i];
    gp[i] = gv * (pv - t[i]) / fmaxf((1.f - pv) * pv, 1e-12f);
  }
}

// ------------------------------------------------------------------ separable resampling with tap tables
__global__ void __launch_bounds__(256) resample2d_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t NC,
                                                         int Hi, int Wi, int Ho, int Wo, const int* __restrict__ ih,
                                                         const float* __restrict__ wh, int Th,
                                                         const int* __restrict__ iw, const float* __restrict__ ww, int Tw) {
  int64_t total = NC * Ho * Wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int wo = (int)(i % Wo);
    int ho = (int)((i / Wo) % Ho);
    int64_t nc = i / ((int64_t)Wo * Ho);
    const float* xp = x + nc * Hi * Wi;
    float acc = 0.f;
    for (int a = 0; a < Th; ++a) {
      float wa = wh[ho * Th + a];
      if (wa == 0.f) continue;
      const float* xr = xp + (int64_t)ih[ho * Th + a] * Wi;
      float racc = 0.f;
      for (int b = 0; b < Tw; ++b) racc = fmaf(ww[wo * Tw + b], xr[iw[wo * Tw + b]], racc);
      acc = fmaf(wa, racc, acc);
    }
    y[i] = acc;
  }
}

// Separable form for whole planes that fit in shared memory (every bicubic / Smooth use of the models: planes up to
// 64x64): one block walks planes; the plane is staged with coalesced (float4) loads, the horizontal pass writes an
// Hi x Wo intermediate to shared memory, the vertical pass streams the output rows.  Tap tables live in shared memory
// too.  (Th + Tw) multiply-adds per output instead of Th*Tw and no global gathers: HBM-bound, (in + out) * 4 B.
// Summation order per output equals resample2d_kernel's (taps b inside, taps a outside).
// The first version was issue-bound (ncu: IPC 2.96, DRAM 20 %): an integer division and three shared-memory reads per
// multiply-add.  When Wo is a power of two <= 256 (every hot shape) a thread now keeps ONE output column for the
// horizontal pass (its T tap weights / indices in registers, loaded once per kernel, T a template parameter) and one
// group of four columns for the vertical pass (one 16-byte shared-memory read and four FMAs per tap, 16-byte stores).
template <int T>   // taps per output of both passes when > 0 (exact unrolling); 0 = run-time Th / Tw
__global__ void __launch_bounds__(256) resample2d_sep_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                             int64_t NC, int Hi, int Wi, int Ho, int Wo,
                                                             const int* __restrict__ ih, const float* __restrict__ wh,
                                                             int Th_, const int* __restrict__ iw,
                                                             const float* __restrict__ ww, int Tw_, int fast) {
  extern __shared__ float sm[];
  const int Th = T > 0 ? T : Th_, Tw = T > 0 ? T : Tw_;
  const int HWi = Hi * Wi, HWo = Ho * Wo, HiWo = Hi * Wo;
  float* pin = sm;                               // [Hi][Wi]
  float* tmp = pin + ((HWi + 3) & ~3);           // [Hi][Wo]
  float* s_ww = tmp + ((HiWo + 3) & ~3);         // [Tw][Wo]
  float* s_wh = s_ww + Wo * Tw;                  // [Ho][Th]
  int* s_iw = reinterpret_cast<int*>(s_wh + Ho * Th);
  int* s_ih = s_iw + Wo * Tw;
  for (int i = threadIdx.x; i < Wo * Tw; i += 256) {  // transposed [Tw][Wo]: lanes (consecutive wo) hit distinct banks
    const int wo = i / Tw, b = i - wo * Tw;
    s_ww[b * Wo + wo] = ww[i];
    s_iw[b * Wo + wo] = iw[i];
  }
  for (int i = threadIdx.x; i < Ho * Th; i += 256) { s_wh[i] = wh[i]; s_ih[i] = ih[i] * Wo; }
  const bool vec = (HWi & 3) == 0 && aligned16(x);
  constexpr int TR = T > 0 ? T : 1;
  // horizontal pass: column wo0, rows h0, h0 + hstep, ...
  const int wo0 = fast ? (int)threadIdx.x % Wo : 0, h0 = fast ? (int)threadIdx.x / Wo : 0, hstep = fast ? 256 / Wo : 1;
  // vertical pass: columns 4*q0 .. 4*q0+3, rows v0, v0 + vstep, ...
  const int Wq = Wo >> 2;
  const int q0 = fast ? (int)threadIdx.x % Wq : 0, v0 = fast ? (int)threadIdx.x / Wq : 0, vstep = fast ? 256 / Wq : 1;
  float rw[TR];
  int ri[TR];
#pragma unroll
  for (int b = 0; b < TR; ++b) {
    rw[b] = (fast && T > 0) ? ww[wo0 * Tw + b] : 0.f;
    ri[b] = (fast && T > 0) ? iw[wo0 * Tw + b] : 0;
  }
  for (int64_t nc = blockIdx.x; nc < NC; nc += gridDim.x) {
    const float* xp = x + nc * HWi;
    __syncthreads();
    if (vec) {
      for (int i = threadIdx.x; i < (HWi >> 2); i += 256)
        reinterpret_cast<float4*>(pin)[i] = __ldg(reinterpret_cast<const float4*>(xp) + i);
    } else {
      for (int i = threadIdx.x; i < HWi; i += 256) pin[i] = __ldg(xp + i);
    }
    __syncthreads();
    float* yp = y + nc * HWo;
    if (fast) {
      for (int h = h0; h < Hi; h += hstep) {
        const float* row = pin + h * Wi;
        float acc = 0.f;
        if (T > 0) {
#pragma unroll
          for (int b = 0; b < TR; ++b) acc = fmaf(rw[b], row[ri[b]], acc);
        } else {
          for (int b = 0; b < Tw; ++b) acc = fmaf(s_ww[b * Wo + wo0], row[s_iw[b * Wo + wo0]], acc);
        }
        tmp[h * Wo + wo0] = acc;
      }
      __syncthreads();
      for (int ho = v0; ho < Ho; ho += vstep) {
        const float* wrow = s_wh + ho * Th;
        const int* irow = s_ih + ho * Th;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int a = 0; a < (T > 0 ? T : 1); ++a) {
          if (T > 0) {
            const float wa = wrow[a];
            if (wa != 0.f) {   // skipped exactly like the scalar kernel: -0.0 / NaN handling stays identical
              const float4 t = *reinterpret_cast<const float4*>(tmp + irow[a] + 4 * q0);
              acc.x = fmaf(wa, t.x, acc.x); acc.y = fmaf(wa, t.y, acc.y);
              acc.z = fmaf(wa, t.z, acc.z); acc.w = fmaf(wa, t.w, acc.w);
            }
          }
        }
        if (T == 0) {
          for (int a = 0; a < Th; ++a) {
            const float wa = wrow[a];
            if (wa == 0.f) continue;
            const float4 t = *reinterpret_cast<const float4*>(tmp + irow[a] + 4 * q0);
            acc.x = fmaf(wa, t.x, acc.x); acc.y = fmaf(wa, t.y, acc.y);
            acc.z = fmaf(wa, t.z, acc.z); acc.w = fmaf(wa, t.w, acc.w);
          }
        }
        *reinterpret_cast<float4*>(yp + ho * Wo + 4 * q0) = acc;
      }
    } else {
      for (int i = threadIdx.x; i < HiWo; i += 256) {
        const int h = i / Wo, wo = i - h * Wo;
        const float* row = pin + h * Wi;
        float acc = 0.f;
        for (int b = 0; b < Tw; ++b) acc = fmaf(s_ww[b * Wo + wo], row[s_iw[b * Wo + wo]], acc);
        tmp[i] = acc;
      }
      __syncthreads();
      for (int i = threadIdx.x; i < HWo; i += 256) {
        const int ho = i / Wo, wo = i - ho * Wo;
        float acc = 0.f;
        for (int a = 0; a < Th; ++a) {
          const float wa = s_wh[ho * Th + a];
          if (wa == 0.f) continue;
          acc = fmaf(wa, tmp[s_ih[ho * Th + a] + wo], acc);
        }
        yp[i] = acc;
      }
    }
  }
}

// Same operator for "few outputs, many taps" (AdaptiveAvgPool2d(5) of a 64x64 map: 25 outputs x 169 taps per plane):
// one block per plane, the plane staged in shared memory with coalesced loads, one warp per output reducing its taps.
__global__ void __launch_bounds__(256) resample2d_plane_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                               int64_t NC, int Hi, int Wi, int Ho, int Wo,
                                                               const int* __restrict__ ih, const float* __restrict__ wh,
                                                               int Th, const int* __restrict__ iw,
                                                               const float* __restrict__ ww, int Tw) {
  extern __shared__ float plane[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int HWi = Hi * Wi, T = Th * Tw;
  for (int64_t nc = blockIdx.x; nc < NC; nc += gridDim.x) {
    const float* xp = x + nc * HWi;
    __syncthreads();
    for (int i = threadIdx.x; i < HWi; i += 256) plane[i] = xp[i];
    __syncthreads();
    for (int o = warp; o < Ho * Wo; o += 8) {
      const int ho = o / Wo, wo = o - ho * Wo;
      float acc = 0.f;
      for (int t = lane; t < T; t += 32) {
        const int a = t / Tw, b = t - a * Tw;
        acc = fmaf(wh[ho * Th + a] * ww[wo * Tw + b], plane[ih[ho * Th + a] * Wi + iw[wo * Tw + b]], acc);
      }
      acc = warp_sum(acc);
      if (lane == 0) y[nc * Ho * Wo + o] = acc;
    }
  }
}

__global__ void __launch_bounds__(256) mixup_kernel(const float* __restrict__ t, const float* __restrict__ f, float lam,
                                                    float* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = lam * t[i] + (1.f - lam) * f[i];
}

// units/mixup.py:26  label:(B,1) f64 broadcast against (B,3,H,W): needs W==1 or B==W; label index = h (row!) when B==H.. 
// torch broadcasting aligns trailing dims: label (B,1) -> (1,1,B,1): dim -2 (H) indexes label, dim -1 broadcast.
__global__ void __launch_bounds__(256) mixup_wgan_kernel(const double* __restrict__ label, const float* __restrict__ t,
                                                         const float* __restrict__ f, double* __restrict__ out, int B,
                                                         int H, int W, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int h = (int)((i / W) % H);
    double l = label[B == 1 ? 0 : h];
    out[i] = l * (double)t[i] + (1.0 - l) * (double)f[i];
  }
}

__global__ void __launch_bounds__(256) clamp_kernel(float* __restrict__ x, int64_t n, float lo, float hi) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] = fminf(fmaxf(x[i], lo), hi);
}

}  // namespace ganb200

extern "C" int ganb200_softmax_m_fwd(const float* x, float* y, int B, int M, int C, void* stream) {
  softmax_m_fwd_kernel<<<ew_grid((int64_t)B * C, 256, 1), 256, 0, (cudaStream_t)stream>>>(x, y, B, M, C);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_softmax_m_bwd(const float* gy, const float* y, float* gx, int B, int M, int C, void* stream) {
  softmax_m_bwd_kernel<<<ew_grid((int64_t)B * C, 256, 1), 256, 0, (cudaStream_t)stream>>>(gy, y, gx, B, M, C);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_mbstd_fwd(const float* x, float* out, int G, int64_t L, float eps, void* stream) {
  GB_REQUIRE(G >= 2 && L > 0, GANB200_BAD_SHAPE, "mbstd: need G>=2");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = ew_grid(L, 256, 1), region = det_region_for(s, 1, grid);
  if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float), s));
  if (G <= 16) mbstd_kernel<0, true><<<grid, 256, 0, s>>>(nullptr, x, nullptr, out, nullptr, G, L, eps, region);
  else mbstd_kernel<0, false><<<grid, 256, 0, s>>>(nullptr, x, nullptr, out, nullptr, G, L, eps, region);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_mbstd_bwd(const float* g, const float* x, float* gx, int G, int64_t L, float eps, void* stream) {
  GB_REQUIRE(G >= 2 && L > 0, GANB200_BAD_SHAPE, "mbstd: need G>=2");
  if (G <= 16) mbstd_kernel<1, true><<<ew_grid(L, 256, 1), 256, 0, (cudaStream_t)stream>>>(g, x, nullptr, nullptr, gx, G, L, eps, 0);
  else mbstd_kernel<1, false><<<ew_grid(L, 256, 1), 256, 0, (cudaStream_t)stream>>>(g, x, nullptr, nullptr, gx, G, L, eps, 0);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_mbstd_bwd_bwd(const float* g, const float* x, const float* ggx, float* out_gg, float* out_gx,
                                     int G, int64_t L, float eps, void* stream) {
  GB_REQUIRE(G >= 2 && L > 0, GANB200_BAD_SHAPE, "mbstd: need G>=2");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = ew_grid(L, 256, 1), region = det_region_for(s, 1, grid);
  if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(out_gg, 0, sizeof(float), s));
  if (G <= 16) mbstd_kernel<2, true><<<grid, 256, 0, s>>>(g, x, ggx, out_gg, out_gx, G, L, eps, region);
  else mbstd_kernel<2, false><<<grid, 256, 0, s>>>(g, x, ggx, out_gg, out_gx, G, L, eps, region);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_bce_fwd(const float* p, const float* t, float* out, int* range_flag, int64_t n, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  GB_REQUIRE(n > 0, GANB200_BAD_SHAPE, "bce: empty");
  const int grid = ew_grid(n, 256, 4), region = det_region_for(s, 1, grid);
  if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float), s));
  if (range_flag) GB_CHECK_CUDA(cudaMemsetAsync(range_flag, 0, sizeof(int), s));
  bce_fwd_kernel<<<grid, 256, 0, s>>>(p, t, out, range_flag, n, region);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_bce_fwd_sticky(const float* p, const float* t, float* out, int* sticky_flag, int64_t n,
                                      void* stream) {
  // as ganb200_bce_fwd, but the flag is only ever SET (never cleared): inside a replayed CUDA graph the host cannot read a
  // per-call flag, so the caller keeps one flag per device, polls it between replays and clears it itself
  cudaStream_t s = (cudaStream_t)stream;
  GB_REQUIRE(n > 0, GANB200_BAD_SHAPE, "bce: empty");
  GB_REQUIRE(sticky_flag != nullptr, GANB200_BAD_ARG, "bce_fwd_sticky: null flag");
  const int grid = ew_grid(n, 256, 4), region = det_region_for(s, 1, grid);
  if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float), s));
  bce_fwd_kernel<<<grid, 256, 0, s>>>(p, t, out, sticky_flag, n, region);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_bce_bwd(const float* g, const float* p, const float* t, float* gp, int64_t n, void* stream) {
  bce_bwd_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(g, p, t, gp, n);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_resample2d(const float* x, float* y, int NC, int Hi, int Wi, int Ho, int Wo, const int* ih,
                                  const float* wh, int Th, const int* iw, const float* ww, int Tw, void* stream) {
  int64_t total = (int64_t)NC * Ho * Wo;
  if (total == 0) return 0;
  GB_REQUIRE(Th > 0 && Tw > 0, GANB200_BAD_SHAPE, "resample2d: empty tap table");
  const size_t sep_bytes = sizeof(float) * (((size_t)Hi * Wi + 3) / 4 * 4 + ((size_t)Hi * Wo + 3) / 4 * 4) +
                           (sizeof(float) + sizeof(int)) * ((size_t)Wo * Tw + (size_t)Ho * Th);
  static const bool plane_first = [] { const char* e = getenv("GANB200_RESAMPLE_PLANE"); return e && e[0] == '1'; }();
  // AdaptiveAvgPool2d(5) of a 64x64 plane: 25 outputs x 169 taps.  Separable it is 320 x 13 + 25 x 13 multiply-adds from
  // shared-memory tables (the plane kernel re-reads four global tap tables per multiply-add): prefer it when it fits.
  const bool sep_fits = sep_bytes <= 48 * 1024;
  if (Ho * Wo <= 64 && Th * Tw >= 16 && (size_t)Hi * Wi * sizeof(float) <= 40 * 1024 && (plane_first || !sep_fits)) {
    int grid = (int)std::min<int64_t>(NC, (int64_t)kNumSMs * 8);
    resample2d_plane_kernel<<<grid, 256, sizeof(float) * Hi * Wi, (cudaStream_t)stream>>>(x, y, NC, Hi, Wi, Ho, Wo, ih,
                                                                                         wh, Th, iw, ww, Tw);
  } else {
    const size_t sep_smem = sep_bytes;
    if (sep_smem <= 48 * 1024 && (Ho * Wo >= 64 || Th * Tw >= 16)) {
      int grid = (int)std::min<int64_t>(NC, (int64_t)kNumSMs * 16);
      // fast path: Wo a power of two in [4, 256] (thread <-> column mappings without division, 16-byte stores)
      const int fast = (Wo >= 4 && Wo <= 256 && (Wo & (Wo - 1)) == 0 && ((uintptr_t)y & 15) == 0) ? 1 : 0;
      const int T = (fast && Th == Tw && Th <= 6) ? Th : 0;
      cudaStream_t st = (cudaStream_t)stream;
#define GB_SEP(TT) resample2d_sep_kernel<TT><<<grid, 256, sep_smem, st>>>(x, y, NC, Hi, Wi, Ho, Wo, ih, wh, Th, iw, ww, Tw, fast)
      switch (T) {
        case 1: GB_SEP(1); break;
        case 2: GB_SEP(2); break;
        case 3: GB_SEP(3); break;
        case 4: GB_SEP(4); break;
        case 5: GB_SEP(5); break;
        case 6: GB_SEP(6); break;
        default: GB_SEP(0); break;
      }
#undef GB_SEP
    } else {
      resample2d_kernel<<<ew_grid(total, 256, 2), 256, 0, (cudaStream_t)stream>>>(x, y, NC, Hi, Wi, Ho, Wo, ih, wh, Th,
                                                                                 iw, ww, Tw);
    }
  }
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_mixup(const float* t, const float* f, float lam, float* out, int64_t n, void* stream) {
  if (n <= 0) return 0;
  mixup_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(t, f, lam, out, n);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_mixup_wgan_f64(const double* label, const float* t, const float* f, double* out, int B, int C,
                                      int H, int W, void* stream) {
  GB_REQUIRE(B == 1 || B == H, GANB200_BAD_SHAPE, "mixup_wgan: (B,1) label only broadcasts against (B,C,H,W) when B==1 or B==H");
  int64_t n = (int64_t)B * C * H * W;
  mixup_wgan_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(label, t, f, out, B, H, W, n);
  GB_CHECK_LAUNCH();
  return 0;
}
extern "C" int ganb200_clamp_(float* x, int64_t n, float lo, float hi, void* stream) {
  if (n <= 0) return 0;
  clamp_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(x, n, lo, hi);
  GB_CHECK_LAUNCH();
  return 0;
}
