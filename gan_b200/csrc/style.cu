// style.cu — fused kernels of the generator's StyleConv layer (generators/generator_13_5.py:219-266 in the reference):
// demodulation, the noise + PReLU tail (noise regenerated in-kernel from a Philox counter, never stored), the fused
// tail backward (PReLU mask-scale + the three row reductions the layer needs) and the demodulation adjoints.
#include "common.cuh"
#include <algorithm>

namespace ganb200 {

// ------------------------------------------------------------------------------------------------ Philox4x32-10
struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  __device__ static __forceinline__ uint4 rounds(uint4 c, uint2 k) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
      uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
      c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
      k.x += W0;
      k.y += W1;
    }
    return c;
  }
};

// four N(0,1) draws for the float4 group `g` of layer `layer`; state = {seed, forward counter}
__device__ __forceinline__ float4 normal4(const int64_t* __restrict__ state, uint32_t layer, uint64_t g) {
  const uint64_t seed = (uint64_t)state[0], ctr = (uint64_t)state[1];
  uint4 r = Philox::rounds(make_uint4((uint32_t)g, (uint32_t)(g >> 32), layer, (uint32_t)ctr),
                           make_uint2((uint32_t)seed, (uint32_t)(seed >> 32) ^ (uint32_t)(ctr >> 32)));
  const float k = 2.3283064365386963e-10f;  // 2^-32
  float u0 = ((float)r.x + 0.5f) * k, u1 = ((float)r.y + 0.5f) * k;
  float u2 = ((float)r.z + 0.5f) * k, u3 = ((float)r.w + 0.5f) * k;
  u0 = fminf(u0, 0.99999994f);
  u2 = fminf(u2, 0.99999994f);
  float r0 = sqrtf(-2.f * __logf(u0)), r1 = sqrtf(-2.f * __logf(u2));
  float s0, c0, s1, c1;
  __sincosf(6.283185307179586f * u1, &s0, &c0);
  __sincosf(6.283185307179586f * u3, &s1, &c1);
  return make_float4(r0 * c0, r0 * s0, r1 * c1, r1 * s1);
}

// z[g] = the noise the fused kernels use (tests, and the unfused reference path)
__global__ void __launch_bounds__(256) noise_fill_kernel(float4* __restrict__ z, int64_t groups,
                                                         const int64_t* __restrict__ state, uint32_t layer) {
  for (int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x; g < groups; g += (int64_t)gridDim.x * 256)
    z[g] = normal4(state, layer, (uint64_t)g);
}

// ------------------------------------------------------------------------------------------------ demodulation
// sinv[b,o] = rsqrt(alpha^2 * sum_i s[b,i]^2 * wsq[o,i] + eps).  A [B,C]x[C,K] product: one block = DEMOD_BT samples
// (s^2 staged in shared memory) x a slice of the output channels, so a wsq row is read once per DEMOD_BT samples.
constexpr int DEMOD_BT = 8;
__global__ void __launch_bounds__(256) demod_kernel(const float* __restrict__ s, const float* __restrict__ wsq,
                                                    float* __restrict__ sinv, int B, int C, int K, float alpha2,
                                                    float eps) {
  extern __shared__ float s2[];  // [DEMOD_BT][C]
  const int b0 = blockIdx.x * DEMOD_BT;
  for (int i = threadIdx.x; i < DEMOD_BT * C; i += 256) {
    const int bl = i / C, c = i - bl * C;
    const float v = (b0 + bl < B) ? s[(int64_t)(b0 + bl) * C + c] : 0.f;
    s2[i] = v * v;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int o = blockIdx.y * 8 + warp; o < K; o += gridDim.y * 8) {
    const float* wr = wsq + (int64_t)o * C;
    float acc[DEMOD_BT];
#pragma unroll
    for (int j = 0; j < DEMOD_BT; ++j) acc[j] = 0.f;
    for (int i = lane; i < C; i += 32) {
      const float w = __ldg(wr + i);
#pragma unroll
      for (int j = 0; j < DEMOD_BT; ++j) acc[j] = fmaf(s2[j * C + i], w, acc[j]);
    }
#pragma unroll
    for (int j = 0; j < DEMOD_BT; ++j) {
      const float t = warp_sum(acc[j]);
      if (lane == 0 && b0 + j < B) sinv[(int64_t)(b0 + j) * K + o] = rsqrtf(alpha2 * t + eps);
    }
  }
}

// gs[b,i] = gs1[b,i] + 2*s[b,i]*alpha^2 * sum_o dd[b,o]*wsq[o,i],  dd[b,o] = -0.5*Ad[b,o]*sinv[b,o]^2
// One block = DEMOD_BS samples x 128 input channels: every wsq element a thread loads serves DEMOD_BS samples (one block
// per sample re-read the whole wsq matrix B times through L2: 600 MB for a 768-channel layer at B = 256).
constexpr int DEMOD_BS = 8;
__global__ void __launch_bounds__(128) demod_bwd_s_kernel(const float* __restrict__ gs1, const float* __restrict__ s,
                                                          const float* __restrict__ Ad, const float* __restrict__ sinv,
                                                          const float* __restrict__ wsq, float* __restrict__ gs, int B,
                                                          int C, int K, float alpha2) {
  extern __shared__ float dd[];   // [DEMOD_BS][K]
  const int b0 = blockIdx.x * DEMOD_BS;
  for (int e = threadIdx.x; e < DEMOD_BS * K; e += 128) {
    const int j = e / K, o = e - j * K;
    float v = 0.f;
    if (b0 + j < B) {
      const float si = sinv[(int64_t)(b0 + j) * K + o];
      v = -0.5f * Ad[(int64_t)(b0 + j) * K + o] * si * si;
    }
    dd[e] = v;
  }
  __syncthreads();
  const int i = blockIdx.y * 128 + threadIdx.x;
  if (i >= C) return;
  float acc[DEMOD_BS];
#pragma unroll
  for (int j = 0; j < DEMOD_BS; ++j) acc[j] = 0.f;
  for (int o = 0; o < K; ++o) {
    const float w = __ldg(wsq + (int64_t)o * C + i);
#pragma unroll
    for (int j = 0; j < DEMOD_BS; ++j) acc[j] = fmaf(dd[j * K + o], w, acc[j]);
  }
#pragma unroll
  for (int j = 0; j < DEMOD_BS; ++j) {
    if (b0 + j < B) {
      const int64_t idx = (int64_t)(b0 + j) * C + i;
      gs[idx] = (gs1 ? gs1[idx] : 0.f) + 2.f * s[idx] * alpha2 * acc[j];
    }
  }
}

// gw[o,i,t] += 2*w[o,i,t]*alpha^2 * sum_b dd[b,o]*s[b,i]^2 ; one thread per (o,i)
__global__ void __launch_bounds__(256) demod_bwd_w_kernel(float* __restrict__ gw, const float* __restrict__ w,
                                                          const float* __restrict__ Ad, const float* __restrict__ sinv,
                                                          const float* __restrict__ s, int B, int C, int K, int RS,
                                                          float alpha2) {
  const int64_t KC = (int64_t)K * C;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < KC; idx += (int64_t)gridDim.x * 256) {
    const int o = (int)(idx / C), i = (int)(idx - (int64_t)o * C);
    float acc = 0.f;
    for (int b = 0; b < B; ++b) {
      float si = sinv[(int64_t)b * K + o], sv = s[(int64_t)b * C + i];
      acc = fmaf(-0.5f * Ad[(int64_t)b * K + o] * si * si, sv * sv, acc);
    }
    const float f = 2.f * alpha2 * acc;
    for (int t = 0; t < RS; ++t) gw[idx * RS + t] = fmaf(f, w[idx * RS + t], gw[idx * RS + t]);
  }
}

// ------------------------------------------------------------------------------------------------ tail forward
// u = y + ns[k]*z (in place on y when u == y);  a = PReLU(u) when slope != null.  Tensors are [N][K][HW], HW % 4 == 0.
__global__ void __launch_bounds__(256) style_tail_fwd_kernel(const float4* y, float4* u,
                                                             float4* __restrict__ a, const float* __restrict__ ns,
                                                             const float* __restrict__ slope, int slope_n, int K,
                                                             int hw4, int64_t groups,
                                                             const int64_t* __restrict__ state, uint32_t layer) {
  for (int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x; g < groups; g += (int64_t)gridDim.x * 256) {
    const int k = (int)((g / hw4) % K);
    float4 v = y[g];
    if (ns) {
      const float sc = ns[k];
      const float4 z = normal4(state, layer, (uint64_t)g);
      v.x = fmaf(sc, z.x, v.x); v.y = fmaf(sc, z.y, v.y); v.z = fmaf(sc, z.z, v.z); v.w = fmaf(sc, z.w, v.w);
      u[g] = v;
    }
    if (slope) {
      const float sl = slope[slope_n == 1 ? 0 : k];
      v.x = v.x > 0.f ? v.x : sl * v.x; v.y = v.y > 0.f ? v.y : sl * v.y;
      v.z = v.z > 0.f ? v.z : sl * v.z; v.w = v.w > 0.f ? v.w : sl * v.w;
      a[g] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------------ tail backward
// per row (n,k):  g_u = g_a * (u>0 ? 1 : slope)          (written when slope != null)
//                 Ad  = sum g_u * (u - ns*z)              = sum over hw of g_u * y_conv  (demodulation adjoint)
//                 Z   = sum g_u * z                       (d/d noise scale, summed over n by the caller)
//                 Pm  = sum g_a * min(u, 0)               (d/d PReLU slope, summed over n by the caller)
template <bool WARP_ROWS>
__global__ void __launch_bounds__(256) style_tail_bwd_kernel(const float4* __restrict__ ga, const float4* __restrict__ u,
                                                             float4* __restrict__ gu, const float* __restrict__ ns,
                                                             const float* __restrict__ slope, int slope_n,
                                                             float* __restrict__ Ad, float* __restrict__ Z,
                                                             float* __restrict__ Pm, int K, int hw4, int64_t rows,
                                                             const int64_t* __restrict__ state, uint32_t layer) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row_step = WARP_ROWS ? (int64_t)gridDim.x * 8 : gridDim.x;
  for (int64_t row = WARP_ROWS ? (int64_t)blockIdx.x * 8 + warp : blockIdx.x; row < rows; row += row_step) {
    const int k = (int)(row % K);
    const float sc = ns ? ns[k] : 0.f;
    const float sl = slope ? slope[slope_n == 1 ? 0 : k] : 1.f;
    float accA = 0.f, accZ = 0.f, accP = 0.f;
    const int t0 = WARP_ROWS ? lane : threadIdx.x, tstep = WARP_ROWS ? 32 : 256;
    for (int j = t0; j < hw4; j += tstep) {
      const int64_t g = row * hw4 + j;
      const float4 gv = ga[g], uv = u[g];
      float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ns) z = normal4(state, layer, (uint64_t)g);
      float4 o;
#define GB_TAIL(c)                                         \
  {                                                        \
    float m = uv.c > 0.f ? 1.f : sl;                       \
    o.c = gv.c * m;                                        \
    accA = fmaf(o.c, uv.c - sc * z.c, accA);               \
    accZ = fmaf(o.c, z.c, accZ);                           \
    accP = fmaf(gv.c, fminf(uv.c, 0.f), accP);             \
  }
      GB_TAIL(x) GB_TAIL(y) GB_TAIL(z) GB_TAIL(w)
#undef GB_TAIL
      if (slope) gu[g] = o;
    }
    if (WARP_ROWS) {
      accA = warp_sum(accA); accZ = warp_sum(accZ); accP = warp_sum(accP);
      if (lane == 0) {
        if (Ad) Ad[row] = accA;
        if (Z) Z[row] = accZ;
        if (Pm) Pm[row] = accP;
      }
    } else {
      accA = block_sum<256>(accA); accZ = block_sum<256>(accZ); accP = block_sum<256>(accP);
      if (threadIdx.x == 0) {
        if (Ad) Ad[row] = accA;
        if (Z) Z[row] = accZ;
        if (Pm) Pm[row] = accP;
      }
    }
  }
}

// gx = dxs * s[n,c] ;  gs1[n,c] = sum_hw dxs * x      (rows of HW, HW % 4 == 0)
template <bool WARP_ROWS>
__global__ void __launch_bounds__(256) scale_reduce_kernel(const float4* __restrict__ dxs, const float* __restrict__ s,
                                                           const float4* __restrict__ x, float4* __restrict__ gx,
                                                           float* __restrict__ gs1, int hw4, int64_t rows) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row_step = WARP_ROWS ? (int64_t)gridDim.x * 8 : gridDim.x;
  for (int64_t row = WARP_ROWS ? (int64_t)blockIdx.x * 8 + warp : blockIdx.x; row < rows; row += row_step) {
    const float sv = s[row];
    float acc = 0.f;
    const int t0 = WARP_ROWS ? lane : threadIdx.x, tstep = WARP_ROWS ? 32 : 256;
    for (int j = t0; j < hw4; j += tstep) {
      const int64_t g = row * hw4 + j;
      const float4 d = dxs[g];
      if (gs1) {
        const float4 xv = x[g];
        acc = fmaf(d.x, xv.x, fmaf(d.y, xv.y, fmaf(d.z, xv.z, fmaf(d.w, xv.w, acc))));
      }
      if (gx) gx[g] = make_float4(d.x * sv, d.y * sv, d.z * sv, d.w * sv);
    }
    if (gs1) {
      if (WARP_ROWS) {
        acc = warp_sum(acc);
        if (lane == 0) gs1[row] = acc;
      } else {
        acc = block_sum<256>(acc);
        if (threadIdx.x == 0) gs1[row] = acc;
      }
    }
  }
}

}  // namespace ganb200
using namespace ganb200;

#define GB_S(st) ((cudaStream_t)(st))

extern "C" int ganb200_noise_fill(float* z, int64_t n, const int64_t* rng_state, int layer, void* stream) {
  GB_REQUIRE(n % 4 == 0, GANB200_BAD_SHAPE, "noise_fill: n must be a multiple of 4");
  noise_fill_kernel<<<ew_grid(n / 4, 256, 2), 256, 0, GB_S(stream)>>>((float4*)z, n / 4, rng_state, (uint32_t)layer);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_demod(const float* s, const float* wsq, float* sinv, int B, int C, int K, float alpha, float eps,
                             void* stream) {
  const int bx = (B + DEMOD_BT - 1) / DEMOD_BT;
  int by = (2 * kNumSMs + bx - 1) / bx;
  by = std::max(1, std::min(by, (K + 7) / 8));
  const size_t smem = sizeof(float) * DEMOD_BT * C;
  constexpr size_t kDemodSmemMax = 160 * 1024;  // C <= 5120 (the widest concat of the final pair is 1590)
  GB_REQUIRE(smem <= kDemodSmemMax, GANB200_BAD_SHAPE, "demod: C too large for the shared-memory stage");
  static bool configured = false;
  if (!configured) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(demod_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDemodSmemMax));
    configured = true;
  }
  demod_kernel<<<dim3(bx, by), 256, smem, GB_S(stream)>>>(s, wsq, sinv, B, C, K, alpha * alpha, eps);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_demod_bwd_s(const float* gs1, const float* s, const float* Ad, const float* sinv,
                                   const float* wsq, float* gs, int B, int C, int K, float alpha, void* stream) {
  GB_REQUIRE((size_t)DEMOD_BS * K * sizeof(float) <= 48 * 1024, GANB200_BAD_SHAPE, "demod_bwd_s: more than 1536 output channels");
  dim3 grid((B + DEMOD_BS - 1) / DEMOD_BS, (C + 127) / 128);
  demod_bwd_s_kernel<<<grid, 128, sizeof(float) * DEMOD_BS * K, GB_S(stream)>>>(gs1, s, Ad, sinv, wsq, gs, B, C, K,
                                                                                alpha * alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_demod_bwd_w(float* gw, const float* w, const float* Ad, const float* sinv, const float* s, int B,
                                   int C, int K, int RS, float alpha, void* stream) {
  demod_bwd_w_kernel<<<ew_grid((int64_t)K * C, 256, 1), 256, 0, GB_S(stream)>>>(gw, w, Ad, sinv, s, B, C, K, RS,
                                                                                alpha * alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_style_tail_fwd(const float* y, float* u, float* a, const float* ns, const float* slope,
                                      int slope_n, int N, int K, int HW, const int64_t* rng_state, int layer,
                                      void* stream) {
  GB_REQUIRE(HW % 4 == 0, GANB200_BAD_SHAPE, "style_tail_fwd: H*W must be a multiple of 4");
  GB_REQUIRE(!ns || rng_state, GANB200_BAD_ARG, "style_tail_fwd: noise needs an rng state");
  const int64_t groups = (int64_t)N * K * (HW / 4);
  style_tail_fwd_kernel<<<ew_grid(groups, 256, 2), 256, 0, GB_S(stream)>>>(
      (const float4*)y, (float4*)u, (float4*)a, ns, slope, slope_n, K, HW / 4, groups, rng_state, (uint32_t)layer);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_style_tail_bwd(const float* ga, const float* u, float* gu, const float* ns, const float* slope,
                                      int slope_n, float* Ad, float* Z, float* Pm, int N, int K, int HW,
                                      const int64_t* rng_state, int layer, void* stream) {
  GB_REQUIRE(HW % 4 == 0, GANB200_BAD_SHAPE, "style_tail_bwd: H*W must be a multiple of 4");
  const int64_t rows = (int64_t)N * K;
  const int hw4 = HW / 4;
  if (hw4 >= 256) {
    int grid = (int)std::min<int64_t>(rows, (int64_t)kNumSMs * 32);
    style_tail_bwd_kernel<false><<<grid, 256, 0, GB_S(stream)>>>((const float4*)ga, (const float4*)u, (float4*)gu, ns,
                                                                 slope, slope_n, Ad, Z, Pm, K, hw4, rows, rng_state,
                                                                 (uint32_t)layer);
  } else {
    int grid = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)kNumSMs * 32);
    style_tail_bwd_kernel<true><<<grid, 256, 0, GB_S(stream)>>>((const float4*)ga, (const float4*)u, (float4*)gu, ns,
                                                                slope, slope_n, Ad, Z, Pm, K, hw4, rows, rng_state,
                                                                (uint32_t)layer);
  }
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_scale_reduce(const float* dxs, const float* s, const float* x, float* gx, float* gs1, int N,
                                    int C, int HW, void* stream) {
  GB_REQUIRE(HW % 4 == 0, GANB200_BAD_SHAPE, "scale_reduce: H*W must be a multiple of 4");
  const int64_t rows = (int64_t)N * C;
  const int hw4 = HW / 4;
  if (hw4 >= 256) {
    int grid = (int)std::min<int64_t>(rows, (int64_t)kNumSMs * 32);
    scale_reduce_kernel<false><<<grid, 256, 0, GB_S(stream)>>>((const float4*)dxs, s, (const float4*)x, (float4*)gx, gs1,
                                                               hw4, rows);
  } else {
    int grid = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)kNumSMs * 32);
    scale_reduce_kernel<true><<<grid, 256, 0, GB_S(stream)>>>((const float4*)dxs, s, (const float4*)x, (float4*)gx, gs1,
                                                              hw4, rows);
  }
  GB_CHECK_LAUNCH();
  return 0;
}
