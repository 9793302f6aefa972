// pointwise.cu — HBM-bound elementwise / broadcast / reduction kernels (fp32, 128-bit accesses when aligned).
// Roofline: HBM.  Algorithmic bytes = 4 B per element read + 4 B per element written, each tensor once.
#include "common.cuh"
#include "det_reduce.cuh"

namespace ganb200 {

thread_local char g_last_error[512] = {0};

// ------------------------------------------------------------------ unary
__device__ __forceinline__ float u_apply(int op, float x, float p0, float p1) {
  switch (op) {
    case GANB200_U_SIGMOID: return 1.f / (1.f + expf(-x));
    case GANB200_U_TANH: return tanhf(x);
    case GANB200_U_SQUARE: return x * x;
    case GANB200_U_AFFINE: return fmaf(p0, x, p1);
    case GANB200_U_RSQRT_EPS: return 1.f / sqrtf(x + p0);
    case GANB200_U_SQRT: return sqrtf(x);
    case GANB200_U_LEAKY: return x > 0.f ? x : p0 * x;
    case GANB200_U_RECIP: return 1.f / x;
    case GANB200_U_LEAKY_MASK: return x > 0.f ? 1.f : p0;
    case GANB200_U_COPY: return x;
    case GANB200_U_POW: return powf(x, p0);
    // exact GELU x * Phi(x) (nn.GELU(), generators/generator_6.py:12) and its first two derivatives
    case GANB200_U_GELU: return 0.5f * x * (1.f + erff(x * 0.70710678118654752f));
    case GANB200_U_GELU_GRAD:
      return 0.5f * (1.f + erff(x * 0.70710678118654752f)) + x * 0.39894228040143268f * expf(-0.5f * x * x);
    case GANB200_U_GELU_GRAD2: return (2.f - x * x) * 0.39894228040143268f * expf(-0.5f * x * x);
  }
  return x;
}

template <int OP>
__global__ void __launch_bounds__(256) ew_unary_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n,
                                                       float p0, float p1, bool vec) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (vec) {
    int64_t n4 = n >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    float4* y4 = reinterpret_cast<float4*>(y);
    for (int64_t i = tid; i < n4; i += stride) {
      float4 a = x4[i];
      a.x = u_apply(OP, a.x, p0, p1); a.y = u_apply(OP, a.y, p0, p1);
      a.z = u_apply(OP, a.z, p0, p1); a.w = u_apply(OP, a.w, p0, p1);
      y4[i] = a;
    }
    for (int64_t i = (n4 << 2) + tid; i < n; i += stride) y[i] = u_apply(OP, x[i], p0, p1);
  } else {
    for (int64_t i = tid; i < n; i += stride) y[i] = u_apply(OP, x[i], p0, p1);
  }
}

// ------------------------------------------------------------------ binary
__device__ __forceinline__ float b_apply(int op, float a, float b, float p0) {
  switch (op) {
    case GANB200_B_ADD: return a + b;
    case GANB200_B_SUB: return a - b;
    case GANB200_B_MUL: return a * b;
    case GANB200_B_DIV: return a / b;
    case GANB200_B_SIGMOID_BWD: return a * b * (1.f - b);
    case GANB200_B_TANH_BWD: return a * (1.f - b * b);
    case GANB200_B_LEAKY_BWD: return b > 0.f ? a : a * p0;
  }
  return a;
}

template <int OP>
__global__ void __launch_bounds__(256) ew_binary_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                        float* __restrict__ y, int64_t n, float p0, bool vec) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (vec) {
    int64_t n4 = n >> 2;
    const float4* a4 = reinterpret_cast<const float4*>(a);
    const float4* b4 = reinterpret_cast<const float4*>(b);
    float4* y4 = reinterpret_cast<float4*>(y);
    for (int64_t i = tid; i < n4; i += stride) {
      float4 u = a4[i], v = b4[i], r;
      r.x = b_apply(OP, u.x, v.x, p0); r.y = b_apply(OP, u.y, v.y, p0);
      r.z = b_apply(OP, u.z, v.z, p0); r.w = b_apply(OP, u.w, v.w, p0);
      y4[i] = r;
    }
    for (int64_t i = (n4 << 2) + tid; i < n; i += stride) y[i] = b_apply(OP, a[i], b[i], p0);
  } else {
    for (int64_t i = tid; i < n; i += stride) y[i] = b_apply(OP, a[i], b[i], p0);
  }
}

// ------------------------------------------------------------------ broadcast ops on [N,C,HW]
__device__ __forceinline__ float v_load(const float* __restrict__ v, int vmode, int n, int c, int C) {
  switch (vmode) {
    case GANB200_V_NC: return v[(int64_t)n * C + c];
    case GANB200_V_C: return v[c];
    case GANB200_V_N: return v[n];
    default: return v[0];
  }
}

// one block row per (n,c): grid-stride over rows, threads stride over HW.  z == nullptr -> plain op.
// mode: 0 add, 1 mul, 2 fma (x + v*z)
template <int MODE>
__global__ void __launch_bounds__(256) ew_bcast_kernel(const float* __restrict__ x, const float* __restrict__ v,
                                                       const float* __restrict__ z, float* __restrict__ y,
                                                       int N, int C, int HW, int vmode) {
  int64_t rows = (int64_t)N * C;
  if (HW >= 64) {
    bool vec = ((HW & 3) == 0) && aligned16(x) && aligned16(y) && (MODE != 2 || aligned16(z));
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
      int n = (int)(row / C), c = (int)(row % C);
      float s = v_load(v, vmode, n, c, C);
      const float* xr = x + row * HW;
      float* yr = y + row * HW;
      const float* zr = (MODE == 2) ? z + row * HW : nullptr;
      if (vec) {
        int hw4 = HW >> 2;
        for (int i = threadIdx.x; i < hw4; i += blockDim.x) {
          float4 a = reinterpret_cast<const float4*>(xr)[i], r;
          if (MODE == 0) { r.x = a.x + s; r.y = a.y + s; r.z = a.z + s; r.w = a.w + s; }
          else if (MODE == 1) { r.x = a.x * s; r.y = a.y * s; r.z = a.z * s; r.w = a.w * s; }
          else {
            float4 q = reinterpret_cast<const float4*>(zr)[i];
            r.x = fmaf(s, q.x, a.x); r.y = fmaf(s, q.y, a.y); r.z = fmaf(s, q.z, a.z); r.w = fmaf(s, q.w, a.w);
          }
          reinterpret_cast<float4*>(yr)[i] = r;
        }
      } else {
        for (int i = threadIdx.x; i < HW; i += blockDim.x) {
          float a = xr[i];
          yr[i] = MODE == 0 ? a + s : (MODE == 1 ? a * s : fmaf(s, zr[i], a));
        }
      }
    }
  } else {
    // small HW (BatchNorm1d-like, SE vectors): flat mapping
    int64_t total = rows * HW;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
      int64_t row = i / HW;
      int n = (int)(row / C), c = (int)(row % C);
      float s = v_load(v, vmode, n, c, C);
      float a = x[i];
      y[i] = MODE == 0 ? a + s : (MODE == 1 ? a * s : fmaf(s, z[i], a));
    }
  }
}

// ------------------------------------------------------------------ reductions of x*(y or 1) over [N,C,HW]
// rmode 0: one warp per (n,c) row.
__global__ void __launch_bounds__(256) reduce_rows_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                          float* __restrict__ out, int64_t rows, int HW, float scale) {
  int warps_per_block = blockDim.x >> 5;
  int lane = threadIdx.x & 31;
  for (int64_t row = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * warps_per_block) {
    const float* xr = x + row * HW;
    const float* yr = y ? y + row * HW : nullptr;
    float acc = 0.f;
    if (((HW & 3) == 0) && aligned16(x) && (!y || aligned16(y))) {
      int hw4 = HW >> 2;
      for (int i = lane; i < hw4; i += 32) {
        float4 a = reinterpret_cast<const float4*>(xr)[i];
        if (yr) {
          float4 b = reinterpret_cast<const float4*>(yr)[i];
          acc += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
        } else acc += a.x + a.y + a.z + a.w;
      }
    } else {
      for (int i = lane; i < HW; i += 32) acc += yr ? xr[i] * yr[i] : xr[i];
    }
    acc = warp_sum(acc);
    if (lane == 0) out[row] = acc * scale;
  }
}

// rmode 1 ([C] over n,hw) / rmode 2 ([N] over c,hw) / rmode 3 (all): grid (groups, splits); det_reduce.cuh sums the splits in order.
// element (g, j): rmode1: g=c, j over (n,hw): idx = (n*C + c)*HW + hw ;  rmode2: g=n, contiguous C*HW ; rmode3: g=0 all.
__global__ void __launch_bounds__(256) reduce_groups_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                            float* __restrict__ out, int N, int C, int HW, int rmode,
                                                            float scale, int region) {
  int g = blockIdx.x;
  int split = blockIdx.y, nsplit = gridDim.y;
  float acc = 0.f;
  if (rmode == 1) {
    // iterate over n in [split::nsplit], hw by threads
    if (HW >= 1024) {
      for (int n = split; n < N; n += nsplit) {
        const float* xr = x + ((int64_t)n * C + g) * HW;
        const float* yr = y ? y + ((int64_t)n * C + g) * HW : nullptr;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) acc += yr ? xr[i] * yr[i] : xr[i];
      }
    } else {
      // small planes: threads over (n, hw) pairs so that all lanes have work
      const int total = N * HW;
      for (int j = split * 256 + threadIdx.x; j < total; j += nsplit * 256) {
        const int n = j / HW, i = j - n * HW;
        const int64_t idx = ((int64_t)n * C + g) * HW + i;
        acc += y ? x[idx] * y[idx] : x[idx];
      }
    }
  } else {
    int64_t len = rmode == 2 ? (int64_t)C * HW : (int64_t)N * C * HW;
    const float* xr = x + (rmode == 2 ? (int64_t)g * len : 0);
    const float* yr = y ? y + (rmode == 2 ? (int64_t)g * len : 0) : nullptr;
    int64_t chunk = (len + nsplit - 1) / nsplit;
    int64_t lo = (int64_t)split * chunk, hi = lo + chunk < len ? lo + chunk : len;
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) acc += yr ? xr[i] * yr[i] : xr[i];
  }
  float t = block_sum<256>(acc);
  det_group_store<256, 1>(region, g, split, nsplit, scale, t, out + g);
}

}  // namespace ganb200

using namespace ganb200;

extern "C" int ganb200_abi_version(void) { return 1; }
extern "C" const char* ganb200_last_error(void) { return g_last_error; }

extern "C" int ganb200_device_check(int device) {
  cudaDeviceProp prop;
  GB_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(GANB200_UNSUPPORTED_ARCH, "device is sm_%lld%s, libganb200 needs sm_100", "", prop.major * 10 + prop.minor);
  return 0;
}

#define UNARY_CASE(OP) \
  case OP: ew_unary_kernel<OP><<<grid, 256, 0, s>>>(x, y, n, p0, p1, vec); break;

extern "C" int ganb200_ew_unary(int op, const float* x, float* y, int64_t n, float p0, float p1, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  bool vec = (((uintptr_t)x | (uintptr_t)y) & 15) == 0;
  int grid = ew_grid(n, 256, 8);
  switch (op) {
    UNARY_CASE(GANB200_U_SIGMOID) UNARY_CASE(GANB200_U_TANH) UNARY_CASE(GANB200_U_SQUARE)
    UNARY_CASE(GANB200_U_AFFINE) UNARY_CASE(GANB200_U_RSQRT_EPS) UNARY_CASE(GANB200_U_SQRT)
    UNARY_CASE(GANB200_U_LEAKY) UNARY_CASE(GANB200_U_RECIP) UNARY_CASE(GANB200_U_LEAKY_MASK)
    UNARY_CASE(GANB200_U_COPY) UNARY_CASE(GANB200_U_POW)
    UNARY_CASE(GANB200_U_GELU) UNARY_CASE(GANB200_U_GELU_GRAD) UNARY_CASE(GANB200_U_GELU_GRAD2)
    default: return fail(GANB200_BAD_ARG, "ew_unary: unknown op%s %lld", "", op);
  }
  GB_CHECK_LAUNCH();
  return 0;
}

#define BINARY_CASE(OP) \
  case OP: ew_binary_kernel<OP><<<grid, 256, 0, s>>>(a, b, y, n, p0, vec); break;

extern "C" int ganb200_ew_binary(int op, const float* a, const float* b, float* y, int64_t n, float p0, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  bool vec = (((uintptr_t)a | (uintptr_t)b | (uintptr_t)y) & 15) == 0;
  int grid = ew_grid(n, 256, 8);
  switch (op) {
    BINARY_CASE(GANB200_B_ADD) BINARY_CASE(GANB200_B_SUB) BINARY_CASE(GANB200_B_MUL) BINARY_CASE(GANB200_B_DIV)
    BINARY_CASE(GANB200_B_SIGMOID_BWD) BINARY_CASE(GANB200_B_TANH_BWD) BINARY_CASE(GANB200_B_LEAKY_BWD)
    default: return fail(GANB200_BAD_ARG, "ew_binary: unknown op%s %lld", "", op);
  }
  GB_CHECK_LAUNCH();
  return 0;
}

static int bcast_grid(int N, int C, int HW) {
  int64_t rows = (int64_t)N * C;
  if (HW >= 64) return (int)(rows < (int64_t)kNumSMs * 16 ? rows : (int64_t)kNumSMs * 16);
  return ew_grid(rows * HW, 256, 4);
}

extern "C" int ganb200_ew_bcast(int op, const float* x, const float* v, float* y, int N, int C, int HW, int vmode,
                                void* stream) {
  if ((int64_t)N * C * HW <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  int grid = bcast_grid(N, C, HW);
  if (op == GANB200_B_ADD) ew_bcast_kernel<0><<<grid, 256, 0, s>>>(x, v, nullptr, y, N, C, HW, vmode);
  else if (op == GANB200_B_MUL) ew_bcast_kernel<1><<<grid, 256, 0, s>>>(x, v, nullptr, y, N, C, HW, vmode);
  else return fail(GANB200_BAD_ARG, "ew_bcast: op must be ADD or MUL%s", "");
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_ew_fma_bcast(const float* x, const float* v, const float* z, float* y, int N, int C, int HW,
                                    int vmode, void* stream) {
  if ((int64_t)N * C * HW <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  ew_bcast_kernel<2><<<bcast_grid(N, C, HW), 256, 0, s>>>(x, v, z, y, N, C, HW, vmode);
  GB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ selective-kernel merge of two branches, fused
// out = a0 * t0 + a1 * t1 with per-(sample, channel) attention weights att[n][i][c] (generator_13_5.py:169,201,294 for
// M = 2).  As mul_bc + fma_bc the forward is 5 passes over a feature map and the backward 8 (g is read four times); here
// 3 and 5.  One block (planes of >= 1024 values) or one warp per (n, c) plane; the attention gradients are whole-plane sums
// taken inside the block / warp in a fixed order.
namespace ganb200 {

template <bool BWD>
__global__ void __launch_bounds__(256) sk_merge2_kernel(const float* __restrict__ g, const float* __restrict__ t0,
                                                        const float* __restrict__ t1, const float* __restrict__ att,
                                                        float* __restrict__ o0, float* __restrict__ o1,
                                                        float* __restrict__ gatt, int64_t rows, int C, int HW,
                                                        bool block_rows) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t step = block_rows ? gridDim.x : (int64_t)gridDim.x * 8;
  const int t_first = block_rows ? threadIdx.x : lane, t_step = block_rows ? 256 : 32;
  const bool vec = (HW & 3) == 0;
  for (int64_t row = block_rows ? blockIdx.x : (int64_t)blockIdx.x * 8 + warp; row < rows; row += step) {
    const int64_t n = row / C;
    const int c = (int)(row - n * C);
    const float a0 = __ldg(att + (n * 2 + 0) * C + c), a1 = __ldg(att + (n * 2 + 1) * C + c);
    const int64_t base = row * HW;
    float s0 = 0.f, s1 = 0.f;
    if (vec) {
      const float4* p0 = reinterpret_cast<const float4*>(t0 + base);
      const float4* p1 = reinterpret_cast<const float4*>(t1 + base);
      const float4* pg = BWD ? reinterpret_cast<const float4*>(g + base) : nullptr;
      float4* q0 = reinterpret_cast<float4*>(o0 + base);
      float4* q1 = BWD ? reinterpret_cast<float4*>(o1 + base) : nullptr;
      for (int i = t_first; i < (HW >> 2); i += t_step) {
        const float4 u = __ldg(p0 + i), v = __ldg(p1 + i);
        if (BWD) {
          const float4 w = __ldg(pg + i);
          s0 += w.x * u.x + w.y * u.y + w.z * u.z + w.w * u.w;
          s1 += w.x * v.x + w.y * v.y + w.z * v.z + w.w * v.w;
          if (o0) {   // null: attention gradient only (the feature-map gradients are formed by sk_bwd_finish_kernel)
            q0[i] = make_float4(a0 * w.x, a0 * w.y, a0 * w.z, a0 * w.w);
            q1[i] = make_float4(a1 * w.x, a1 * w.y, a1 * w.z, a1 * w.w);
          }
        } else {
          q0[i] = make_float4(fmaf(a1, v.x, a0 * u.x), fmaf(a1, v.y, a0 * u.y), fmaf(a1, v.z, a0 * u.z),
                              fmaf(a1, v.w, a0 * u.w));
        }
      }
    } else {
      for (int i = t_first; i < HW; i += t_step) {
        const float u = t0[base + i], v = t1[base + i];
        if (BWD) {
          const float w = g[base + i];
          s0 = fmaf(w, u, s0);
          s1 = fmaf(w, v, s1);
          if (o0) {
            o0[base + i] = a0 * w;
            o1[base + i] = a1 * w;
          }
        } else {
          o0[base + i] = fmaf(a1, v, a0 * u);
        }
      }
    }
    if (BWD) {
      if (block_rows) {
        s0 = block_sum<256>(s0);
        s1 = block_sum<256>(s1);
        if (threadIdx.x == 0) { gatt[(n * 2 + 0) * C + c] = s0; gatt[(n * 2 + 1) * C + c] = s1; }
      } else {
        s0 = warp_sum(s0);
        s1 = warp_sum(s1);
        if (lane == 0) { gatt[(n * 2 + 0) * C + c] = s0; gatt[(n * 2 + 1) * C + c] = s1; }
      }
    }
  }
}

}  // namespace ganb200

static bool sk_aligned(const void* a, const void* b, const void* c, const void* d, const void* e) {
  return ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)c) | ((uintptr_t)d) | ((uintptr_t)e)) & 15) == 0;
}

extern "C" int ganb200_sk_merge2_fwd(const float* t0, const float* t1, const float* att, float* out, int N, int C, int HW,
                                     void* stream) {
  const int64_t rows = (int64_t)N * C;
  if (rows * HW <= 0) return 0;
  GB_REQUIRE(t0 && t1 && att && out, GANB200_BAD_ARG, "sk_merge2_fwd: null argument");
  GB_REQUIRE((HW & 3) != 0 || sk_aligned(t0, t1, out, nullptr, nullptr), GANB200_BAD_ARG, "sk_merge2_fwd: unaligned tensor");
  const bool block_rows = HW >= 1024;
  const int grid = (int)std::min<int64_t>(block_rows ? rows : (rows + 7) / 8, (int64_t)kNumSMs * 32);
  sk_merge2_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(nullptr, t0, t1, att, out, nullptr, nullptr, rows, C, HW,
                                                                  block_rows);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_sk_merge2_bwd(const float* g, const float* t0, const float* t1, const float* att, float* g0,
                                     float* g1, float* gatt, int N, int C, int HW, void* stream) {
  const int64_t rows = (int64_t)N * C;
  if (rows * HW <= 0) return 0;
  GB_REQUIRE(g && t0 && t1 && att && gatt && ((g0 == nullptr) == (g1 == nullptr)), GANB200_BAD_ARG,
             "sk_merge2_bwd: null argument");
  GB_REQUIRE((HW & 3) != 0 || sk_aligned(g, t0, t1, g0, g1), GANB200_BAD_ARG, "sk_merge2_bwd: unaligned tensor");
  const bool block_rows = HW >= 1024;
  const int grid = (int)std::min<int64_t>(block_rows ? rows : (rows + 7) / 8, (int64_t)kNumSMs * 32);
  sk_merge2_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(g, t0, t1, att, g0, g1, gatt, rows, C, HW, block_rows);
  GB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ SK attention input: AdaptiveAvgPool2d(OUT) of t0 + t1
// (generator_13_5.py:77-80: fea_U = sum of the branches, then the 5x5 pooling).  As add + pooling it is 3 passes over a
// feature map plus a pooling launch that re-reads the sum; here both branches are read once.  One thread per column of a
// plane (W threads per plane, 256 / W planes per block): per output row bin the column sums of that bin's input rows, then
// the OUT x OUT outputs from the column sums in shared memory.  Bins as torch: [floor(o n / OUT), ceil((o + 1) n / OUT)).
namespace ganb200 {

template <int OUT>
__global__ void __launch_bounds__(256) pool_sum2_kernel(const float* __restrict__ t0, const float* __restrict__ t1,
                                                        float* __restrict__ out, int64_t planes, int H, int W) {
  __shared__ float colsum[256 * OUT];          // [plane in block][OUT][W]
  const int ppb = 256 / W;
  const int pl = threadIdx.x / W, w = threadIdx.x - pl * W;
  const int HW = H * W;
  for (int64_t p0 = (int64_t)blockIdx.x * ppb; p0 < planes; p0 += (int64_t)gridDim.x * ppb) {
    const int64_t plane = p0 + pl;
    const bool ok = plane < planes;
    const float* a = t0 + plane * HW + w;
    const float* b = t1 + plane * HW + w;
#pragma unroll
    for (int o = 0; o < OUT; ++o) {
      const int hs = (o * H) / OUT, he = ((o + 1) * H + OUT - 1) / OUT;
      float acc = 0.f;
      if (ok) {
#pragma unroll 4
        for (int h = hs; h < he; ++h) acc += __ldg(a + h * W) + __ldg(b + h * W);
      }
      colsum[(pl * OUT + o) * W + w] = acc / (float)(he - hs);
    }
    __syncthreads();
    if (ok && w < OUT * OUT) {
      const int o = w / OUT, j = w - o * OUT;
      const int ws = (j * W) / OUT, we = ((j + 1) * W + OUT - 1) / OUT;
      const float* row = colsum + (pl * OUT + o) * W;
      float acc = 0.f;
      for (int c = ws; c < we; ++c) acc += row[c];
      out[plane * (OUT * OUT) + w] = acc / (float)(we - ws);
    }
    __syncthreads();
  }
}

}  // namespace ganb200

// g_i = att[n][i][c] * g + P^T(gp) for both branches in one pass: the merge's feature-map gradients and the adjoint of
// pool_sum2 (the same for both branches) summed where they are produced, instead of two launches plus two accumulations
// by the autograd engine (12 passes over a feature map -> 6 with the attention-gradient pass of sk_merge2_kernel).
// One block per plane; the OUT x OUT pooled gradients, pre-divided by their bin areas, sit in shared memory.
namespace ganb200 {
template <int OUT>
__global__ void __launch_bounds__(256) sk_bwd_finish_kernel(const float* __restrict__ g, const float* __restrict__ att,
                                                            const float* __restrict__ gp, float* __restrict__ g0,
                                                            float* __restrict__ g1, int64_t planes, int C, int H, int W) {
  __shared__ float sgp[OUT * OUT];
  __shared__ int lo[2][64], hi[2][64];       // first / last bin of every row (0) and column (1); H, W <= 64
  const int HW = H * W, Wq = W >> 2;
  for (int i = threadIdx.x; i < H + W; i += 256) {
    const int d = i < H ? 0 : 1, n = d ? W : H, p = d ? i - H : i;
    int a = OUT, b = -1;
    for (int o = 0; o < OUT; ++o) {
      const int s = (o * n) / OUT, e = ((o + 1) * n + OUT - 1) / OUT;
      if (p >= s && p < e) { a = min(a, o); b = max(b, o); }
    }
    lo[d][p] = a; hi[d][p] = b;
  }
  for (int64_t plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int64_t n = plane / C;
    const int c = (int)(plane - n * C);
    const float a0 = __ldg(att + (n * 2 + 0) * C + c), a1 = __ldg(att + (n * 2 + 1) * C + c);
    __syncthreads();
    if (threadIdx.x < OUT * OUT) {
      const int o = threadIdx.x / OUT, j = threadIdx.x - o * OUT;
      const int hs = (o * H) / OUT, he = ((o + 1) * H + OUT - 1) / OUT;
      const int ws = (j * W) / OUT, we = ((j + 1) * W + OUT - 1) / OUT;
      sgp[threadIdx.x] = __ldg(gp + plane * (OUT * OUT) + threadIdx.x) / (float)((he - hs) * (we - ws));
    }
    __syncthreads();
    const float4* pg = reinterpret_cast<const float4*>(g + plane * HW);
    float4* q0 = reinterpret_cast<float4*>(g0 + plane * HW);
    float4* q1 = reinterpret_cast<float4*>(g1 + plane * HW);
    for (int q = threadIdx.x; q < (HW >> 2); q += 256) {
      const int h = q / Wq, w0 = (q - h * Wq) << 2;
      const float4 v = __ldg(pg + q);
      float up[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float t = 0.f;
        for (int o = lo[0][h]; o <= hi[0][h]; ++o)
          for (int j = lo[1][w0 + k]; j <= hi[1][w0 + k]; ++j) t += sgp[o * OUT + j];
        up[k] = t;
      }
      q0[q] = make_float4(fmaf(a0, v.x, up[0]), fmaf(a0, v.y, up[1]), fmaf(a0, v.z, up[2]), fmaf(a0, v.w, up[3]));
      q1[q] = make_float4(fmaf(a1, v.x, up[0]), fmaf(a1, v.y, up[1]), fmaf(a1, v.z, up[2]), fmaf(a1, v.w, up[3]));
    }
  }
}
}  // namespace ganb200

extern "C" int ganb200_sk_bwd_finish(const float* g, const float* att, const float* gp, float* g0, float* g1, int N, int C,
                                     int H, int W, int out_size, void* stream) {
  const int64_t planes = (int64_t)N * C;
  if (planes <= 0) return 0;
  GB_REQUIRE(g && att && gp && g0 && g1, GANB200_BAD_ARG, "sk_bwd_finish: null argument");
  GB_REQUIRE(out_size == 5 && H == W && (W == 32 || W == 64), GANB200_BAD_SHAPE, "sk_bwd_finish: 32x32 or 64x64 planes, 5x5 pooling");
  GB_REQUIRE(((((uintptr_t)g) | ((uintptr_t)g0) | ((uintptr_t)g1)) & 15) == 0, GANB200_BAD_ARG, "sk_bwd_finish: unaligned tensor");
  const int grid = (int)std::min<int64_t>(planes, (int64_t)kNumSMs * 32);
  sk_bwd_finish_kernel<5><<<grid, 256, 0, (cudaStream_t)stream>>>(g, att, gp, g0, g1, planes, C, H, W);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_pool_sum2_eligible(int H, int W, int out_size) {
  return out_size == 5 && H == W && (W == 32 || W == 64);   // W threads per plane must cover the 25 outputs
}

extern "C" int ganb200_pool_sum2(const float* t0, const float* t1, float* out, int64_t planes, int H, int W, int out_size,
                                 void* stream) {
  if (planes <= 0) return 0;
  GB_REQUIRE(t0 && t1 && out, GANB200_BAD_ARG, "pool_sum2: null argument");
  GB_REQUIRE(ganb200_pool_sum2_eligible(H, W, out_size), GANB200_BAD_SHAPE, "pool_sum2: 32x32 or 64x64 planes to 5x5");
  const int ppb = 256 / W;
  const int grid = (int)std::min<int64_t>((planes + ppb - 1) / ppb, (int64_t)kNumSMs * 32);
  pool_sum2_kernel<5><<<grid, 256, 0, (cudaStream_t)stream>>>(t0, t1, out, planes, H, W);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_reduce(const float* x, const float* y, float* out, int N, int C, int HW, int rmode, float scale,
                              void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  GB_REQUIRE(N > 0 && C > 0 && HW > 0, GANB200_BAD_SHAPE, "reduce: empty shape");
  if (rmode == 0) {
    int64_t rows = (int64_t)N * C;
    int grid = (int)((rows + 7) / 8);
    if (grid > kNumSMs * 16) grid = kNumSMs * 16;
    reduce_rows_kernel<<<grid, 256, 0, s>>>(x, y, out, rows, HW, scale);
  } else if (rmode >= 1 && rmode <= 3) {
    int groups = rmode == 1 ? C : (rmode == 2 ? N : 1);
    int64_t per_group = (int64_t)N * C * HW / groups;
    int max_split = rmode == 1 ? (HW >= 1024 ? N : (int)(((int64_t)N * HW + 255) / 256)) : (int)((per_group + 2047) / 2048);
    int want = (kNumSMs * 4 + groups - 1) / groups;
    int nsplit = want < max_split ? want : max_split;
    if (nsplit < 1) nsplit = 1;
    const int region = det_region_for(s, groups, nsplit);
    if (region < 0) GB_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * groups, s));
    reduce_groups_kernel<<<dim3(groups, nsplit), 256, 0, s>>>(x, y, out, N, C, HW, rmode, scale, region);
  } else {
    return fail(GANB200_BAD_ARG, "reduce: bad rmode%s %lld", "", rmode);
  }
  GB_CHECK_LAUNCH();
  return 0;
}
