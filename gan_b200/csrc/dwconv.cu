// dwconv.cu — depthwise R x R (R = 3, 5, 7) convolution family (fprop / dgrad / wgrad) and channel shuffle, fp32 NCHW.
// Used by discriminators/discriminator_7_wgangpsn.py:55,61,72 (ShuffleNetV2 units: Conv2d(C,C,3,stride,padding=1,
// groups=C)), discriminator_6_wgangp.py:12-13 (bottlenecks with depthwise 3x3 / 5x5 / 7x7, reflect-padded by the caller)
// and :27-41 (channel_shuffle).  One multiply-add per 4 bytes moved: these are HBM-bound streaming kernels
// (algorithmic bytes = 4 B read per input element + 4 B written per output element; the 3x3 window re-reads hit L1),
// they stay on CUDA cores.  The three kernels are each other's derivatives (bilinear map), so the autograd
// Functions in gan_b200/ops.py are closed under double backward exactly as for the dense convs.
#include "common.cuh"

namespace ganb200 {

// y[n,c,p,q] = alpha * sum_{r,s} w[c,r,s] * x[n,c,p*st-pad+r, q*st-pad+s] + bias[c]
template <int ST, int R>
__global__ void __launch_bounds__(256) dw3_fprop_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ bias, float* __restrict__ y, int NC,
                                                        int C, int H, int W, int P, int Q, int pad, float alpha) {
  grid_stride((int64_t)NC * P * Q, [&](auto i) {
    int q = (int)(i % Q);
    auto t = i / Q;
    int p = (int)(t % P);
    int nc = (int)(t / P);
    int c = nc % C;
    const float* xp = x + (int64_t)nc * H * W;
    const float* wp = w + c * (R * R);
    int h0 = p * ST - pad, w0 = q * ST - pad;
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      int h = h0 + r;
      if ((unsigned)h >= (unsigned)H) continue;
#pragma unroll
      for (int s = 0; s < R; ++s) {
        int ww = w0 + s;
        if ((unsigned)ww < (unsigned)W) acc = fmaf(__ldg(wp + r * R + s), __ldg(xp + h * W + ww), acc);
      }
    }
    y[i] = alpha * acc + (bias ? __ldg(bias + c) : 0.f);
  });
}

// dx[n,c,h,w] = alpha * sum_{r,s} w[c,r,s] * dy[n,c,(h+pad-r)/st,(w+pad-s)/st]   (where divisible and in range)
template <int ST, int R>
__global__ void __launch_bounds__(256) dw3_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                        float* __restrict__ dx, int NC, int C, int H, int W, int P,
                                                        int Q, int pad, float alpha) {
  grid_stride((int64_t)NC * H * W, [&](auto i) {
    int ww = (int)(i % W);
    auto t = i / W;
    int h = (int)(t % H);
    int nc = (int)(t / H);
    int c = nc % C;
    const float* gp = dy + (int64_t)nc * P * Q;
    const float* wp = w + c * (R * R);
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      int ph = h + pad - r;
      if (ph < 0 || (ST == 2 && (ph & 1))) continue;
      int p = ph / ST;
      if (p >= P) continue;
#pragma unroll
      for (int s = 0; s < R; ++s) {
        int pw = ww + pad - s;
        if (pw < 0 || (ST == 2 && (pw & 1))) continue;
        int q = pw / ST;
        if (q < Q) acc = fmaf(__ldg(wp + r * R + s), __ldg(gp + p * Q + q), acc);
      }
    }
    dx[i] = alpha * acc;
  });
}

// partial[split][c][R*R] = sum over the split's samples and all (p,q) of dy[n,c,p,q] * x[n,c,p*st-pad+r,q*st-pad+s]
// grid (C, splits); a warp walks output rows so both reads are coalesced.
template <int ST, int R>
__global__ void __launch_bounds__(256) dw3_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                        float* __restrict__ partial, int N, int C, int H, int W, int P,
                                                        int Q, int pad) {
  const int c = blockIdx.x, split = blockIdx.y, splits = gridDim.y;
  float acc[R * R];
#pragma unroll
  for (int k = 0; k < R * R; ++k) acc[k] = 0.f;
  const int PQ = P * Q;
  for (int n = split; n < N; n += splits) {
    const float* xp = x + ((int64_t)n * C + c) * H * W;
    const float* gp = dy + ((int64_t)n * C + c) * PQ;
    for (int i = threadIdx.x; i < PQ; i += 256) {
      int p = i / Q, q = i - p * Q;
      float g = __ldg(gp + i);
      int h0 = p * ST - pad, w0 = q * ST - pad;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        int h = h0 + r;
        if ((unsigned)h >= (unsigned)H) continue;
#pragma unroll
        for (int s = 0; s < R; ++s) {
          int ww = w0 + s;
          if ((unsigned)ww < (unsigned)W) acc[r * R + s] = fmaf(g, __ldg(xp + h * W + ww), acc[r * R + s]);
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < R * R; ++k) {
    float t = block_sum<256>(acc[k]);
    if (threadIdx.x == 0) partial[((int64_t)split * C + c) * (R * R) + k] = t;
  }
}

// dw[c][k] = alpha * sum_split partial[split][c][k]   (fixed order: deterministic)
__global__ void __launch_bounds__(256) dw3_wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw,
                                                               int C9, int splits, float alpha) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C9) return;
  float acc = 0.f;
  for (int s = 0; s < splits; ++s) acc += partial[(int64_t)s * C9 + i];
  dw[i] = alpha * acc;
}

// channel shuffle: y[n, j*G + i] = x[n, i*(C/G) + j]
__global__ void __launch_bounds__(256) chan_shuffle_kernel(const float* __restrict__ x, float* __restrict__ y, int N,
                                                           int C, int HW4, int G, bool vec) {
  int cpg = C / G;
  grid_stride((int64_t)N * C * HW4, [&](auto i) {
    int e = (int)(i % HW4);
    auto t = i / HW4;
    int co = (int)(t % C);
    int n = (int)(t / C);
    int ci = (co % G) * cpg + co / G;
    int64_t src = ((int64_t)n * C + ci) * HW4 + e;
    if (vec)
      reinterpret_cast<float4*>(y)[i] = __ldg(reinterpret_cast<const float4*>(x) + src);
    else
      y[i] = __ldg(x + src);
  });
}

static int dw_check(int N, int C, int H, int W, int R, int S, int stride, int pad, int* P, int* Q) {
  GB_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0, GANB200_BAD_SHAPE, "dwconv: empty tensor");
  GB_REQUIRE(R == S && (R == 3 || R == 5 || R == 7), GANB200_BAD_SHAPE, "dwconv: square 3x3, 5x5 or 7x7 kernels only");
  GB_REQUIRE(stride == 1 || stride == 2, GANB200_BAD_SHAPE, "dwconv: stride must be 1 or 2");
  GB_REQUIRE(pad >= 0 && pad <= R / 2 + 1, GANB200_BAD_SHAPE, "dwconv: pad must be 0 .. R/2 + 1");
  GB_REQUIRE(H + 2 * pad >= R && W + 2 * pad >= S, GANB200_BAD_SHAPE, "dwconv: empty output");
  *P = (H + 2 * pad - R) / stride + 1;
  *Q = (W + 2 * pad - S) / stride + 1;
  GB_REQUIRE(*P > 0 && *Q > 0, GANB200_BAD_SHAPE, "dwconv: empty output");
  return 0;
}

}  // namespace ganb200
using namespace ganb200;

// kernel<stride, R> for stride in {1, 2}, R in {3, 5, 7} (dw_check has admitted nothing else)
#define GB_DW_DISPATCH(KERNEL, GRID, STREAM, ...)                                                      \
  do {                                                                                                 \
    if (stride == 1) {                                                                                 \
      if (R == 3) KERNEL<1, 3><<<GRID, 256, 0, STREAM>>>(__VA_ARGS__);                                  \
      else if (R == 5) KERNEL<1, 5><<<GRID, 256, 0, STREAM>>>(__VA_ARGS__);                             \
      else KERNEL<1, 7><<<GRID, 256, 0, STREAM>>>(__VA_ARGS__);                                         \
    } else {                                                                                           \
      if (R == 3) KERNEL<2, 3><<<GRID, 256, 0, STREAM>>>(__VA_ARGS__);                                  \
      else if (R == 5) KERNEL<2, 5><<<GRID, 256, 0, STREAM>>>(__VA_ARGS__);                             \
      else KERNEL<2, 7><<<GRID, 256, 0, STREAM>>>(__VA_ARGS__);                                         \
    }                                                                                                  \
  } while (0)

static int dw_splits(int N, int C) {
  int splits = (4 * kNumSMs + C - 1) / C;
  if (splits > N) splits = N;
  return splits < 1 ? 1 : splits;
}

extern "C" int ganb200_dwconv2d_fprop(const float* x, const float* w, const float* bias, float* y, int N, int C, int H,
                                      int W, int R, int S, int stride, int pad, float alpha, void* stream) {
  int P, Q;
  if (int rc = dw_check(N, C, H, W, R, S, stride, pad, &P, &Q)) return rc;
  int64_t total = (int64_t)N * C * P * Q;
  int grid = ew_grid(total, 256, 2);
  cudaStream_t s = (cudaStream_t)stream;
  GB_DW_DISPATCH(dw3_fprop_kernel, grid, s, x, w, bias, y, N * C, C, H, W, P, Q, pad, alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_dwconv2d_dgrad(const float* dy, const float* w, float* dx, int N, int C, int H, int W, int R,
                                      int S, int stride, int pad, float alpha, void* stream) {
  int P, Q;
  if (int rc = dw_check(N, C, H, W, R, S, stride, pad, &P, &Q)) return rc;
  int64_t total = (int64_t)N * C * H * W;
  int grid = ew_grid(total, 256, 2);
  cudaStream_t s = (cudaStream_t)stream;
  GB_DW_DISPATCH(dw3_dgrad_kernel, grid, s, dy, w, dx, N * C, C, H, W, P, Q, pad, alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int64_t ganb200_dwconv2d_wgrad_workspace_bytes(int N, int C) {   // 3x3 kernels
  return (int64_t)dw_splits(N, C) * C * 9 * sizeof(float);
}
extern "C" int64_t ganb200_dwconv2d_wgrad_workspace_bytes_rs(int N, int C, int R, int S) {
  return (int64_t)dw_splits(N, C) * C * R * S * sizeof(float);
}

extern "C" int ganb200_dwconv2d_wgrad(const float* x, const float* dy, float* dw, void* workspace,
                                      int64_t workspace_bytes, int N, int C, int H, int W, int R, int S, int stride,
                                      int pad, float alpha, void* stream) {
  int P, Q;
  if (int rc = dw_check(N, C, H, W, R, S, stride, pad, &P, &Q)) return rc;
  int64_t need = ganb200_dwconv2d_wgrad_workspace_bytes_rs(N, C, R, S);
  GB_REQUIRE(workspace != nullptr && workspace_bytes >= need, GANB200_BAD_SHAPE, "dwconv wgrad: workspace too small");
  int splits = dw_splits(N, C);
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid(C, splits);
  float* partial = (float*)workspace;
  GB_DW_DISPATCH(dw3_wgrad_kernel, grid, s, x, dy, partial, N, C, H, W, P, Q, pad);
  dw3_wgrad_reduce_kernel<<<ceil_div(C * R * S, 256), 256, 0, s>>>(partial, dw, C * R * S, splits, alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the depthwise wgrad (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA-graph capture, tight serving loops, harnesses that
// forbid allocation).  The split partials live in a file-scope device array instead of a caller-provided workspace.
// Every GPU that loads the library holds the array (256 KiB) until the process exits.  All calls share it: two calls
// that overlap on the GPU (different streams) race; calls on one stream are ordered and safe.  Nothing has to start at
// zero: dw3_wgrad_kernel writes every partial[split][c][k] that dw3_wgrad_reduce_kernel of the same call reads, so the
// previous call's contents are simply overwritten.
constexpr int64_t kDwWgradStaticFloats = 65536;   // 256 KiB: splits * C * 9 floats, i.e. C up to ~6 600 channels
static __device__ __align__(256) float g_dw_wgrad_partials[kDwWgradStaticFloats];

// The array's name is a host-side stand-in in host code, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static int dw_static_partials(float** out) {
  constexpr int kMaxDev = 64;
  static float* ptr[kMaxDev] = {nullptr};
  int dev = -1;
  GB_CHECK_CUDA(cudaGetDevice(&dev));
  GB_REQUIRE(dev >= 0 && dev < kMaxDev, GANB200_BAD_ARG, "dwconv wgrad (static scratch): device index out of range");
  if (ptr[dev] == nullptr) {
    void* p = nullptr;
    GB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_dw_wgrad_partials));
    GB_REQUIRE(p != nullptr, GANB200_CUDA_ERROR, "dwconv wgrad (static scratch): null symbol address");
    ptr[dev] = (float*)p;   // same value from any thread: a benign race
  }
  *out = ptr[dev];
  return 0;
}

extern "C" int64_t ganb200_dwconv2d_wgrad_static_max_bytes(void) { return kDwWgradStaticFloats * (int64_t)sizeof(float); }

extern "C" int ganb200_dwconv2d_wgrad_static(const float* x, const float* dy, float* dw, int N, int C, int H, int W,
                                             int R, int S, int stride, int pad, float alpha, void* stream) {
  int P, Q;
  if (int rc = dw_check(N, C, H, W, R, S, stride, pad, &P, &Q)) return rc;
  const int64_t need = ganb200_dwconv2d_wgrad_workspace_bytes_rs(N, C, R, S);
  if (need > kDwWgradStaticFloats * (int64_t)sizeof(float))
    return fail(GANB200_WORKSPACE_TOO_SMALL,
                "dwconv wgrad (static scratch): this shape needs %s%lld bytes of split partials, the built-in scratch "
                "holds %lld; use ganb200_dwconv2d_wgrad with a caller workspace", "", (long long)need,
                (long long)(kDwWgradStaticFloats * sizeof(float)));
  float* partial = nullptr;
  if (int rc = dw_static_partials(&partial)) return rc;
  const int splits = dw_splits(N, C);
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid(C, splits);
  GB_DW_DISPATCH(dw3_wgrad_kernel, grid, s, x, dy, partial, N, C, H, W, P, Q, pad);
  dw3_wgrad_reduce_kernel<<<ceil_div(C * R * S, 256), 256, 0, s>>>(partial, dw, C * R * S, splits, alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

extern "C" int ganb200_chan_shuffle(const float* x, float* y, int N, int C, int HW, int groups, void* stream) {
  if ((int64_t)N * C * HW <= 0) return 0;
  GB_REQUIRE(groups >= 1 && C % groups == 0, GANB200_BAD_SHAPE, "chan_shuffle: C must be a multiple of groups");
  bool vec = ((HW & 3) == 0) && ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0;
  int HW4 = vec ? HW / 4 : HW;
  int64_t total = (int64_t)N * C * HW4;
  chan_shuffle_kernel<<<ew_grid(total, 256, 2), 256, 0, (cudaStream_t)stream>>>(x, y, N, C, HW4, groups, vec);
  GB_CHECK_LAUNCH();
  return 0;
}
