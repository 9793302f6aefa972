// conv_simt.cu — fp32 CUDA-core implicit-GEMM convolution: fprop / dgrad / wgrad (engine 1).
// This is the exact-fp32 engine (parity <= 1e-5 tests, odd shapes, strided dgrad); the bf16 tcgen05 engine
// lives in conv_tc.cu.  One tiled GEMM core, three "problem" functors that gather operands on the fly
// (no im2col buffer, replicate padding folded into the index clamp).
#include "common.cuh"
#include "conv_common.cuh"
#include "det_reduce.cuh"

namespace ganb200 {

constexpr int BM = 128, BN = 64, BK = 16;
constexpr int PADA = 1, PADB = 4;

// thread layout for compute: tr = tid % 16 owns rows tr + 16*i (i<8); tc = tid / 16 owns cols tc*4 + j (j<4)
template <class P>
__global__ void __launch_bounds__(256) simt_gemm_kernel(P p) {
  __shared__ float As[BK][BM + PADA];
  __shared__ __align__(16) float Bs[BK][BN + PADB];
  const int tid = threadIdx.x;
  const int tile_m0 = blockIdx.x * BM, tile_n0 = blockIdx.y * BN;
  const int Kd = p.Kdim();
  // split-K over blockIdx.z (only wgrad uses gridDim.z > 1)
  int k_tiles = (Kd + BK - 1) / BK;
  int per = (k_tiles + gridDim.z - 1) / gridDim.z;
  int kbeg = blockIdx.z * per * BK;
  int kend = min(Kd, (int)(blockIdx.z + 1) * per * BK);

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  typename P::ACtx actx;
  typename P::BCtx bctx;
  p.init_a(actx, tile_m0, tid);
  p.init_b(bctx, tile_n0, tid);

  float ra[8], rb[4];
  if (kbeg < kend) {
    p.load_a(ra, actx, kbeg, tid);
    p.load_b(rb, bctx, kbeg, tid);
  }
  const int tr = tid & 15, tc = tid >> 4;
  for (int k0 = kbeg; k0 < kend; k0 += BK) {
    // registers -> smem
    if (P::A_KFAST) {
      int kl = tid & 15, ml = tid >> 4;
#pragma unroll
      for (int j = 0; j < 8; ++j) As[kl][ml + 16 * j] = ra[j];
    } else {
      int ml = tid & 127, kl = tid >> 7;
#pragma unroll
      for (int j = 0; j < 8; ++j) As[kl + 2 * j][ml] = ra[j];
    }
    {
      int kl = tid & 15, nl = tid >> 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) Bs[kl][nl + 16 * j] = rb[j];
    }
    __syncthreads();
    if (k0 + BK < kend) {
      p.load_a(ra, actx, k0 + BK, tid);
      p.load_b(rb, bctx, k0 + BK, tid);
    }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = As[k][tr + 16 * i];
      float4 b = *reinterpret_cast<const float4*>(&Bs[k][tc * 4]);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[i][0] = fmaf(a[i], b.x, acc[i][0]);
        acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
        acc[i][2] = fmaf(a[i], b.z, acc[i][2]);
        acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) p.store(tile_m0 + tr + 16 * i, tile_n0 + tc * 4 + j, acc[i][j]);
}

// ------------------------------------------------------------------ fprop: M = N*P*Q, Ncols = K, Kdim = C*R*S
struct FpropP : ConvGeom {
  const float* x; const float* w; const float* bias; const float* iscale; const float* oscale; float* y;
  static constexpr bool A_KFAST = false;
  struct ACtx { const float* xn; const float* sc; int ih0, iw0; bool valid; };
  struct BCtx { int dummy; };
  __device__ int Kdim() const { return CRS; }
  __device__ void init_a(ACtx& c, int m0, int tid) const {
    int m = m0 + (tid & 127);
    c.valid = m < N * PQ;
    int n = c.valid ? m / PQ : 0;
    int pq = c.valid ? m % PQ : 0;
    c.ih0 = (pq / Q) * stride - pad;
    c.iw0 = (pq % Q) * stride - pad;
    c.xn = x + (int64_t)n * C * HW;
    c.sc = iscale ? iscale + (int64_t)n * C : nullptr;
  }
  __device__ void init_b(BCtx&, int, int) const {}
  __device__ void load_a(float (&r)[8], const ACtx& c, int k0, int tid) const {
    int kb = k0 + (tid >> 7);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int k = kb + 2 * j;
      float v = 0.f;
      if (c.valid && k < CRS) {
        int ci = k / RS, rs = k - ci * RS;
        int r_ = rs / S, s_ = rs - r_ * S;
        int ih = c.ih0 + r_, iw = c.iw0 + s_;
        bool ok = true;
        if (pad_mode) { ih = min(max(ih, 0), H - 1); iw = min(max(iw, 0), W - 1); }
        else ok = (ih >= 0) & (ih < H) & (iw >= 0) & (iw < W);
        if (ok) {
          v = c.xn[(int64_t)ci * HW + ih * W + iw];
          if (c.sc) v *= c.sc[ci];
        }
      }
      r[j] = v;
    }
  }
  __device__ void load_b(float (&r)[4], const BCtx&, int k0, int tid) const {
    int k = k0 + (tid & 15);
    int nb = blockIdx.y * BN + (tid >> 4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = nb + 16 * j;
      r[j] = (k < CRS && n < K) ? w[(int64_t)n * CRS + k] : 0.f;
    }
  }
  __device__ void store(int m, int n, float v) const {
    if (m >= N * PQ || n >= K) return;
    int img = m / PQ, pq = m - img * PQ;
    v *= alpha;
    if (oscale) v *= oscale[(int64_t)img * K + n];
    if (bias) v += bias[n];
    y[((int64_t)img * K + n) * PQ + pq] = v;
  }
};

// ------------------------------------------------------------------ dgrad over a domain Hd x Wd with zero-pad `pe`
// M = N*Hd*Wd, Ncols = C, Kdim = K*R*S ; out[n,c,h,w] (domain layout) = iscale*alpha*sum
struct DgradP : ConvGeom {
  const float* dy; const float* w; const float* iscale; const float* oscale; float* out;
  int Hd, Wd, pe;
  static constexpr bool A_KFAST = false;
  struct ACtx { const float* dyn; const float* sc; int h, w; bool valid; };
  struct BCtx { int dummy; };
  __device__ int Kdim() const { return K * RS; }
  __device__ void init_a(ACtx& c, int m0, int tid) const {
    int m = m0 + (tid & 127);
    int HWd = Hd * Wd;
    c.valid = m < N * HWd;
    int n = c.valid ? m / HWd : 0;
    int hw = c.valid ? m % HWd : 0;
    c.h = hw / Wd + pe;
    c.w = hw % Wd + pe;
    c.dyn = dy + (int64_t)n * K * PQ;
    c.sc = oscale ? oscale + (int64_t)n * K : nullptr;
  }
  __device__ void init_b(BCtx&, int, int) const {}
  __device__ void load_a(float (&r)[8], const ACtx& c, int k0, int tid) const {
    int kb = k0 + (tid >> 7);
    int Kd = K * RS;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int k = kb + 2 * j;
      float v = 0.f;
      if (c.valid && k < Kd) {
        int ko = k / RS, rs = k - ko * RS;
        int r_ = rs / S, s_ = rs - r_ * S;
        int hh = c.h - r_, ww = c.w - s_;
        if (hh >= 0 && ww >= 0) {
          int p_ = hh / stride, q_ = ww / stride;
          if (p_ * stride == hh && q_ * stride == ww && p_ < P && q_ < Q) {
            v = c.dyn[(int64_t)ko * PQ + p_ * Q + q_];
            if (c.sc) v *= c.sc[ko];
          }
        }
      }
      r[j] = v;
    }
  }
  __device__ void load_b(float (&r)[4], const BCtx&, int k0, int tid) const {
    int k = k0 + (tid & 15);
    int nb = blockIdx.y * BN + (tid >> 4);
    int ko = k / RS, rs = k - ko * RS;
    bool kok = k < K * RS;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int c = nb + 16 * j;
      r[j] = (kok && c < C) ? w[((int64_t)ko * C + c) * RS + rs] : 0.f;
    }
  }
  __device__ void store(int m, int c, float v) const {
    int HWd = Hd * Wd;
    if (m >= N * HWd || c >= C) return;
    int img = m / HWd, hw = m - img * HWd;
    v *= alpha;
    if (iscale) v *= iscale[(int64_t)img * C + c];
    out[((int64_t)img * C + c) * HWd + hw] = v;
  }
};

// ------------------------------------------------------------------ wgrad: M = K, Ncols = C*R*S, Kdim = N*P*Q (split-K)
// Split-K partials go to slice blockIdx.z of a scratch array and simt_splitk_sum_kernel adds the slices in order
// (`slice` = K*CRS).  slice == 0: no scratch for this launch (see simt_wgrad) — atomics into a zeroed dw.
constexpr int kWgradRegions = 4;
constexpr size_t kWgradScratch = (size_t)1 << 22;   // floats per region (16 MB)
static __device__ __align__(16) float g_wgrad_part[kWgradRegions * kWgradScratch];

__global__ void __launch_bounds__(256) simt_splitk_sum_kernel(float* __restrict__ dw, int region, int64_t slice,
                                                              int splits) {
  const float* part = g_wgrad_part + (size_t)region * kWgradScratch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slice; i += (int64_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int z = 0; z < splits; ++z) acc += __ldcg(part + z * slice + i);
    dw[i] = acc;
  }
}

struct WgradP : ConvGeom {
  const float* x; const float* dy; const float* iscale; const float* oscale; float* dw;
  int64_t slice; int region;
  static constexpr bool A_KFAST = true;
  struct ACtx { int dummy; };
  struct BCtx { int ci[4], r_[4], s_[4]; bool ok[4]; };
  __device__ int Kdim() const { return N * PQ; }
  __device__ void init_a(ACtx&, int, int) const {}
  __device__ void init_b(BCtx& b, int n0, int tid) const {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int col = n0 + (tid >> 4) + 16 * j;
      b.ok[j] = col < CRS;
      int cc = b.ok[j] ? col : 0;
      b.ci[j] = cc / RS;
      int rs = cc - b.ci[j] * RS;
      b.r_[j] = rs / S;
      b.s_[j] = rs - b.r_[j] * S;
    }
  }
  __device__ void load_a(float (&r)[8], const ACtx&, int k0, int tid) const {
    int k = k0 + (tid & 15);
    bool kok = k < N * PQ;
    int n = kok ? k / PQ : 0, pq = kok ? k - n * PQ : 0;
    int mb = blockIdx.x * BM + (tid >> 4);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int ko = mb + 16 * j;
      float v = 0.f;
      if (kok && ko < K) {
        v = dy[((int64_t)n * K + ko) * PQ + pq];
        if (oscale) v *= oscale[(int64_t)n * K + ko];
      }
      r[j] = v;
    }
  }
  __device__ void load_b(float (&r)[4], const BCtx& b, int k0, int tid) const {
    int k = k0 + (tid & 15);
    bool kok = k < N * PQ;
    int n = kok ? k / PQ : 0, pq = kok ? k - n * PQ : 0;
    int p_ = pq / Q, q_ = pq - p_ * Q;
    int ih0 = p_ * stride - pad, iw0 = q_ * stride - pad;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v = 0.f;
      if (kok && b.ok[j]) {
        int ih = ih0 + b.r_[j], iw = iw0 + b.s_[j];
        bool ok = true;
        if (pad_mode) { ih = min(max(ih, 0), H - 1); iw = min(max(iw, 0), W - 1); }
        else ok = (ih >= 0) & (ih < H) & (iw >= 0) & (iw < W);
        if (ok) {
          v = x[((int64_t)n * C + b.ci[j]) * HW + ih * W + iw];
          if (iscale) v *= iscale[(int64_t)n * C + b.ci[j]];
        }
      }
      r[j] = v;
    }
  }
  __device__ void store(int ko, int col, float v) const {
    if (ko >= K || col >= CRS) return;
    if (gridDim.z == 1) dw[(int64_t)ko * CRS + col] = v * alpha;
    else if (slice) g_wgrad_part[(size_t)region * kWgradScratch + blockIdx.z * slice + (int64_t)ko * CRS + col] = v * alpha;
    else atomicAdd(dw + (int64_t)ko * CRS + col, v * alpha);
  }
};

// adjoint of replicate padding: dx[n,c,h,w] = sum of dxp over the padded positions that clamp to (h,w).
// VEC = 4 outputs per thread along w (W % 4 == 0, float4 store); HBM-bound: 4 B read + 4 B written per element.
template <int VEC>
__global__ void __launch_bounds__(256) fold_replicate_kernel(const float* __restrict__ dxp, float* __restrict__ dx,
                                                             int64_t NC, int H, int W, int pad) {
  const int Hp = H + 2 * pad, Wp = W + 2 * pad, Wv = W / VEC;
  grid_stride(NC * H * Wv, [&](auto i) {
    const int w0 = (int)(i % Wv) * VEC;
    const auto t = i / Wv;
    const int h = (int)(t % H);
    const int64_t nc = (int64_t)(t / H);
    int h_lo = h == 0 ? 0 : h + pad, h_hi = h == H - 1 ? Hp - 1 : h + pad;
    if (H == 1) { h_lo = 0; h_hi = Hp - 1; }
    const float* src = dxp + nc * Hp * Wp;
    float acc[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const int w = w0 + j;
      int w_lo = w == 0 ? 0 : w + pad, w_hi = w == W - 1 ? Wp - 1 : w + pad;
      if (W == 1) { w_lo = 0; w_hi = Wp - 1; }
      float a_ = 0.f;
      for (int a = h_lo; a <= h_hi; ++a)
        for (int b = w_lo; b <= w_hi; ++b) a_ += __ldg(src + a * Wp + b);
      acc[j] = a_;
    }
    float* dst = dx + (nc * H + h) * W + w0;
    if (VEC == 4) *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    else dst[0] = acc[0];
  });
}

static void launch_fold(const float* dxp, float* dx, int64_t NC, int H, int W, int pad, cudaStream_t s) {
  if ((W & 3) == 0 && ((uintptr_t)dx & 15) == 0)
    fold_replicate_kernel<4><<<ew_grid(NC * H * (W / 4), 256, 1), 256, 0, s>>>(dxp, dx, NC, H, W, pad);
  else
    fold_replicate_kernel<1><<<ew_grid(NC * H * W, 256, 2), 256, 0, s>>>(dxp, dx, NC, H, W, pad);
}

}  // namespace ganb200
using namespace ganb200;

int ganb200::fold_replicate(const float* dxp, float* dx, int64_t NC, int H, int W, int pad, cudaStream_t s) {
  launch_fold(dxp, dx, NC, H, W, pad, s);
  GB_CHECK_LAUNCH();
  return 0;
}

int ganb200::simt_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias,
                        const float* iscale, const float* oscale, float* y, cudaStream_t s) {
  FpropP p;
  fill_geom(p, d);
  p.x = x; p.w = w; p.bias = bias; p.iscale = iscale; p.oscale = oscale; p.y = y;
  dim3 grid(ceil_div((int64_t)p.N * p.PQ, BM), ceil_div(p.K, BN), 1);
  simt_gemm_kernel<FpropP><<<grid, 256, 0, s>>>(p);
  GB_CHECK_LAUNCH();
  return 0;
}

size_t ganb200::simt_dgrad_workspace(const ganb200_conv_desc* d) {
  if (d->pad_mode == 1 && d->pad > 0)
    return sizeof(float) * (size_t)d->N * d->C * (d->H + 2 * d->pad) * (d->W + 2 * d->pad);
  return 0;
}

int ganb200::simt_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale,
                        const float* oscale, float* dx, void* ws, size_t ws_bytes, cudaStream_t s) {
  DgradP p;
  fill_geom(p, d);
  p.dy = dy; p.w = w; p.iscale = iscale; p.oscale = oscale;
  bool rep = d->pad_mode == 1 && d->pad > 0;
  if (rep) {
    GB_REQUIRE(ws && ws_bytes >= simt_dgrad_workspace(d), GANB200_WORKSPACE_TOO_SMALL, "dgrad: workspace too small");
    p.Hd = d->H + 2 * d->pad; p.Wd = d->W + 2 * d->pad; p.pe = 0; p.out = (float*)ws;
  } else {
    p.Hd = d->H; p.Wd = d->W; p.pe = d->pad; p.out = dx;
  }
  dim3 grid(ceil_div((int64_t)p.N * p.Hd * p.Wd, BM), ceil_div(p.C, BN), 1);
  simt_gemm_kernel<DgradP><<<grid, 256, 0, s>>>(p);
  GB_CHECK_LAUNCH();
  if (rep) {
    launch_fold((const float*)ws, dx, (int64_t)d->N * d->C, d->H, d->W, d->pad, s);
    GB_CHECK_LAUNCH();
  }
  return 0;
}

int ganb200::simt_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale,
                        const float* oscale, float* dw, cudaStream_t s) {
  WgradP p;
  fill_geom(p, d);
  p.x = x; p.dy = dy; p.iscale = iscale; p.oscale = oscale; p.dw = dw;
  int tiles = ceil_div(p.K, BM) * ceil_div(p.CRS, BN);
  int64_t Kd = (int64_t)p.N * p.PQ;
  int k_tiles = ceil_div(Kd, BK);
  int splits = (kNumSMs * 4 + tiles - 1) / tiles;
  int max_splits = (k_tiles + 7) / 8;  // at least 8 k-tiles per split
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  if (splits > 65535) splits = 65535;
  const int64_t slice = (int64_t)p.K * p.CRS;
  const int region = det_region(s);
  const bool det = region >= 0 && region < kWgradRegions;
  if (det && splits > 1 && (size_t)slice * splits > kWgradScratch)   // fewer, longer splits rather than atomics
    splits = (int)std::max<int64_t>(1, (int64_t)(kWgradScratch / (size_t)slice));
  p.slice = det && splits > 1 ? slice : 0;
  p.region = det ? region : 0;
  if (splits > 1 && !p.slice) GB_CHECK_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)slice, s));
  dim3 grid(ceil_div(p.K, BM), ceil_div(p.CRS, BN), splits);
  simt_gemm_kernel<WgradP><<<grid, 256, 0, s>>>(p);
  GB_CHECK_LAUNCH();
  if (p.slice) {
    simt_splitk_sum_kernel<<<ew_grid(slice, 256, 1), 256, 0, s>>>(dw, region, slice, splits);
    GB_CHECK_LAUNCH();
  }
  return 0;
}
