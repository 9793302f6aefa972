// conv_tc.cu — bf16 tcgen05/TMEM implicit-GEMM convolution engine (engine 2), sm_100a only.
//
// fprop  y[n,k,p,q] = oscale[n,k]*alpha*sum_{c,r,s} w[k,c,r,s]*iscale[n,c]*xpad[n,c,p+r,q+s] + bias[k]   (stride 1)
//   1. pack_act    : x fp32 NCHW -> bf16 NHWC [N][Hp][Wp][Cp] (iscale folded in, replicate padding materialised,
//                    zero padding left to TMA out-of-bounds fill)
//   2. pack_weight : w fp32 KCRS -> bf16 [R*S][K][Cp]
//   3. tc_conv_kernel : one CTA = 128 output pixels (a bn x bh x bw box) x BN output channels.
//        warp 0 : TMA producer — per (tap, 64-channel chunk) one 4-D box of activations (the box shifted by the tap:
//                 that is the whole im2col) and one 3-D box of weights into a SWIZZLE_128B ring
//        warp 1 : tcgen05.mma issuer (one thread), fp32 accumulator 128 lanes x BN columns in TMEM
//        warps 2-5 : epilogue — tcgen05.ld, alpha / oscale / bias, fp32 NCHW stores
// dgrad (stride 1) = the same kernel on dy with the flipped, transposed weights and zero padding R-1, written to the
//   padded domain and folded (adjoint of replicate padding) when needed.
// wgrad = conv_tc_wgrad.cu.
#include "common.cuh"
#include "conv_common.cuh"
#include "tc_ptx.cuh"
#include <algorithm>

namespace ganb200 {

using namespace ptx;

constexpr int TILE_M = 128;
constexpr int BK = 64;                 // bf16 elements per k-chunk = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = TILE_M * BK * 2;
constexpr int TC_THREADS = 192;

// ------------------------------------------------------------------------------------------------ packing
// x [N][C][H][W] fp32 (* scale[n][c]) -> out [N][Hp][Wp][Cp] bf16, (hp,wp) reading clamp(hp-pad), clamp(wp-pad).
// One block = one (n, hp) row x 64 channels; loads coalesced along w, stores coalesced along c.
__global__ void __launch_bounds__(256) pack_act_kernel(const float* __restrict__ x, const float* __restrict__ scale,
                                                       __nv_bfloat16* __restrict__ out, int C, int H, int W, int Cp,
                                                       int pad) {
  extern __shared__ float tile[];  // [64][W+1]
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const int n = blockIdx.x / Hp, hp = blockIdx.x % Hp;
  const int c0 = blockIdx.y * 64;
  const int h = min(max(hp - pad, 0), H - 1);
  const int ld = W + 1;
  for (int i = threadIdx.x; i < 64 * W; i += 256) {
    int cl = i / W, w = i - cl * W;
    int c = c0 + cl;
    float v = 0.f;
    if (c < C) {
      v = x[(((int64_t)n * C + c) * H + h) * W + w];
      if (scale) v *= scale[(int64_t)n * C + c];
    }
    tile[cl * ld + w] = v;
  }
  __syncthreads();
  __nv_bfloat16* orow = out + ((int64_t)n * Hp + hp) * Wp * Cp;
  const int cvalid = min(64, Cp - c0);  // Cp is a multiple of 8
  for (int i = threadIdx.x; i < 32 * Wp; i += 256) {
    int wp = i >> 5, cl = (i & 31) * 2;
    if (cl >= cvalid) continue;
    int w = min(max(wp - pad, 0), W - 1);
    __nv_bfloat162 v = __floats2bfloat162_rn(tile[cl * ld + w], tile[(cl + 1) * ld + w]);
    *reinterpret_cast<__nv_bfloat162*>(orow + (int64_t)wp * Cp + c0 + cl) = v;
  }
}

// w fp32 -> out[rs][ko][cp] bf16 with out(rs,ko,ci) = w[ko*sk + ci*sc + rs'] , rs' = flip ? RS-1-rs : rs
// (fprop: sk = C*RS, sc = RS; dgrad: roles of K and C swapped: sk = RS, sc = Cw*RS, flip = 1)
__global__ void __launch_bounds__(256) pack_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out,
                                                          int Ko, int Ci, int Cp, int RS, int64_t sk, int64_t sc,
                                                          int flip) {
  int64_t total = (int64_t)RS * Ko * Cp;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    int ci = (int)(i % Cp);
    int64_t t = i / Cp;
    int ko = (int)(t % Ko), rs = (int)(t / Ko);
    float v = 0.f;
    if (ci < Ci) v = w[ko * sk + ci * sc + (flip ? RS - 1 - rs : rs)];
    out[i] = __float2bfloat16(v);
  }
}

// ------------------------------------------------------------------------------------------------ main kernel
struct TcConvParams {
  int N, K, P, Q;       // output [N][K][P][Q]
  int S, taps;          // filter width, R*S
  int C;                // reduction channels
  int bw, bh, bn;       // tile box (bw*bh*bn == 128)
  int tiles_q, tiles_p; // ceil(Q/bw), ceil(P/bh)
  int off_h, off_w;     // tap coordinate offsets: 0 when the padding is materialised, -pad for TMA zero fill
  int n_umma;           // UMMA N = columns of the accumulator actually computed (multiple of 16, <= BN)
  float alpha;
  const float* oscale;  // [N][K] or null
  const float* bias;    // [K] or null
  float* y;
};

template <int BN, int STAGES>
struct TcSmem {
  static constexpr int B_STAGE_BYTES = BN * BK * 2;
  static constexpr int BYTES = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + 1024 /*align*/ + 256 /*barriers*/;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(TC_THREADS) tc_conv_kernel(const __grid_constant__ CUtensorMap tmA,
                                                             const __grid_constant__ CUtensorMap tmB,
                                                             const TcConvParams p) {
  constexpr int B_STAGE_BYTES = BN * BK * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + STAGES * A_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smB + STAGES * B_STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // tile coordinates
  int t = blockIdx.x;
  const int tq = t % p.tiles_q; t /= p.tiles_q;
  const int tp = t % p.tiles_p; t /= p.tiles_p;
  const int q0 = tq * p.bw, p0 = tp * p.bh, n0 = t * p.bn;
  const int k0 = blockIdx.y * BN;
  const int k_chunks = (p.C + BK - 1) / BK;
  const int iters = p.taps * k_chunks;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        mbar_wait(&empty[stage], phase ^ 1);
        const int tap = it / k_chunks, kc = it - tap * k_chunks;
        const int r = tap / p.S, s = tap - r * p.S;
        mbar_expect_tx(&full[stage], A_STAGE_BYTES + B_STAGE_BYTES);
        tma_load_4d(smA + stage * A_STAGE_BYTES, &tmA, &full[stage], kc * BK, q0 + s + p.off_w, p0 + r + p.off_h, n0);
        tma_load_3d(smB + stage * B_STAGE_BYTES, &tmB, &full[stage], kc * BK, k0, tap);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(TILE_M, p.n_umma, 0, 0);
      for (int it = 0; it < iters; ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const int kc = it % k_chunks;
        const int ksub = min(BK / 16, (p.C - kc * BK + 15) / 16);
        const uint64_t adesc = make_smem_desc_sw128(smem_u32(smA + stage * A_STAGE_BYTES), 16, 1024);
        const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB + stage * B_STAGE_BYTES), 16, 1024);
        for (int k = 0; k < ksub; ++k)  // +32 bytes (= 2 in the >>4 address field) per 16-element k step
          mma_bf16_ss(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
        mma_commit(&empty[stage]);
      }
      mma_commit(tmem_full);
    }
    __syncwarp();
  } else {
    // epilogue: warp w may touch TMEM lanes [32*(w%4), +32)
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const int wl = m % p.bw, hl = (m / p.bw) % p.bh, nl = m / (p.bw * p.bh);
    const int q = q0 + wl, pp = p0 + hl, n = n0 + nl;
    const bool row_ok = (q < p.Q) && (pp < p.P) && (n < p.N);
    const int64_t PQ = (int64_t)p.P * p.Q;
    float* ybase = p.y + ((int64_t)n * p.K) * PQ + (int64_t)pp * p.Q + q;
    const float* osc = p.oscale ? p.oscale + (int64_t)n * p.K : nullptr;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < p.n_umma; c0 += 32) {
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
      if (row_ok) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int k = k0 + c0 + j;
          if (c0 + j < p.n_umma && k < p.K) {
            float o = v[j] * p.alpha;
            if (osc) o *= osc[k];
            if (p.bias) o += p.bias[k];
            ybase[(int64_t)k * PQ] = o;
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// bf16 tensor, `rank` dims (innermost first), byte strides for dims 1..rank-1, SWIZZLE_128B, zero OOB fill
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
                   const uint32_t* box) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail(GANB200_CUDA_ERROR, "cuTensorMapEncodeTiled not available%s", "");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides[i];
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(GANB200_CUDA_ERROR, "cuTensorMapEncodeTiled failed (%s code %lld)", "", (long long)r);
  return 0;
}

static inline int pow2ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline int cpad(int c) { return (c + 7) / 8 * 8; }
static inline void* align_ws(void* ws) { return (void*)(((uintptr_t)ws + 1023) & ~uintptr_t(1023)); }

// the generic problem both fprop and stride-1 dgrad map to
struct TcProblem {
  int N, Ci, Hi, Wi;      // input  fp32 NCHW
  int Ko, R, S;           // filter
  int Po, Qo;             // output fp32 NCHW [N][Ko][Po][Qo]
  int mat_pad;            // replicate padding materialised while packing
  int off_h, off_w;       // tap coordinate offsets (<= 0): zero padding through TMA OOB fill
  int64_t w_sk, w_sc;     // weight strides for (ko, ci); taps innermost
  int w_flip;
  float alpha;
};

static size_t problem_act_bytes(const TcProblem& q) {
  return align_up((size_t)q.N * (q.Hi + 2 * q.mat_pad) * (q.Wi + 2 * q.mat_pad) * cpad(q.Ci) * 2, 1024);
}
static size_t problem_w_bytes(const TcProblem& q) { return align_up((size_t)q.R * q.S * q.Ko * cpad(q.Ci) * 2, 1024); }

template <int BN, int STAGES>
static int launch_tc(const CUtensorMap& ta, const CUtensorMap& tb, const TcConvParams& p, dim3 grid, cudaStream_t s) {
  static bool configured = false;
  constexpr int SMEM = TcSmem<BN, STAGES>::BYTES;
  if (!configured) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(tc_conv_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    configured = true;
  }
  tc_conv_kernel<BN, STAGES><<<grid, TC_THREADS, SMEM, s>>>(ta, tb, p);
  GB_CHECK_LAUNCH();
  return 0;
}

static int run_problem(const TcProblem& q, const float* in, const float* in_scale, const float* w, float* out,
                       const float* out_scale, const float* bias, void* ws, cudaStream_t s) {
  const int Cp = cpad(q.Ci), Hp = q.Hi + 2 * q.mat_pad, Wp = q.Wi + 2 * q.mat_pad, RS = q.R * q.S;
  __nv_bfloat16* act = (__nv_bfloat16*)ws;
  __nv_bfloat16* wb = (__nv_bfloat16*)((uint8_t*)ws + problem_act_bytes(q));
  {
    dim3 grid(q.N * Hp, (Cp + 63) / 64);
    size_t sm = sizeof(float) * 64 * (q.Wi + 1);
    pack_act_kernel<<<grid, 256, sm, s>>>(in, in_scale, act, q.Ci, q.Hi, q.Wi, Cp, q.mat_pad);
    GB_CHECK_LAUNCH();
    int64_t total = (int64_t)RS * q.Ko * Cp;
    pack_weight_kernel<<<ew_grid(total, 256, 1), 256, 0, s>>>(w, wb, q.Ko, q.Ci, Cp, RS, q.w_sk, q.w_sc, q.w_flip);
    GB_CHECK_LAUNCH();
  }
  TcConvParams p;
  p.N = q.N; p.K = q.Ko; p.P = q.Po; p.Q = q.Qo; p.S = q.S; p.taps = RS; p.C = q.Ci;
  p.bw = std::min(pow2ceil(q.Qo), TILE_M);
  p.bh = std::min(pow2ceil(q.Po), TILE_M / p.bw);
  p.bn = TILE_M / (p.bw * p.bh);
  p.tiles_q = ceil_div(q.Qo, p.bw); p.tiles_p = ceil_div(q.Po, p.bh);
  p.alpha = q.alpha; p.oscale = out_scale; p.bias = bias; p.y = out;
  const int BN = q.Ko >= 128 ? 128 : 64;
  p.n_umma = q.Ko >= BN ? BN : (q.Ko + 15) / 16 * 16;
  CUtensorMap ta, tb;
  {
    uint64_t dims[4] = {(uint64_t)q.Ci, (uint64_t)Wp, (uint64_t)Hp, (uint64_t)q.N};
    uint64_t str[3] = {(uint64_t)Cp * 2, (uint64_t)Wp * Cp * 2, (uint64_t)Hp * Wp * Cp * 2};
    uint32_t box[4] = {BK, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
    int rc = make_tmap_bf16(&ta, act, 4, dims, str, box);
    if (rc) return rc;
    uint64_t wd[3] = {(uint64_t)q.Ci, (uint64_t)q.Ko, (uint64_t)RS};
    uint64_t wst[2] = {(uint64_t)Cp * 2, (uint64_t)q.Ko * Cp * 2};
    uint32_t wbox[3] = {BK, (uint32_t)BN, 1};
    rc = make_tmap_bf16(&tb, wb, 3, wd, wst, wbox);
    if (rc) return rc;
  }
  p.off_h = q.off_h; p.off_w = q.off_w;
  dim3 grid(p.tiles_q * p.tiles_p * ceil_div(q.N, p.bn), ceil_div(q.Ko, BN));
  if (BN == 128) return launch_tc<128, 3>(ta, tb, p, grid, s);
  return launch_tc<64, 4>(ta, tb, p, grid, s);
}

// fold_replicate_kernel lives in conv_simt.cu


static bool shape_ok(const ganb200_conv_desc* d) {
  return d->stride == 1 && d->H <= 128 && d->W <= 128 && d->R <= 7 && d->S <= 7;
}

int tc_eligible(const ganb200_conv_desc* d, int op) {
  if (!shape_ok(d)) return 0;
  const int64_t M = (int64_t)d->N * d->P * d->Q;
  if (op == GANB200_CONV_FPROP) return d->C >= 8 && d->K >= 8 && M >= 64;
  if (op == GANB200_CONV_DGRAD) return d->C >= 8 && d->K >= 8 && (int64_t)d->N * d->H * d->W >= 64;
  return d->C >= 8 && d->K >= 8 && M >= 256;
}

static size_t wgrad_workspace(const ganb200_conv_desc* d);

static TcProblem fprop_problem(const ganb200_conv_desc* d) {
  TcProblem q;
  q.N = d->N; q.Ci = d->C; q.Hi = d->H; q.Wi = d->W; q.Ko = d->K; q.R = d->R; q.S = d->S; q.Po = d->P; q.Qo = d->Q;
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  q.mat_pad = rep ? d->pad : 0;
  q.off_h = q.off_w = rep ? 0 : -d->pad;
  q.w_sk = (int64_t)d->C * d->R * d->S; q.w_sc = (int64_t)d->R * d->S; q.w_flip = 0;
  q.alpha = d->alpha;
  return q;
}

// dgrad as a convolution of dy: output domain = padded input (replicate, folded afterwards) or the input itself (zeros)
static TcProblem dgrad_problem(const ganb200_conv_desc* d) {
  TcProblem q;
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  q.N = d->N; q.Ci = d->K; q.Hi = d->P; q.Wi = d->Q; q.Ko = d->C; q.R = d->R; q.S = d->S;
  q.Po = rep ? d->H + 2 * d->pad : d->H;
  q.Qo = rep ? d->W + 2 * d->pad : d->W;
  q.mat_pad = 0;
  q.off_h = -(d->R - 1 - (rep ? 0 : d->pad));
  q.off_w = -(d->S - 1 - (rep ? 0 : d->pad));
  q.w_sk = (int64_t)d->R * d->S; q.w_sc = (int64_t)d->C * d->R * d->S; q.w_flip = 1;
  q.alpha = d->alpha;
  return q;
}

static size_t dgrad_fold_bytes(const ganb200_conv_desc* d) {
  if (d->pad_mode == 1 && d->pad > 0)
    return align_up(sizeof(float) * (size_t)d->N * d->C * (d->H + 2 * d->pad) * (d->W + 2 * d->pad), 1024);
  return 0;
}

size_t tc_workspace(const ganb200_conv_desc* d, int op) {
  if (op == GANB200_CONV_FPROP) {
    TcProblem q = fprop_problem(d);
    return problem_act_bytes(q) + problem_w_bytes(q) + 1024;
  }
  if (op == GANB200_CONV_DGRAD) {
    TcProblem q = dgrad_problem(d);
    return problem_act_bytes(q) + problem_w_bytes(q) + dgrad_fold_bytes(d) + 1024;
  }
  return wgrad_workspace(d);
}

int tc_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias, const float* iscale,
             const float* oscale, float* y, void* ws, size_t ws_bytes, cudaStream_t s) {
  GB_REQUIRE(ws && ws_bytes >= tc_workspace(d, GANB200_CONV_FPROP), GANB200_WORKSPACE_TOO_SMALL,
             "tc fprop: workspace too small");
  ws = align_ws(ws);
  TcProblem q = fprop_problem(d);
  return run_problem(q, x, iscale, w, y, oscale, bias, ws, s);
}

int tc_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale, const float* oscale,
             float* dx, void* ws, size_t ws_bytes, cudaStream_t s) {
  GB_REQUIRE(ws && ws_bytes >= tc_workspace(d, GANB200_CONV_DGRAD), GANB200_WORKSPACE_TOO_SMALL,
             "tc dgrad: workspace too small");
  ws = align_ws(ws);
  TcProblem q = dgrad_problem(d);
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  float* out = dx;
  if (rep) out = (float*)((uint8_t*)ws + problem_act_bytes(q) + problem_w_bytes(q));
  int rc = run_problem(q, dy, oscale, w, out, iscale, nullptr, ws, s);
  if (rc) return rc;
  if (rep) return fold_replicate(out, dx, (int64_t)d->N * d->C, d->H, d->W, d->pad, s);
  return 0;
}


// ================================================================================================ wgrad
// dw[k,c,r,s] = alpha * sum_{n,p,q} (oscale*dy)[n,k,p,q] * (iscale*xpad)[n,c,p+r,q+s]
// GEMM per tap: D[128 kout x BN c] += A^T B over pixels, A = dy NHWC bf16, B = x NHWC bf16 shifted by the tap.
// Both operands are MN-major in shared memory (channels contiguous, pixels = the GEMM K dimension): a TMA box of
// (64 channels x 64 pixels) is exactly eight SWIZZLE_128B MN-major atoms stacked along K (SBO = 1024 B); the second
// 64-channel group of the tile sits LBO = 8192 B further.  One CTA = one (kout tile, c tile, tap, pixel split);
// partial sums go to a workspace and are reduced deterministically (no atomics).
constexpr int WG_PIX = 64;                        // pixels per pipeline stage
constexpr int WG_GROUP_BYTES = WG_PIX * 128;      // one (64 channels x 64 pixels) box

struct TcWgradParams {
  int K, C, S, taps;
  int bw, bh, bn;            // pixel box, bw*bh*bn == 64
  int tiles_q, tiles_p, tiles_n;
  int chunks, chunks_per_split, splits;
  int off_h, off_w;
  int n_umma;                // multiple of 16 <= BN
  float* part;               // [splits][taps][K][C]
};

template <int BN, int STAGES>
struct TcWgSmem {
  static constexpr int A_BYTES = 2 * WG_GROUP_BYTES;
  static constexpr int B_BYTES = (BN / 64) * WG_GROUP_BYTES;
  static constexpr int BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 + 256;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(TC_THREADS) tc_wgrad_kernel(const __grid_constant__ CUtensorMap tmA,
                                                              const __grid_constant__ CUtensorMap tmB,
                                                              const TcWgradParams p) {
  constexpr int A_BYTES = TcWgSmem<BN, STAGES>::A_BYTES, B_BYTES = TcWgSmem<BN, STAGES>::B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + STAGES * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smB + STAGES * B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * TILE_M, c0 = blockIdx.y * BN;
  const int tap = blockIdx.z % p.taps, split = blockIdx.z / p.taps;
  const int r = tap / p.S, s = tap - r * p.S;
  const int ch_beg = split * p.chunks_per_split;
  const int ch_end = min(p.chunks, ch_beg + p.chunks_per_split);
  const int iters = ch_end - ch_beg;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        mbar_wait(&empty[stage], phase ^ 1);
        int t = ch_beg + it;
        const int tq = t % p.tiles_q; t /= p.tiles_q;
        const int tp = t % p.tiles_p; t /= p.tiles_p;
        const int q0 = tq * p.bw, p0 = tp * p.bh, n0 = t * p.bn;
        mbar_expect_tx(&full[stage], A_BYTES + B_BYTES);
        uint8_t* a = smA + stage * A_BYTES;
        uint8_t* b = smB + stage * B_BYTES;
        tma_load_4d(a, &tmA, &full[stage], k0, q0, p0, n0);
        tma_load_4d(a + WG_GROUP_BYTES, &tmA, &full[stage], k0 + 64, q0, p0, n0);
#pragma unroll
        for (int g = 0; g < BN / 64; ++g)
          tma_load_4d(b + g * WG_GROUP_BYTES, &tmB, &full[stage], c0 + g * 64, q0 + s + p.off_w, p0 + r + p.off_h, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(TILE_M, p.n_umma, 1, 1);
      for (int it = 0; it < iters; ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        // MN-major SWIZZLE_128B: LBO = next 64-channel group, SBO = next 8 pixels
        const uint64_t adesc = make_smem_desc_sw128(smem_u32(smA + stage * A_BYTES), WG_GROUP_BYTES, 1024);
        const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB + stage * B_BYTES), WG_GROUP_BYTES, 1024);
#pragma unroll
        for (int k = 0; k < WG_PIX / 16; ++k)  // 16 pixels = 2048 bytes = 128 in the >>4 address field
          mma_bf16_ss(tmem_base, adesc + 128 * k, bdesc + 128 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
        mma_commit(&empty[stage]);
      }
      mma_commit(tmem_full);
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int k = k0 + quarter * 32 + lane;
    float* row = p.part + (((int64_t)split * p.taps + tap) * p.K + k) * p.C + c0;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    for (int j0 = 0; j0 < p.n_umma; j0 += 32) {
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)j0, v);
      if (k < p.K) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j0 + j < p.n_umma && c0 + j0 + j < p.C) row[j0 + j] = iters > 0 ? v[j] : 0.f;
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// dw[k][c][rs] = alpha * sum_split part[split][rs][k][c]
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ dw, int K,
                                                           int C, int RS, int splits, float alpha) {
  const int64_t KC = (int64_t)K * C;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < KC; i += (int64_t)gridDim.x * 256) {
    for (int rs = 0; rs < RS; ++rs) {
      float acc = 0.f;
      for (int sp = 0; sp < splits; ++sp) acc += part[((int64_t)sp * RS + rs) * KC + i];
      dw[i * RS + rs] = acc * alpha;
    }
  }
}

struct WgradPlan {
  int bw, bh, bn, tiles_q, tiles_p, tiles_n, chunks, splits, per, BN;
  size_t dy_bytes, x_bytes, part_bytes;
};

static WgradPlan wgrad_plan(const ganb200_conv_desc* d) {
  WgradPlan w;
  w.bw = std::min(pow2ceil(d->Q), WG_PIX);
  w.bh = std::min(pow2ceil(d->P), WG_PIX / w.bw);
  w.bn = WG_PIX / (w.bw * w.bh);
  w.tiles_q = ceil_div(d->Q, w.bw); w.tiles_p = ceil_div(d->P, w.bh); w.tiles_n = ceil_div(d->N, w.bn);
  w.chunks = w.tiles_q * w.tiles_p * w.tiles_n;
  w.BN = d->C > 64 ? 128 : 64;
  const int tiles = ceil_div(d->K, TILE_M) * ceil_div(d->C, w.BN) * d->R * d->S;
  int splits = std::max(1, (2 * kNumSMs + tiles - 1) / tiles);
  splits = std::min(splits, std::max(1, w.chunks / 4));
  w.per = ceil_div(w.chunks, splits);
  w.splits = ceil_div(w.chunks, w.per);
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  const int mp = rep ? d->pad : 0;
  w.dy_bytes = align_up((size_t)d->N * d->P * d->Q * cpad(d->K) * 2, 1024);
  w.x_bytes = align_up((size_t)d->N * (d->H + 2 * mp) * (d->W + 2 * mp) * cpad(d->C) * 2, 1024);
  w.part_bytes = align_up((size_t)w.splits * d->R * d->S * d->K * d->C * 4, 1024);
  return w;
}

static size_t wgrad_workspace(const ganb200_conv_desc* d) {
  WgradPlan w = wgrad_plan(d);
  return w.dy_bytes + w.x_bytes + w.part_bytes + 1024;
}

template <int BN, int STAGES>
static int launch_wg(const CUtensorMap& ta, const CUtensorMap& tb, const TcWgradParams& p, dim3 grid, cudaStream_t s) {
  static bool configured = false;
  constexpr int SMEM = TcWgSmem<BN, STAGES>::BYTES;
  if (!configured) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(tc_wgrad_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    configured = true;
  }
  tc_wgrad_kernel<BN, STAGES><<<grid, TC_THREADS, SMEM, s>>>(ta, tb, p);
  GB_CHECK_LAUNCH();
  return 0;
}

int tc_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale, const float* oscale,
             float* dw, void* ws, size_t ws_bytes, cudaStream_t s) {
  GB_REQUIRE(ws && ws_bytes >= wgrad_workspace(d), GANB200_WORKSPACE_TOO_SMALL, "tc wgrad: workspace too small");
  ws = align_ws(ws);
  const WgradPlan w = wgrad_plan(d);
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  const int mp = rep ? d->pad : 0;
  const int Kp = cpad(d->K), Cp = cpad(d->C), Hp = d->H + 2 * mp, Wp = d->W + 2 * mp, RS = d->R * d->S;
  __nv_bfloat16* dyh = (__nv_bfloat16*)ws;
  __nv_bfloat16* xh = (__nv_bfloat16*)((uint8_t*)ws + w.dy_bytes);
  float* part = (float*)((uint8_t*)ws + w.dy_bytes + w.x_bytes);
  {
    dim3 g1(d->N * d->P, (Kp + 63) / 64);
    pack_act_kernel<<<g1, 256, sizeof(float) * 64 * (d->Q + 1), s>>>(dy, oscale, dyh, d->K, d->P, d->Q, Kp, 0);
    GB_CHECK_LAUNCH();
    dim3 g2(d->N * Hp, (Cp + 63) / 64);
    pack_act_kernel<<<g2, 256, sizeof(float) * 64 * (d->W + 1), s>>>(x, iscale, xh, d->C, d->H, d->W, Cp, mp);
    GB_CHECK_LAUNCH();
  }
  TcWgradParams p;
  p.K = d->K; p.C = d->C; p.S = d->S; p.taps = RS;
  p.bw = w.bw; p.bh = w.bh; p.bn = w.bn; p.tiles_q = w.tiles_q; p.tiles_p = w.tiles_p; p.tiles_n = w.tiles_n;
  p.chunks = w.chunks; p.chunks_per_split = w.per; p.splits = w.splits;
  p.off_h = p.off_w = rep ? 0 : -d->pad;
  p.n_umma = d->C >= w.BN ? w.BN : (d->C + 15) / 16 * 16;
  p.part = part;
  CUtensorMap ta, tb;
  {
    uint64_t dims[4] = {(uint64_t)d->K, (uint64_t)d->Q, (uint64_t)d->P, (uint64_t)d->N};
    uint64_t str[3] = {(uint64_t)Kp * 2, (uint64_t)d->Q * Kp * 2, (uint64_t)d->P * d->Q * Kp * 2};
    uint32_t box[4] = {64, (uint32_t)w.bw, (uint32_t)w.bh, (uint32_t)w.bn};
    int rc = make_tmap_bf16(&ta, dyh, 4, dims, str, box);
    if (rc) return rc;
    uint64_t xd[4] = {(uint64_t)d->C, (uint64_t)Wp, (uint64_t)Hp, (uint64_t)d->N};
    uint64_t xs[3] = {(uint64_t)Cp * 2, (uint64_t)Wp * Cp * 2, (uint64_t)Hp * Wp * Cp * 2};
    rc = make_tmap_bf16(&tb, xh, 4, xd, xs, box);
    if (rc) return rc;
  }
  dim3 grid(ceil_div(d->K, TILE_M), ceil_div(d->C, w.BN), RS * w.splits);
  int rc = w.BN == 128 ? launch_wg<128, 4>(ta, tb, p, grid, s) : launch_wg<64, 4>(ta, tb, p, grid, s);
  if (rc) return rc;
  wgrad_reduce_kernel<<<ew_grid((int64_t)d->K * d->C, 256, 1), 256, 0, s>>>(part, dw, d->K, d->C, RS, w.splits, d->alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

}  // namespace ganb200
