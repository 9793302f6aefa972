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
//      tc_conv_halo_kernel (stride-1 problems on planes >= 256 pixels): one box per filter COLUMN, the filter rows and
//        two stacked sub-tiles are row-shifted views of it (see the kernel's comment).
// dgrad (stride 1) = the same kernels on dy with the flipped, transposed weights and zero padding R-1, written to the
//   padded domain and folded (adjoint of replicate padding) when needed; stride 2 = one problem per output phase.
// wgrad = tc_wgrad_kernel (one CTA per tap and pixel split) / tc_wgrad_ring_kernel (one CTA per filter column, rows of
//   the input in a shared-memory ring), both below, + a deterministic split reduction.
// Producer and MMA warps run warp-uniform loops and predicate the issue on elect.sync (tc_ptx.cuh: elect_one).
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

// Vectorised variant for W % 4 == 0 (every hot shape): one block = one INPUT row (n, h) x 64 channels.  float4 loads along
// w, the row is transposed in shared memory as bf16x2 words [w][33] (conflict-free both ways), 16-byte stores along c;
// the replicated border rows / columns are written from the same tile, so every input byte is read once.
__device__ __forceinline__ uint32_t pack_bf162(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__global__ void __launch_bounds__(256) pack_act_vec_kernel(const float* __restrict__ x, const float* __restrict__ scale,
                                                           __nv_bfloat16* __restrict__ out, int C, int H, int W, int Cp,
                                                           int pad) {
  extern __shared__ uint32_t tw[];  // [W][33]: word (w, cp) = channels (c0 + 2cp, c0 + 2cp + 1) of pixel w
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const int n = blockIdx.x / H, h = blockIdx.x % H;
  const int c0 = blockIdx.y * 64;
  const int G4 = W >> 2;
  const bool tiled = (G4 & 7) == 0;  // warp = 8 w-groups x 4 channel pairs -> 32 distinct banks
  for (int i = threadIdx.x; i < 32 * G4; i += 256) {
    int g, cp;
    if (tiled) {
      const int l = i & 31, t = i >> 5, tiles_g = G4 >> 3;
      g = (t % tiles_g) * 8 + (l >> 2);
      cp = (t / tiles_g) * 4 + (l & 3);
    } else {
      g = i % G4;
      cp = i / G4;
    }
    const int c = c0 + 2 * cp;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (c < C) {
      a = __ldg(reinterpret_cast<const float4*>(x + (((int64_t)n * C + c) * H + h) * W) + g);
      if (scale) { float sc = __ldg(scale + (int64_t)n * C + c); a.x *= sc; a.y *= sc; a.z *= sc; a.w *= sc; }
    }
    if (c + 1 < C) {
      b = __ldg(reinterpret_cast<const float4*>(x + (((int64_t)n * C + c + 1) * H + h) * W) + g);
      if (scale) { float sc = __ldg(scale + (int64_t)n * C + c + 1); b.x *= sc; b.y *= sc; b.z *= sc; b.w *= sc; }
    }
    uint32_t* d = tw + (4 * g) * 33 + cp;
    d[0] = pack_bf162(a.x, b.x);
    d[33] = pack_bf162(a.y, b.y);
    d[66] = pack_bf162(a.z, b.z);
    d[99] = pack_bf162(a.w, b.w);
  }
  __syncthreads();
  const int chunks = min(64, Cp - c0) >> 3;  // 16-byte chunks of valid (padded) channels
  const int hp_lo = (h == 0) ? 0 : h + pad, hp_hi = (h == H - 1) ? Hp - 1 : h + pad;
  for (int hp = hp_lo; hp <= hp_hi; ++hp) {
    __nv_bfloat16* orow = out + ((int64_t)n * Hp + hp) * Wp * Cp + c0;
    for (int i = threadIdx.x; i < 8 * Wp; i += 256) {
      const int k = i & 7, wp = i >> 3;
      if (k >= chunks) continue;
      const int w = min(max(wp - pad, 0), W - 1);
      const uint32_t* sp = tw + w * 33 + 4 * k;
      uint4 v = make_uint4(sp[0], sp[1], sp[2], sp[3]);
      *reinterpret_cast<uint4*>(orow + (int64_t)wp * Cp + 8 * k) = v;
    }
  }
}

// Second generation of the vectorised pack (used for 16 x 16 planes).  W is a template parameter (shifts instead of
// divisions), a warp loads a 4 w-group x 8 channel-pair tile, the transposed tile has a 34-word pitch — stores hit 32
// distinct banks (8*g + 2*j + cp) and the store phase reads two conflict-free 64-bit words per 16-byte store.
// (ncu on the first generation at 64 x 64: 4.26 TB/s with 0.69 issued warp-instructions per scheduler and cycle; cutting
// its instruction count this way did NOT make it faster there, so it stays in use for the wide planes.)
// ROWS input rows per block: every channel then contributes ROWS * W * 4 contiguous bytes (1 KB at W = 64) instead of
// one 256-byte row — the first generation read 64 separate 256-byte pieces 16 KB apart per block, one DRAM page each.
template <int LOG_G4, int ROWS>
__global__ void __launch_bounds__(256) pack_act_vec2_kernel(const float* __restrict__ x, const float* __restrict__ scale,
                                                            __nv_bfloat16* __restrict__ out, int C, int H, int Cp,
                                                            int pad) {
  constexpr int G4 = 1 << LOG_G4, W = 4 * G4, PITCH = 34;
  __shared__ __align__(16) uint32_t tw[ROWS][W * PITCH];
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const int hb = (H + ROWS - 1) / ROWS;
  const int n = blockIdx.x / hb, h0 = (blockIdx.x - n * hb) * ROWS;
  const int c0 = blockIdx.y * 64;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gl = lane >> 3, cl = lane & 7;
  constexpr int TILES_G = G4 / 4, TILES = TILES_G * 4;       // tiles of 4 w-groups x 8 channel pairs, per row
  const float* xrow = x + (((int64_t)n * C + c0) * H + h0) * W;
  const float* srow = scale ? scale + (int64_t)n * C + c0 : nullptr;
  const int64_t cstride = (int64_t)H * W;
  const int rows = min(ROWS, H - h0);
#pragma unroll
  for (int t = warp; t < TILES * ROWS; t += 8) {
    const int r = t / TILES, tt = t - r * TILES;
    if (r >= rows) break;
    const int g = (tt % TILES_G) * 4 + gl, cp = (tt / TILES_G) * 8 + cl;
    const int c = c0 + 2 * cp;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (c < C) {
      a = __ldg(reinterpret_cast<const float4*>(xrow + (2 * cp) * cstride + r * W) + g);
      if (srow) { const float sc = __ldg(srow + 2 * cp); a.x *= sc; a.y *= sc; a.z *= sc; a.w *= sc; }
    }
    if (c + 1 < C) {
      b = __ldg(reinterpret_cast<const float4*>(xrow + (2 * cp + 1) * cstride + r * W) + g);
      if (srow) { const float sc = __ldg(srow + 2 * cp + 1); b.x *= sc; b.y *= sc; b.z *= sc; b.w *= sc; }
    }
    uint32_t* d = tw[r] + (4 * g) * PITCH + cp;
    d[0] = pack_bf162(a.x, b.x);
    d[PITCH] = pack_bf162(a.y, b.y);
    d[2 * PITCH] = pack_bf162(a.z, b.z);
    d[3 * PITCH] = pack_bf162(a.w, b.w);
  }
  __syncthreads();
  const int chunks = min(64, Cp - c0) >> 3;   // 16-byte chunks of valid (padded) channels
  const int k = threadIdx.x & 7;
  if (k < chunks) {
    for (int r = 0; r < rows; ++r) {
      const int h = h0 + r;
      const int hp_lo = (h == 0) ? 0 : h + pad, hp_hi = (h == H - 1) ? Hp - 1 : h + pad;
      for (int hp = hp_lo; hp <= hp_hi; ++hp) {
        __nv_bfloat16* orow = out + ((int64_t)n * Hp + hp) * Wp * Cp + c0 + 8 * k;
        for (int wp = threadIdx.x >> 3; wp < Wp; wp += 32) {
          const int w = min(max(wp - pad, 0), W - 1);
          const uint2* sp = reinterpret_cast<const uint2*>(tw[r] + w * PITCH + 4 * k);
          const uint2 lo = sp[0], hi = sp[1];
          *reinterpret_cast<uint4*>(orow + (int64_t)wp * Cp) = make_uint4(lo.x, lo.y, hi.x, hi.y);
        }
      }
    }
  }
}

constexpr int MAX_TAPS = 49;
struct TapTable {
  int n;
  int src[MAX_TAPS];          // index of the tap inside the R*S filter
  signed char dh[MAX_TAPS];   // input coordinate offset of the tap
  signed char dw[MAX_TAPS];
};

// w fp32 -> out[t][ko][cp] bf16 with out(t,ko,ci) = w[ko*sk + ci*sc + taps.src[t]]
// (fprop: sk = C*RS, sc = RS; dgrad: the roles of K and C swap: sk = RS, sc = C*RS)
__global__ void __launch_bounds__(256) pack_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out,
                                                          int Ko, int Ci, int Cp, int64_t sk, int64_t sc,
                                                          const TapTable taps) {
  int64_t total = (int64_t)taps.n * Ko * Cp;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    int ci = (int)(i % Cp);
    int64_t t = i / Cp;
    int ko = (int)(t % Ko), tp = (int)(t / Ko);
    float v = 0.f;
    if (ci < Ci) v = w[ko * sk + ci * sc + taps.src[tp]];
    out[i] = __float2bfloat16(v);
  }
}

// ------------------------------------------------------------------------------------------------ main kernel
struct TcConvParams {
  int N, K, P, Q;       // output tile grid [N][K][P][Q] (one phase of the output when o_step > 1)
  int taps;             // number of filter taps
  int C;                // reduction channels
  int bw, bh, bn;       // tile box (bw*bh*bn == 128)
  int tiles_q, tiles_p; // ceil(Q/bw), ceil(P/bh)
  int tiles_m, tiles_n; // pixel tiles (tiles_q*tiles_p*ceil(N/bn)), output-channel tiles (ceil(K/BN))
  int in_step;          // input coordinate of output (p,q) = (p*in_step + dh[tap], q*in_step + dw[tap])
  int OH, OW;           // full output plane; element (p,q) is stored at (p*o_step + o_off_h, q*o_step + o_off_w)
  int o_step, o_off_h, o_off_w;
  signed char dh[MAX_TAPS], dw[MAX_TAPS];
  int n_umma;           // UMMA N = columns of the accumulator actually computed (multiple of 16, <= BN)
  float alpha;
  const float* oscale;  // [N][K] or null
  const float* bias;    // [K] or null
  float* y;
  // scatter epilogue (dgrad of a replicate-padded convolution, see epi_fold32): the tile grid covers the PADDED input
  // domain, the value of padded pixel (ph, pw) is ADDED to pixel (clamp(ph - pad), clamp(pw - pad)) of the FH x FW plane
  int fold, fold_pad, FH, FW;
  const float* red_x;   // [N][K][FH][FW] or null: red_out[n][k] += sum over pixels of (alpha * acc) * red_x
  float* red_out;       // [N][K]
};

template <int BN, int STAGES>
struct TcSmem {
  static constexpr int B_STAGE_BYTES = BN * BK * 2;
  static constexpr int BYTES = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + 1024 /*align*/ + 256 /*barriers*/;
};

// ------------------------------------------------------------------------------------------------ epilogue
// One thread = one output pixel, v[j] = accumulator of output channel kc + j (32 consecutive channels out of TMEM).
// y[k] = v * alpha * oscale[k] + bias[k], stored to the NCHW plane of channel k (plane stride `plane_bytes`); a warp
// writes 32 consecutive pixels of one channel per store instruction (128 B).  The epilogue warps are the only ones
// that touch global memory from registers, and with one or two CTAs per SM there are 4-8 of them per SM: every
// instruction per element counts (the first version spent ~25 on 64-bit index math and per-element null checks and
// was the kernel's critical path), so the pointer checks are template parameters, the channel guard only exists in
// the tail variant and the address advances by one 32-bit add per channel.
template <bool HAS_OSC, bool HAS_BIAS, bool FULL>
__device__ __forceinline__ void epi_store32(const float (&v)[32], char* yk, uint32_t plane_bytes, int nvalid, float alpha,
                                            const float* __restrict__ osc, const float* __restrict__ bias) {
  uint32_t off = 0;
#pragma unroll
  for (int j = 0; j < 32; ++j, off += plane_bytes) {
    if (FULL || j < nvalid) {
      float sc = alpha;
      if (HAS_OSC) sc *= __ldg(osc + j);
      float o = v[j] * sc;
      if (HAS_BIAS) o += __ldg(bias + j);
      *reinterpret_cast<float*>(yk + off) = o;
    }
  }
}

__device__ __forceinline__ void epi_chunk(const float (&v)[32], char* yk, uint32_t plane_bytes, int nvalid, float alpha,
                                          const float* osc, const float* bias) {
  if (nvalid >= 32) {
    if (osc) {
      if (bias) epi_store32<true, true, true>(v, yk, plane_bytes, 32, alpha, osc, bias);
      else epi_store32<true, false, true>(v, yk, plane_bytes, 32, alpha, osc, bias);
    } else {
      if (bias) epi_store32<false, true, true>(v, yk, plane_bytes, 32, alpha, osc, bias);
      else epi_store32<false, false, true>(v, yk, plane_bytes, 32, alpha, osc, bias);
    }
  } else {
    if (osc) {
      if (bias) epi_store32<true, true, false>(v, yk, plane_bytes, nvalid, alpha, osc, bias);
      else epi_store32<true, false, false>(v, yk, plane_bytes, nvalid, alpha, osc, bias);
    } else {
      if (bias) epi_store32<false, true, false>(v, yk, plane_bytes, nvalid, alpha, osc, bias);
      else epi_store32<false, false, false>(v, yk, plane_bytes, nvalid, alpha, osc, bias);
    }
  }
}

// Scatter epilogue: dgrad of a replicate-padded convolution WITHOUT the padded fp32 intermediate and the fold pass.
// The accumulator of a thread belongs to pixel (ph, pw) of the padded input; the adjoint of replicate padding adds it to
// the clamped pixel of the unpadded plane.  Only pixels on the border of the plane receive more than one contribution
// (`multi`): those go through red.global.add into a border that the host zeroed, every other pixel is a plain store.
// With red_x the per-(sample, channel) sum of (alpha * acc) * x — the style gradient of the modulated convolution,
// sum_hw dxs * x (generator_13_5.py:239 backward) — is formed here as well: by linearity the sum over the padded
// domain against x at the clamped position equals the sum over the folded plane.  32 channel values per lane are
// reduced over the warp's 32 pixels by recursive halving (31 shuffles, lane j ends with channel j) and added to
// red_out[n][k] with one coalesced atomic per warp and chunk.
template <bool HAS_OSC, bool HAS_RED>
__device__ __forceinline__ void epi_fold32(const float (&v)[32], char* yk, const char* xk, uint32_t plane_bytes, int nvalid,
                                           float alpha, const float* __restrict__ osc, bool ok, bool multi,
                                           float* __restrict__ red, int lane, float* racc = nullptr) {
  float r[32];
  uint32_t off = 0;
  if (HAS_RED) {
    // all 32 loads first (independent, ~one memory latency for the batch): interleaved with the stores below the compiler
    // must assume aliasing and serialises load -> store -> load, ~30 latencies per chunk (measured: 4x the whole kernel)
#pragma unroll
    for (int j = 0; j < 32; ++j, off += plane_bytes)
      r[j] = (ok && j < nvalid) ? __ldg(reinterpret_cast<const float*>(xk + off)) : 0.f;
    off = 0;
  }
#pragma unroll
  for (int j = 0; j < 32; ++j, off += plane_bytes) {
    const bool okj = ok && j < nvalid;
    const float t = v[j] * alpha;
    if (HAS_RED) r[j] *= t;
    if (okj) {
      const float o = HAS_OSC ? t * __ldg(osc + j) : t;
      float* dst = reinterpret_cast<float*>(yk + off);
      if (multi) atomicAdd(dst, o); else *dst = o;
    }
  }
  if (HAS_RED) {
#pragma unroll
    for (int w = 16; w >= 1; w >>= 1) {
      const bool upper = (lane & w) != 0;
#pragma unroll
      for (int j = 0; j < w; ++j) {
        const float keep = upper ? r[j + w] : r[j];
        const float send = upper ? r[j] : r[j + w];
        r[j] = keep + __shfl_xor_sync(0xffffffffu, send, w);
      }
    }
    if (racc) *racc += r[0];                       // the caller flushes when the (image, channel tile) changes
    else if (lane < nvalid) atomicAdd(red + lane, r[0]);
  }
}

__device__ __forceinline__ void epi_fold_chunk(const float (&v)[32], char* yk, const char* xk, uint32_t plane_bytes,
                                               int nvalid, float alpha, const float* osc, bool ok, bool multi, float* red,
                                               int lane, float* racc = nullptr) {
  if (red) {
    if (osc) epi_fold32<true, true>(v, yk, xk, plane_bytes, nvalid, alpha, osc, ok, multi, red, lane, racc);
    else epi_fold32<false, true>(v, yk, xk, plane_bytes, nvalid, alpha, osc, ok, multi, red, lane, racc);
  } else {
    if (osc) epi_fold32<true, false>(v, yk, xk, plane_bytes, nvalid, alpha, osc, ok, multi, red, lane);
    else epi_fold32<false, false>(v, yk, xk, plane_bytes, nvalid, alpha, osc, ok, multi, red, lane);
  }
}

// Persistent: each CTA walks tiles blockIdx.x, +gridDim.x, ... (n-tile fastest so that concurrent CTAs share the
// activation tile in L2).  The accumulator is double-buffered in TMEM (2 x BN columns): while the epilogue warps drain
// tile i, the producer and the MMA issuer are already in the main loop of tile i+1.
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
  uint64_t* tmem_full = empty + STAGES;    // [2]
  uint64_t* tmem_empty = tmem_full + 2;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_chunks = (p.C + BK - 1) / BK;
  const int total_tiles = p.tiles_m * p.tiles_n;
  constexpr uint32_t TMEM_COLS = 2 * BN;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);   // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // producer: the whole warp walks the loop (uniform control flow), the elected lane issues
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int t = tile / p.tiles_n;
      const int k0 = (tile - t * p.tiles_n) * BN;
      const int tq = t % p.tiles_q; t /= p.tiles_q;
      const int tp = t % p.tiles_p; t /= p.tiles_p;
      const int q0 = tq * p.bw, p0 = tp * p.bh, n0 = t * p.bn;
      for (int tap = 0; tap < p.taps; ++tap)
        for (int kc = 0; kc < k_chunks; ++kc) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (leader) {
            mbar_expect_tx(&full[stage], A_STAGE_BYTES + (uint32_t)p.n_umma * 128u);   // weight box = n_umma rows
            tma_load_4d(smA + stage * A_STAGE_BYTES, &tmA, &full[stage], kc * BK, q0 * p.in_step + p.dw[tap],
                        p0 * p.in_step + p.dh[tap], n0);
            tma_load_3d(smB + stage * B_STAGE_BYTES, &tmB, &full[stage], kc * BK, k0, tap);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
    }
    __syncwarp();
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(TILE_M, p.n_umma, 0, 0);
    int stage = 0, local = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const int acc = local & 1;
      mbar_wait(&tmem_empty[acc], ((local >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
      for (int tap = 0; tap < p.taps; ++tap)
        for (int kc = 0; kc < k_chunks; ++kc) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const int ksub = min(BK / 16, (p.C - kc * BK + 15) / 16);
          const uint64_t adesc = make_smem_desc_sw128(smem_u32(smA + stage * A_STAGE_BYTES), 16, 1024);
          const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB + stage * B_STAGE_BYTES), 16, 1024);
          if (leader) {
            // +32 bytes (= 2 in the >>4 address field) per 16-element k step
            mma_chain_n(ksub, tmem_d, adesc, bdesc, idesc, (tap > 0 || kc > 0) ? 1u : 0u, 2u);
            mma_commit(&empty[stage]);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      if (leader) mma_commit(&tmem_full[acc]);
    }
    __syncwarp();
  } else {
    // epilogue: warp w may touch TMEM lanes [32*(w%4), +32)
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const int wl = m % p.bw, hl = (m / p.bw) % p.bh, nl = m / (p.bw * p.bh);
    const int64_t PQ = (int64_t)p.OH * p.OW;
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      int t = tile / p.tiles_n;
      const int k0 = (tile - t * p.tiles_n) * BN;
      const int tq = t % p.tiles_q; t /= p.tiles_q;
      const int tp = t % p.tiles_p; t /= p.tiles_p;
      const int q = tq * p.bw + wl, pp = tp * p.bh + hl, n = t * p.bn + nl;
      const bool row_ok = (q < p.Q) && (pp < p.P) && (n < p.N);
      int64_t pix = (int64_t)(pp * p.o_step + p.o_off_h) * p.OW + q * p.o_step + p.o_off_w;
      int64_t plane = PQ;
      bool multi = false;
      if (p.fold) {
        const int fh = min(max(pp * p.o_step + p.o_off_h - p.fold_pad, 0), p.FH - 1);
        const int fw = min(max(q * p.o_step + p.o_off_w - p.fold_pad, 0), p.FW - 1);
        multi = p.fold_pad > 0 && (fh == 0 || fh == p.FH - 1 || fw == 0 || fw == p.FW - 1);
        pix = (int64_t)fh * p.FW + fw;
        plane = (int64_t)p.FH * p.FW;
      }
      const int64_t nbase = row_ok ? ((int64_t)n * p.K) * plane + pix : 0;
      float* ybase = p.y + nbase;
      const float* osc = p.oscale ? p.oscale + (int64_t)(row_ok ? n : 0) * p.K : nullptr;
      const int acc = local & 1;
      mbar_wait(&tmem_full[acc], (local >> 1) & 1);
      tc_fence_after();
      const uint32_t plane_bytes = (uint32_t)plane * 4u;
      for (int c0 = 0; c0 < p.n_umma; c0 += 32) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + c0), v);
        const int kc = k0 + c0;
        const int nvalid = min(p.n_umma - c0, p.K - kc);
        if (p.fold) {
          // bn == 1 whenever red_out is set (host side): a warp's 32 pixels belong to one image
          if (nvalid > 0)
            epi_fold_chunk(v, reinterpret_cast<char*>(ybase + (int64_t)kc * plane),
                           reinterpret_cast<const char*>(p.red_x ? p.red_x + nbase + (int64_t)kc * plane : nullptr),
                           plane_bytes, nvalid, p.alpha, osc ? osc + kc : nullptr, row_ok, multi,
                           p.red_out ? p.red_out + (int64_t)min(t * p.bn, p.N - 1) * p.K + kc : nullptr, lane);
        } else if (row_ok && nvalid > 0)
          epi_chunk(v, reinterpret_cast<char*>(ybase + (int64_t)kc * PQ), plane_bytes, nvalid, p.alpha,
                    osc ? osc + kc : nullptr, p.bias ? p.bias + kc : nullptr);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}


// ------------------------------------------------------------------------------------------------ halo kernel
// Stride-1 problems on planes of >= 256 output pixels whose taps form a dense R x S grid.  tc_conv_kernel fetches one
// activation box per filter tap, so every input pixel crosses L2 -> shared memory R*S times and the narrow 64x64
// convolutions of the generator are bound by that fill, not by the tensor pipe.  Here one CTA tile is MT stacked
// sub-tiles (MT*bh output rows x bw columns, bw*bh = 128) of ONE image, and per (filter column s, 64-channel chunk) ONE
// box of MT*bh + R - 1 input rows is loaded: the R taps of that column are the same box read from a start address
// shifted by whole rows (bw * 128 bytes, a multiple of the 1024-byte swizzle atom, so the SWIZZLE_128B phase is
// unchanged), and the MT sub-tiles are two more row shifts feeding MT accumulators.  Activation fill per output pixel
// drops from R*S to S * (MT*bh + R - 1) / (MT*bh) boxes, and each weight box now serves MT*128 pixels.
// Rings: A (halo boxes; as many stages as fit beside the B ring, >= 3) and B (weight boxes, HALO_B_STAGES x BN*128 B)
// with their own barriers.  Measured (ncu, r01): L2 -> shared-memory fill runs at ~(bytes in flight per SM) / 2.3 us,
// so the win comes from fewer bytes per tile AND as many of them in flight as 227 KB allow.
constexpr int HALO_SMEM_MAX = 227 * 1024;   // opt-in shared memory of one CTA (one CTA per SM)
constexpr int HALO_MAX_STAGES = 8;          // per ring (barrier slots)
constexpr int HALO_B_STAGES = 4;
constexpr int HALO_MIN_A_STAGES = 3;        // fill is latency-bound: at least two halo boxes in flight behind the one in use

struct TcHaloParams {
  int N, K, P, Q, C;
  int R, S;             // dense tap grid; tap (r, s) reads input (p + dh0 + r, q + dw0 + s)
  int dh0, dw0;
  int bw, bh, mt;       // sub-tile box (bw * bh == 128) and sub-tiles per CTA tile (stacked along rows)
  int a_bytes;          // bytes of one halo box = bw * (mt*bh + R - 1) * 128 (a multiple of 1024)
  int a_stages;         // stages of the A ring (3..8)
  int tiles_q, tiles_p, tiles_m, tiles_n;
  int OH, OW, o_step, o_off_h, o_off_w;
  int n_umma;
  float alpha;
  const float* oscale;
  const float* bias;
  float* y;
  int fold, fold_pad, FH, FW;       // scatter epilogue, as in TcConvParams
  const float* red_x;
  float* red_out;
  // range_mode (set with red_out): a CTA walks a CONTIGUOUS range of tiles, pixel tiles fastest, so that consecutive tiles
  // belong to one (image, channel tile) and the epilogue keeps the reduction in registers until that key changes
  int range_mode, tiles_per_cta;
  unsigned char tap_idx[MAX_TAPS];  // packed-weight index of grid point r * S + s
};

template <int BN>
struct TcHaloSmem {
  static constexpr int B_STAGE_BYTES = BN * BK * 2;
  static constexpr int FIXED = HALO_B_STAGES * B_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr int A_REGION = HALO_SMEM_MAX - FIXED;
};

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_conv_halo_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                  const __grid_constant__ CUtensorMap tmB,
                                                                  const TcHaloParams p) {
  constexpr int B_STAGE_BYTES = BN * BK * 2;
  constexpr int MT_MAX = 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int SA = p.a_stages;
  uint8_t* smA = smem;
  uint8_t* smB = smem + SA * p.a_bytes;
  uint64_t* fullA = reinterpret_cast<uint64_t*>(smB + HALO_B_STAGES * B_STAGE_BYTES);
  uint64_t* emptyA = fullA + HALO_MAX_STAGES;
  uint64_t* fullB = emptyA + HALO_MAX_STAGES;
  uint64_t* emptyB = fullB + HALO_B_STAGES;
  uint64_t* tmem_full = emptyB + HALO_B_STAGES;  // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_chunks = (p.C + BK - 1) / BK;
  const int total_tiles = p.tiles_m * p.tiles_n;
  const int rows_per_tile = p.mt * p.bh;
  constexpr uint32_t TMEM_COLS = 2 * MT_MAX * BN;  // double-buffered, MT accumulators each
  // tile walk: strided over the grid (default: concurrent CTAs share activation rows in L2) or a contiguous range
  const int t_first = p.range_mode ? (int)blockIdx.x * p.tiles_per_cta : (int)blockIdx.x;
  const int t_end = p.range_mode ? min(total_tiles, t_first + p.tiles_per_cta) : total_tiles;
  const int t_step = p.range_mode ? 1 : (int)gridDim.x;
  // tile -> (k0, tq, tp, n): channel tile fastest by default, pixel tiles fastest in range mode
  auto decode = [&](int tile, int& k0, int& tq, int& tp, int& n) {
    if (p.range_mode) {
      int t = tile;
      tq = t % p.tiles_q; t /= p.tiles_q;
      tp = t % p.tiles_p; t /= p.tiles_p;
      k0 = (t % p.tiles_n) * BN;
      n = t / p.tiles_n;
    } else {
      int t = tile / p.tiles_n;
      k0 = (tile - t * p.tiles_n) * BN;
      tq = t % p.tiles_q; t /= p.tiles_q;
      tp = t % p.tiles_p; t /= p.tiles_p;
      n = t;
    }
  };

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < SA; ++i) { mbar_init(&fullA[i], 1); mbar_init(&emptyA[i], 1); }
    for (int i = 0; i < HALO_B_STAGES; ++i) { mbar_init(&fullB[i], 1); mbar_init(&emptyB[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // producer: warp-uniform loops, the elected lane issues
    const bool leader = elect_one();
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0;
    for (int tile = t_first; tile < t_end; tile += t_step) {
      int k0, tq, tp, n0;
      decode(tile, k0, tq, tp, n0);
      const int q0 = tq * p.bw, p0 = tp * rows_per_tile;
      for (int kc = 0; kc < k_chunks; ++kc)
        for (int s = 0; s < p.S; ++s) {
          mbar_wait(&emptyA[sa], pa ^ 1);
          if (leader) {
            mbar_expect_tx(&fullA[sa], (uint32_t)p.a_bytes);
            tma_load_4d(smA + sa * p.a_bytes, &tmA, &fullA[sa], kc * BK, q0 + p.dw0 + s, p0 + p.dh0, n0);
          }
          if (++sa == SA) { sa = 0; pa ^= 1; }
          for (int r = 0; r < p.R; ++r) {
            mbar_wait(&emptyB[sb], pb ^ 1);
            if (leader) {
              mbar_expect_tx(&fullB[sb], (uint32_t)p.n_umma * 128u);   // weight box = n_umma rows
              tma_load_3d(smB + sb * B_STAGE_BYTES, &tmB, &fullB[sb], kc * BK, k0, (int)p.tap_idx[r * p.S + s]);
            }
            if (++sb == HALO_B_STAGES) { sb = 0; pb ^= 1; }
          }
        }
    }
    __syncwarp();
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(TILE_M, p.n_umma, 0, 0);
    const uint32_t row_bytes = (uint32_t)p.bw * 128u;  // one input row of the halo box
    const uint32_t sub_bytes = (uint32_t)p.bh * row_bytes;
    int sa = 0, sb = 0, local = 0;
    uint32_t pa = 0, pb = 0;
    for (int tile = t_first; tile < t_end; tile += t_step, ++local) {
      const int acc = local & 1;
      mbar_wait(&tmem_empty[acc], ((local >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * MT_MAX * BN);
      for (int kc = 0; kc < k_chunks; ++kc) {
        const int ksub = min(BK / 16, (p.C - kc * BK + 15) / 16);
        for (int s = 0; s < p.S; ++s) {
          mbar_wait(&fullA[sa], pa);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smA + sa * p.a_bytes);
          for (int r = 0; r < p.R; ++r) {
            mbar_wait(&fullB[sb], pb);
            tc_fence_after();
            const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB + sb * B_STAGE_BYTES), 16, 1024);
            const uint64_t adesc0 = make_smem_desc_sw128(a_base + (uint32_t)r * row_bytes, 16, 1024);
            const uint32_t fresh = (kc == 0 && s == 0 && r == 0) ? 0u : 1u;
            if (leader) {
              mma_chain_n(ksub, tmem_d, adesc0, bdesc, idesc, fresh, 2u);
              if (p.mt == 2)
                mma_chain_n(ksub, tmem_d + (uint32_t)BN, adesc0 + (uint64_t)(sub_bytes >> 4), bdesc, idesc, fresh, 2u);
              mma_commit(&emptyB[sb]);
            }
            if (++sb == HALO_B_STAGES) { sb = 0; pb ^= 1; }
          }
          if (leader) mma_commit(&emptyA[sa]);
          if (++sa == SA) { sa = 0; pa ^= 1; }
        }
      }
      if (leader) mma_commit(&tmem_full[acc]);
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const int wl = m % p.bw, hl = m / p.bw;
    const int64_t PQ = (int64_t)p.OH * p.OW;
    const uint32_t plane_bytes = (uint32_t)PQ * 4u;
    int local = 0;
    float racc[BN / 32];            // range mode: this lane's running sums, one per 32-channel chunk
    int racc_n = -1, racc_k0 = 0;
#pragma unroll
    for (int c = 0; c < BN / 32; ++c) racc[c] = 0.f;
    auto flush = [&]() {
      if (racc_n >= 0) {
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) {
          const int k = racc_k0 + c * 32 + lane;
          if (c * 32 < p.n_umma && k < p.K) atomicAdd(p.red_out + (int64_t)racc_n * p.K + k, racc[c]);
          racc[c] = 0.f;
        }
      }
    };
    for (int tile = t_first; tile < t_end; tile += t_step, ++local) {
      int k0, tq, tp, n;
      decode(tile, k0, tq, tp, n);
      if (p.range_mode && p.red_out && (n != racc_n || k0 != racc_k0)) { flush(); racc_n = n; racc_k0 = k0; }
      const int q = tq * p.bw + wl;
      const float* osc = p.oscale ? p.oscale + (int64_t)n * p.K : nullptr;
      const int acc = local & 1;
      mbar_wait(&tmem_full[acc], (local >> 1) & 1);
      tc_fence_after();
      for (int j = 0; j < p.mt; ++j) {
        const int pp = tp * rows_per_tile + j * p.bh + hl;
        const bool row_ok = (q < p.Q) && (pp < p.P);
        int64_t pix = (int64_t)(pp * p.o_step + p.o_off_h) * p.OW + q * p.o_step + p.o_off_w;
        int64_t plane = PQ;
        bool multi = false;
        if (p.fold) {
          const int fh = min(max(pp * p.o_step + p.o_off_h - p.fold_pad, 0), p.FH - 1);
          const int fw = min(max(q * p.o_step + p.o_off_w - p.fold_pad, 0), p.FW - 1);
          multi = p.fold_pad > 0 && (fh == 0 || fh == p.FH - 1 || fw == 0 || fw == p.FW - 1);
          pix = (int64_t)fh * p.FW + fw;
          plane = (int64_t)p.FH * p.FW;
        }
        const int64_t nbase = ((int64_t)n * p.K) * plane + (row_ok ? pix : 0);
        float* ybase = p.y + nbase;
        const uint32_t pb = p.fold ? (uint32_t)plane * 4u : plane_bytes;
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 32) {   // static trip count: racc[c0 >> 5] stays in registers
          if (c0 >= p.n_umma) break;
          float v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)((acc * MT_MAX + j) * BN + c0), v);
          const int kc = k0 + c0;
          const int nvalid = min(p.n_umma - c0, p.K - kc);
          if (p.fold) {
            if (nvalid > 0)
              epi_fold_chunk(v, reinterpret_cast<char*>(ybase + (int64_t)kc * plane),
                             reinterpret_cast<const char*>(p.red_x ? p.red_x + nbase + (int64_t)kc * plane : nullptr), pb,
                             nvalid, p.alpha, osc ? osc + kc : nullptr, row_ok, multi,
                             p.red_out ? p.red_out + (int64_t)n * p.K + kc : nullptr, lane,
                             p.range_mode ? &racc[c0 >> 5] : nullptr);
          } else if (row_ok && nvalid > 0)
            epi_chunk(v, reinterpret_cast<char*>(ybase + (int64_t)kc * PQ), plane_bytes, nvalid, p.alpha,
                      osc ? osc + kc : nullptr, p.bias ? p.bias + kc : nullptr);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
    if (p.range_mode && p.red_out) flush();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ flat kernel
// Stride-1 forward problems whose taps all fall inside the MATERIALISED padding of the packed input (replicate-padded
// convolutions: every layer of the generator and the critic).  The padded NHWC input is read as what it is in memory, a
// 2-D matrix [N*Hp*Wp flat pixels][C], and the convolution as a 1-D one over flat pixels: output o = (n*Hp + y)*Wp + x
// needs input rows o + (dh0 + r)*Wp + (dw0 + s).  One CTA tile is MT*128 CONSECUTIVE flat pixels of one image, and per
// 64-channel chunk ONE contiguous range of MT*128 + (R-1)*Wp + (S-1) flat rows is loaded (as 64-row TMA boxes): every
// tap of the filter is that same shared-memory range read from a start address shifted by (r*Wp + s) rows of 128 bytes.
// The shifts are not multiples of the 8-row SWIZZLE_128B atom; the tensor core applies the swizzle to the absolute
// shared-memory address, so any 128-byte shift addresses the rows TMA wrote (csrc/probe.cu / ganb200_tc_shift_probe is the
// hardware check, tests/test_gpu_kernels.py).  Activation fill per output pixel drops from S * (rows + R - 1) / rows boxes
// (halo kernel; measured 48 -> 48 5x5 at 64x64: 1.64 GB L2 -> shared memory per launch at 6.3 TB/s, tensor pipe 23 %) to
// (MT*128 + (R-1)*Wp + S-1) / (MT*128) rows — 2.1 instead of 7.5 input rows per output row on that layer.
// Outputs at x >= Q (the Wp - Q columns of row padding) and y >= P are computed and dropped: (Wp - Q) / Wp of the MMA work.
// Same tap order and k order per output element as the halo kernel: identical results.
// Warps: 0 = activation producer, 1 = MMA issuer, 2-5 = epilogue, 6 = weight producer (its own ring, so that the next
// tile's activations are requested while the current tile's weight boxes still stream).
constexpr int FLAT_THREADS = 224;
constexpr int FLAT_BOX_ROWS = 32;                           // rows per small TMA box (4 KB): the tail of a stage
constexpr int FLAT_BOX_BYTES = FLAT_BOX_ROWS * 128;
constexpr int FLAT_BIG_ROWS = 256;                          // rows per large TMA box (32 KB, the largest box there is):
constexpr int FLAT_BIG_BYTES = FLAT_BIG_ROWS * 128;         // every TMA issue costs the producer ~0.1 us (measured)
constexpr int FLAT_MAX_B_STAGES = 16;                       // weight ring (barrier slots)
constexpr int FLAT_FIXED_BYTES = 1024 /*align*/ + 512 /*barriers*/;

struct TcFlatParams {
  int N, K, P, Q, C;
  int R, S, dh0, dw0;
  int Hp, Wp;            // padded input plane
  int mt;                // sub-tiles (accumulators) of 128 flat pixels per CTA tile: 1, 2 or (64-wide tiles) 4
  int nbig, nbox;        // 256-row and 32-row boxes per activation stage
  int a_stages;
  int b_stages;          // weight ring: boxes of n_umma * 128 bytes; the fill is latency-bound, so as deep as fits
  int tiles_img;         // tiles per image = ceil(P * Wp / (mt * 128))
  int tiles_m, tiles_n;
  int OH, OW, o_step, o_off_h, o_off_w;
  int n_umma;
  float alpha;
  const float* oscale;
  const float* bias;
  float* y;
  // diagnosis (ganb200_conv_flat_trace): clock64 stamps of CTA 0's pipeline roles, [role][tile < 32][4]; null = off
  unsigned long long* trace;
  unsigned char tap_idx[MAX_TAPS];  // packed-weight index of grid point r * S + s
};
constexpr int FLAT_TRACE_TILES = 32;
__device__ __forceinline__ void flat_stamp(unsigned long long* trace, int role, int tile_local, int k) {
  if (trace && blockIdx.x == 0 && tile_local < FLAT_TRACE_TILES)
    trace[(role * FLAT_TRACE_TILES + tile_local) * 4 + k] = (unsigned long long)clock64();
}

template <int BN>
__global__ void __launch_bounds__(FLAT_THREADS, 1) tc_conv_flat_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                       const __grid_constant__ CUtensorMap tmA32,
                                                                       const __grid_constant__ CUtensorMap tmB,
                                                                       const TcFlatParams p) {
  constexpr int MT_MAX = 256 / BN;             // accumulators per tile: 2 x MT_MAX x BN = 512 TMEM columns
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int SA = p.a_stages, SB = p.b_stages;
  const int a_bytes = p.nbig * FLAT_BIG_BYTES + p.nbox * FLAT_BOX_BYTES;
  const int b_bytes = p.n_umma * 128;            // a multiple of 1024 (n_umma is a multiple of 8)
  uint8_t* smA = smem;
  uint8_t* smB = smem + SA * a_bytes;
  uint64_t* fullA = reinterpret_cast<uint64_t*>(smB + SB * b_bytes);
  uint64_t* emptyA = fullA + HALO_MAX_STAGES;
  uint64_t* fullB = emptyA + HALO_MAX_STAGES;
  uint64_t* emptyB = fullB + FLAT_MAX_B_STAGES;
  uint64_t* tmem_full = emptyB + FLAT_MAX_B_STAGES;  // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_chunks = (p.C + BK - 1) / BK;
  const int total_tiles = p.tiles_m * p.tiles_n;
  const int tile_px = p.mt * TILE_M;
  constexpr uint32_t TMEM_COLS = 2 * MT_MAX * BN;
  // tile -> (k0, position of the tile inside its image, image): channel tile fastest
  auto decode = [&](int tile, int& k0, int& ti, int& n) {
    const int t = tile / p.tiles_n;
    k0 = (tile - t * p.tiles_n) * BN;
    n = t / p.tiles_img;
    ti = t - n * p.tiles_img;
  };

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmA32);
    prefetch_tmap(&tmB);
    for (int i = 0; i < SA; ++i) { mbar_init(&fullA[i], 1); mbar_init(&emptyA[i], 1); }
    for (int i = 0; i < SB; ++i) { mbar_init(&fullB[i], 1); mbar_init(&emptyB[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // activation producer
    const bool leader = elect_one();
    int sa = 0, local = 0;
    uint32_t pa = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      int k0, ti, n;
      decode(tile, k0, ti, n);
      const int row0 = n * p.Hp * p.Wp + ti * tile_px + p.dh0 * p.Wp + p.dw0;
      for (int kc = 0; kc < k_chunks; ++kc) {
        if (leader && kc == 0) flat_stamp(p.trace, 0, local, 0);
        mbar_wait(&emptyA[sa], pa ^ 1);
        if (leader && kc == 0) flat_stamp(p.trace, 0, local, 1);
        if (leader) {
          mbar_expect_tx(&fullA[sa], (uint32_t)a_bytes);
          uint8_t* dst = smA + sa * a_bytes;
          int row = row0;
          for (int b = 0; b < p.nbig; ++b, dst += FLAT_BIG_BYTES, row += FLAT_BIG_ROWS)
            tma_load_3d(dst, &tmA, &fullA[sa], kc * BK, row, 0);
          for (int b = 0; b < p.nbox; ++b, dst += FLAT_BOX_BYTES, row += FLAT_BOX_ROWS)
            tma_load_3d(dst, &tmA32, &fullA[sa], kc * BK, row, 0);
        }
        if (++sa == SA) { sa = 0; pa ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 6) {
    // weight producer: one box per (chunk, tap), in the order the MMA warp consumes them
    const bool leader = elect_one();
    int sb = 0;
    uint32_t pb = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int k0, ti, n;
      decode(tile, k0, ti, n);
      for (int kc = 0; kc < k_chunks; ++kc)
        for (int s = 0; s < p.S; ++s)
          for (int r = 0; r < p.R; ++r) {
            mbar_wait(&emptyB[sb], pb ^ 1);
            if (leader) {
              mbar_expect_tx(&fullB[sb], (uint32_t)p.n_umma * 128u);
              tma_load_3d(smB + sb * b_bytes, &tmB, &fullB[sb], kc * BK, k0, (int)p.tap_idx[r * p.S + s]);
            }
            if (++sb == SB) { sb = 0; pb ^= 1; }
          }
    }
    __syncwarp();
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(TILE_M, p.n_umma, 0, 0);
    int sa = 0, sb = 0, local = 0;
    uint32_t pa = 0, pb = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const int acc = local & 1;
      if (leader) flat_stamp(p.trace, 1, local, 0);
      mbar_wait(&tmem_empty[acc], ((local >> 1) & 1) ^ 1);
      tc_fence_after();
      if (leader) flat_stamp(p.trace, 1, local, 1);
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * MT_MAX * BN);
      for (int kc = 0; kc < k_chunks; ++kc) {
        const int ksub = min(BK / 16, (p.C - kc * BK + 15) / 16);
        mbar_wait(&fullA[sa], pa);
        tc_fence_after();
        if (leader && kc == 0) flat_stamp(p.trace, 1, local, 2);
        const uint32_t a_base = smem_u32(smA + sa * a_bytes);
        for (int s = 0; s < p.S; ++s)
          for (int r = 0; r < p.R; ++r) {
            mbar_wait(&fullB[sb], pb);
            tc_fence_after();
            const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB + sb * b_bytes), 16, 1024);
            const uint64_t adesc0 = make_smem_desc_sw128(a_base + (uint32_t)(r * p.Wp + s) * 128u, 16, 1024);
            const uint32_t fresh = (kc == 0 && s == 0 && r == 0) ? 0u : 1u;
            if (leader) {
#pragma unroll
              for (int j = 0; j < MT_MAX; ++j)
                if (j < p.mt)
                  mma_chain_n(ksub, tmem_d + (uint32_t)(j * BN), adesc0 + (uint64_t)((j * TILE_M * 128) >> 4), bdesc, idesc,
                              fresh, 2u);
              mma_commit(&emptyB[sb]);
            }
            if (++sb == SB) { sb = 0; pb ^= 1; }
          }
        if (leader) mma_commit(&emptyA[sa]);
        if (++sa == SA) { sa = 0; pa ^= 1; }
      }
      if (leader) mma_commit(&tmem_full[acc]);
      if (leader) flat_stamp(p.trace, 1, local, 3);
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const int64_t PQ = (int64_t)p.OH * p.OW;
    const uint32_t plane_bytes = (uint32_t)PQ * 4u;
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      int k0, ti, n;
      decode(tile, k0, ti, n);
      const float* osc = p.oscale ? p.oscale + (int64_t)n * p.K : nullptr;
      const int acc = local & 1;
      if (warp == 2 && lane == 0) flat_stamp(p.trace, 2, local, 0);
      mbar_wait(&tmem_full[acc], (local >> 1) & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) flat_stamp(p.trace, 2, local, 1);
      for (int j = 0; j < p.mt; ++j) {
        const int o = ti * tile_px + j * TILE_M + m;      // flat position inside the image's padded plane
        const int yy = o / p.Wp, xx = o - yy * p.Wp;
        const bool row_ok = (xx < p.Q) && (yy < p.P);
        const int64_t pix = (int64_t)(yy * p.o_step + p.o_off_h) * p.OW + xx * p.o_step + p.o_off_w;
        float* ybase = p.y + ((int64_t)n * p.K) * PQ + (row_ok ? pix : 0);
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 32) {
          if (c0 >= p.n_umma) break;
          float v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)((acc * MT_MAX + j) * BN + c0), v);
          const int kc = k0 + c0;
          const int nvalid = min(p.n_umma - c0, p.K - kc);
          if (row_ok && nvalid > 0)
            epi_chunk(v, reinterpret_cast<char*>(ybase + (int64_t)kc * PQ), plane_bytes, nvalid, p.alpha,
                      osc ? osc + kc : nullptr, p.bias ? p.bias + kc : nullptr);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (warp == 2 && lane == 0) flat_stamp(p.trace, 2, local, 2);
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
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
                   const uint32_t* box, const uint32_t* estr = nullptr) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail(GANB200_CUDA_ERROR, "cuTensorMapEncodeTiled not available%s", "");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = estr ? estr[i] : 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides[i];
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(GANB200_CUDA_ERROR, "cuTensorMapEncodeTiled failed (%s code %lld)", "", (long long)r);
  return 0;
}

static inline int pow2ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
// Channel pitch of the packed (bf16, channels-last) operands.  Tensors of more than 32 channels get whole 128-byte rows
// (64 channels): TMA fills shared memory from rows whose global pitch is a multiple of 128 bytes at about twice the
// rate of the 96-byte rows a 48-channel tensor would have (tools/flat_bench.py, profiles/r02_flat_kernel_experiments.txt:
// the activation stage of a 64-channel layer arrives at 7.2 TB/s chip-wide, that of a 48-channel layer at 3.7 TB/s).
// GANB200_CPAD64=0: multiples of 8 as in round 1.
static inline int cpad(int c) {
  static int wide = -1;
  if (wide < 0) { const char* e = getenv("GANB200_CPAD64"); wide = (e && e[0] == '0') ? 0 : 1; }
  return (wide && c > 32) ? (c + 63) / 64 * 64 : (c + 7) / 8 * 8;
}
static inline void* align_ws(void* ws) { return (void*)(((uintptr_t)ws + 1023) & ~uintptr_t(1023)); }

// the generic problem fprop and dgrad map to: out[n,ko,(p,q)] = alpha*oscale*sum_{tap,ci} W[tap][ko][ci] *
// in[n, ci, p*in_step + dh[tap], q*in_step + dw[tap]] (+bias), stored at (p*o_step + o_off_h, q*o_step + o_off_w)
struct TcProblem {
  int N, Ci, Hi, Wi;      // input  fp32 NCHW
  int Ko;
  int Po, Qo;             // output tile grid
  int OH, OW, o_step, o_off_h, o_off_w;
  int mat_pad;            // replicate padding materialised while packing
  int in_step;
  int64_t w_sk, w_sc;     // weight strides for (ko, ci); taps innermost
  TapTable taps;
  float alpha;
  // scatter epilogue (dgrad of replicate padding fused, optional style-gradient reduction): see epi_fold32
  int fold = 0, fold_pad = 0, FH = 0, FW = 0;
  const float* red_x = nullptr;
  float* red_out = nullptr;
};

static size_t act_bytes(int N, int Ci, int Hi, int Wi, int mat_pad) {
  return align_up((size_t)N * (Hi + 2 * mat_pad) * (Wi + 2 * mat_pad) * cpad(Ci) * 2, 1024);
}
static size_t w_bytes(int taps, int Ko, int Ci) { return align_up((size_t)taps * Ko * cpad(Ci) * 2, 1024); }

// 1 = use the halo kernel where it applies (default), 0 = always the per-tap kernel.  GANB200_CONV_HALO=0 in the
// environment or ganb200_conv_halo_mode(0) switch it off (parity tests run both).
static int g_halo_mode = -1;
static int halo_mode() {
  if (g_halo_mode < 0) {
    const char* e = getenv("GANB200_CONV_HALO");
    g_halo_mode = (e && e[0] == '0') ? 0 : (e && e[0] == '2') ? 2 : 1;
  }
  return g_halo_mode;
}

// Fills `h` and returns true when problem `q` can run on the halo kernel.
static bool halo_plan(const TcProblem& q, int BN, TcHaloParams& h) {
  if (q.in_step != 1 || q.taps.n < 1 || q.taps.n > MAX_TAPS) return false;
  if ((int64_t)q.Po * q.Qo < 256 || q.Qo < 8) return false;
  // Measured (tools/conv_bench.py, B = 256): 3-tap-wide filters over channel counts that are multiples of 64 (the
  // critic's 64/128/256-channel 3x3 convs) run 5-10 % faster on the per-tap kernel, whose 128-byte-pitch boxes the TMA
  // streams at full rate with two CTAs per SM; everything else (k5, 48/96/108-channel, 1x1, strided-dgrad phases) is
  // 1.1-2.2x faster here.
  if (halo_mode() != 2 && q.taps.n == 9 && q.Ci % 64 == 0 && q.Ko % 64 == 0) return false;
  int dh_lo = 127, dh_hi = -128, dw_lo = 127, dw_hi = -128;
  for (int t = 0; t < q.taps.n; ++t) {
    dh_lo = std::min(dh_lo, (int)q.taps.dh[t]); dh_hi = std::max(dh_hi, (int)q.taps.dh[t]);
    dw_lo = std::min(dw_lo, (int)q.taps.dw[t]); dw_hi = std::max(dw_hi, (int)q.taps.dw[t]);
  }
  const int R = dh_hi - dh_lo + 1, S = dw_hi - dw_lo + 1;
  if (R * S != q.taps.n) return false;
  bool seen[MAX_TAPS] = {false};
  for (int t = 0; t < q.taps.n; ++t) {
    const int g = (q.taps.dh[t] - dh_lo) * S + (q.taps.dw[t] - dw_lo);
    if (seen[g]) return false;
    seen[g] = true;
    h.tap_idx[g] = (unsigned char)t;
  }
  // tile: the (bw, bh, mt) with the least L2 -> shared-memory fill over the whole problem
  double best = -1.0;
  // bw >= 32 keeps the epilogue's warp stores at 128 contiguous bytes (planes narrower than 32: the whole row)
  const int a_region = BN == 128 ? TcHaloSmem<128>::A_REGION : TcHaloSmem<64>::A_REGION;
  for (int bw = std::min(32, pow2ceil(q.Qo)); bw <= TILE_M; bw <<= 1) {
    const int bh = TILE_M / bw;
    if (bw > pow2ceil(q.Qo)) continue;
    for (int mt = 1; mt <= 2; ++mt) {
      const int rows = mt * bh + R - 1;
      const int64_t a_bytes = (int64_t)bw * rows * 128;
      if (a_bytes * HALO_MIN_A_STAGES > a_region || rows > 256) continue;
      const int64_t tiles = (int64_t)ceil_div(q.Qo, bw) * ceil_div(q.Po, mt * bh) * q.N;
      const double cost = (double)tiles * ((double)S * a_bytes + (double)R * S * std::min(q.Ko, BN) * 128.0);
      if (best < 0 || cost < best) { best = cost; h.bw = bw; h.bh = bh; h.mt = mt; }
    }
  }
  if (best < 0) return false;
  h.N = q.N; h.K = q.Ko; h.P = q.Po; h.Q = q.Qo; h.C = q.Ci;
  h.R = R; h.S = S; h.dh0 = dh_lo; h.dw0 = dw_lo;
  h.a_bytes = h.bw * (h.mt * h.bh + R - 1) * 128;
  h.a_stages = std::min(HALO_MAX_STAGES, a_region / h.a_bytes);
  h.tiles_q = ceil_div(q.Qo, h.bw); h.tiles_p = ceil_div(q.Po, h.mt * h.bh);
  h.tiles_m = h.tiles_q * h.tiles_p * q.N; h.tiles_n = ceil_div(q.Ko, BN);
  h.OH = q.OH; h.OW = q.OW; h.o_step = q.o_step; h.o_off_h = q.o_off_h; h.o_off_w = q.o_off_w;
  return true;
}

// 1 (default) = the flat kernel where it measured faster than the halo / per-tap kernels (the rule is at the call site in
// run_problem); 2 = wherever a forward problem qualifies (parity tests); 0 = never.  GANB200_CONV_FLAT or
// ganb200_conv_flat_mode().
// What the experiments showed (profiles/r02_flat_kernel_experiments.txt): the 48-channel layers are NOT bound by the bytes
// they move from L2 — with a quarter of the activation fill (this kernel), with the weight ring 13 deep instead of 4, with
// no weight traffic at all, with 2, 3 or 4 activation stages in flight, with 32-, 64- or 256-row TMA boxes, the launch
// takes the same time to 5 %; with the MMAs removed it still takes 60 % of it.  What it does follow is the NUMBER OF CTA
// TILES: time = (shared-memory-bound MMA time: 4 KB of A per 24-cycle N = 48 instruction) + ~4 us per tile and CTA.  Four
// accumulators per tile (512 pixels; 64-wide channel tiles leave the TMEM columns for it) instead of two is what made this
// kernel 10-14 % faster than the halo kernel on those layers.
static unsigned long long* g_flat_trace = nullptr;   // ganb200_conv_flat_trace
static int g_flat_mode = -1;
static int flat_mode() {
  if (g_flat_mode < 0) {
    const char* e = getenv("GANB200_CONV_FLAT");
    g_flat_mode = (e && e[0] == '0') ? 0 : (e && e[0] == '2') ? 2 : 1;
  }
  return g_flat_mode;
}

// Fills `f` and returns true when problem `q` can run on the flat kernel.
static bool flat_plan(const TcProblem& q, int BN, TcFlatParams& f) {
  if (q.in_step != 1 || q.taps.n < 1 || q.taps.n > MAX_TAPS || q.fold || q.red_out) return false;
  if (q.Qo < 32 || q.Po < 8) return false;     // narrow planes: the row padding would be too large a share of the work
  const int Hp = q.Hi + 2 * q.mat_pad, Wp = q.Wi + 2 * q.mat_pad;
  int dh_lo = 127, dh_hi = -128, dw_lo = 127, dw_hi = -128;
  for (int t = 0; t < q.taps.n; ++t) {
    dh_lo = std::min(dh_lo, (int)q.taps.dh[t]); dh_hi = std::max(dh_hi, (int)q.taps.dh[t]);
    dw_lo = std::min(dw_lo, (int)q.taps.dw[t]); dw_hi = std::max(dw_hi, (int)q.taps.dw[t]);
  }
  const int R = dh_hi - dh_lo + 1, S = dw_hi - dw_lo + 1;
  if (R * S != q.taps.n) return false;
  // every tap of every valid output must read a pixel of the SAME image's padded plane: the flat view has no border
  if (dh_lo < 0 || dw_lo < 0 || q.Po - 1 + dh_hi > Hp - 1 || q.Qo - 1 + dw_hi > Wp - 1) return false;
  bool seen[MAX_TAPS] = {false};
  for (int t = 0; t < q.taps.n; ++t) {
    const int g = (q.taps.dh[t] - dh_lo) * S + (q.taps.dw[t] - dw_lo);
    if (seen[g]) return false;
    seen[g] = true;
    f.tap_idx[g] = (unsigned char)t;
  }
  f.n_umma = q.Ko >= BN ? BN : (q.Ko + 15) / 16 * 16;
  const int b_bytes = f.n_umma * 128;
  const int room = HALO_SMEM_MAX - FLAT_FIXED_BYTES;
  int best_mt = 0;
  double best_cost = 0;
  for (int mt = 256 / BN; mt >= 1; mt >>= 1) {
    const int rows = mt * TILE_M + (R - 1) * Wp + (S - 1);
    const int nbig = rows / FLAT_BIG_ROWS;
    const int nbox = ceil_div(rows - nbig * FLAT_BIG_ROWS, FLAT_BOX_ROWS);
    const int a_bytes = nbig * FLAT_BIG_BYTES + nbox * FLAT_BOX_BYTES;
    // two activation stages; the rest goes to the weight ring (its boxes are small and the ring is latency-bound: 4
    // stages of 6 KB kept the 48-channel 5x5 layers at 23 % tensor pipe whatever the activation traffic was); what is
    // left after 16 weight stages buys a third activation stage
    int a_stages = 2;
    int b_stages = std::min(FLAT_MAX_B_STAGES, (room - a_stages * a_bytes) / b_bytes);
    if (b_stages < 4) continue;
    if (b_stages == FLAT_MAX_B_STAGES && room - a_stages * a_bytes - b_stages * b_bytes >= a_bytes) a_stages = 3;
    const int tiles_img = ceil_div(q.Po * Wp, mt * TILE_M);
    // measured (tools/flat_bench.py with 1, 2 and 4 sub-tiles per tile and 2, 3 and 4 activation stages; 48 and 96
    // channels; profiles/r02_flat_kernel_experiments.txt): a launch costs its MMA time plus ~4 us per CTA tile, whatever
    // the number of stages in flight, the box size or the number of TMA issues — so the fewest, largest tiles win
    const double cost = (double)tiles_img * (4.0 + 1.6 * mt);
    if (!best_mt || cost < best_cost) {
      best_mt = mt; best_cost = cost; f.nbig = nbig; f.nbox = nbox; f.a_stages = a_stages; f.b_stages = b_stages; f.tiles_img = tiles_img;
    }
  }
  if (!best_mt) return false;
  f.mt = best_mt;
  f.N = q.N; f.K = q.Ko; f.P = q.Po; f.Q = q.Qo; f.C = q.Ci;
  f.R = R; f.S = S; f.dh0 = dh_lo; f.dw0 = dw_lo; f.Hp = Hp; f.Wp = Wp;
  f.tiles_m = f.tiles_img * q.N; f.tiles_n = ceil_div(q.Ko, BN);
  f.OH = q.OH; f.OW = q.OW; f.o_step = q.o_step; f.o_off_h = q.o_off_h; f.o_off_w = q.o_off_w;
  if ((int64_t)q.N * Hp * Wp + 4096 >= ((int64_t)1 << 31)) return false;
  return true;
}

template <int BN>
static int launch_flat(const CUtensorMap& ta, const CUtensorMap& ta32, const CUtensorMap& tb, const TcFlatParams& p,
                       dim3 grid, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(tc_conv_flat_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       HALO_SMEM_MAX));
    configured = true;
  }
  const int smem = FLAT_FIXED_BYTES + p.b_stages * p.n_umma * 128 +
                   p.a_stages * (p.nbig * FLAT_BIG_BYTES + p.nbox * FLAT_BOX_BYTES);
  tc_conv_flat_kernel<BN><<<grid, FLAT_THREADS, smem, s>>>(ta, ta32, tb, p);
  GB_CHECK_LAUNCH();
  return 0;
}

static int pack_input(const float* in, const float* in_scale, __nv_bfloat16* act, int N, int Ci, int Hi, int Wi,
                      int mat_pad, cudaStream_t s) {
  const int Cp = cpad(Ci), Hp = Hi + 2 * mat_pad;
  static int pack_v2 = -1;
  if (pack_v2 < 0) { const char* e = getenv("GANB200_PACK_V2"); pack_v2 = (e && e[0] == '0') ? 0 : 1; }
  // measured at N = 256 (tools/pack_bench.py): the 4-rows-per-block kernel wins on 16 x 16 planes (2.9 -> 3.9 TB/s: the
  // one-row kernel's blocks are too small there), ties at 32 x 32 and loses at 64 x 64 (4.4 -> 4.1 TB/s)
  if (pack_v2 && Wi == 16 && ((uintptr_t)in & 15) == 0) {
    constexpr int ROWS = 4;
    dim3 grid(N * ((Hi + ROWS - 1) / ROWS), (Cp + 63) / 64);
    pack_act_vec2_kernel<2, ROWS><<<grid, 256, 0, s>>>(in, in_scale, act, Ci, Hi, Cp, mat_pad);
    GB_CHECK_LAUNCH();
    return 0;
  }
  if ((Wi & 3) == 0 && ((uintptr_t)in & 15) == 0 && Wi <= 1024) {
    dim3 grid(N * Hi, (Cp + 63) / 64);
    pack_act_vec_kernel<<<grid, 256, sizeof(uint32_t) * 33 * Wi, s>>>(in, in_scale, act, Ci, Hi, Wi, Cp, mat_pad);
    GB_CHECK_LAUNCH();
    return 0;
  }
  dim3 grid(N * Hp, (Cp + 63) / 64);
  pack_act_kernel<<<grid, 256, sizeof(float) * 64 * (Wi + 1), s>>>(in, in_scale, act, Ci, Hi, Wi, Cp, mat_pad);
  GB_CHECK_LAUNCH();
  return 0;
}

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

template <int BN>
static int launch_halo(const CUtensorMap& ta, const CUtensorMap& tb, const TcHaloParams& p, dim3 grid, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    GB_CHECK_CUDA(cudaFuncSetAttribute(tc_conv_halo_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       HALO_SMEM_MAX));
    configured = true;
  }
  const int smem = TcHaloSmem<BN>::FIXED + p.a_stages * p.a_bytes;
  tc_conv_halo_kernel<BN><<<grid, TC_THREADS, smem, s>>>(ta, tb, p);
  GB_CHECK_LAUNCH();
  return 0;
}

// `act`: the packed input; `wb`: scratch for this problem's packed weights
static int run_problem(const TcProblem& q, const __nv_bfloat16* act, const float* w, __nv_bfloat16* wb, float* out,
                       const float* out_scale, const float* bias, cudaStream_t s, bool w_packed = false) {
  const int Cp = cpad(q.Ci), Hp = q.Hi + 2 * q.mat_pad, Wp = q.Wi + 2 * q.mat_pad, NT = q.taps.n;
  if (!w_packed) {
    int64_t total = (int64_t)NT * q.Ko * Cp;
    pack_weight_kernel<<<ew_grid(total, 256, 1), 256, 0, s>>>(w, wb, q.Ko, q.Ci, Cp, q.w_sk, q.w_sc, q.taps);
    GB_CHECK_LAUNCH();
  }
  const int BN = q.Ko > 64 ? 128 : 64;   // 64 < Ko < 128 (96, 108): one 128-wide tile, the activations are read once
  {
    TcFlatParams f;
    // where it measured faster (tools/flat_bench.py, B = 256): 64-wide channel tiles on 64 x 64 planes (four sub-tiles per
    // CTA tile: 48 -> 48 5x5 1.10x, 3x3 1.14x, 48 -> 54 1.10x, ToRGB 108 -> 3 1.14x, the critic's 64 -> 64 1.02x) and
    // the 3x3 layers of up to 96 channels
    const bool flat_wins = q.Qo >= 64 && (q.Ko <= 64 || (q.taps.n == 9 && q.Ci <= 96 && q.Ko <= 96));
    if (halo_mode() >= 1 && (flat_mode() == 2 || (flat_mode() == 1 && flat_wins)) && flat_plan(q, BN, f)) {
      f.alpha = q.alpha; f.oscale = out_scale; f.bias = bias; f.y = out; f.trace = g_flat_trace;
      CUtensorMap ta, ta32, tb;
      uint64_t dims[3] = {(uint64_t)q.Ci, (uint64_t)q.N * Hp * Wp, 1};
      uint64_t str[2] = {(uint64_t)Cp * 2, (uint64_t)q.N * Hp * Wp * Cp * 2};
      uint32_t box[3] = {BK, FLAT_BIG_ROWS, 1}, box32[3] = {BK, FLAT_BOX_ROWS, 1};
      int rc = make_tmap_bf16(&ta, act, 3, dims, str, box);
      if (rc) return rc;
      rc = make_tmap_bf16(&ta32, act, 3, dims, str, box32);
      if (rc) return rc;
      uint64_t wd[3] = {(uint64_t)q.Ci, (uint64_t)q.Ko, (uint64_t)NT};
      uint64_t wst[2] = {(uint64_t)Cp * 2, (uint64_t)q.Ko * Cp * 2};
      uint32_t wbox[3] = {BK, (uint32_t)f.n_umma, 1};
      rc = make_tmap_bf16(&tb, wb, 3, wd, wst, wbox);
      if (rc) return rc;
      const int64_t total = (int64_t)f.tiles_m * f.tiles_n;
      dim3 grid((unsigned)std::min<int64_t>(total, kNumSMs));
      return BN == 128 ? launch_flat<128>(ta, ta32, tb, f, grid, s) : launch_flat<64>(ta, ta32, tb, f, grid, s);
    }
  }
  {
    TcHaloParams h;
    if (halo_mode() >= 1 && halo_plan(q, BN, h)) {
      h.n_umma = q.Ko >= BN ? BN : (q.Ko + 15) / 16 * 16;
      h.alpha = q.alpha; h.oscale = out_scale; h.bias = bias; h.y = out;
      h.fold = q.fold; h.fold_pad = q.fold_pad; h.FH = q.FH; h.FW = q.FW; h.red_x = q.red_x; h.red_out = q.red_out;
      h.range_mode = q.red_out ? 1 : 0;
      h.tiles_per_cta = 0;
      CUtensorMap ta, tb;
      uint64_t dims[4] = {(uint64_t)q.Ci, (uint64_t)Wp, (uint64_t)Hp, (uint64_t)q.N};
      uint64_t str[3] = {(uint64_t)Cp * 2, (uint64_t)Wp * Cp * 2, (uint64_t)Hp * Wp * Cp * 2};
      uint32_t box[4] = {BK, (uint32_t)h.bw, (uint32_t)(h.mt * h.bh + h.R - 1), 1};
      int rc = make_tmap_bf16(&ta, act, 4, dims, str, box);
      if (rc) return rc;
      uint64_t wd[3] = {(uint64_t)q.Ci, (uint64_t)q.Ko, (uint64_t)NT};
      uint64_t wst[2] = {(uint64_t)Cp * 2, (uint64_t)q.Ko * Cp * 2};
      // only the n_umma filter rows the MMA reads (K = 96: 12 KB instead of 16 KB per tap and chunk; the weights are
      // 60 % of this kernel's L2 -> shared-memory traffic on the k5 layers)
      uint32_t wbox[3] = {BK, (uint32_t)h.n_umma, 1};
      rc = make_tmap_bf16(&tb, wb, 3, wd, wst, wbox);
      if (rc) return rc;
      const int64_t total = (int64_t)h.tiles_m * h.tiles_n;
      dim3 grid((unsigned)std::min<int64_t>(total, kNumSMs));  // one CTA per SM (shared memory and TMEM are per-SM here)
      if (h.range_mode) {
        h.tiles_per_cta = ceil_div(total, grid.x);
        grid.x = (unsigned)ceil_div(total, h.tiles_per_cta);
      }
      return BN == 128 ? launch_halo<128>(ta, tb, h, grid, s) : launch_halo<64>(ta, tb, h, grid, s);
    }
  }
  TcConvParams p;
  p.N = q.N; p.K = q.Ko; p.P = q.Po; p.Q = q.Qo; p.taps = NT; p.C = q.Ci;
  int64_t best = -1;
  {
    // tile box (bw x bh x bn = 128 pixels, powers of two): the shape that covers [N][Po][Qo] with the fewest tiles
    // (padded dgrad domains such as 68x68 would waste half of a 128-wide row tile); ties go to the wider box.
    for (int bw = std::min(pow2ceil(q.Qo), TILE_M); bw >= 1; bw >>= 1) {
      const int bh = std::min(pow2ceil(q.Po), TILE_M / bw), bn = TILE_M / (bw * bh);
      if (q.red_out && (bn != 1 || (bw * bh) % 32 != 0)) continue;   // the warp reduction assumes one image per warp
      const int64_t tiles = (int64_t)ceil_div(q.Qo, bw) * ceil_div(q.Po, bh) * ceil_div(q.N, bn);
      if (best < 0 || tiles < best) { best = tiles; p.bw = bw; p.bh = bh; p.bn = bn; }
    }
  }
  p.tiles_q = ceil_div(q.Qo, p.bw); p.tiles_p = ceil_div(q.Po, p.bh);
  GB_REQUIRE(best >= 0, GANB200_BAD_SHAPE, "tc conv: no tile box for the fused reduction");
  p.alpha = q.alpha; p.oscale = out_scale; p.bias = bias; p.y = out;
  p.fold = q.fold; p.fold_pad = q.fold_pad; p.FH = q.FH; p.FW = q.FW; p.red_x = q.red_x; p.red_out = q.red_out;
  p.in_step = q.in_step; p.OH = q.OH; p.OW = q.OW; p.o_step = q.o_step; p.o_off_h = q.o_off_h; p.o_off_w = q.o_off_w;
  for (int t = 0; t < NT; ++t) { p.dh[t] = q.taps.dh[t]; p.dw[t] = q.taps.dw[t]; }
  p.n_umma = q.Ko >= BN ? BN : (q.Ko + 15) / 16 * 16;
  CUtensorMap ta, tb;
  {
    const uint32_t st = (uint32_t)q.in_step;
    uint64_t dims[4] = {(uint64_t)q.Ci, (uint64_t)Wp, (uint64_t)Hp, (uint64_t)q.N};
    uint64_t str[3] = {(uint64_t)Cp * 2, (uint64_t)Wp * Cp * 2, (uint64_t)Hp * Wp * Cp * 2};
    uint32_t box[4] = {BK, (uint32_t)p.bw * st, (uint32_t)p.bh * st, (uint32_t)p.bn};
    uint32_t es[4] = {1, st, st, 1};
    int rc = make_tmap_bf16(&ta, act, 4, dims, str, box, es);
    if (rc) return rc;
    uint64_t wd[3] = {(uint64_t)q.Ci, (uint64_t)q.Ko, (uint64_t)NT};
    uint64_t wst[2] = {(uint64_t)Cp * 2, (uint64_t)q.Ko * Cp * 2};
    uint32_t wbox[3] = {BK, (uint32_t)p.n_umma, 1};
    rc = make_tmap_bf16(&tb, wb, 3, wd, wst, wbox);
    if (rc) return rc;
  }
  p.tiles_m = p.tiles_q * p.tiles_p * ceil_div(q.N, p.bn);
  p.tiles_n = ceil_div(q.Ko, BN);
  const int64_t total = (int64_t)p.tiles_m * p.tiles_n;
  // persistent grid: 2 CTAs per SM (shared memory: 4 x 24 KB resp. 3 x 32 KB stages; TMEM: 2 x 2*BN <= 512 columns)
  dim3 grid((unsigned)std::min<int64_t>(total, 2 * kNumSMs));
  if (BN == 128) return launch_tc<128, 3>(ta, tb, p, grid, s);
  return launch_tc<64, 4>(ta, tb, p, grid, s);
}

// fold_replicate_kernel lives in conv_simt.cu


static bool shape_ok(const ganb200_conv_desc* d) {
  return (d->stride == 1 || d->stride == 2) && d->H <= 128 && d->W <= 128 && d->R <= 7 && d->S <= 7 &&
         d->R >= d->stride && d->S >= d->stride && d->pad <= 8;
}

int tc_eligible(const ganb200_conv_desc* d, int op) {
  if (!shape_ok(d)) return 0;
  const int64_t M = (int64_t)d->N * d->P * d->Q;
  if (op == GANB200_CONV_FPROP) return M >= 64;
  if (op == GANB200_CONV_DGRAD) return (int64_t)d->N * d->H * d->W >= 64;
  return M >= 256;
}

static size_t wgrad_workspace(const ganb200_conv_desc* d);

static TcProblem fprop_problem(const ganb200_conv_desc* d) {
  TcProblem q;
  q.N = d->N; q.Ci = d->C; q.Hi = d->H; q.Wi = d->W; q.Ko = d->K; q.Po = d->P; q.Qo = d->Q;
  q.OH = d->P; q.OW = d->Q; q.o_step = 1; q.o_off_h = q.o_off_w = 0;
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  q.mat_pad = rep ? d->pad : 0;
  q.in_step = d->stride;
  const int off = rep ? 0 : -d->pad;
  q.taps.n = d->R * d->S;
  for (int r = 0; r < d->R; ++r)
    for (int c = 0; c < d->S; ++c) {
      const int t = r * d->S + c;
      q.taps.src[t] = t; q.taps.dh[t] = (signed char)(r + off); q.taps.dw[t] = (signed char)(c + off);
    }
  q.w_sk = (int64_t)d->C * d->R * d->S; q.w_sc = (int64_t)d->R * d->S;
  q.alpha = d->alpha;
  return q;
}

// dgrad as stride*stride convolutions of dy, one per output phase (a,b) = (h % stride, w % stride) of the output
// domain (the padded input for replicate padding — folded afterwards — or the input itself for zero padding):
//   dx[st*i + a] = sum over taps r with (a + pz - r) % st == 0 of w[r] * dy[i + (a + pz - r)/st]
// pz = pad for zero padding (domain = input), 0 for replicate padding (domain = padded input).
static TcProblem dgrad_problem(const ganb200_conv_desc* d, int a, int b) {
  TcProblem q;
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  const int st = d->stride, pz = rep ? 0 : d->pad;
  const int DH = rep ? d->H + 2 * d->pad : d->H, DW = rep ? d->W + 2 * d->pad : d->W;
  q.N = d->N; q.Ci = d->K; q.Hi = d->P; q.Wi = d->Q; q.Ko = d->C;
  q.OH = DH; q.OW = DW; q.o_step = st; q.o_off_h = a; q.o_off_w = b;
  q.Po = (DH - a + st - 1) / st; q.Qo = (DW - b + st - 1) / st;
  q.mat_pad = 0; q.in_step = 1;
  q.taps.n = 0;
  for (int r = 0; r < d->R; ++r) {
    if (((a + pz - r) % st + st) % st != 0) continue;
    for (int c = 0; c < d->S; ++c) {
      if (((b + pz - c) % st + st) % st != 0) continue;
      const int t = q.taps.n++;
      q.taps.src[t] = r * d->S + c;
      q.taps.dh[t] = (signed char)((a + pz - r) / st);   // exact: the numerator is a multiple of st
      q.taps.dw[t] = (signed char)((b + pz - c) / st);
    }
  }
  q.w_sk = (int64_t)d->R * d->S; q.w_sc = (int64_t)d->C * d->R * d->S;
  q.alpha = d->alpha;
  return q;
}

// packed weights of all stride*stride phases, each region rounded up to 1024 bytes
static size_t dgrad_w_bytes(const ganb200_conv_desc* d) {
  return w_bytes(d->R * d->S, d->C, d->K) + (size_t)1024 * d->stride * d->stride;
}

static size_t dgrad_fold_bytes(const ganb200_conv_desc* d) {
  if (d->pad_mode == 1 && d->pad > 0)
    return align_up(sizeof(float) * (size_t)d->N * d->C * (d->H + 2 * d->pad) * (d->W + 2 * d->pad), 1024);
  return 0;
}

size_t tc_workspace(const ganb200_conv_desc* d, int op) {
  if (op == GANB200_CONV_FPROP) {
    const bool rep = d->pad_mode == 1 && d->pad > 0;
    return act_bytes(d->N, d->C, d->H, d->W, rep ? d->pad : 0) + w_bytes(d->R * d->S, d->K, d->C) + 1024;
  }
  if (op == GANB200_CONV_DGRAD)
    return act_bytes(d->N, d->K, d->P, d->Q, 0) + dgrad_w_bytes(d) + dgrad_fold_bytes(d) + 1024;
  return wgrad_workspace(d);
}

int tc_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias, const float* iscale,
             const float* oscale, float* y, void* ws, size_t ws_bytes, cudaStream_t s, const void* xp, const void* wp) {
  GB_REQUIRE((xp && wp) || (ws && ws_bytes >= tc_workspace(d, GANB200_CONV_FPROP)), GANB200_WORKSPACE_TOO_SMALL,
             "tc fprop: workspace too small");
  ws = align_ws(ws);
  TcProblem q = fprop_problem(d);
  __nv_bfloat16* act = xp ? (__nv_bfloat16*)xp : (__nv_bfloat16*)ws;
  __nv_bfloat16* wb = wp ? (__nv_bfloat16*)wp
                         : (__nv_bfloat16*)((uint8_t*)ws + act_bytes(q.N, q.Ci, q.Hi, q.Wi, q.mat_pad));
  if (!xp) {
    int rc = pack_input(x, iscale, act, q.N, q.Ci, q.Hi, q.Wi, q.mat_pad, s);
    if (rc) return rc;
  }
  return run_problem(q, act, w, wb, y, oscale, bias, s, wp != nullptr);
}

// ---- packed operands as first-class buffers (cached by the caller across the passes that share them) ----------------
size_t tc_packed_act_bytes(int N, int C, int H, int W, int pad) { return act_bytes(N, C, H, W, pad) + 1024; }

int tc_pack_act(const float* x, const float* scale, void* out, int N, int C, int H, int W, int pad, cudaStream_t s) {
  GB_REQUIRE(((uintptr_t)out & 1023) == 0, GANB200_BAD_ARG, "pack_act: output must be 1024-byte aligned");
  return pack_input(x, scale, (__nv_bfloat16*)out, N, C, H, W, pad, s);
}

size_t tc_packed_weight_bytes(const ganb200_conv_desc* d, int op) {
  if (op == GANB200_CONV_FPROP) return w_bytes(d->R * d->S, d->K, d->C) + 1024;
  if (op == GANB200_CONV_DGRAD) return dgrad_w_bytes(d) + 1024;
  return 0;
}

int tc_pack_weight(const ganb200_conv_desc* d, int op, const float* w, void* out, cudaStream_t s) {
  GB_REQUIRE(((uintptr_t)out & 1023) == 0, GANB200_BAD_ARG, "pack_weight: output must be 1024-byte aligned");
  auto pack = [&](const TcProblem& q, __nv_bfloat16* dst) {
    const int Cp = cpad(q.Ci);
    int64_t total = (int64_t)q.taps.n * q.Ko * Cp;
    pack_weight_kernel<<<ew_grid(total, 256, 1), 256, 0, s>>>(w, dst, q.Ko, q.Ci, Cp, q.w_sk, q.w_sc, q.taps);
  };
  if (op == GANB200_CONV_FPROP) {
    pack(fprop_problem(d), (__nv_bfloat16*)out);
  } else if (op == GANB200_CONV_DGRAD) {
    size_t woff = 0;   // the same phase order and offsets as dgrad_run
    for (int a = 0; a < d->stride; ++a)
      for (int b = 0; b < d->stride; ++b) {
        TcProblem q = dgrad_problem(d, a, b);
        if (q.Po <= 0 || q.Qo <= 0) continue;
        GB_REQUIRE(q.taps.n > 0, GANB200_BAD_SHAPE, "tc dgrad: an output phase has no filter tap");
        pack(q, (__nv_bfloat16*)((uint8_t*)out + woff));
        woff += w_bytes(q.taps.n, d->C, d->K);
      }
  } else {
    return fail(GANB200_BAD_ARG, "pack_weight: wgrad has no weight operand%s", "");
  }
  GB_CHECK_LAUNCH();
  return 0;
}

static int dgrad_run(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale,
                     const float* oscale, float* dx, void* ws, cudaStream_t s, bool fold, float** padded_out,
                     const void* dyp = nullptr, const void* wp = nullptr, const float* red_x = nullptr,
                     float* red_out = nullptr);

int tc_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale, const float* oscale,
             float* dx, void* ws, size_t ws_bytes, cudaStream_t s, const void* dyp, const void* wp) {
  GB_REQUIRE(ws && ws_bytes >= tc_workspace(d, GANB200_CONV_DGRAD), GANB200_WORKSPACE_TOO_SMALL,
             "tc dgrad: workspace too small");
  return dgrad_run(d, dy, w, iscale, oscale, dx, align_ws(ws), s, true, nullptr, dyp, wp);
}

// 0 (default): dgrad of replicate-padded convolutions writes the padded fp32 domain to the workspace and a fold pass adds
// the border back — every sum in a fixed order.  1: scatter straight into the unpadded plane from the MMA epilogue
// (epi_fold32; no intermediate, ~0.2 % of the step at batch 256) — the border pixels then take red.global.add from
// several CTAs, so their low bits depend on the schedule.  GANB200_DGRAD_SCATTER=1 or ganb200_conv_scatter_mode(1).
static int g_scatter_mode = -1;
static int scatter_mode() {
  if (g_scatter_mode < 0) {
    const char* e = getenv("GANB200_DGRAD_SCATTER");
    g_scatter_mode = (e && e[0] == '1') ? 1 : 0;
  }
  return g_scatter_mode;
}

// zero the border pixels of every H x W plane (the only pixels the scatter epilogue accumulates into)
__global__ void __launch_bounds__(256) zero_border_kernel(float* __restrict__ x, int64_t planes, int H, int W) {
  const int per = (H <= 2 || W <= 2) ? H * W : 2 * W + 2 * (H - 2);
  grid_stride(planes * per, [&](auto i) {
    const int e = (int)(i % per);
    const int64_t pl = (int64_t)(i / per);
    int h, w;
    if (H <= 2 || W <= 2) { h = e / W; w = e - h * W; }
    else if (e < W) { h = 0; w = e; }
    else if (e < 2 * W) { h = H - 1; w = e - W; }
    else { const int r = e - 2 * W; h = 1 + (r >> 1); w = (r & 1) ? W - 1 : 0; }
    x[(pl * H + h) * W + w] = 0.f;
  });
}

// ws (1024-aligned): [dy packed][phase weights][padded fp32 output (replicate padding only)].  With fold == false the
// padded-domain result is left in the workspace (*padded_out) for a fused consumer.
static int dgrad_run(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale,
                     const float* oscale, float* dx, void* ws, cudaStream_t s, bool fold, float** padded_out,
                     const void* dyp, const void* wp, const float* red_x, float* red_out) {
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  const __nv_bfloat16* act = dyp ? (const __nv_bfloat16*)dyp : (const __nv_bfloat16*)ws;
  uint8_t* ws_w = (uint8_t*)ws + act_bytes(d->N, d->K, d->P, d->Q, 0);
  uint8_t* wbase = wp ? (uint8_t*)wp : ws_w;
  float* out = dx;
  // scatter epilogue: straight into dx (callers that pass fold == true, or a reduction target)
  const bool scatter = (fold || red_out) && dx && (rep || red_out) && scatter_mode() == 1;
  if (rep && !scatter) out = (float*)(ws_w + dgrad_w_bytes(d));
  if (padded_out) *padded_out = scatter ? nullptr : out;
  if (scatter) {
    if (rep) {
      const int64_t planes = (int64_t)d->N * d->C;
      const int per = (d->H <= 2 || d->W <= 2) ? d->H * d->W : 2 * d->W + 2 * (d->H - 2);
      zero_border_kernel<<<ew_grid(planes * per, 256, 1), 256, 0, s>>>(dx, planes, d->H, d->W);
      GB_CHECK_LAUNCH();
    }
    if (red_out) GB_CHECK_CUDA(cudaMemsetAsync(red_out, 0, sizeof(float) * (size_t)d->N * d->C, s));
  }
  int rc = 0;
  if (!dyp) {
    rc = pack_input(dy, oscale, (__nv_bfloat16*)ws, d->N, d->K, d->P, d->Q, 0, s);
    if (rc) return rc;
  }
  size_t woff = 0;
  for (int a = 0; a < d->stride; ++a)
    for (int b = 0; b < d->stride; ++b) {
      TcProblem q = dgrad_problem(d, a, b);
      if (q.Po <= 0 || q.Qo <= 0) continue;
      GB_REQUIRE(q.taps.n > 0, GANB200_BAD_SHAPE, "tc dgrad: an output phase has no filter tap");
      if (scatter) {
        q.fold = 1; q.fold_pad = rep ? d->pad : 0; q.FH = d->H; q.FW = d->W; q.red_x = red_x; q.red_out = red_out;
      }
      rc = run_problem(q, act, w, (__nv_bfloat16*)(wbase + woff), out, iscale, nullptr, s, wp != nullptr);
      if (rc) return rc;
      woff += w_bytes(q.taps.n, d->C, d->K);
    }
  if (rep && fold && !scatter) return fold_replicate(out, dx, (int64_t)d->N * d->C, d->H, d->W, d->pad, s);
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
  int off_h, off_w, in_step;
  int n_umma;                // multiple of 16 <= BN
  int m;                     // rows of the MMA: 128, or 64 when K <= 64 (see wgrad_m)
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
    const bool leader = elect_one();   // warp-uniform loop, the elected lane issues (see tc_ptx.cuh)
    for (int it = 0; it < iters; ++it) {
      const int stage = it % STAGES;
      const uint32_t phase = (it / STAGES) & 1;
      mbar_wait(&empty[stage], phase ^ 1);
      int t = ch_beg + it;
      const int tq = t % p.tiles_q; t /= p.tiles_q;
      const int tp = t % p.tiles_p; t /= p.tiles_p;
      const int q0 = tq * p.bw, p0 = tp * p.bh, n0 = t * p.bn;
      if (leader) {
        mbar_expect_tx(&full[stage], (p.m == 64 ? A_BYTES / 2 : A_BYTES) + B_BYTES);
        uint8_t* a = smA + stage * A_BYTES;
        uint8_t* b = smB + stage * B_BYTES;
        tma_load_4d(a, &tmA, &full[stage], k0, q0, p0, n0);
        if (p.m != 64) tma_load_4d(a + WG_GROUP_BYTES, &tmA, &full[stage], k0 + 64, q0, p0, n0);
#pragma unroll
        for (int g = 0; g < BN / 64; ++g)
          tma_load_4d(b + g * WG_GROUP_BYTES, &tmB, &full[stage], c0 + g * 64, q0 * p.in_step + s + p.off_w,
                      p0 * p.in_step + r + p.off_h, n0);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(p.m, p.n_umma, 1, 1);
    for (int it = 0; it < iters; ++it) {
      const int stage = it % STAGES;
      const uint32_t phase = (it / STAGES) & 1;
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      // MN-major SWIZZLE_128B: LBO = next 64-channel group, SBO = next 8 pixels
      const uint64_t adesc = make_smem_desc_sw128(smem_u32(smA + stage * A_BYTES), WG_GROUP_BYTES, 1024);
      const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB + stage * B_BYTES), WG_GROUP_BYTES, 1024);
      if (leader) {
        // 16 pixels = 2048 bytes = 128 in the >>4 address field
        mma_chain<WG_PIX / 16>(tmem_base, adesc, bdesc, idesc, it > 0 ? 1u : 0u, 128u);
        mma_commit(&empty[stage]);
      }
    }
    if (leader) mma_commit(tmem_full);
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    // M = 64: accumulator row r sits in TMEM lane 32 * (r / 16) + r % 16 (csrc/probe.cu, profiles/r02_tc_m64_lane_probe.txt)
    const int k = k0 + (p.m == 64 ? quarter * 16 + lane : quarter * 32 + lane);
    const bool k_ok = k < p.K && (p.m != 64 || lane < 16);
    float* row = p.part + (((int64_t)split * p.taps + tap) * p.K + k) * p.C + c0;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    for (int j0 = 0; j0 < p.n_umma; j0 += 32) {
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)j0, v);
      if (k_ok) {
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

// ------------------------------------------------------------------------------------------------ wgrad, row ring
// tc_wgrad_kernel gives every filter tap its own CTAs, so dy and x cross L2 -> shared memory R*S times each and the
// narrow 64x64 layers (K = C = 48: an N = 48 MMA per 16 pixels) run at 80-100 TFLOP/s, all fill.  Here one CTA owns a
// filter COLUMN s and a range of output rows: per output row (n, p) it loads the dy row once and ONE new x row; the R
// rows p .. p+R-1 of the padded input sit in a ring of shared-memory slots (each 1024-aligned, so every slot is a valid
// MN-major SWIZZLE_128B operand) and feed R MMAs chains into R accumulators (R * n_umma <= 512 TMEM columns).
// Fill per output row drops from R * (dy + x) to (dy + x).  Rows are global indices g = n * P + p; a CTA's range may
// span images: at an image boundary the R-1 rows left in the ring are released and the ring is primed again.
constexpr int WGR_MAX_D = 4;    // dy stages
constexpr int WGR_MAX_X = 12;   // x row slots

struct TcWgRingParams {
  int K, C, R, S, P, Q;
  int rows_total, rows_per_split;
  int off_h, off_w;
  int a_groups, b_groups;   // 64-channel groups of the dy tile (128 filters: 2) and of the x tile (BN / 64)
  int BN, n_umma, acc_pitch;
  int d_stages, x_slots;
  int row_bytes;            // Q * 128: one image row of one 64-channel group
  int m;                    // rows of the MMA: 128, or 64 when K <= 64 (see wgrad_m)
  uint32_t tmem_cols;
  float* part;              // [splits][R*S][K][C]
};

__global__ void __launch_bounds__(TC_THREADS, 1) tc_wgrad_ring_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                      const __grid_constant__ CUtensorMap tmB,
                                                                      const TcWgRingParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int d_bytes = p.a_groups * p.row_bytes, x_bytes = p.b_groups * p.row_bytes;
  uint8_t* smD = smem;
  uint8_t* smX = smem + p.d_stages * d_bytes;
  uint64_t* full_d = reinterpret_cast<uint64_t*>(smX + p.x_slots * x_bytes);
  uint64_t* empty_d = full_d + WGR_MAX_D;
  uint64_t* full_x = empty_d + WGR_MAX_D;
  uint64_t* empty_x = full_x + WGR_MAX_X;
  uint64_t* tmem_full = empty_x + WGR_MAX_X;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * TILE_M, c0 = blockIdx.y * p.BN;
  const int s = blockIdx.z % p.S, split = blockIdx.z / p.S;
  const int g_beg = min(split * p.rows_per_split, p.rows_total);
  const int g_end = min(g_beg + p.rows_per_split, p.rows_total);
  const int NS = p.x_slots, ND = p.d_stages, R = p.R;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < ND; ++i) { mbar_init(&full_d[i], 1); mbar_init(&empty_d[i], 1); }
    for (int i = 0; i < NS; ++i) { mbar_init(&full_x[i], 1); mbar_init(&empty_x[i], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    const bool leader = elect_one();
    int sd = 0, sx = 0;
    uint32_t pd = 0, px = 0;
    for (int g = g_beg; g < g_end;) {
      const int n = g / p.P;
      int pr = g - n * p.P;
      const int seg_end = min(g_end, (n + 1) * p.P);
      // x rows of this segment in load order: the R-1 rows that prime the ring, then one per output row
      const int first_row = pr + p.off_h;
      const int n_x = (seg_end - g) + R - 1;
      int xi = 0;
      for (; g < seg_end; ++g, ++pr) {
        const int want = (pr + p.off_h + R - 1) - first_row + 1;   // x rows needed up to this output row
        for (; xi < want && xi < n_x; ++xi) {
          mbar_wait(&empty_x[sx], px ^ 1);
          if (leader) {
            mbar_expect_tx(&full_x[sx], (uint32_t)x_bytes);
            for (int gi = 0; gi < p.b_groups; ++gi)
              tma_load_4d(smX + sx * x_bytes + gi * p.row_bytes, &tmB, &full_x[sx], c0 + gi * 64, s + p.off_w,
                          first_row + xi, n);
          }
          if (++sx == NS) { sx = 0; px ^= 1; }
        }
        mbar_wait(&empty_d[sd], pd ^ 1);
        if (leader) {
          mbar_expect_tx(&full_d[sd], (uint32_t)d_bytes);
          for (int gi = 0; gi < p.a_groups; ++gi)
            tma_load_4d(smD + sd * d_bytes + gi * p.row_bytes, &tmA, &full_d[sd], k0 + gi * 64, 0, pr, n);
        }
        if (++sd == ND) { sd = 0; pd ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(p.m, p.n_umma, 1, 1);
    int sd = 0, sx = 0;          // sx / px: slot and fill parity of the OLDEST row in the ring window
    uint32_t pd = 0, px = 0;
    bool fresh = true;
    for (int g = g_beg; g < g_end;) {
      const int n = g / p.P;
      const int seg_end = min(g_end, (n + 1) * p.P);
      bool primed = false;
      for (; g < seg_end; ++g) {
        mbar_wait(&full_d[sd], pd);
        {
          // the first output row of a segment needs all R ring rows, every later one only the newest
          int sl = sx;
          uint32_t par = px;
          for (int r = 0; r < R; ++r) {
            if (!primed || r == R - 1) mbar_wait(&full_x[sl], par);
            if (++sl == NS) { sl = 0; par ^= 1; }
          }
          primed = true;
        }
        tc_fence_after();
        // MN-major SWIZZLE_128B: LBO = next 64-channel group, SBO = next 8 pixels
        const uint64_t adesc = make_smem_desc_sw128(smem_u32(smD + sd * d_bytes), (uint32_t)p.row_bytes, 1024);
        if (leader) {
          int sl = sx;
          const uint32_t acc = fresh ? 0u : 1u;
          for (int r = 0; r < R; ++r) {
            const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smX + sl * x_bytes), (uint32_t)p.row_bytes, 1024);
            // 16 pixels = 2048 bytes = 128 in the >>4 address field; rows are 64 pixels = 4 k steps
            mma_chain<4>(tmem_base + (uint32_t)(r * p.acc_pitch), adesc, bdesc, idesc, acc, 128u);
            if (++sl == NS) sl = 0;
          }
          mma_commit(&empty_d[sd]);
          mma_commit(&empty_x[sx]);   // the oldest row is not needed by any later output row
        }
        fresh = false;
        if (++sd == ND) { sd = 0; pd ^= 1; }
        if (++sx == NS) { sx = 0; px ^= 1; }
      }
      // image (or range) boundary: release the R-1 rows still in the ring
      for (int r = 0; r + 1 < R; ++r) {
        if (leader) mma_commit(&empty_x[sx]);
        if (++sx == NS) { sx = 0; px ^= 1; }
      }
    }
    if (leader) mma_commit(tmem_full);
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int k = k0 + (p.m == 64 ? quarter * 16 + lane : quarter * 32 + lane);   // M = 64: 16 lanes per quarter
    const bool k_ok = k < p.K && (p.m != 64 || lane < 16);
    const bool any = g_end > g_beg;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    for (int r = 0; r < R; ++r) {
      float* row = p.part + (((int64_t)split * (R * p.S) + (r * p.S + s)) * p.K + k) * p.C + c0;
      for (int j0 = 0; j0 < p.n_umma; j0 += 32) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(r * p.acc_pitch + j0), v);
        if (k_ok) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j0 + j < p.n_umma && c0 + j0 + j < p.C) row[j0 + j] = any ? v[j] : 0.f;
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// dw[k][c][rs] = alpha * sum_split part[split][rs][k][c]
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ dw, int K,
                                                           int C, int RS, int splits, float alpha) {
  // one thread per (rs, k*C + c): reads coalesced along (k,c); the split sum runs in a fixed order (deterministic)
  const uint32_t KC = (uint32_t)K * C, total = KC * RS;
  for (uint32_t j = blockIdx.x * 256 + threadIdx.x; j < total; j += gridDim.x * 256) {
    const uint32_t rs = j / KC, i = j - rs * KC;
    const float* src = part + (size_t)rs * KC + i;
    const size_t step = (size_t)RS * KC;
    float acc = 0.f;
    int sp = 0;
    for (; sp + 4 <= splits; sp += 4) {
      const float a0 = __ldg(src + (size_t)sp * step), a1 = __ldg(src + (size_t)(sp + 1) * step);
      const float a2 = __ldg(src + (size_t)(sp + 2) * step), a3 = __ldg(src + (size_t)(sp + 3) * step);
      acc += a0; acc += a1; acc += a2; acc += a3;
    }
    for (; sp < splits; ++sp) acc += __ldg(src + (size_t)sp * step);
    dw[(size_t)i * RS + rs] = acc * alpha;
  }
}

// Rows of the wgrad MMAs.  The tensor core fetches its shared-memory operands at ~70 bytes per clock and SM (flat-kernel
// trace, profiles/r02_flat_kernel_trace.txt: a 128 x 48 x 16 MMA takes 77 cycles, a 128 x 96 x 16 one 107), so a narrow MMA
// costs what its operands weigh: with K <= 64 filters an M = 128 instruction reads 4 KB of dy of which half is padding;
// M = 64 reads 2 KB.  GANB200_WGRAD_M64=0: always 128.
static int wgrad_m(const ganb200_conv_desc* d) {
  static int on = -1;
  if (on < 0) { const char* e = getenv("GANB200_WGRAD_M64"); on = (e && e[0] == '0') ? 0 : 1; }
  return (on && d->K <= 64) ? 64 : TILE_M;
}

struct WgradPlan {
  int bw, bh, bn, tiles_q, tiles_p, tiles_n, chunks, splits, per, BN;
  size_t dy_bytes, x_bytes, part_bytes;
  // row-ring kernel (ring.ok): stride 1, rows of 32 or 64 pixels, more than one filter row
  struct { bool ok; int BN, n_umma, acc_pitch, a_groups, b_groups, d_stages, x_slots, splits, rows_per_split, smem;
           uint32_t tmem_cols; } ring;
};

static void wgrad_ring_plan(const ganb200_conv_desc* d, WgradPlan& w) {
  w.ring.ok = false;
  if (d->stride != 1 || d->Q != 64 || d->R < 2 || d->R > 7 || d->S > 7) return;
  const int n16 = (d->C + 15) / 16 * 16;
  const int BN = d->C > 64 ? 128 : 64;
  const int n_umma = std::min(n16, BN);
  const int pitch = (n_umma + 31) / 32 * 32;
  if (d->R * pitch > 512) return;
  w.ring.BN = BN; w.ring.n_umma = n_umma; w.ring.acc_pitch = pitch;
  w.ring.tmem_cols = (uint32_t)std::max(32, pow2ceil(d->R * pitch));
  w.ring.a_groups = d->K > 64 ? 2 : 1;
  w.ring.b_groups = BN / 64;
  const int row_bytes = d->Q * 128;
  int ds = WGR_MAX_D, xs = std::min(WGR_MAX_X, d->R + 4);
  auto bytes = [&](int a, int b) { return a * w.ring.a_groups * row_bytes + b * w.ring.b_groups * row_bytes + 1024 + 512; };
  while (bytes(ds, xs) > HALO_SMEM_MAX && xs > d->R + 1) --xs;
  while (bytes(ds, xs) > HALO_SMEM_MAX && ds > 2) --ds;
  if (bytes(ds, xs) > HALO_SMEM_MAX) return;
  w.ring.d_stages = ds; w.ring.x_slots = xs; w.ring.smem = bytes(ds, xs);
  const int units = ceil_div(d->K, TILE_M) * ceil_div(d->C, BN) * d->S;
  const int rows = d->N * d->P;
  // whole waves: the TMEM allocation (and usually shared memory) admits one CTA per SM, two when both halve
  const int per_sm = (w.ring.tmem_cols <= 256 && 2 * (w.ring.smem + 1024) <= HALO_SMEM_MAX) ? 2 : 1;
  int splits = std::max(1, per_sm * kNumSMs / units);
  splits = std::min(splits, std::max(1, rows / 16));     // at least 16 rows behind every ring priming
  w.ring.rows_per_split = ceil_div(rows, splits);
  w.ring.splits = ceil_div(rows, w.ring.rows_per_split);
  w.ring.ok = true;
}

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
  wgrad_ring_plan(d, w);
  const int max_splits = w.ring.ok ? std::max(w.splits, w.ring.splits) : w.splits;   // either kernel may run
  w.part_bytes = align_up((size_t)max_splits * d->R * d->S * d->K * d->C * 4, 1024);
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

static int wgrad_run(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale,
                     const float* oscale, float* dw, void* ws, cudaStream_t s, const __nv_bfloat16* dy_prepacked,
                     const __nv_bfloat16* x_prepacked = nullptr);

int tc_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale, const float* oscale,
             float* dw, void* ws, size_t ws_bytes, cudaStream_t s, const void* xp, const void* dyp) {
  GB_REQUIRE(ws && ws_bytes >= wgrad_workspace(d), GANB200_WORKSPACE_TOO_SMALL, "tc wgrad: workspace too small");
  return wgrad_run(d, x, dy, iscale, oscale, dw, align_ws(ws), s, (const __nv_bfloat16*)dyp, (const __nv_bfloat16*)xp);
}

// ws (1024-aligned): [dy packed (unused when dy_prepacked)][x packed][partials]
static int wgrad_run(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale,
                     const float* oscale, float* dw, void* ws, cudaStream_t s, const __nv_bfloat16* dy_prepacked,
                     const __nv_bfloat16* x_prepacked) {
  const WgradPlan w = wgrad_plan(d);
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  const int mp = rep ? d->pad : 0;
  const int Kp = cpad(d->K), Cp = cpad(d->C), Hp = d->H + 2 * mp, Wp = d->W + 2 * mp, RS = d->R * d->S;
  const __nv_bfloat16* dyh = dy_prepacked ? dy_prepacked : (const __nv_bfloat16*)ws;
  const __nv_bfloat16* xh = x_prepacked ? x_prepacked : (const __nv_bfloat16*)((uint8_t*)ws + w.dy_bytes);
  float* part = (float*)((uint8_t*)ws + w.dy_bytes + w.x_bytes);
  {
    if (!dy_prepacked) {
      if (int rc = pack_input(dy, oscale, (__nv_bfloat16*)ws, d->N, d->K, d->P, d->Q, 0, s)) return rc;
    }
    if (!x_prepacked) {
      if (int rc = pack_input(x, iscale, (__nv_bfloat16*)((uint8_t*)ws + w.dy_bytes), d->N, d->C, d->H, d->W, mp, s))
        return rc;
    }
  }
  if (w.ring.ok && halo_mode() >= 1) {
    TcWgRingParams r;
    r.K = d->K; r.C = d->C; r.R = d->R; r.S = d->S; r.P = d->P; r.Q = d->Q;
    r.rows_total = d->N * d->P; r.rows_per_split = w.ring.rows_per_split;
    r.off_h = r.off_w = rep ? 0 : -d->pad;
    r.a_groups = w.ring.a_groups; r.b_groups = w.ring.b_groups;
    r.BN = w.ring.BN; r.n_umma = w.ring.n_umma; r.acc_pitch = w.ring.acc_pitch;
    r.d_stages = w.ring.d_stages; r.x_slots = w.ring.x_slots;
    r.row_bytes = d->Q * 128; r.tmem_cols = w.ring.tmem_cols; r.part = part; r.m = wgrad_m(d);
    CUtensorMap ta, tb;
    uint64_t dims[4] = {(uint64_t)d->K, (uint64_t)d->Q, (uint64_t)d->P, (uint64_t)d->N};
    uint64_t str[3] = {(uint64_t)Kp * 2, (uint64_t)d->Q * Kp * 2, (uint64_t)d->P * d->Q * Kp * 2};
    uint32_t box[4] = {64, (uint32_t)d->Q, 1, 1};
    int rc = make_tmap_bf16(&ta, dyh, 4, dims, str, box);
    if (rc) return rc;
    uint64_t xd[4] = {(uint64_t)d->C, (uint64_t)Wp, (uint64_t)Hp, (uint64_t)d->N};
    uint64_t xs[3] = {(uint64_t)Cp * 2, (uint64_t)Wp * Cp * 2, (uint64_t)Hp * Wp * Cp * 2};
    rc = make_tmap_bf16(&tb, xh, 4, xd, xs, box);
    if (rc) return rc;
    static bool configured = false;
    if (!configured) {
      GB_CHECK_CUDA(cudaFuncSetAttribute(tc_wgrad_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         HALO_SMEM_MAX));
      configured = true;
    }
    dim3 grid(ceil_div(d->K, TILE_M), ceil_div(d->C, w.ring.BN), d->S * w.ring.splits);
    tc_wgrad_ring_kernel<<<grid, TC_THREADS, w.ring.smem, s>>>(ta, tb, r);
    GB_CHECK_LAUNCH();
    wgrad_reduce_kernel<<<ew_grid((int64_t)d->K * d->C * RS, 256, 1), 256, 0, s>>>(part, dw, d->K, d->C, RS,
                                                                                 w.ring.splits, d->alpha);
    GB_CHECK_LAUNCH();
    return 0;
  }
  TcWgradParams p;
  p.K = d->K; p.C = d->C; p.S = d->S; p.taps = RS;
  p.bw = w.bw; p.bh = w.bh; p.bn = w.bn; p.tiles_q = w.tiles_q; p.tiles_p = w.tiles_p; p.tiles_n = w.tiles_n;
  p.chunks = w.chunks; p.chunks_per_split = w.per; p.splits = w.splits;
  p.off_h = p.off_w = rep ? 0 : -d->pad;
  p.in_step = d->stride;
  p.n_umma = d->C >= w.BN ? w.BN : (d->C + 15) / 16 * 16;
  p.part = part; p.m = wgrad_m(d);
  CUtensorMap ta, tb;
  {
    uint64_t dims[4] = {(uint64_t)d->K, (uint64_t)d->Q, (uint64_t)d->P, (uint64_t)d->N};
    uint64_t str[3] = {(uint64_t)Kp * 2, (uint64_t)d->Q * Kp * 2, (uint64_t)d->P * d->Q * Kp * 2};
    uint32_t box[4] = {64, (uint32_t)w.bw, (uint32_t)w.bh, (uint32_t)w.bn};
    int rc = make_tmap_bf16(&ta, dyh, 4, dims, str, box);
    if (rc) return rc;
    uint64_t xd[4] = {(uint64_t)d->C, (uint64_t)Wp, (uint64_t)Hp, (uint64_t)d->N};
    uint64_t xs[3] = {(uint64_t)Cp * 2, (uint64_t)Wp * Cp * 2, (uint64_t)Hp * Wp * Cp * 2};
    const uint32_t st = (uint32_t)d->stride;
    uint32_t xbox[4] = {64, (uint32_t)w.bw * st, (uint32_t)w.bh * st, (uint32_t)w.bn};
    uint32_t es[4] = {1, st, st, 1};
    rc = make_tmap_bf16(&tb, xh, 4, xd, xs, xbox, es);
    if (rc) return rc;
  }
  dim3 grid(ceil_div(d->K, TILE_M), ceil_div(d->C, w.BN), RS * w.splits);
  int rc = w.BN == 128 ? launch_wg<128, 4>(ta, tb, p, grid, s) : launch_wg<64, 4>(ta, tb, p, grid, s);
  if (rc) return rc;
  wgrad_reduce_kernel<<<ew_grid((int64_t)d->K * d->C * RS, 256, 1), 256, 0, s>>>(part, dw, d->K, d->C, RS, w.splits, d->alpha);
  GB_CHECK_LAUNCH();
  return 0;
}

// ================================================================================================ fused modconv backward
// plane (n,c): v(h,w) = sum of dxp over the padded positions that clamp to (h,w) (adjoint of replicate padding);
//              gx = v * s[n,c];  gs1[n,c] = sum_hw v * x
template <bool WARP_ROWS>
__global__ void __launch_bounds__(256) fold_scale_reduce_kernel(const float* __restrict__ dxp, const float* __restrict__ s,
                                                                const float* __restrict__ x, float* __restrict__ gx,
                                                                float* __restrict__ gs1, int H, int W, int pad,
                                                                int64_t rows) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int Hp = H + 2 * pad, Wp = W + 2 * pad, HW = H * W;
  const int64_t row_step = WARP_ROWS ? (int64_t)gridDim.x * 8 : gridDim.x;
  for (int64_t row = WARP_ROWS ? (int64_t)blockIdx.x * 8 + warp : blockIdx.x; row < rows; row += row_step) {
    const float sv = s[row];
    const float* src = dxp + row * Hp * Wp;
    float acc = 0.f;
    const int t0 = WARP_ROWS ? lane : threadIdx.x, tstep = WARP_ROWS ? 32 : 256;
    for (int i = t0; i < HW; i += tstep) {
      const int h = i / W, w = i - h * W;
      int h_lo = h == 0 ? 0 : h + pad, h_hi = h == H - 1 ? Hp - 1 : h + pad;
      int w_lo = w == 0 ? 0 : w + pad, w_hi = w == W - 1 ? Wp - 1 : w + pad;
      if (H == 1) { h_lo = 0; h_hi = Hp - 1; }
      if (W == 1) { w_lo = 0; w_hi = Wp - 1; }
      float v = 0.f;
      for (int a = h_lo; a <= h_hi; ++a)
        for (int b = w_lo; b <= w_hi; ++b) v += src[a * Wp + b];
      if (gs1) acc = fmaf(v, x[row * HW + i], acc);
      if (gx) gx[row * HW + i] = v * sv;
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

// The same for planes of >= 1024 pixels with W % 4 == 0 (the generator's 32 x 32 and 64 x 64 layers — where the time is):
// ncu on the scalar kernel above, 96 channels at 64 x 64, N = 256: 592 us for 1.24 GB = 2.1 TB/s — one dependent 4-byte
// load per thread and iteration cannot cover the HBM latency.  Here a thread owns UNROLL float4 quads per trip: all centre
// loads of the padded plane (the only ones an interior quad needs) and the float4 loads of x are issued first; quads on
// the border ring of the plane then add what the replicate padding folds onto them.  One block per (n, c) plane.
// WARP_PLANES: one WARP per plane (no block-wide barrier between planes: the eight warps of a block stream eight planes
// independently and a warp's loads of its next trip are not held back by the others' reduction) instead of one block.
template <int UNROLL, bool WARP_PLANES>
__global__ void __launch_bounds__(256) fold_scale_reduce_vec_kernel(const float* __restrict__ dxp, const float* __restrict__ s,
                                                                     const float* __restrict__ x, float* __restrict__ gx,
                                                                     float* __restrict__ gs1, int H, int W, int pad,
                                                                     int64_t rows) {
  const int Hp = H + 2 * pad, Wp = W + 2 * pad, HW = H * W, Wq = W >> 2, quads = HW >> 2;
  constexpr int TS = WARP_PLANES ? 32 : 256;                   // threads that share a plane
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row_first = WARP_PLANES ? (int64_t)blockIdx.x * 8 + warp : blockIdx.x;
  const int64_t row_step = WARP_PLANES ? (int64_t)gridDim.x * 8 : gridDim.x;
  for (int64_t row = row_first; row < rows; row += row_step) {
    const float sv = s[row];
    const float* src = dxp + row * Hp * Wp;
    const float4* xr = reinterpret_cast<const float4*>(x + row * HW);
    float4* gr = reinterpret_cast<float4*>(gx + row * HW);
    float acc = 0.f;
    for (int q0 = WARP_PLANES ? lane : (int)threadIdx.x; q0 < quads; q0 += TS * UNROLL) {
      float4 c[UNROLL], xv[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int q = q0 + u * TS;
        c[u] = xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q < quads) {
          const int h = q / Wq, w0 = (q - h * Wq) << 2;
          const float* p = src + (h + pad) * Wp + w0 + pad;
          c[u] = make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3));
          if (gs1) xv[u] = __ldg(xr + q);
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int q = q0 + u * TS;
        if (q >= quads) continue;
        const int h = q / Wq, w0 = (q - h * Wq) << 2;
        if (h == 0 || h == H - 1 || w0 == 0 || w0 + 4 == W) {   // a quad on the ring: the general clamp sums
          int h_lo = h == 0 ? 0 : h + pad, h_hi = h == H - 1 ? Hp - 1 : h + pad;
          if (H == 1) { h_lo = 0; h_hi = Hp - 1; }
          float v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int w = w0 + j;
            const int w_lo = w == 0 ? 0 : w + pad, w_hi = w == W - 1 ? Wp - 1 : w + pad;
            float t = 0.f;
            for (int a = h_lo; a <= h_hi; ++a)
              for (int b = w_lo; b <= w_hi; ++b) t += __ldg(src + a * Wp + b);
            v[j] = t;
          }
          c[u] = make_float4(v[0], v[1], v[2], v[3]);
        }
        if (gs1) acc = fmaf(c[u].x, xv[u].x, fmaf(c[u].y, xv[u].y, fmaf(c[u].z, xv[u].z, fmaf(c[u].w, xv[u].w, acc))));
        if (gx) gr[q] = make_float4(c[u].x * sv, c[u].y * sv, c[u].z * sv, c[u].w * sv);
      }
    }
    if (gs1) {
      if (WARP_PLANES) {
        acc = warp_sum(acc);
        if (lane == 0) gs1[row] = acc;
      } else {
        acc = block_sum<256>(acc);
        if (threadIdx.x == 0) gs1[row] = acc;
      }
    }
  }
}

static size_t modconv_bwd_bytes(const ganb200_conv_desc* d) {
  const WgradPlan w = wgrad_plan(d);
  const bool rep = d->pad_mode == 1 && d->pad > 0;
  const size_t direct = rep ? 0 : align_up(sizeof(float) * (size_t)d->N * d->C * d->H * d->W, 1024);
  return tc_workspace(d, GANB200_CONV_DGRAD) + w.dy_bytes + w.x_bytes + w.part_bytes + direct + 2048;
}

}  // namespace ganb200

using namespace ganb200;

extern "C" int ganb200_conv_scatter_mode(int mode) {
  const int prev = scatter_mode();
  if (mode == 0 || mode == 1) g_scatter_mode = mode;
  return prev;
}

extern "C" int ganb200_conv_flat_trace(unsigned long long* device_buf) {
  g_flat_trace = device_buf;
  return 0;
}

extern "C" int ganb200_conv_flat_mode(int mode) {
  const int prev = flat_mode();
  if (mode >= 0 && mode <= 2) g_flat_mode = mode;
  return prev;
}

extern "C" int ganb200_conv_halo_mode(int mode) {
  const int prev = halo_mode();
  if (mode >= 0 && mode <= 2) g_halo_mode = mode;
  return prev;
}

extern "C" size_t ganb200_modconv_bwd_workspace_bytes(const ganb200_conv_desc* d) {
  if (validate_desc(d) != 0) return 0;
  return modconv_bwd_bytes(d);
}

extern "C" int ganb200_modconv_bwd_eligible(const ganb200_conv_desc* d) {
  if (validate_desc(d) != 0) return 0;
  return d->stride == 1 && d->pad_mode == 1 && tc_eligible(d, GANB200_CONV_DGRAD) && tc_eligible(d, GANB200_CONV_WGRAD);
}

// Backward of y = sinv*alpha*conv(w, s*xpad) given gu = dL/dy (generators/generator_13_5.py:234-248, shared-weight form):
//   gx = s * dxs, gs1 = sum_hw dxs * x with dxs = fold(alpha * sum w * (sinv*gu));  gw = alpha * wgrad(s*x, sinv*gu).
// gu is packed to bf16 NHWC once and feeds both the dgrad and the wgrad kernels; the fold of the replicate padding is
// fused with the style scaling and the row reduction.  gx / gs1 (together) or gw may be NULL.
static int modconv_bwd_impl(const ganb200_conv_desc* d, const float* x, const float* w, const float* s,
                            const float* sinv, const float* gu, float* gx, float* gs1, float* gw, void* ws,
                            size_t ws_bytes, void* stream, const void* wp_dgrad);

extern "C" int ganb200_modconv_bwd(const ganb200_conv_desc* d, const float* x, const float* w, const float* s,
                                   const float* sinv, const float* gu, float* gx, float* gs1, float* gw, void* ws,
                                   size_t ws_bytes, void* stream) {
  return modconv_bwd_impl(d, x, w, s, sinv, gu, gx, gs1, gw, ws, ws_bytes, stream, nullptr);
}

// same, with the dgrad-layout packed weights supplied by the caller (ganb200_conv2d_pack_weight(op = DGRAD), cached per
// optimiser step)
extern "C" int ganb200_modconv_bwd_pk(const ganb200_conv_desc* d, const float* x, const float* w, const void* wp_dgrad,
                                      const float* s, const float* sinv, const float* gu, float* gx, float* gs1,
                                      float* gw, void* ws, size_t ws_bytes, void* stream) {
  return modconv_bwd_impl(d, x, w, s, sinv, gu, gx, gs1, gw, ws, ws_bytes, stream, wp_dgrad);
}

static int modconv_bwd_impl(const ganb200_conv_desc* d, const float* x, const float* w, const float* s,
                            const float* sinv, const float* gu, float* gx, float* gs1, float* gw, void* ws,
                            size_t ws_bytes, void* stream, const void* wp_dgrad) {
  int rc = validate_desc(d);
  if (rc) return rc;
  GB_REQUIRE(ganb200_modconv_bwd_eligible(d), GANB200_BAD_SHAPE, "modconv_bwd: shape not eligible");
  GB_REQUIRE(ws && ws_bytes >= modconv_bwd_bytes(d), GANB200_WORKSPACE_TOO_SMALL, "modconv_bwd: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* base = (uint8_t*)align_ws(ws);
  float* dxp = nullptr;
  const WgradPlan wp = wgrad_plan(d);
  uint8_t* wws = base + (tc_workspace(d, GANB200_CONV_DGRAD) / 1024) * 1024;
  // pad == 0 (1x1 convs): no padded domain, the dgrad kernels write a plain [N,C,H,W] buffer behind the wgrad scratch
  float* direct = (float*)(wws + wp.dy_bytes + wp.x_bytes + wp.part_bytes);
  // dgrad part of the workspace first: its head is the packed gu, shared with wgrad
  // GANB200_MODCONV_RED=1 (experimental): the dgrad epilogue also scales by s and reduces sum_hw dxs * x into gs1
  // (epi_fold32 with red_x; contiguous tile ranges so the sums stay in registers per image).  Measured at B = 256
  // (tools/scatter_bench.py): it wins where the main loop is long enough to hide the four epilogue warps' extra loads
  // (96->96 5x5 @64x64: 1.54 -> 1.38 ms) and loses where it is not (48->48 3x3: 0.54 -> 0.60 ms; 96->96 3x3: 1.06 -> 1.25)
  // — ~0 over the generator's mix, so the default keeps the padded domain + fold_scale_reduce pass for the modulated
  // convolution and uses the scatter epilogue for the plain dgrads (the critic's: 0.45 -> 0.31 ms at 64 channels).
  static int red_mode = -1;
  if (red_mode < 0) { const char* e = getenv("GANB200_MODCONV_RED"); red_mode = (e && e[0] == '1') ? 1 : 0; }
  const bool rep0 = d->pad_mode == 1 && d->pad > 0;
  const bool fused = red_mode == 1 && scatter_mode() == 1 && (gx || gs1) && (int64_t)d->H * d->W >= 256;
  bool folded = false;   // dxp holds the folded (unpadded) dxs
  if (fused) {
    // gs1 alone (gx not wanted): the scaled scatter still needs a target; the padded-output region of the workspace serves
    float* tgt = gx ? gx : (rep0 ? (float*)(base + act_bytes(d->N, d->K, d->P, d->Q, 0) + dgrad_w_bytes(d)) : direct);
    rc = dgrad_run(d, gu, w, s, sinv, tgt, base, st, true, nullptr, nullptr, wp_dgrad, gs1 ? x : nullptr, gs1);
    if (rc) return rc;
  } else {
    rc = dgrad_run(d, gu, w, nullptr, sinv, direct, base, st, false, &dxp, nullptr, wp_dgrad);
    if (rc) return rc;
  }
  if (!fused && (gx || gs1)) {
    const int64_t rows = (int64_t)d->N * d->C;
    const int HW = d->H * d->W;
    const int pad = (d->pad_mode == 1 && !folded) ? d->pad : 0;
    const bool vec = HW >= 1024 && (d->W & 3) == 0 && ((((uintptr_t)x) | ((uintptr_t)gx)) & 15) == 0;
    if (vec) {
      static int warp_planes = -1;
      if (warp_planes < 0) { const char* e = getenv("GANB200_FOLD_WARP_PLANES"); warp_planes = (e && e[0] == '0') ? 0 : 1; }
      if (warp_planes) {
        int grid = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)kNumSMs * 32);
        fold_scale_reduce_vec_kernel<4, true><<<grid, 256, 0, st>>>(dxp, s, x, gx, gs1, d->H, d->W, pad, rows);
      } else {
        int grid = (int)std::min<int64_t>(rows, (int64_t)kNumSMs * 32);
        if (HW >= 4096) fold_scale_reduce_vec_kernel<4, false><<<grid, 256, 0, st>>>(dxp, s, x, gx, gs1, d->H, d->W, pad, rows);
        else fold_scale_reduce_vec_kernel<1, false><<<grid, 256, 0, st>>>(dxp, s, x, gx, gs1, d->H, d->W, pad, rows);
      }
    } else if (HW >= 1024) {
      int grid = (int)std::min<int64_t>(rows, (int64_t)kNumSMs * 32);
      fold_scale_reduce_kernel<false><<<grid, 256, 0, st>>>(dxp, s, x, gx, gs1, d->H, d->W, pad, rows);
    } else {
      int grid = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)kNumSMs * 32);
      fold_scale_reduce_kernel<true><<<grid, 256, 0, st>>>(dxp, s, x, gx, gs1, d->H, d->W, pad, rows);
    }
    GB_CHECK_LAUNCH();
  }
  if (gw) {
    rc = wgrad_run(d, x, gu, s, sinv, gw, wws, st, (const __nv_bfloat16*)base);
    if (rc) return rc;
  }
  return 0;
}
