/*
 * ganb200.h — C-ABI of libganb200.so: the B200 (sm_100a) kernels under the GAN training step.
 *
 * The reference (xxxxx23124/-GAN-) has no FFI of its own: every FLOP of its hot path is an ATen call
 * made from Python.  The entry points below are what a binding for that path binds
 * instead of ATen; each one cites the reference call site(s) it replaces (file:line under
 * /root/reference).  INTEGRATION.md shows the ctypes stub a maintainer adds on the reference side.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes.  No torch types.  All tensors are DEVICE pointers to
 *    contiguous float32 unless stated; layouts are the reference's (NCHW activations, KCRS weights).
 *  - `stream` is a cudaStream_t passed as void*.  Nothing here synchronises the stream or allocates
 *    device memory, so every call is CUDA-graph capturable.  Scratch comes from the caller (`ws`).
 *  - Return value: 0 = ok, negative = ganb200_status.  ganb200_last_error() gives a thread-local text.
 *  - Ownership: the library borrows pointers for the duration of the enqueue only.
 *  - There is no CPU path: the library refuses to run on a device that is not sm_100.
 */
#ifndef GANB200_H
#define GANB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  GANB200_OK = 0,
  GANB200_BAD_SHAPE = -1,
  GANB200_BAD_ALIGN = -2,
  GANB200_UNSUPPORTED_ARCH = -3,
  GANB200_CUDA_ERROR = -4,
  GANB200_WORKSPACE_TOO_SMALL = -5,
  GANB200_BAD_ARG = -6
} ganb200_status;

int ganb200_abi_version(void);
const char* ganb200_last_error(void);
/* 0 when `device` is a compute-capability 10.x GPU, GANB200_UNSUPPORTED_ARCH otherwise. */
int ganb200_device_check(int device);

/* ------------------------------------------------------------------------------------------------
 * Convolution family.  Replaces nn.functional.conv2d + nn.ReplicationPad2d / zero padding and
 * nn.ConvTranspose2d / nn.functional.linear and all their autograd backward formulas:
 *   EqualizedConv2d.forward      discriminators/discriminator_9_4.py:38-40, generators/generator_13_5.py:36-38
 *   Conv2dWeightModulate.forward generators/generator_13_5.py:234-248 (shared-weight form)
 *   EqualizedLinear.forward      discriminator_9_4.py:26-27, generator_13_5.py:25-26   (H=W=R=S=1)
 *   nn.ConvTranspose2d           generator_13_5.py:156,594  (= dgrad of a strided conv)
 *
 *   fprop: y[n,k,p,q]  = oscale[n,k] * alpha * sum_{c,r,s} w[k,c,r,s] * iscale[n,c] * xpad[n,c,p*stride+r,q*stride+s] + bias[k]
 *   dgrad: dx[n,c,h,w] = iscale[n,c] * fold( alpha * sum_{k,r,s} w[k,c,r,s] * oscale[n,k] * dy[n,k,p,q] )
 *   wgrad: dw[k,c,r,s] = alpha * sum_{n,p,q} oscale[n,k]*dy[n,k,p,q] * iscale[n,c]*xpad[n,c,p*stride+r,q*stride+s]
 * xpad = x padded by `pad` on each side with zeros (pad_mode 0) or edge replication (pad_mode 1);
 * fold = adjoint of that padding.  iscale/oscale/bias may be NULL.
 * P = (H + 2*pad - R)/stride + 1, Q likewise.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int N, C, H, W;   /* input  (x / dx)            */
  int K, R, S;      /* filter (w / dw): K,C,R,S   */
  int stride, pad;
  int pad_mode;     /* 0 zeros, 1 replicate       */
  int P, Q;         /* output (y / dy) spatial    */
  float alpha;
  int engine;       /* 0 auto, 1 fp32 CUDA-core implicit GEMM, 2 bf16 tcgen05/TMEM implicit GEMM */
} ganb200_conv_desc;

enum { GANB200_CONV_FPROP = 0, GANB200_CONV_DGRAD = 1, GANB200_CONV_WGRAD = 2 };

size_t ganb200_conv2d_workspace_bytes(const ganb200_conv_desc* d, int op);
/* 1 if the tcgen05 engine can run this (desc, op); engine 0 picks it when this returns 1. */
int ganb200_conv2d_tc_eligible(const ganb200_conv_desc* d, int op);
/* Which implementation ganb200_conv2d_{fprop,dgrad,wgrad} runs for (desc, op): 1 fp32 CUDA-core implicit GEMM,
 * 2 bf16 tcgen05 implicit GEMM, 3 fp32 one-launch linear kernel (1x1 filter on a 1x1 map, N <= 1024, N*C*K <= 2^26, engines 0 and 1:
 * nn.Linear / EqualizedLinear, generators/generator_13_5.py:19-26).  0 for an invalid descriptor. */
int ganb200_conv2d_route(const ganb200_conv_desc* d, int op);
/* tcgen05 engine, stride-1 fprop / dgrad on planes of >= 256 pixels: 1 (default) loads one halo box per filter
 * column and feeds the filter rows from row-shifted views of it; 0 loads one box per filter tap (and runs wgrad per tap); 2 uses the halo kernel even for
 * the shapes the measured heuristic leaves on the per-tap kernel (tuning).  Same results
 * (identical operands and accumulation order per output element).  Returns the previous mode; any other value
 * only queries.  Process-wide; also GANB200_CONV_HALO=0 in the environment. */
int ganb200_conv_halo_mode(int mode);
/* tcgen05 engine, forward problems with stride 1 whose taps stay inside the materialised (replicate) padding can run on
 * the flat kernel — the padded NHWC input read as a [flat pixel][channel] matrix, one contiguous row range per tile and
 * 64-channel chunk, every tap a 128-byte-row shift of the A descriptor.  1 (default) = where it measured faster than the
 * halo / per-tap kernels (64 x 64 planes: layers of up to 64 output channels, 3x3 layers of up to 96 channels), 2 =
 * wherever a problem qualifies, 0 = never.  Same operands and accumulation order per output element as the halo kernel.  Returns the previous mode; any
 * other value only queries.  Process-wide; also GANB200_CONV_FLAT in the environment. */
int ganb200_conv_flat_mode(int mode);
/* Diagnosis of the flat kernel's pipeline: while `device_buf` (3 * 32 * 4 uint64 on the device, zeroed by the caller) is
 * set, CTA 0 of every flat launch writes clock64() stamps for its first 32 tiles — role 0 activation producer
 * {before, after the wait for a free stage}, role 1 MMA issuer {before / after the wait for a free accumulator, after the
 * wait for the activation stage, after the last MMA of the tile is issued}, role 2 epilogue warp {before / after the wait
 * for the accumulator, after its last store}.  NULL switches it off (default).  tools/flat_trace.py prints the timeline. */
int ganb200_conv_flat_trace(unsigned long long* device_buf);
/* Hardware self-test behind the tcgen05 engine's tap addressing (csrc/probe.cu; no reference counterpart): loads a
 * [rows][64] bf16 box (SWIZZLE_128B) with TMA and returns out[128][n] = a[shift : shift + 128] * b^T with the A
 * descriptor's start address shifted by `shift` 128-byte rows — any shift, not only multiples of the 8-row swizzle atom.
 * base_offset_mode 0 leaves the descriptor's base_offset field 0, 1 sets it to (start >> 7) & 7.  rows <= 256,
 * n in {32, 64}. */
int ganb200_tc_shift_probe(const void* a, const void* b, float* out, int rows, int n, int shift, int base_offset_mode,
                           void* stream);
/* the same with an M = 64 (m = 64) or M = 128 instruction; `out` always receives all 128 TMEM lanes */
int ganb200_tc_shift_probe_m(const void* a, const void* b, float* out, int rows, int n, int shift, int base_offset_mode,
                             int m, void* stream);
/* tcgen05 engine, dgrad of replicate-padded convolutions: 0 (default) = padded fp32 domain in the workspace, then a fold
 * pass — every sum in a fixed order, the result is reproducible bit for bit; 1 = the MMA epilogue scatters each
 * padded-domain value straight into the clamped pixel of the unpadded plane (plain stores inside, red.global.add on the
 * border ring, which the host zeroes first): no intermediate, but the border pixels' low bits depend on the schedule.
 * Returns the previous mode; any other value only queries.  Process-wide; also GANB200_DGRAD_SCATTER=1 in the
 * environment. */
int ganb200_conv_scatter_mode(int mode);
int ganb200_conv2d_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias,
                         const float* iscale, const float* oscale, float* y,
                         void* ws, size_t ws_bytes, void* stream);
int ganb200_conv2d_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w,
                         const float* iscale, const float* oscale, float* dx,
                         void* ws, size_t ws_bytes, void* stream);
int ganb200_conv2d_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy,
                         const float* iscale, const float* oscale, float* dw,
                         void* ws, size_t ws_bytes, void* stream);

/* Packed operands of the tcgen05 engine as caller-owned buffers.  Every tcgen05-routed pass first re-packs its fp32 NCHW /
 * KCRS operands to bf16 ([N][H+2p][W+2p][C8] activations with the per-sample scale folded in and replicate padding
 * materialised; [tap][K][C8] weights per output phase).  The same packed tensor serves several passes — x: fprop and the
 * wgrad of the same layer; dy: dgrad and wgrad; the weights: every pass until the next optimiser step — so the host side
 * may pack once, keep the buffer (1024-byte aligned, sizes below) and hand it to the *_pk entry points
 * (gan_b200/kernels.py caches them on the tensors).  A NULL packed pointer means "pack the fp32 operand as before".
 *   pack_act(x, scale[N,C]|NULL, pad): pad = desc.pad for a replicate-padded x (fprop / wgrad), 0 for dy and zero padding
 *   pack_weight(desc, op): op = FPROP or DGRAD layout (wgrad has no weight operand); 0 bytes when (desc, op) is not
 *   routed to the tcgen05 engine.  iscale / oscale given to a *_pk call must be the ones folded into the packed operand
 *   (they are only applied to operands packed inside the call and, for dgrad, to the output). */
size_t ganb200_tc_packed_act_bytes(int N, int C, int H, int W, int pad);
int ganb200_tc_pack_act(const float* x, const float* scale, void* out, int N, int C, int H, int W, int pad, void* stream);
size_t ganb200_conv2d_packed_weight_bytes(const ganb200_conv_desc* d, int op);
int ganb200_conv2d_pack_weight(const ganb200_conv_desc* d, int op, const float* w, void* out, void* stream);
int ganb200_conv2d_fprop_pk(const ganb200_conv_desc* d, const void* xp, const void* wp, const float* bias,
                            const float* oscale, float* y, void* stream);
int ganb200_conv2d_dgrad_pk(const ganb200_conv_desc* d, const float* dy, const void* dyp, const float* w, const void* wp,
                            const float* iscale, const float* oscale, float* dx, void* ws, size_t ws_bytes, void* stream);
int ganb200_conv2d_wgrad_pk(const ganb200_conv_desc* d, const float* x, const void* xp, const float* dy, const void* dyp,
                            const float* iscale, const float* oscale, float* dw, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Separable linear resampling with per-output tap tables:
 *   y[nc,ho,wo] = sum_{a<Th} sum_{b<Tw} wh[ho,a]*ww[wo,b] * x[nc, ih[ho,a], iw[wo,b]]
 * One kernel for Smooth (discriminator_9_4.py:56-72, generator_13_5.py:134-150), bicubic x2 / x0.5
 * (generator_13_5.py:160, discriminator_9_4.py:81), AdaptiveAvgPool2d(5) (discriminator_9_4.py:86),
 * ReplicationPad2d, their compositions and all their adjoints (the host builds the tables).
 * ---------------------------------------------------------------------------------------------- */
int ganb200_resample2d(const float* x, float* y, int NC, int Hi, int Wi, int Ho, int Wo,
                       const int* ih, const float* wh, int Th,
                       const int* iw, const float* ww, int Tw, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Pointwise ops.  Replaces the aten elementwise calls of the model files and loops
 * (sigmoid: discriminator_9_4.py:109; gate product :160; residual add :161; noise add
 * generator_13_5.py:265; tanh generator_1.py; LeakyReLU generator_1.py / discriminator_1.py ...).
 * ---------------------------------------------------------------------------------------------- */
enum {
  GANB200_U_SIGMOID = 0, GANB200_U_TANH = 1, GANB200_U_SQUARE = 2, GANB200_U_AFFINE = 3, /* p0*x+p1 */
  GANB200_U_RSQRT_EPS = 4 /* rsqrt(x+p0) */, GANB200_U_SQRT = 5, GANB200_U_LEAKY = 6 /* slope p0 */,
  GANB200_U_RECIP = 7, GANB200_U_LEAKY_MASK = 8 /* x>0 ? 1 : p0 */, GANB200_U_COPY = 9,
  GANB200_U_POW = 10 /* x^p0 */,
  GANB200_U_GELU = 11 /* x*Phi(x), exact erf form: nn.GELU(), generators/generator_6.py:12 */,
  GANB200_U_GELU_GRAD = 12 /* Phi(x) + x*phi(x) */, GANB200_U_GELU_GRAD2 = 13 /* (2 - x^2)*phi(x) */
};
enum {
  GANB200_B_ADD = 0, GANB200_B_SUB = 1, GANB200_B_MUL = 2, GANB200_B_DIV = 3,
  GANB200_B_SIGMOID_BWD = 4 /* a*b*(1-b) */, GANB200_B_TANH_BWD = 5 /* a*(1-b*b) */,
  GANB200_B_LEAKY_BWD = 6 /* a * (b>0 ? 1 : p0) */
};
/* v-broadcast modes for a logical [N,C,HW] tensor */
enum { GANB200_V_NC = 0 /* v[n*C+c] */, GANB200_V_C = 1 /* v[c] */, GANB200_V_N = 2 /* v[n] */, GANB200_V_SCALAR = 3 };

int ganb200_ew_unary(int op, const float* x, float* y, int64_t n, float p0, float p1, void* stream);
int ganb200_ew_binary(int op, const float* a, const float* b, float* y, int64_t n, float p0, void* stream);
/* y = x + v (op ADD) or x * v (op MUL) with v broadcast per vmode */
int ganb200_ew_bcast(int op, const float* x, const float* v, float* y, int N, int C, int HW, int vmode, void* stream);
/* y = x + v*z, v broadcast per vmode (noise injection generator_13_5.py:265) */
int ganb200_ew_fma_bcast(const float* x, const float* v, const float* z, float* y, int N, int C, int HW, int vmode, void* stream);

/* out = sum x*(y or 1) over: rmode 0: HW -> [N,C];  1: N,HW -> [C];  2: C,HW -> [N];  3: all -> [1].
 * Replaces the bias/scale gradient reductions, AdaptiveAvgPool2d(1) (scaled by the caller), compute_grad2's
 * pow(2).view(B,-1).sum(1) (train/ganR2.py:36-38), torch.mean (train/wgangpR2.py:21). */
/* Self-attention of the discriminator_9 family (discriminators/discriminator_9_3.py:58-80: torch.bmm, nn.Softmax).
 * ganb200_bmm: c[b][m][n] = alpha * sum_k A[b][m][k] * B[b][k][n], fp32, c contiguous [batch][M][N]; the operands are
 * addressed through element strides (A[b][m][k] = a[b * sab + m * sam + k * sak], B[b][k][n] = b[b * sbb + k * sbk +
 * n * sbn]), so the transposes of the forward, backward and double-backward products need no copy.
 * ganb200_softmax_rows_fwd / _bwd: softmax over the last dimension of [rows][n]; gx = y * (gy - sum(gy * y)). */
int ganb200_bmm(const float* a, const float* b, float* c, int batch, int M, int N, int K, int64_t sab, int64_t sam,
                int64_t sak, int64_t sbb, int64_t sbk, int64_t sbn, float alpha, void* stream);
int ganb200_softmax_rows_fwd(const float* x, float* y, int64_t rows, int n, void* stream);
int ganb200_softmax_rows_bwd(const float* gy, const float* y, float* gx, int64_t rows, int n, void* stream);

/* Pixelwise feature-vector normalisation (generators/generator_11.py:9-16, `Pixnorm`): x:[N,C,HW]
 *   fwd: rnorm[n,p] = 1 / sqrt(eps + mean_c x[n,c,p]^2),  y = x * rnorm
 *   bwd: gx = rnorm * (gy - y * mean_c(gy * y)) */
int ganb200_pixnorm_fwd(const float* x, float* y, float* rnorm, int N, int C, int64_t HW, float eps, void* stream);
int ganb200_pixnorm_bwd(const float* gy, const float* y, const float* rnorm, float* gx, int N, int C, int64_t HW,
                        void* stream);
/* AdaptiveAvgPool2d(out_size) of t0 + t1 without materialising the sum (the input of the SK attention,
 * generators/generator_13_5.py:77-80): t0, t1 [planes, H, W] -> out [planes, out_size, out_size].  Eligible shapes
 * (ganb200_pool_sum2_eligible): 32 x 32 and 64 x 64 planes to 5 x 5; torch's bins. */
int ganb200_pool_sum2_eligible(int H, int W, int out_size);
int ganb200_pool_sum2(const float* t0, const float* t1, float* out, int64_t planes, int H, int W, int out_size,
                      void* stream);
/* Backward of the SK block's two uses of its branches in one pass: g_i = att[n][i][c] * g + P^T(gp), i = 0, 1, with P the
 * AdaptiveAvgPool2d(out_size) of ganb200_pool_sum2, g the gradient of the merged map [N, C, H, W] and gp the gradient of
 * the pooled sum [N, C, out_size, out_size] (ganb200_sk_merge2_bwd with g0 = g1 = NULL supplies the attention gradient
 * alone).  32 x 32 and 64 x 64 planes, out_size 5. */
int ganb200_sk_bwd_finish(const float* g, const float* att, const float* gp, float* g0, float* g1, int N, int C, int H,
                          int W, int out_size, void* stream);
/* Selective-kernel merge of two branches in one pass each way: out = att[n][0][c] * t0 + att[n][1][c] * t1
 * ((feas * attention).sum(dim=1) with M = 2: generators/generator_13_5.py:169,201,294); backward g0 = att0 * g,
 * g1 = att1 * g, gatt[n][i][c] = sum_hw g * t_i.  t0, t1, out, g* : [N, C, HW]; att, gatt: [N, 2, C].  g0 = g1 = NULL:
 * gatt only. */
int ganb200_sk_merge2_fwd(const float* t0, const float* t1, const float* att, float* out, int N, int C, int HW,
                          void* stream);
int ganb200_sk_merge2_bwd(const float* g, const float* t0, const float* t1, const float* att, float* g0, float* g1,
                          float* gatt, int N, int C, int HW, void* stream);
int ganb200_reduce(const float* x, const float* y, float* out, int N, int C, int HW, int rmode, float scale, void* stream);

/* PReLU (nn.PReLU(C) discriminator_9_4.py:91,136; nn.PReLU() scalar generator_3_progan.py): a_n = C or 1.
 *   prelu_fwd : y = x>0 ? x : a*x
 *   maskscale : y = g * (x>0 ? 1 : a)         (PReLU input-gradient; linear in g, so also its own double-backward)
 *   prelu_bwd : gx = maskscale(gy,x,a), ga[c] = sum gy*min(x,0)   (ga overwritten)
 *   masked_dot: out[c] = sum u*v*[x<=0]        (d/da of maskscale, needed by the R1/R2/GP double backward) */
int ganb200_prelu_fwd(const float* x, const float* a, float* y, int N, int C, int HW, int a_n, void* stream);
int ganb200_maskscale(const float* g, const float* x, const float* a, float* y, int N, int C, int HW, int a_n, void* stream);
int ganb200_prelu_bwd(const float* gy, const float* x, const float* a, float* gx, float* ga, int N, int C, int HW, int a_n, void* stream);
int ganb200_masked_dot(const float* u, const float* v, const float* x, float* out, int N, int C, int HW, int a_n, void* stream);

/* BatchNorm, training mode (nn.BatchNorm2d / nn.BatchNorm1d in generator_13_5.py:49,56,157,...):
 * batch statistics (biased variance, eps inside sqrt), affine, running stats with `momentum`
 * (running_var gets the unbiased variance), x viewed as [N,C,HW] (HW=1 for BatchNorm1d).
 * If sum_ws != NULL the kernel only produces per-channel (sum, sumsq-about-mean) partials for a
 * cross-rank reduction (data-parallel sync-BN) — see bn_stats / bn_apply. */
int ganb200_bn_stats(const float* x, float* mean, float* var_biased, int N, int C, int HW, void* stream);
int ganb200_bn_apply(const float* x, const float* mean, const float* var_biased, const float* gamma, const float* beta,
                     float* y, float* save_invstd, float* running_mean, float* running_var,
                     int64_t* num_batches_tracked /* += 1 when non-NULL */,
                     int N, int C, int HW, float eps, float momentum, int64_t count_total, void* stream);
/* gx, ggamma, gbeta from (gy, x, mean, invstd, gamma).  If reduce_ws != NULL (2*C floats, holding
 * already-reduced sum(gy) and sum(gy*xhat)) the reduction phase is skipped (sync-BN). */
int ganb200_bn_bwd_reduce(const float* gy, const float* x, const float* mean, const float* invstd,
                          float* sum_gy, float* sum_gy_xhat, int N, int C, int HW, void* stream);
int ganb200_bn_bwd_apply(const float* gy, const float* x, const float* mean, const float* invstd, const float* gamma,
                         const float* sum_gy, const float* sum_gy_xhat, float* gx, float* ggamma, float* gbeta,
                         int N, int C, int HW, int64_t count_total, void* stream);

/* Fused training-mode batch norm for small per-channel populations (N*HW <= 8192: every BatchNorm1d and the
 * BatchNorm2d of the 5x5 / 3x3 / 1x1 attention maps, generator_13_5.py:60-66,100-118,212-214): ONE launch computes the
 * batch mean and biased variance (two-pass), y, save_mean / save_invstd, the running statistics (unbiased variance,
 * momentum) and num_batches_tracked += 1; the backward ONE launch computes dx, dgamma, dbeta.  Same results as
 * ganb200_bn_stats + ganb200_bn_apply resp. ganb200_bn_bwd_reduce + ganb200_bn_bwd_apply; deterministic.
 * prelu_slope ([C] or NULL) fuses the nn.PReLU(C) that follows the norm in _fc_bn_act / _conv_bn_act
 * (generator_13_5.py:60-66): y = prelu(bn(x)); the backward then takes gy = dL/dy, recomputes bn(x) from x, mean,
 * invstd, gamma, beta, and also returns gslope[c] = sum gy * min(bn(x), 0). */
int ganb200_bn_small_eligible(int N, int C, int HW);
int ganb200_bn_small_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean, float* invstd,
                         float* running_mean, float* running_var, int64_t* num_batches_tracked, int N, int C, int HW,
                         float eps, float momentum, const float* prelu_slope, void* stream);
int ganb200_bn_small_bwd(const float* gy, const float* x, const float* mean, const float* invstd, const float* gamma,
                         float* gx, float* ggamma, float* gbeta, int N, int C, int HW, const float* beta,
                         const float* prelu_slope, float* gslope, void* stream);

/* softmax over dim=1 of a [B,M,C] tensor (nn.Softmax(dim=1), generator_13_5.py:72,87) */
int ganb200_softmax_m_fwd(const float* x, float* y, int B, int M, int C, void* stream);
int ganb200_softmax_m_bwd(const float* gy, const float* y, float* gx, int B, int M, int C, void* stream);

/* MiniBatchStdDev scalar (discriminator_9_4.py:42-54): x viewed as [G, L]; s = mean_j sqrt(var_unbiased_j + eps).
 *   fwd    : out[0] = s
 *   bwd    : gx = g[0] * ds/dx
 *   bwd_bwd: given ggx: out_gg[0] = <ggx, ds/dx>,  out_gx = g[0] * (d2s/dx2 . ggx) */
int ganb200_mbstd_fwd(const float* x, float* out, int G, int64_t L, float eps, void* stream);
int ganb200_mbstd_bwd(const float* g, const float* x, float* gx, int G, int64_t L, float eps, void* stream);
int ganb200_mbstd_bwd_bwd(const float* g, const float* x, const float* ggx, float* out_gg, float* out_gx,
                          int G, int64_t L, float eps, void* stream);

/* Grouped launches: a device table of jobs describes a whole family of small independent layers and ONE launch runs
 * them all (gan_b200/csrc/group.cu).  Built for the 519 per-layer style MLPs of generators/generator_13_5.py:205-216,
 * 223-227 (to_style = EqualizedLinear, BatchNorm1d, PReLU, EqualizedLinear, BatchNorm1d on the same w), which layer by
 * layer are ~12 000 launches of 8 us per training iteration.
 *
 * linear_group: out = alpha * A B^T (+ bias) per job, fp32.  An operand is a fixed pointer (a parameter) or, when the
 * pointer is NULL, an element offset into the base pointer given at launch (activations: new memory every call, same
 * table).  Element (i, r) of A sits at a[r*lda + i] when a_red_is_row else a[i*lda + r]; B likewise with (j, r).
 * first_block = index of the job's first CTA in the flattened list of 32 x 32 output tiles (prefix sum). */
typedef struct {
  const float* a; const float* b; float* out; const float* bias;
  int64_t a_off, b_off, out_off;
  int ni, nj, nr, lda, ldb, ldo, a_red_is_row, b_red_is_row;
  float alpha;
  int first_block;
} ganb200_glin_job;
int ganb200_linear_group(const ganb200_glin_job* jobs, int n_jobs, int n_blocks, const float* a_base,
                         const float* b_base, float* out_base, void* stream);

/* Training-mode BatchNorm1d (+ the per-channel PReLU that follows it when slope != NULL) on row-major [N][C] matrices,
 * one job per norm (torch.nn.BatchNorm1d + nn.PReLU of generator_13_5.py:60-62, 189-193).  x / y / gy / gx share the layout
 * base + x_off; per-channel arrays (mean, var, invstd, sums, bias gradient) are flat with the job's C entries at stat_off.
 *   stats     : mean, biased variance over the N rows
 *   apply     : y = PReLU((x - mean) / sqrt(var + eps) * gamma + beta), invstd out, running statistics with momentum
 *               (unbiased variance from count_total = N * world_size), num_batches_tracked += 1
 *   bwd_reduce: gu = gy * PReLU'(u);  s1 = sum_n gu (= dbeta),  s2 = sum_n gu * xhat (= dgamma),  gslope = sum_n gy*min(u,0)
 *   bwd_apply : gx = gamma * invstd * (gu - s1/count - xhat * s2/count) with the (all-reduced) sums; gbias = sum_n gx
 * first_block = prefix sum of ceil(C / 32) (one CTA per job and 32 channels). */
typedef struct {
  const float* gamma; const float* beta; const float* slope;
  float* running_mean; float* running_var; int64_t* num_batches_tracked;
  int64_t x_off, stat_off;
  int C, first_block;
} ganb200_gbn_job;
int ganb200_bn1d_group_stats(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* x_base, int N,
                             float* mean, float* var, void* stream);
int ganb200_bn1d_group_apply(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* x_base, float* y_base,
                             int N, const float* mean, const float* var, float* invstd, float eps, float momentum,
                             int64_t count_total, void* stream);
int ganb200_bn1d_group_bwd_reduce(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* gy_base,
                                  const float* x_base, int N, const float* mean, const float* invstd, float* s1,
                                  float* s2, float* gslope, void* stream);
int ganb200_bn1d_group_bwd_apply(const ganb200_gbn_job* jobs, int n_jobs, int n_blocks, const float* gy_base,
                                 const float* x_base, float* gx_base, int N, const float* mean, const float* invstd,
                                 const float* s1, const float* s2, int64_t count_total, float* gbias, void* stream);

/* Input pipeline with the data set cached in HBM (gan_b200/data.py; reference units/dataloader.py:5-32 decodes, flips,
 * resizes and normalises every image again every epoch on the host): out[b] = cache[idx[b]] ([C,H,W] fp32 images, already
 * resized and normalised once), columns reversed where flip[b] != 0 (RandomHorizontalFlip).  idx: int64[B], flip: uint8[B],
 * both on the device. */
int ganb200_gather_flip(const float* cache, const int64_t* idx, const unsigned char* flip, float* out, int B, int C,
                        int H, int W, void* stream);

/* Preview grid of the sampling path (train/trainunits.py:52-56: generator(fixed_noise) -> torchvision
 * make_grid(padding=2, normalize=True) -> save_image).  minmax: out2 = {min(x), max(x)} over all n elements.
 * make_grid_u8: x [N,C,H,W] fp32 (C = 1 is shown as grey, like torchvision) -> out uint8 [GH][GW][3], the bytes torchvision's
 * save_image would hand to the PNG encoder: v = (clamp(x,min,max) - min) / max(max - min, 1e-5), pixel =
 * (uint8)clamp(v*255 + 0.5, 0, 255), padding pixels 0; GH = (H+padding)*ceil(N/nrow) + padding, GW = (W+padding)*min(nrow,N)
 * + padding, and the bare image when N == 1. */
int ganb200_minmax(const float* x, int64_t n, float* out2, void* stream);
int ganb200_make_grid_u8(const float* x, const float* minmax, unsigned char* out, int N, int C, int H, int W, int nrow,
                         int padding, void* stream);

/* BCELoss, mean reduction, log clamped at -100 (torch.nn.BCELoss, train/gan.py:21, ganR2.py:17).
 * range_flag[0] is set to 1 if any p is outside [0,1] (the caller raises, as the reference does). */
int ganb200_bce_fwd(const float* p, const float* t, float* out, int* range_flag, int64_t n, void* stream);
/* Same loss, but sticky_flag[0] is only ever set to 1, never cleared: the form used inside a captured CUDA graph, where the
 * host cannot read a per-call flag; the caller polls and clears the flag between replays (gan_b200/graphed.py) and raises
 * the reference's RuntimeError (torch.nn.BCELoss input range check) when it tripped. */
int ganb200_bce_fwd_sticky(const float* p, const float* t, float* out, int* sticky_flag, int64_t n, void* stream);
int ganb200_bce_bwd(const float* g, const float* p, const float* t, float* gp, int64_t n, void* stream);

/* Fused multi-tensor Adam / AdamW (torch.optim.Adam train/trainunits.py:18-19; AdamW train/wgangp.py:17-18).
 * Table entries are device pointers; `tab` is a DEVICE array.
 * grad_scale multiplies every gradient first (1/world_size for data-parallel averaging). */
typedef struct {
  float* p; const float* g; float* m; float* v; int64_t n;
} ganb200_adam_entry;
/* block_tab: DEVICE array of n_blocks (tensor index, chunk index) int32 pairs; each CTA updates elements
 * [chunk_idx*chunk, min(n, (chunk_idx+1)*chunk)) of its tensor.  step is the 1-based Adam step count; when
 * step_dev != NULL the count is read from that DEVICE int64 instead (so a captured CUDA graph can be replayed:
 * bump it with ganb200_counter_add before the update). */
int ganb200_multi_tensor_adam_blocks(const ganb200_adam_entry* tab, const int32_t* block_tab, int n_blocks,
                                     int64_t chunk, float lr, float beta1, float beta2, float eps,
                                     float weight_decay, int decoupled, int64_t step, const int64_t* step_dev,
                                     float grad_scale, void* stream);

/* dst = src for every table entry (entry.p = destination, entry.g = source, entry.n elements; m / v ignored), same block
 * table as the Adam launch: the gradients of a data-parallel bucket gathered into the flat buffer that ncclAllReduce
 * sums in place and the fused Adam then reads — one launch per bucket (gan_b200/dist.py; no reference counterpart: the
 * reference is single-device, main.py:23-25). */
int ganb200_multi_tensor_copy(const ganb200_adam_entry* tab, const int32_t* block_tab, int n_blocks, int64_t chunk,
                              void* stream);
int ganb200_counter_add(int64_t* counter, int64_t inc, void* stream);

/* Small all-reduces of the data-parallel step over NVLink peer memory (csrc/peer.cu; no reference counterpart: the
 * reference is single-device, main.py:23-25).  Every rank allocates one exchange buffer (ganb200_peer_alloc: cudaMalloc of
 * ganb200_peer_bytes(cap_floats), zeroed), exports it (64-byte cudaIpc handle), opens the other ranks' handles, and keeps a
 * DEVICE array of the `world` buffer addresses (its own at index `rank`).  ganb200_peer_allreduce is then one single-CTA
 * launch: stage -> signal every rank -> wait for every rank -> add the staged vectors in rank order -> epilogue:
 *   GANB200_PEER_SUM       out0 = sum_r in0, out1 = sum_r in1             (BatchNorm backward: sum dy, sum dy * xhat)
 *   GANB200_PEER_MEAN      out0 = mean_r in0 (, out1 = mean_r in1)        (MiniBatchStdDev scalar and its adjoint)
 *   GANB200_PEER_BN_STATS  in0 = local mean, in1 = local biased variance of equal shards ->
 *                          out0 = global mean, out1 = global biased variance
 * in1 / out1 may both be null (one vector).  n elements per vector, n * (1 or 2) <= cap_floats.  All ranks must issue the
 * same calls in the same order; the call is stream-ordered and CUDA-graph capturable (the sequence number is on the
 * device).  A wait longer than ~20 s sets the buffer's error word (ganb200_peer_error) instead of hanging. */
#define GANB200_PEER_SUM 0
#define GANB200_PEER_MEAN 1
#define GANB200_PEER_BN_STATS 2
int64_t ganb200_peer_bytes(int64_t cap_floats);
int ganb200_peer_alloc(int64_t cap_floats, void** buf);
int ganb200_peer_free(void* buf);
int ganb200_peer_export(void* buf, unsigned char* handle64);
int ganb200_peer_open(const unsigned char* handle64, void** buf);
int ganb200_peer_close(void* buf);
int ganb200_peer_allreduce(const void* bufs_dev, int rank, int world, int64_t cap_floats, const float* in0,
                           const float* in1, float* out0, float* out1, int64_t n, int op, void* stream);
int ganb200_peer_error(const void* buf, int* error);
/* The fused small-population BatchNorm launches (ganb200_bn_small_fwd / _bwd) under data parallelism: every block swaps
 * its two statistics with the same channel's block on the other ranks in the middle of the kernel (8-byte {value, sequence
 * number} stores over NVLink, peer.cuh), so sync-BN stays one launch each way.  mean / invstd returned are the global ones,
 * running_var is unbiased with the global count; ggamma / gbeta / gslope stay rank-local partial sums.
 * C <= ganb200_bn_small_sync_max_channels() (all blocks must be co-resident). */
int ganb200_bn_small_sync_max_channels(void);
int ganb200_bn_small_fwd_sync(const float* x, const float* gamma, const float* beta, float* y, float* mean, float* invstd,
                              float* running_mean, float* running_var, int64_t* num_batches_tracked, int N, int C, int HW,
                              float eps, float momentum, const float* prelu_slope, const void* peer_bufs_dev, int rank,
                              int world, void* stream);
int ganb200_bn_small_bwd_sync(const float* gy, const float* x, const float* mean, const float* invstd, const float* gamma,
                              float* gx, float* ggamma, float* gbeta, int N, int C, int HW, const float* beta,
                              const float* prelu_slope, float* gslope, const void* peer_bufs_dev, int rank, int world,
                              void* stream);

/* Channel-slice copy on [N,C,HW] tensors: dst[:, dst_c0:dst_c0+ncopy] (+)= src[:, src_c0:src_c0+ncopy].
 * Replaces torch.cat(dim=1) / channel slicing and their adjoints (generator_13_5.py:451-467,549-560,
 * discriminator_9_4.py:54). */
int ganb200_chan_copy(const float* src, float* dst, int N, int src_C, int src_c0, int dst_C, int dst_c0,
                      int ncopy, int HW, int accumulate, void* stream);

/* units/mixup.py:9-17 and :19-27 (with its (B,1)x(B,3,64,64) float64 broadcast) */
int ganb200_mixup(const float* t, const float* f, float lam, float* out, int64_t n, void* stream);
int ganb200_mixup_wgan_f64(const double* label, const float* t, const float* f, double* out,
                           int B, int C, int H, int W, void* stream);

/* Clamp all elements to [lo,hi] in place (train/wgan.py:47-48 weight clipping). */
int ganb200_clamp_(float* x, int64_t n, float lo, float hi, void* stream);

/* Spectral norm (torch.nn.utils.parametrizations.spectral_norm as used by
 * discriminators/discriminator_7_wgangpsn.py:14,53-76,112,121): W viewed as [rows, cols] (rows = out channels).
 *   power_iter       : n_iter x { u = normalize(W v); v = normalize(W^T u) } (eps 1e-12, torch's order), in place;
 *                      sigma[0] = u . (W v)
 *   power_iter_batched: the same for a device table of n_jobs matrices in ONE launch (one CTA per matrix).  `jobs`
 *                      is a device array of ganb200_spectral_job; max_rows2_plus_cols = max over jobs of 2*rows+cols.
 *   scale            : out = W / sigma[0]                  (the parametrised weight)
 *   scale_bwd        : gW = G/sigma - <G,W>/sigma^2 * u v^T (u, v constants, as in torch); dot_scratch: 1 float. */
typedef struct ganb200_spectral_job {
  const float* W;
  float* u;
  float* v;
  float* sigma;
  int rows, cols;
} ganb200_spectral_job;
int ganb200_spectral_power_iter(const float* W, float* u, float* v, float* sigma, int rows, int cols,
                                int n_iter, float eps, void* stream);
int ganb200_spectral_power_iter_batched(const void* jobs, int n_jobs, int max_rows2_plus_cols, int n_iter,
                                        float eps, void* stream);
int ganb200_spectral_scale(const float* W, const float* sigma, float* out, int64_t n, void* stream);
int ganb200_spectral_scale_bwd(const float* G, const float* W, const float* u, const float* v,
                               const float* sigma, float* dot_scratch, float* gW, int rows, int cols,
                               void* stream);

/* Depthwise R x R convolution (R = S in {3, 5, 7}), zero padding 0 .. R/2 + 1, stride 1|2
 * (discriminators/discriminator_7_wgangpsn.py:55,61,72: nn.Conv2d(C, C, 3, stride, padding=1, groups=C);
 * discriminators/discriminator_6_wgangp.py:12-13: kernel 3 / 5 / 7, reflect padding applied by the caller).
 * x:[N,C,H,W]  w:[C,1,R,R]  y:[N,C,P,Q].
 *   fprop: y  = alpha * dwconv(x, w) + bias        (bias may be NULL)
 *   dgrad: dx = alpha * dwconv^T(dy, w)
 *   wgrad: dw = alpha * sum_{n,p,q} dy * x_window   (workspace: ganb200_dwconv2d_wgrad_workspace_bytes for 3x3,
 *          ganb200_dwconv2d_wgrad_workspace_bytes_rs for any admitted kernel) */
int ganb200_dwconv2d_fprop(const float* x, const float* w, const float* bias, float* y, int N, int C, int H,
                           int W, int R, int S, int stride, int pad, float alpha, void* stream);
int ganb200_dwconv2d_dgrad(const float* dy, const float* w, float* dx, int N, int C, int H, int W, int R,
                           int S, int stride, int pad, float alpha, void* stream);
int64_t ganb200_dwconv2d_wgrad_workspace_bytes(int N, int C);
int64_t ganb200_dwconv2d_wgrad_workspace_bytes_rs(int N, int C, int R, int S);
int ganb200_dwconv2d_wgrad(const float* x, const float* dy, float* dw, void* workspace,
                           int64_t workspace_bytes, int N, int C, int H, int W, int R, int S, int stride,
                           int pad, float alpha, void* stream);

/* Allocation-free depthwise wgrad.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same result as ganb200_dwconv2d_wgrad, but the split partials live in a 256 KiB file-scope device array inside the
 * library (held by every GPU that loads it, until exit) instead of a caller workspace; the call performs no cudaMalloc /
 * cudaFree / allocating library call.  Shapes whose partials exceed ganb200_dwconv2d_wgrad_static_max_bytes()
 * (splits * C * R * S * 4 bytes > 256 KiB; for 3x3 kernels C above ~6 600) are rejected with GANB200_WORKSPACE_TOO_SMALL.  All calls share
 * the array: calls on one stream are ordered; calls that overlap on the GPU (different streams) race. */
int64_t ganb200_dwconv2d_wgrad_static_max_bytes(void);
int ganb200_dwconv2d_wgrad_static(const float* x, const float* dy, float* dw, int N, int C, int H, int W, int R, int S,
                                  int stride, int pad, float alpha, void* stream);

/* channel_shuffle (discriminator_7_wgangpsn.py:27-41): y[n, j*groups + i] = x[n, i*(C/groups) + j].
 * Its adjoint / inverse is the same call with groups' = C/groups. */
int ganb200_chan_shuffle(const float* x, float* y, int N, int C, int HW, int groups, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused StyleConv layer of the final generator (generators/generator_13_5.py:219-266):
 *   y = sinv[b,o] * alpha * conv(w, s[b,i]*xpad)      Conv2dWeightModulate.forward :234-248 (shared-weight form)
 *   u = y + scale_noise[o] * z,  z ~ N(0,1)            StyleConv.forward :264-265
 *   a = PReLU(u)                                        the nn.PReLU that follows it (:315-321, :291)
 * The noise is never stored: z is a pure function of (rng_state = {seed, forward counter}, layer id, element index)
 * — Philox4x32-10 + Box-Muller — and is regenerated by the backward kernel.  Tensors are [N][K][HW], HW % 4 == 0.
 * ---------------------------------------------------------------------------------------------- */
/* z[i] for i < n (n % 4 == 0): the exact noise the fused kernels draw for (rng_state, layer) */
int ganb200_noise_fill(float* z, int64_t n, const int64_t* rng_state, int layer, void* stream);
/* sinv[b,o] = rsqrt(alpha^2 * sum_i s[b,i]^2 * wsq[o,i] + eps), wsq[o,i] = sum_taps w[o,i,:]^2   (:241-243) */
int ganb200_demod(const float* s, const float* wsq, float* sinv, int B, int C, int K, float alpha, float eps, void* stream);
/* adjoints of the demodulation: Ad[b,o] = sum_hw g_y*y;  dd = -0.5*Ad*sinv^2
 *   gs[b,i] = gs1[b,i] + 2*s[b,i]*alpha^2*sum_o dd[b,o]*wsq[o,i]          (gs1 may be NULL)
 *   gw[o,i,t] += 2*w[o,i,t]*alpha^2*sum_b dd[b,o]*s[b,i]^2 */
int ganb200_demod_bwd_s(const float* gs1, const float* s, const float* Ad, const float* sinv, const float* wsq,
                        float* gs, int B, int C, int K, float alpha, void* stream);
int ganb200_demod_bwd_w(float* gw, const float* w, const float* Ad, const float* sinv, const float* s, int B, int C,
                        int K, int RS, float alpha, void* stream);
/* u = y + ns[k]*z (u may alias y; written only when ns != NULL);  a = PReLU(u) (written only when slope != NULL) */
int ganb200_style_tail_fwd(const float* y, float* u, float* a, const float* ns, const float* slope, int slope_n,
                           int N, int K, int HW, const int64_t* rng_state, int layer, void* stream);
/* g_u = g_a*(u>0?1:slope) (written when slope != NULL); per row (n,k): Ad = sum g_u*(u - ns*z), Z = sum g_u*z,
 * Pm = sum g_a*min(u,0).  Ad / Z / Pm may be NULL. */
int ganb200_style_tail_bwd(const float* ga, const float* u, float* gu, const float* ns, const float* slope, int slope_n,
                           float* Ad, float* Z, float* Pm, int N, int K, int HW, const int64_t* rng_state, int layer,
                           void* stream);
/* Backward of the modulated convolution y = sinv*alpha*conv(w, s*xpad) given gu = dL/dy (stride 1, replicate padding):
 *   gx = s*dxs, gs1[n,c] = sum_hw dxs*x with dxs = fold(alpha*sum w*(sinv*gu));   gw = alpha*wgrad(s*x, sinv*gu)
 * One bf16 pack of gu feeds the dgrad and the wgrad tensor-core kernels; the padding adjoint, the style scaling and the
 * row reduction are one kernel.  gx/gs1 or gw may be NULL.  Only for shapes ganb200_modconv_bwd_eligible() accepts. */
size_t ganb200_modconv_bwd_workspace_bytes(const ganb200_conv_desc* d);
int ganb200_modconv_bwd_eligible(const ganb200_conv_desc* d);
int ganb200_modconv_bwd(const ganb200_conv_desc* d, const float* x, const float* w, const float* s, const float* sinv,
                        const float* gu, float* gx, float* gs1, float* gw, void* ws, size_t ws_bytes, void* stream);
/* the same with the dgrad-layout packed weights supplied by the caller (ganb200_conv2d_pack_weight, op = DGRAD) */
int ganb200_modconv_bwd_pk(const ganb200_conv_desc* d, const float* x, const float* w, const void* wp_dgrad,
                           const float* s, const float* sinv, const float* gu, float* gx, float* gs1, float* gw,
                           void* ws, size_t ws_bytes, void* stream);
/* gx = dxs * s[n,c] (gx may be NULL);  gs1[n,c] = sum_hw dxs*x (gs1 may be NULL) */
int ganb200_scale_reduce(const float* dxs, const float* s, const float* x, float* gx, float* gs1, int N, int C, int HW,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GANB200_H */
