// conv_common.cuh — geometry shared by the two convolution engines.
#pragma once
#include "common.cuh"

namespace ganb200 {

struct ConvGeom {
  int N, C, H, W, K, R, S, stride, pad, pad_mode, P, Q;
  int RS, CRS, PQ, HW;
  float alpha;
};

template <class G>
inline void fill_geom(G& g, const ganb200_conv_desc* d) {
  g.N = d->N; g.C = d->C; g.H = d->H; g.W = d->W; g.K = d->K; g.R = d->R; g.S = d->S;
  g.stride = d->stride; g.pad = d->pad; g.pad_mode = d->pad_mode; g.P = d->P; g.Q = d->Q;
  g.RS = d->R * d->S; g.CRS = d->C * g.RS; g.PQ = d->P * d->Q; g.HW = d->H * d->W;
  g.alpha = d->alpha;
}

inline int validate_desc(const ganb200_conv_desc* d) {
  if (!d) return fail(GANB200_BAD_ARG, "conv: null desc%s", "");
  if (d->N <= 0 || d->C <= 0 || d->H <= 0 || d->W <= 0 || d->K <= 0 || d->R <= 0 || d->S <= 0 || d->stride <= 0 ||
      d->pad < 0)
    return fail(GANB200_BAD_SHAPE, "conv: non-positive dimension%s", "");
  if (d->pad_mode != 0 && d->pad_mode != 1) return fail(GANB200_BAD_ARG, "conv: pad_mode must be 0 or 1%s", "");
  int P = (d->H + 2 * d->pad - d->R) / d->stride + 1, Q = (d->W + 2 * d->pad - d->S) / d->stride + 1;
  if (d->H + 2 * d->pad < d->R || d->W + 2 * d->pad < d->S || P != d->P || Q != d->Q)
    return fail(GANB200_BAD_SHAPE, "conv: P/Q inconsistent with H,W,R,S,stride,pad (expected P=%s%lld Q=%lld)", "", P, Q);
  return 0;
}

// fp32 CUDA-core engine (conv_simt.cu)
int simt_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias, const float* iscale,
               const float* oscale, float* y, cudaStream_t s);
size_t simt_dgrad_workspace(const ganb200_conv_desc* d);
int simt_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale, const float* oscale,
               float* dx, void* ws, size_t ws_bytes, cudaStream_t s);
int simt_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale, const float* oscale,
               float* dw, cudaStream_t s);

// fp32 one-launch kernels for 1x1 convolutions on 1x1 maps = nn.Linear (conv_linear.cu)
int linear_eligible(const ganb200_conv_desc* d);
int linear_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias, const float* iscale,
                 const float* oscale, float* y, cudaStream_t s);
int linear_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale, const float* oscale,
                 float* dx, cudaStream_t s);
int linear_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale, const float* oscale,
                 float* dw, cudaStream_t s);

int fold_replicate(const float* dxp, float* dx, int64_t NC, int H, int W, int pad, cudaStream_t s);

// bf16 tcgen05 engine (conv_tc.cu)
int tc_eligible(const ganb200_conv_desc* d, int op);
size_t tc_workspace(const ganb200_conv_desc* d, int op);
int tc_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias, const float* iscale,
             const float* oscale, float* y, void* ws, size_t ws_bytes, cudaStream_t s, const void* xp = nullptr,
             const void* wp = nullptr);
int tc_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale, const float* oscale,
             float* dx, void* ws, size_t ws_bytes, cudaStream_t s, const void* dyp = nullptr, const void* wp = nullptr);
int tc_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale, const float* oscale,
             float* dw, void* ws, size_t ws_bytes, cudaStream_t s, const void* xp = nullptr, const void* dyp = nullptr);
// packed bf16 operands of the engine as caller-owned buffers (conv_tc.cu)
size_t tc_packed_act_bytes(int N, int C, int H, int W, int pad);
int tc_pack_act(const float* x, const float* scale, void* out, int N, int C, int H, int W, int pad, cudaStream_t s);
size_t tc_packed_weight_bytes(const ganb200_conv_desc* d, int op);
int tc_pack_weight(const ganb200_conv_desc* d, int op, const float* w, void* out, cudaStream_t s);

}  // namespace ganb200
