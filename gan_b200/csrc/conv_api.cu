// conv_api.cu — C-ABI entry points of the convolution family; picks the engine.
#include "common.cuh"
#include "conv_common.cuh"
using namespace ganb200;

// engines 0 (auto) and 1 (fp32): linear layers at training batch sizes run as ONE fp32 launch (conv_linear.cu) instead
// of pack + pack + tcgen05 kernel (+ memset + split reduce) resp. the generic fp32 implicit GEMM; engine 2 forces the
// tcgen05 implicit GEMM.
static bool use_linear(const ganb200_conv_desc* d) { return d->engine != 2 && linear_eligible(d) == 1; }

static bool use_tc(const ganb200_conv_desc* d, int op) {
  if (d->engine == 1 || use_linear(d)) return false;
  return tc_eligible(d, op) == 1;
}

extern "C" int ganb200_conv2d_tc_eligible(const ganb200_conv_desc* d, int op) {
  if (validate_desc(d) != 0) return 0;
  return tc_eligible(d, op);
}

extern "C" int ganb200_conv2d_route(const ganb200_conv_desc* d, int op) {
  if (validate_desc(d) != 0) return 0;
  if (use_linear(d)) return 3;
  return use_tc(d, op) ? 2 : 1;
}

extern "C" size_t ganb200_conv2d_workspace_bytes(const ganb200_conv_desc* d, int op) {
  if (validate_desc(d) != 0) return 0;
  if (use_linear(d)) return 0;
  if (use_tc(d, op)) return tc_workspace(d, op);
  if (op == GANB200_CONV_DGRAD) return simt_dgrad_workspace(d);
  return 0;
}

extern "C" int ganb200_conv2d_fprop(const ganb200_conv_desc* d, const float* x, const float* w, const float* bias,
                                    const float* iscale, const float* oscale, float* y, void* ws, size_t ws_bytes,
                                    void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  if (d->engine == 2 && !tc_eligible(d, GANB200_CONV_FPROP))
    return fail(GANB200_BAD_SHAPE, "conv fprop: shape not eligible for the tcgen05 engine%s", "");
  if (use_linear(d)) return linear_fprop(d, x, w, bias, iscale, oscale, y, (cudaStream_t)stream);
  if (use_tc(d, GANB200_CONV_FPROP)) return tc_fprop(d, x, w, bias, iscale, oscale, y, ws, ws_bytes, (cudaStream_t)stream);
  return simt_fprop(d, x, w, bias, iscale, oscale, y, (cudaStream_t)stream);
}

extern "C" int ganb200_conv2d_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale,
                                    const float* oscale, float* dx, void* ws, size_t ws_bytes, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  if (d->engine == 2 && !tc_eligible(d, GANB200_CONV_DGRAD))
    return fail(GANB200_BAD_SHAPE, "conv dgrad: shape not eligible for the tcgen05 engine%s", "");
  if (use_linear(d)) return linear_dgrad(d, dy, w, iscale, oscale, dx, (cudaStream_t)stream);
  if (use_tc(d, GANB200_CONV_DGRAD)) return tc_dgrad(d, dy, w, iscale, oscale, dx, ws, ws_bytes, (cudaStream_t)stream);
  return simt_dgrad(d, dy, w, iscale, oscale, dx, ws, ws_bytes, (cudaStream_t)stream);
}

extern "C" int ganb200_conv2d_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale,
                                    const float* oscale, float* dw, void* ws, size_t ws_bytes, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  if (d->engine == 2 && !tc_eligible(d, GANB200_CONV_WGRAD))
    return fail(GANB200_BAD_SHAPE, "conv wgrad: shape not eligible for the tcgen05 engine%s", "");
  if (use_linear(d)) return linear_wgrad(d, x, dy, iscale, oscale, dw, (cudaStream_t)stream);
  if (use_tc(d, GANB200_CONV_WGRAD)) return tc_wgrad(d, x, dy, iscale, oscale, dw, ws, ws_bytes, (cudaStream_t)stream);
  return simt_wgrad(d, x, dy, iscale, oscale, dw, (cudaStream_t)stream);
}

// ---- packed operands (see include/ganb200.h) ---------------------------------------------------------------------------
extern "C" size_t ganb200_tc_packed_act_bytes(int N, int C, int H, int W, int pad) {
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0 || pad < 0) return 0;
  return tc_packed_act_bytes(N, C, H, W, pad);
}

extern "C" int ganb200_tc_pack_act(const float* x, const float* scale, void* out, int N, int C, int H, int W, int pad,
                                   void* stream) {
  GB_REQUIRE(x && out && N > 0 && C > 0 && H > 0 && W > 0 && pad >= 0, GANB200_BAD_ARG, "tc_pack_act: bad arguments");
  return tc_pack_act(x, scale, out, N, C, H, W, pad, (cudaStream_t)stream);
}

extern "C" size_t ganb200_conv2d_packed_weight_bytes(const ganb200_conv_desc* d, int op) {
  if (validate_desc(d) != 0 || !use_tc(d, op)) return 0;
  return tc_packed_weight_bytes(d, op);
}

extern "C" int ganb200_conv2d_pack_weight(const ganb200_conv_desc* d, int op, const float* w, void* out, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  GB_REQUIRE(use_tc(d, op), GANB200_BAD_SHAPE, "conv2d_pack_weight: (desc, op) does not run on the tcgen05 engine");
  return tc_pack_weight(d, op, w, out, (cudaStream_t)stream);
}

extern "C" int ganb200_conv2d_fprop_pk(const ganb200_conv_desc* d, const void* xp, const void* wp, const float* bias,
                                       const float* oscale, float* y, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  GB_REQUIRE(xp && wp && use_tc(d, GANB200_CONV_FPROP), GANB200_BAD_ARG,
             "conv2d_fprop_pk: needs both packed operands and a tcgen05-routed shape");
  return tc_fprop(d, nullptr, nullptr, bias, nullptr, oscale, y, nullptr, 0, (cudaStream_t)stream, xp, wp);
}

extern "C" int ganb200_conv2d_dgrad_pk(const ganb200_conv_desc* d, const float* dy, const void* dyp, const float* w,
                                       const void* wp, const float* iscale, const float* oscale, float* dx, void* ws,
                                       size_t ws_bytes, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  GB_REQUIRE(use_tc(d, GANB200_CONV_DGRAD) && (dy || dyp) && (w || wp), GANB200_BAD_ARG,
             "conv2d_dgrad_pk: tcgen05-routed shapes only; every operand as fp32 or packed");
  return tc_dgrad(d, dy, w, iscale, oscale, dx, ws, ws_bytes, (cudaStream_t)stream, dyp, wp);
}

extern "C" int ganb200_conv2d_wgrad_pk(const ganb200_conv_desc* d, const float* x, const void* xp, const float* dy,
                                       const void* dyp, const float* iscale, const float* oscale, float* dw, void* ws,
                                       size_t ws_bytes, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  GB_REQUIRE(use_tc(d, GANB200_CONV_WGRAD) && (x || xp) && (dy || dyp), GANB200_BAD_ARG,
             "conv2d_wgrad_pk: tcgen05-routed shapes only; every operand as fp32 or packed");
  return tc_wgrad(d, x, dy, iscale, oscale, dw, ws, ws_bytes, (cudaStream_t)stream, xp, dyp);
}
