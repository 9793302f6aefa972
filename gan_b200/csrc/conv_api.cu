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
