// conv_api.cu — C-ABI entry points of the convolution family; picks the engine.
#include "common.cuh"
#include "conv_common.cuh"
using namespace ganb200;

static bool use_tc(const ganb200_conv_desc* d, int op) {
  if (d->engine == 1) return false;
  return tc_eligible(d, op) == 1;
}

extern "C" int ganb200_conv2d_tc_eligible(const ganb200_conv_desc* d, int op) {
  if (validate_desc(d) != 0) return 0;
  return tc_eligible(d, op);
}

extern "C" size_t ganb200_conv2d_workspace_bytes(const ganb200_conv_desc* d, int op) {
  if (validate_desc(d) != 0) return 0;
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
  if (use_tc(d, GANB200_CONV_FPROP)) return tc_fprop(d, x, w, bias, iscale, oscale, y, ws, ws_bytes, (cudaStream_t)stream);
  return simt_fprop(d, x, w, bias, iscale, oscale, y, (cudaStream_t)stream);
}

extern "C" int ganb200_conv2d_dgrad(const ganb200_conv_desc* d, const float* dy, const float* w, const float* iscale,
                                    const float* oscale, float* dx, void* ws, size_t ws_bytes, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  if (d->engine == 2 && !tc_eligible(d, GANB200_CONV_DGRAD))
    return fail(GANB200_BAD_SHAPE, "conv dgrad: shape not eligible for the tcgen05 engine%s", "");
  if (use_tc(d, GANB200_CONV_DGRAD)) return tc_dgrad(d, dy, w, iscale, oscale, dx, ws, ws_bytes, (cudaStream_t)stream);
  return simt_dgrad(d, dy, w, iscale, oscale, dx, ws, ws_bytes, (cudaStream_t)stream);
}

extern "C" int ganb200_conv2d_wgrad(const ganb200_conv_desc* d, const float* x, const float* dy, const float* iscale,
                                    const float* oscale, float* dw, void* ws, size_t ws_bytes, void* stream) {
  int rc = validate_desc(d);
  if (rc) return rc;
  if (d->engine == 2 && !tc_eligible(d, GANB200_CONV_WGRAD))
    return fail(GANB200_BAD_SHAPE, "conv wgrad: shape not eligible for the tcgen05 engine%s", "");
  if (use_tc(d, GANB200_CONV_WGRAD)) return tc_wgrad(d, x, dy, iscale, oscale, dw, ws, ws_bytes, (cudaStream_t)stream);
  return simt_wgrad(d, x, dy, iscale, oscale, dw, (cudaStream_t)stream);
}
