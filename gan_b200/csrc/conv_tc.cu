// conv_tc.cu — bf16 tcgen05/TMEM implicit-GEMM convolution engine (placeholder until the kernel lands).
#include "common.cuh"
#include "conv_common.cuh"
namespace ganb200 {
int tc_eligible(const ganb200_conv_desc*, int) { return 0; }
size_t tc_workspace(const ganb200_conv_desc*, int) { return 0; }
int tc_fprop(const ganb200_conv_desc*, const float*, const float*, const float*, const float*, const float*, float*, void*, size_t, cudaStream_t) { return fail(GANB200_BAD_ARG, "tc engine not built%s", ""); }
int tc_dgrad(const ganb200_conv_desc*, const float*, const float*, const float*, const float*, float*, void*, size_t, cudaStream_t) { return fail(GANB200_BAD_ARG, "tc engine not built%s", ""); }
int tc_wgrad(const ganb200_conv_desc*, const float*, const float*, const float*, const float*, float*, void*, size_t, cudaStream_t) { return fail(GANB200_BAD_ARG, "tc engine not built%s", ""); }
}
