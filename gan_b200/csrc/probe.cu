// probe.cu — hardware self-test: a SWIZZLE_128B K-major operand whose start address is shifted by a whole number of
// 128-byte rows that is NOT a multiple of the 8-row swizzle atom.
//
// The halo kernel (conv_tc.cu) reads the R taps of one filter column from one box by shifting the A descriptor's start
// by whole input rows, which are multiples of 1024 bytes.  Reading the S horizontal taps the same way needs shifts of
// single pixels (128 bytes): the tile TMA wrote is then addressed from the middle of a swizzle atom.  This kernel loads
// one [rows][64] bf16 box with TMA, multiplies rows [shift, shift + 128) of it with a [n][64] weight box on the tensor
// core and returns the accumulator, with the descriptor's base_offset field either 0 or (start >> 7) & 7
// (tests/test_gpu_kernels.py compares both with the expected product).
#include "common.cuh"
#include "conv_common.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>

namespace ganb200 {
using namespace ptx;
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
                   const uint32_t* box, const uint32_t* estr);

__global__ void __launch_bounds__(128, 1) tc_shift_probe_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB, float* out,
                                                                int rows, int n, int shift, int base_offset_mode,
                                                                int m) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;                        // rows * 128 B
  uint8_t* smB = smem + 256 * 128;            // n * 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smB + 64 * 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(&bars[0], (uint32_t)(rows + n) * 128u);
      tma_load_3d(smA, &tmA, &bars[0], 0, 0, 0);
      tma_load_3d(smB, &tmB, &bars[0], 0, 0, 0);
    }
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t a_addr = smem_u32(smA) + (uint32_t)shift * 128u;
      uint64_t adesc = make_smem_desc_sw128(a_addr, 16, 1024);
      if (base_offset_mode) adesc |= (uint64_t)((a_addr >> 7) & 7u) << 49;
      const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smB), 16, 1024);
      mma_chain<4>(tmem_base, adesc, bdesc, make_idesc_bf16(m, n, 0, 0), 0u, 2u);
      mma_commit(&bars[1]);
    }
    __syncwarp();
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();
  for (int c0 = 0; c0 < n; c0 += 32) {
    float v[32];
    tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
    for (int j = 0; j < 32 && c0 + j < n; ++j) out[(warp * 32 + lane) * n + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

}  // namespace ganb200
using namespace ganb200;

// a: bf16 [rows][64] (rows <= 256, rows >= shift + 128), b: bf16 [n][64] (n = 32 or 64), out: fp32 [128][n]
// ganb200_tc_shift_probe_m: the same with an M = 64 instruction (m = 64): `out` still receives all 128 TMEM lanes, which
// shows where the tensor core puts the 64 accumulator rows (lanes the MMA does not write keep what they held).
extern "C" int ganb200_tc_shift_probe_m(const void* a, const void* b, float* out, int rows, int n, int shift,
                                        int base_offset_mode, int m, void* stream);
extern "C" int ganb200_tc_shift_probe(const void* a, const void* b, float* out, int rows, int n, int shift,
                                      int base_offset_mode, void* stream) {
  return ganb200_tc_shift_probe_m(a, b, out, rows, n, shift, base_offset_mode, 128, stream);
}
extern "C" int ganb200_tc_shift_probe_m(const void* a, const void* b, float* out, int rows, int n, int shift,
                                        int base_offset_mode, int m, void* stream) {
  GB_REQUIRE(m == 64 || m == 128, GANB200_BAD_ARG, "tc_shift_probe: m must be 64 or 128");
  GB_REQUIRE(a && b && out, GANB200_BAD_ARG, "tc_shift_probe: null argument");
  GB_REQUIRE(rows >= shift + 128 && rows <= 256 && shift >= 0 && (n == 32 || n == 64), GANB200_BAD_SHAPE,
             "tc_shift_probe: rows in [shift + 128, 256], n in {32, 64}");
  CUtensorMap ta, tb;
  uint64_t ad[3] = {64, (uint64_t)rows, 1}, as[2] = {128, (uint64_t)rows * 128};
  uint32_t abox[3] = {64, (uint32_t)rows, 1};
  int rc = make_tmap_bf16(&ta, a, 3, ad, as, abox, nullptr);
  if (rc) return rc;
  uint64_t bd[3] = {64, (uint64_t)n, 1}, bs[2] = {128, (uint64_t)n * 128};
  uint32_t bbox[3] = {64, (uint32_t)n, 1};
  rc = make_tmap_bf16(&tb, b, 3, bd, bs, bbox, nullptr);
  if (rc) return rc;
  const int smem = 256 * 128 + 64 * 128 + 64 + 1024;
  GB_CHECK_CUDA(cudaFuncSetAttribute(tc_shift_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  tc_shift_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(ta, tb, out, rows, n, shift, base_offset_mode, m);
  GB_CHECK_LAUNCH();
  return 0;
}
