// common.cuh — shared helpers for libganb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/ganb200.h"

namespace ganb200 {

extern thread_local char g_last_error[512];

inline int fail(int code, const char* fmt, const char* a = "", long long b = 0, long long c = 0) {
  snprintf(g_last_error, sizeof(g_last_error), fmt, a, b, c);
  return code;
}

#define GB_CHECK_CUDA(expr)                                                                 \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return ganb200::fail(GANB200_CUDA_ERROR, "CUDA error %s at line %lld", cudaGetErrorString(_e), __LINE__); \
  } while (0)

#define GB_CHECK_LAUNCH() GB_CHECK_CUDA(cudaGetLastError())

#define GB_REQUIRE(cond, code, msg)                                        \
  do {                                                                     \
    if (!(cond)) return ganb200::fail(code, "%s (line %lld)", msg, __LINE__); \
  } while (0)

constexpr int kNumSMs = 148;

inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// grid for a grid-stride elementwise kernel: a multiple of the SM count, capped
inline int ew_grid(int64_t n_items, int threads, int items_per_thread = 4) {
  int64_t blocks = (n_items + (int64_t)threads * items_per_thread - 1) / ((int64_t)threads * items_per_thread);
  int64_t cap = (int64_t)kNumSMs * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// Grid-stride loop over `total` items with 32-bit index arithmetic whenever the count allows it: the body's i / W,
// i % W then compile to 32-bit unsigned divisions (~20 instructions) instead of 64-bit ones (~80), which is what makes a
// flat-index streaming kernel ALU-bound instead of HBM-bound.  `body` is a generic lambda: [&](auto i) { ... }.
template <typename Body>
__device__ __forceinline__ void grid_stride(int64_t total, Body body) {
  if (total < (int64_t)0x7fffffff) {
    const uint32_t n = (uint32_t)total, stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) body(i);
  } else {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) body(i);
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum; result valid in thread 0 (and broadcast to all via smem)
template <int THREADS>
__device__ __forceinline__ float block_sum(float v) {
  __shared__ float red[32];
  __shared__ float total;
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  if (wid == 0) {
    float t = (lane < THREADS / 32) ? red[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) total = t;
  }
  __syncthreads();
  return total;
}

__device__ __forceinline__ bool aligned16(const void* p) { return (((uintptr_t)p) & 15) == 0; }

}  // namespace ganb200
