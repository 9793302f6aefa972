// det_reduce.cuh — deterministic cross-block sums (no floating-point atomics).
//
// Several reductions split one output over many CTAs (BatchNorm statistics over N*H*W, PReLU slope gradients, loss means,
// the MiniBatchStdDev scalar ...).  Round 1 combined the CTA partials with atomicAdd: the sum then depends on the order
// the CTAs happen to finish, the step is not reproducible run to run, and generator_13_5 at small batches (BatchNorm over
// 4-8 samples in every attention head) amplifies that 1e-7 to 5e-3 at its output.  Here every CTA stores its partial in
// a fixed slot of a scratch array, takes a ticket, and the CTA that draws the LAST ticket adds the partials in slot order:
// the result is a fixed function of the inputs and the launch geometry.
//
// The scratch is a file-scope __device__ array per translation unit, cut into kDetRegions regions; a launch uses the
// region that belongs to its STREAM (det_region: first come first served, sticky), so launches that are ordered on one
// stream share a region safely and launches on different streams (gan_b200/streams.py branch streams, graph branches
// captured from them) never share one.  A stream that arrives after the regions are taken, or a launch whose partials
// do not fit a region, gets region -1: the kernel then falls back to atomicAdd into an output the host wrapper zeroed
// (correct, order dependent).  Tickets reset themselves.  A graph keeps the region of the stream it was captured on:
// replay it on one stream at a time.
#pragma once
#include <mutex>
#include "common.cuh"

namespace ganb200 {

constexpr int kDetRegions = 16;
constexpr int kDetPartials = 1 << 17;   // floats per region (512 KB)
constexpr int kDetGroups = 1 << 13;     // ticket counters per region

static __device__ __align__(16) float g_det_part[(size_t)kDetRegions * kDetPartials];
static __device__ unsigned int g_det_count[kDetRegions * kDetGroups];

// region of this stream, or -1 when all regions belong to other streams (shared by every translation unit)
inline int det_region(cudaStream_t s) {
  static std::mutex mu;
  static cudaStream_t seen[kDetRegions];
  static int n = 0;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < n; ++i)
    if (seen[i] == s) return i;
  if (n < kDetRegions) { seen[n] = s; return n++; }
  return -1;
}

// host side: the region to launch with — -1 (atomics into a zeroed output) when the stream has none or the launch's
// (groups x nsplit x sums) partials do not fit
inline int det_region_for(cudaStream_t s, int64_t groups, int64_t nsplit, int sums = 1) {
  if (nsplit <= 1) return 0;   // unused by the kernel
  if (groups > kDetGroups || groups * nsplit * sums > kDetPartials) return -1;
  return det_region(s);
}

// Called by ALL threads of CTA (g, split) with the CTA's partial t (as block_sum returns it).  Writes
// out_a[0] = scale * (sum over the nsplit CTAs of group g), and the same for out_b / tb when SUMS == 2 (one ticket):
//   nsplit == 1   the CTA's own value;
//   region < 0    no scratch for this stream: atomicAdd into out_* (the host wrapper zeroed them) — order dependent;
//   otherwise     partial -> slot `split`, ticket; the CTA with the last ticket adds the slots in order.
template <int THREADS, int SUMS>
__device__ __forceinline__ void det_group_store(int region, int g, int split, int nsplit, float scale, float ta,
                                                float* out_a, float tb = 0.f, float* out_b = nullptr) {
  if (nsplit == 1) {
    if (threadIdx.x == 0) { *out_a = ta * scale; if (SUMS == 2) *out_b = tb * scale; }
    return;
  }
  if (region < 0) {
    if (threadIdx.x == 0) { atomicAdd(out_a, ta * scale); if (SUMS == 2) atomicAdd(out_b, tb * scale); }
    return;
  }
  __shared__ bool last;
  float* part = g_det_part + (size_t)region * kDetPartials + (size_t)g * nsplit * SUMS;
  unsigned int* count = g_det_count + region * kDetGroups + g;
  if (threadIdx.x == 0) {
    part[split] = ta;
    if (SUMS == 2) part[nsplit + split] = tb;
    __threadfence();
    last = atomicAdd(count, 1u) == (unsigned int)nsplit - 1u;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  float a = 0.f, b = 0.f;
  for (int s = threadIdx.x; s < nsplit; s += THREADS) {
    a += __ldcg(part + s);
    if (SUMS == 2) b += __ldcg(part + nsplit + s);
  }
  a = block_sum<THREADS>(a);
  if (SUMS == 2) b = block_sum<THREADS>(b);
  if (threadIdx.x == 0) {
    *count = 0u;
    *out_a = a * scale;
    if (SUMS == 2) *out_b = b * scale;
  }
}

}  // namespace ganb200
