// peer.cuh — layout of a rank's exchange buffer and the device-side pieces of the NVLink peer exchanges (peer.cu has the
// protocol description and the single-CTA all-reduce; layers.cu uses the per-block LL exchange inside the fused
// small-population BatchNorm kernels).
#pragma once
#include "common.cuh"

namespace ganb200 {

constexpr int kPeerMaxWorld = 16;
constexpr int kPeerFlagStride = 32;                        // uint32 per flag line (128 B)
constexpr int kPeerLLSlots = 2048;                         // per-block exchanges of one launch (channels)
constexpr size_t kPeerCtlOffset = kPeerMaxWorld * 128;     // bytes: [0] sequence number, [1] error word, [2] done counter
constexpr size_t kPeerLLOffset = kPeerCtlOffset + 128;     // u64 ll[2 parities][kPeerMaxWorld sources][kPeerLLSlots][2]
constexpr size_t kPeerLLBytes = (size_t)2 * kPeerMaxWorld * kPeerLLSlots * 2 * 8;
constexpr size_t kPeerDataOffset = kPeerLLOffset + kPeerLLBytes;   // float data[2 parities][cap]
constexpr unsigned long long kPeerTimeoutNs = 20000000000ull;

// what a kernel needs to take part in an exchange; bufs == nullptr: single rank, no exchange
struct PeerRef {
  char* const* bufs;   // device array [world] of exchange buffers (own mapping at [rank])
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// sequence number of the exchange this launch performs (every thread of every block reads it before peer_seq_commit of
// the same launch can run: the commit waits for a ticket from every block)
__device__ __forceinline__ unsigned int peer_seq_begin(const PeerRef& p) {
  return reinterpret_cast<const unsigned int*>(p.bufs[p.rank] + kPeerCtlOffset)[0] + 1u;
}

// end of a multi-block launch: the block that draws the last ticket publishes the sequence number.  Call after the
// block's exchange, from all threads.
__device__ __forceinline__ void peer_seq_commit(const PeerRef& p, unsigned int seq) {
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* ctl = reinterpret_cast<unsigned int*>(p.bufs[p.rank] + kPeerCtlOffset);
    if (atomicAdd(ctl + 2, 1u) == gridDim.x - 1) {
      ctl[2] = 0u;
      __threadfence();
      ctl[0] = seq;
    }
  }
}

// Per-block exchange of two floats with the same block of every other rank, NCCL-LL style: each value travels as ONE
// 8-byte store {value, sequence number} into slot (parity, source rank, slot) of the destination's buffer, so the
// receiver needs no fence — it polls its own memory until the sequence number shows.  Returns the sums over ranks, added
// in rank order, in every thread.  Call from all threads of the block (blockDim.x >= world); a and b block-uniform.
__device__ __forceinline__ float2 peer_ll_sum2(const PeerRef& p, unsigned int seq, int slot, float a, float b) {
  __shared__ float sa[kPeerMaxWorld], sb[kPeerMaxWorld];
  const size_t par = (size_t)(seq & 1u) * kPeerMaxWorld;
  if (threadIdx.x < p.world) {
    const int r = threadIdx.x;
    const unsigned long long tag = (unsigned long long)seq << 32;
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(p.bufs[r] + kPeerLLOffset) +
                              ((par + p.rank) * kPeerLLSlots + slot) * 2;
    st_relaxed_sys_u64(dst, tag | __float_as_uint(a));
    st_relaxed_sys_u64(dst + 1, tag | __float_as_uint(b));
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(p.bufs[p.rank] + kPeerLLOffset) +
                                    ((par + r) * kPeerLLSlots + slot) * 2;
    const unsigned long long t0 = global_ns();
    unsigned int spins = 0;
    unsigned long long va, vb;
    for (;;) {
      va = ld_relaxed_sys_u64(src);
      vb = ld_relaxed_sys_u64(src + 1);
      if ((unsigned int)(va >> 32) == seq && (unsigned int)(vb >> 32) == seq) break;
      if ((++spins & 0xfffu) == 0 && global_ns() - t0 > kPeerTimeoutNs) {
        reinterpret_cast<unsigned int*>(p.bufs[p.rank] + kPeerCtlOffset)[1] = 1u;
        break;
      }
    }
    sa[r] = __uint_as_float((unsigned int)va);
    sb[r] = __uint_as_float((unsigned int)vb);
  }
  __syncthreads();
  float2 out = make_float2(0.f, 0.f);
  for (int r = 0; r < p.world; ++r) { out.x += sa[r]; out.y += sb[r]; }
  __syncthreads();   // sa / sb may be rewritten by the next call
  return out;
}

}  // namespace ganb200
