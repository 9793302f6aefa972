// peer.cu — small all-reduces of the data-parallel step over NVLink peer memory, fused with their arithmetic.
//
// Batch-sharded data parallelism has to exchange BatchNorm statistics in the middle of every forward and backward pass
// (gan_b200/dist.py; SURVEY §8(e): the reference's BatchNorm sees the whole batch): 2 575 exchanges of a few hundred
// floats per generator_13_5 + discriminator_9_4 iteration.  Each one as torch.stack -> ncclAllReduce -> 4 elementwise
// launches costs ~50 us of pure latency on the critical path (measured: 1 209 ms -> 1 337 ms per step from 1 to 8 GPUs).
// Here ONE single-CTA launch per exchange does all of it:
//
//   stage   the CTA writes its operands (for BatchNorm statistics: mean and var + mean^2) into its own exchange buffer,
//   signal  fence.sys, then one release-store of the call's sequence number into the flag word it owns in EVERY rank's
//           buffer (NVLink stores through cudaIpc peer mappings),
//   wait    spins (acquire loads, local memory) until every rank's flag word in its own buffer carries the sequence number,
//   reduce  reads the W staged vectors over NVLink, adds them in rank order (every rank gets the same bits, run to run),
//           applies the epilogue (mean over ranks; var = E[x^2] - mean^2) and writes the results.
//
// Buffer of one rank (ganb200_peer_bytes): [flags: one 128-byte line per source rank][control: sequence number, error
// word, done counter][LL slots of the per-block exchanges, peer.cuh][data slot 0][data slot 1].  Slots alternate with the sequence number: a rank can only reach call s+2 (which
// rewrites the slot of call s) after every rank has signalled call s+1, i.e. after every rank has finished reading call s.
// The sequence number lives on the device, so the launch is CUDA-graph capturable and replays keep counting.
// All ranks must issue the same exchanges in the same order on their step stream (they run the same program).
// A wait that exceeds ~20 s (a dead peer) sets the error word and returns, so that a failed rank cannot hang the box.
#include "peer.cuh"

namespace ganb200 {

struct PeerArgs {
  char* const* bufs;   // device array [world] of exchange buffers (own mapping at [rank])
  int rank, world;
  int64_t cap;         // floats per data slot
  const float* in0; const float* in1;
  float* out0; float* out1;
  int n;               // elements per vector
  int op;
};

__global__ void __launch_bounds__(1024) peer_allreduce_kernel(PeerArgs a) {
  char* own = a.bufs[a.rank];
  unsigned int* ctl = reinterpret_cast<unsigned int*>(own + kPeerCtlOffset);
  const unsigned int seq = ctl[0] + 1u;     // every thread reads it before thread 0 advances it (barriers below)
  const int nvec = a.in1 ? 2 : 1;
  const int total = a.n * nvec;
  float* mine = reinterpret_cast<float*>(own + kPeerDataOffset) + (size_t)(seq & 1u) * a.cap;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    float v;
    if (i < a.n) v = a.in0[i];
    else {
      v = a.in1[i - a.n];
      if (a.op == GANB200_PEER_BN_STATS) { const float m = a.in0[i - a.n]; v = fmaf(m, m, v); }
    }
    mine[i] = v;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < a.world) {
    const int r = threadIdx.x;
    st_release_sys(reinterpret_cast<unsigned int*>(a.bufs[r]) + a.rank * kPeerFlagStride, seq);
    const unsigned int* flag = reinterpret_cast<const unsigned int*>(own) + r * kPeerFlagStride;
    const unsigned long long t0 = global_ns();
    unsigned int spins = 0;
    while ((int)(ld_acquire_sys(flag) - seq) < 0) {
      if ((++spins & 0xfffu) == 0 && global_ns() - t0 > kPeerTimeoutNs) { ctl[1] = 1u; break; }
    }
  }
  __syncthreads();
  const float inv = 1.f / (float)a.world;
  const size_t slot = kPeerDataOffset + (size_t)(seq & 1u) * a.cap * sizeof(float);
  for (int i = threadIdx.x; i < a.n; i += blockDim.x) {
    float s0 = 0.f, s1 = 0.f;
    for (int r = 0; r < a.world; ++r) {
      const float* src = reinterpret_cast<const float*>(a.bufs[r] + slot);
      s0 += ld_relaxed_sys(src + i);
      if (nvec == 2) s1 += ld_relaxed_sys(src + a.n + i);
    }
    if (a.op == GANB200_PEER_SUM) {
      a.out0[i] = s0;
      if (nvec == 2) a.out1[i] = s1;
    } else if (a.op == GANB200_PEER_MEAN) {
      a.out0[i] = s0 * inv;
      if (nvec == 2) a.out1[i] = s1 * inv;
    } else {  // GANB200_PEER_BN_STATS: equal shards — global mean, global biased variance
      const float m = s0 * inv;
      a.out0[i] = m;
      a.out1[i] = fmaf(-m, m, s1 * inv);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) ctl[0] = seq;
}

}  // namespace ganb200
using namespace ganb200;

extern "C" int64_t ganb200_peer_bytes(int64_t cap_floats) {
  return (int64_t)kPeerDataOffset + 2 * cap_floats * (int64_t)sizeof(float);
}

extern "C" int ganb200_peer_alloc(int64_t cap_floats, void** buf) {
  GB_REQUIRE(buf != nullptr && cap_floats > 0, GANB200_BAD_ARG, "peer_alloc: bad arguments");
  void* p = nullptr;
  GB_CHECK_CUDA(cudaMalloc(&p, (size_t)ganb200_peer_bytes(cap_floats)));
  GB_CHECK_CUDA(cudaMemset(p, 0, (size_t)ganb200_peer_bytes(cap_floats)));
  GB_CHECK_CUDA(cudaDeviceSynchronize());
  *buf = p;
  return 0;
}

extern "C" int ganb200_peer_free(void* buf) {
  if (buf) GB_CHECK_CUDA(cudaFree(buf));
  return 0;
}

extern "C" int ganb200_peer_export(void* buf, unsigned char* handle64) {
  GB_REQUIRE(buf && handle64, GANB200_BAD_ARG, "peer_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  cudaIpcMemHandle_t h;
  GB_CHECK_CUDA(cudaIpcGetMemHandle(&h, buf));
  memcpy(handle64, &h, 64);
  return 0;
}

extern "C" int ganb200_peer_open(const unsigned char* handle64, void** buf) {
  GB_REQUIRE(buf && handle64, GANB200_BAD_ARG, "peer_open: null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  GB_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *buf = p;
  return 0;
}

extern "C" int ganb200_peer_close(void* buf) {
  if (buf) GB_CHECK_CUDA(cudaIpcCloseMemHandle(buf));
  return 0;
}

extern "C" int ganb200_peer_allreduce(const void* bufs_dev, int rank, int world, int64_t cap_floats, const float* in0,
                                      const float* in1, float* out0, float* out1, int64_t n, int op, void* stream) {
  GB_REQUIRE(bufs_dev && in0 && out0, GANB200_BAD_ARG, "peer_allreduce: null argument");
  GB_REQUIRE(world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, GANB200_BAD_ARG, "peer_allreduce: bad rank / world");
  GB_REQUIRE(op == GANB200_PEER_SUM || op == GANB200_PEER_MEAN || op == GANB200_PEER_BN_STATS, GANB200_BAD_ARG, "peer_allreduce: bad op");
  GB_REQUIRE((in1 == nullptr) == (out1 == nullptr), GANB200_BAD_ARG, "peer_allreduce: in1 and out1 go together");
  GB_REQUIRE(op != GANB200_PEER_BN_STATS || in1 != nullptr, GANB200_BAD_ARG, "peer_allreduce: BN_STATS needs mean and var");
  GB_REQUIRE(n > 0 && n * (in1 ? 2 : 1) <= cap_floats, GANB200_BAD_SHAPE, "peer_allreduce: vector does not fit the exchange slot");
  PeerArgs a;
  a.bufs = (char* const*)bufs_dev; a.rank = rank; a.world = world; a.cap = cap_floats;
  a.in0 = in0; a.in1 = in1; a.out0 = out0; a.out1 = out1; a.n = (int)n; a.op = op;
  int threads = 32;
  while (threads < n && threads < 1024) threads <<= 1;
  if (threads < world) threads = 32;
  peer_allreduce_kernel<<<1, threads, 0, (cudaStream_t)stream>>>(a);
  GB_CHECK_LAUNCH();
  return 0;
}

// error word of the own buffer (set by a wait that timed out); synchronises the device
extern "C" int ganb200_peer_error(const void* buf, int* error) {
  GB_REQUIRE(buf && error, GANB200_BAD_ARG, "peer_error: null argument");
  unsigned int e = 0;
  GB_CHECK_CUDA(cudaMemcpy(&e, (const char*)buf + kPeerCtlOffset + 4, 4, cudaMemcpyDeviceToHost));
  *error = (int)e;
  return 0;
}
