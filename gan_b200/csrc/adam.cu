// adam.cu — fused multi-tensor Adam / AdamW (torch.optim semantics, fp32 state).
// Roofline: HBM.  Algorithmic bytes = 28 B per live parameter per step (read p,g,m,v; write p,m,v).
// One launch updates every tensor: the host supplies a table of (p,g,m,v,n) entries and a block table
// mapping each CTA to (tensor, chunk), so small tensors (G13_5 has 10 653 live ones) cost one CTA each
// instead of one launch each.
#include "common.cuh"

namespace ganb200 {

struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay, grad_scale;
  float step_size;         // lr / (1 - beta1^t)
  float inv_bc2_sqrt;      // 1 / sqrt(1 - beta2^t)
  int decoupled;
  const int64_t* step_dev; // when non-null the 1-based step count lives on the device (CUDA-graph replay)
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamHyper& h) {
  g *= h.grad_scale;
  if (h.weight_decay != 0.f) {
    if (h.decoupled) p *= (1.f - h.lr * h.weight_decay);
    else g = fmaf(h.weight_decay, p, g);
  }
  // torch: exp_avg.lerp_(grad, 1-beta1)
  float w = 1.f - h.beta1;
  m = (w < 0.5f) ? fmaf(w, g - m, m) : g - (g - m) * (1.f - w);
  // torch: exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
  v = fmaf((1.f - h.beta2) * g, g, v * h.beta2);
  float denom = sqrtf(v) * h.inv_bc2_sqrt + h.eps;
  p -= h.step_size * (m / denom);
}

__global__ void __launch_bounds__(256) multi_tensor_adam_kernel(const ganb200_adam_entry* __restrict__ tab,
                                                                const int2* __restrict__ block_tab, int64_t chunk,
                                                                AdamHyper h) {
  if (h.step_dev) {  // bias corrections from the device-side step counter, once per CTA
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
      double t = (double)(*h.step_dev);
      s_bc[0] = (float)((double)h.lr / (1.0 - pow((double)h.beta1, t)));
      s_bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)h.beta2, t)));
    }
    __syncthreads();
    h.step_size = s_bc[0];
    h.inv_bc2_sqrt = s_bc[1];
  }
  int2 bt = block_tab[blockIdx.x];
  ganb200_adam_entry e = tab[bt.x];
  int64_t lo = (int64_t)bt.y * chunk;
  int64_t hi = lo + chunk < e.n ? lo + chunk : e.n;
  float* p = e.p + lo; const float* g = e.g + lo; float* m = e.m + lo; float* v = e.v + lo;
  int64_t n = hi - lo;
  bool vec = ((((uintptr_t)p) | ((uintptr_t)g) | ((uintptr_t)m) | ((uintptr_t)v)) & 15) == 0;
  if (vec) {
    int64_t n4 = n >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
      float4 pv = reinterpret_cast<float4*>(p)[i];
      float4 gv = reinterpret_cast<const float4*>(g)[i];
      float4 mv = reinterpret_cast<float4*>(m)[i];
      float4 vv = reinterpret_cast<float4*>(v)[i];
      adam_one(pv.x, gv.x, mv.x, vv.x, h); adam_one(pv.y, gv.y, mv.y, vv.y, h);
      adam_one(pv.z, gv.z, mv.z, vv.z, h); adam_one(pv.w, gv.w, mv.w, vv.w, h);
      reinterpret_cast<float4*>(p)[i] = pv;
      reinterpret_cast<float4*>(m)[i] = mv;
      reinterpret_cast<float4*>(v)[i] = vv;
    }
    for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) adam_one(p[i], g[i], m[i], v[i], h);
  } else {
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) adam_one(p[i], g[i], m[i], v[i], h);
  }
}

// dst[i] = src[i] for every table entry (p = destination, g = source; m, v unused): gathers the gradient tensors of a
// data-parallel bucket into its flat all-reduce buffer in ONE launch (gan_b200/dist.py) instead of torch.cat + copy_.
__global__ void __launch_bounds__(256) multi_tensor_copy_kernel(const ganb200_adam_entry* __restrict__ tab,
                                                                const int2* __restrict__ block_tab, int64_t chunk) {
  const int2 bt = block_tab[blockIdx.x];
  const ganb200_adam_entry e = tab[bt.x];
  const int64_t lo = (int64_t)bt.y * chunk;
  const int64_t hi = lo + chunk < e.n ? lo + chunk : e.n;
  float* dst = e.p + lo;
  const float* src = e.g + lo;
  const int64_t n = hi - lo;
  if (((((uintptr_t)dst) | ((uintptr_t)src)) & 15) == 0) {
    const int64_t n4 = n >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x)
      reinterpret_cast<float4*>(dst)[i] = __ldg(reinterpret_cast<const float4*>(src) + i);
    for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  } else {
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  }
}

}  // namespace ganb200
using namespace ganb200;

extern "C" int ganb200_multi_tensor_copy(const ganb200_adam_entry* tab, const int32_t* block_tab, int n_blocks,
                                         int64_t chunk, void* stream) {
  if (n_blocks <= 0) return 0;
  GB_REQUIRE(chunk > 0 && (chunk % 4) == 0, GANB200_BAD_ARG, "multi_tensor_copy: chunk must be a positive multiple of 4");
  multi_tensor_copy_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(tab, reinterpret_cast<const int2*>(block_tab), chunk);
  GB_CHECK_LAUNCH();
  return 0;
}

// block_tab: DEVICE array of n_blocks int32 pairs (tensor index, chunk index); chunk = elements per CTA.
extern "C" int ganb200_multi_tensor_adam_blocks(const ganb200_adam_entry* tab, const int32_t* block_tab, int n_blocks,
                                                int64_t chunk, float lr, float beta1, float beta2, float eps,
                                                float weight_decay, int decoupled, int64_t step,
                                                const int64_t* step_dev, float grad_scale, void* stream) {
  if (n_blocks <= 0) return 0;
  GB_REQUIRE(chunk > 0 && (chunk % 4) == 0, GANB200_BAD_ARG, "adam: chunk must be a positive multiple of 4");
  AdamHyper h;
  h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.weight_decay = weight_decay; h.grad_scale = grad_scale;
  double bc1 = 1.0 - pow((double)beta1, (double)step);
  double bc2 = 1.0 - pow((double)beta2, (double)step);
  h.step_size = (float)((double)lr / bc1);
  h.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
  h.decoupled = decoupled;
  h.step_dev = step_dev;
  multi_tensor_adam_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(tab, reinterpret_cast<const int2*>(block_tab),
                                                                       chunk, h);
  GB_CHECK_LAUNCH();
  return 0;
}
