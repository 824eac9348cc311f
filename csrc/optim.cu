// Stand-alone flat OptimMethod update: one pass over (weights, grads, state) that also emits the bf16 compute
// copy of the weights — the single-GPU form of the reference's "optimMethod.optimize on the slice"
// (S/optim/DistriOptimizer.scala:370-389) + "sendWeightPartition" compress (S/parameters/AllReduceParameter.scala:315-328).
// Also: sum-of-squares / abs-max reductions used by L2-norm clipping, LARS and calcScales
// (S/parameters/Util.scala:30-50, S/optim/LarsSGD.scala:289-335, S/nn/MklInt8Convertible.scala:142-172).
#include "bigdl_b200.h"
#include "host_common.h"
#include "optim_update.cuh"
#include "ptx.cuh"

namespace bdl {

__global__ void optim_update_kernel(int kind, float* __restrict__ w, const float* __restrict__ g,
                                    float* __restrict__ s0, float* __restrict__ s1, uint16_t* __restrict__ w16,
                                    size_t n, OptimHP hp, const OptimHP* __restrict__ hp_dev, int truncate,
                                    int zero_grad, float* __restrict__ g_mut) {
  if (hp_dev) hp = *hp_dev;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float a = s0 ? s0[i] : 0.f, b = s1 ? s1[i] : 0.f;
    const float nw = optim_dispatch(kind, w[i], g[i], &a, &b, hp);
    w[i] = nw;
    if (s0) s0[i] = a;
    if (s1) s1[i] = b;
    if (w16) w16[i] = f32_to_bf16_bits(nw, truncate);
    if (zero_grad) g_mut[i] = 0.f;
  }
}

// out[0] += sum x^2 (fp32 input). One atomic per block.
__global__ void sumsq_f32_kernel(const float* __restrict__ x, size_t n, float* __restrict__ out) {
  __shared__ float sh[32];
  float s = 0.f;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) s += x[i] * x[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) atomicAdd(out, t);
  }
}

template <typename T>
__global__ void absmax_kernel(const T* __restrict__ x, size_t n, float* __restrict__ out) {
  __shared__ float sh[32];
  float m = 0.f;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    m = fmaxf(m, fabsf(static_cast<float>(x[i])));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    t = warp_max(t);
    // non-negative floats order like their bit patterns
    if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(out), __float_as_int(t));
  }
}

__global__ void scale_f32_kernel(float* __restrict__ x, size_t n, float alpha, const float* __restrict__ norm_sq,
                                 float max_norm) {
  // alpha is used directly unless norm_sq is given, in which case x *= min(1, max_norm / sqrt(*norm_sq))
  float a = alpha;
  if (norm_sq) {
    const float nrm = sqrtf(*norm_sq);
    a = nrm > max_norm ? max_norm / nrm : 1.f;
  }
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) x[i] *= a;
}

}  // namespace bdl

using namespace bdl;

static int grid_n(size_t n) {
  const long long cap = static_cast<long long>(bdl_num_sms()) * 8;
  long long b = static_cast<long long>((n + 255) / 256);
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

// hp: 14 floats laid out like OptimHP (ints bit-cast) — see bigdl_b200/ops/optim.py::pack_hp
extern "C" int bdl_optim_update(int kind, float* w, float* g, float* s0, float* s1, void* w16, size_t n,
                                const void* hp_host, const void* hp_dev, int truncate, int zero_grad, void* stream) {
  if (n == 0) return BDL_OK;
  if (kind < 0 || kind > OPT_LARS) return bdl_fail("bdl_optim_update: unknown OptimMethod kind");
  OptimHP hp{};
  if (hp_host) hp = *static_cast<const OptimHP*>(hp_host);
  optim_update_kernel<<<grid_n(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      kind, w, g, s0, s1, static_cast<uint16_t*>(w16), n, hp, static_cast<const OptimHP*>(hp_dev), truncate, zero_grad,
      g);
  return bdl_check_launch("optim_update_kernel");
}

extern "C" int bdl_optim_hp_size() { return static_cast<int>(sizeof(OptimHP)); }

extern "C" int bdl_sumsq_f32(const float* x, size_t n, float* out, void* stream) {
  if (n == 0) return BDL_OK;
  sumsq_f32_kernel<<<grid_n(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, out);
  return bdl_check_launch("sumsq_f32_kernel");
}

extern "C" int bdl_absmax(const void* x, size_t n, int is_bf16, float* out, void* stream) {
  if (n == 0) return BDL_OK;
  if (is_bf16)
    absmax_kernel<<<grid_n(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x), n, out);
  else
    absmax_kernel<<<grid_n(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const float*>(x), n, out);
  return bdl_check_launch("absmax_kernel");
}

extern "C" int bdl_scale_f32(float* x, size_t n, float alpha, const float* norm_sq, float max_norm, void* stream) {
  if (n == 0) return BDL_OK;
  scale_f32_kernel<<<grid_n(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, alpha, norm_sq, max_norm);
  return bdl_check_launch("scale_f32_kernel");
}
