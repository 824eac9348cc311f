// NVLS (NVLink SHARP) one-shot all-reduce: the sum over the ranks is computed INSIDE the NVSwitch.
//
// Every rank's bucket lives at the same offset of a symmetric allocation that is also mapped through a multicast object
// (one virtual address that stands for "this offset on every GPU of the group"). `multimem.ld_reduce` on that address
// makes the switch fetch the 16 bytes from all ranks, add them (fp32 accumulation) and return one result, so a rank
// receives bucket-bytes instead of world x bucket-bytes and does no adds itself; `multimem.st` writes one value to all
// ranks. The peer-load kernel of comm_fused.cu stays the default and the fallback (no multicast support, CPU tests).
// Reference: the aggregate step of S/utils/DistriParameterSynchronizer.scala:234-356 (fetch every peer's slice, sum, put
// the result back), S/parameters/AllReduceParameter.scala:236-286.
#include <cuda_bf16.h>
#include <stdint.h>

#include "bigdl_b200.h"
#include "host_common.h"

namespace bdl {

__device__ __forceinline__ void multimem_ld_reduce_bf16x8(const void* mc, uint32_t (&r)[4]) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "l"(mc)
               : "memory");
}
__device__ __forceinline__ void multimem_ld_reduce_f32x4(const void* mc, float (&r)[4]) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3])
               : "l"(mc)
               : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, const uint32_t (&r)[4]) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3])
               : "memory");
}

// out[i] = scale * sum_ranks bucket_r[i] for the whole bucket (one-shot: every rank reduces everything it needs)
__global__ void __launch_bounds__(256) nvls_reduce_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ mc,
                                                                      float* __restrict__ out, long long n8,
                                                                      float scale) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    uint32_t r[4];
    multimem_ld_reduce_bf16x8(mc + i * 8, r);
    float4 lo, hi;
    lo.x = __uint_as_float(r[0] << 16) * scale; lo.y = __uint_as_float(r[0] & 0xFFFF0000u) * scale;
    lo.z = __uint_as_float(r[1] << 16) * scale; lo.w = __uint_as_float(r[1] & 0xFFFF0000u) * scale;
    hi.x = __uint_as_float(r[2] << 16) * scale; hi.y = __uint_as_float(r[2] & 0xFFFF0000u) * scale;
    hi.z = __uint_as_float(r[3] << 16) * scale; hi.w = __uint_as_float(r[3] & 0xFFFF0000u) * scale;
    *reinterpret_cast<float4*>(out + i * 8) = lo;
    *reinterpret_cast<float4*>(out + i * 8 + 4) = hi;
  }
}

// two-shot, in place: rank r reduces slice r and multicasts the sum back into every rank's buffer
__global__ void __launch_bounds__(256) nvls_allreduce_bf16_inplace_kernel(__nv_bfloat16* __restrict__ mc, long long first8,
                                                                          long long count8) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count8; i += stride) {
    uint32_t r[4];
    multimem_ld_reduce_bf16x8(mc + (first8 + i) * 8, r);
    multimem_st_bf16x8(mc + (first8 + i) * 8, r);
  }
}

__global__ void __launch_bounds__(256) nvls_allreduce_f32_kernel(const float* __restrict__ mc, float* __restrict__ out,
                                                                 long long n4, float scale) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float r[4];
    multimem_ld_reduce_f32x4(mc + i * 4, r);
    *reinterpret_cast<float4*>(out + i * 4) = make_float4(r[0] * scale, r[1] * scale, r[2] * scale, r[3] * scale);
  }
}

static unsigned nvls_grid(long long work) {
  long long blocks = (work + 255) / 256;
  const long long cap = static_cast<long long>(bdl_num_sms()) * 4;
  if (blocks > cap) blocks = cap;
  return static_cast<unsigned>(blocks < 1 ? 1 : blocks);
}

}  // namespace bdl

// mc: MULTICAST address of the bucket (16-byte aligned), n elements (multiple of 8 for bf16, 4 for fp32). The callers
// bracket these kernels with a barrier over the group (all ranks have written / all ranks have finished reading).
extern "C" int bdl_nvls_reduce_bf16_to_f32(const void* mc, float* out, long long n, float scale, void* stream) {
  if (mc == nullptr || out == nullptr || n <= 0 || (n % 8) || (reinterpret_cast<uintptr_t>(mc) & 15u) ||
      (reinterpret_cast<uintptr_t>(out) & 15u))
    return bdl_fail("bdl_nvls_reduce_bf16_to_f32: n must be a positive multiple of 8 and pointers 16-byte aligned");
  bdl::nvls_reduce_bf16_to_f32_kernel<<<bdl::nvls_grid(n / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(mc), out, n / 8, scale);
  return bdl_check_launch("nvls_reduce_bf16_to_f32_kernel");
}

extern "C" int bdl_nvls_allreduce_bf16_inplace(void* mc, long long n, int rank, int world, void* stream) {
  if (mc == nullptr || n <= 0 || (n % 8) || world <= 0 || rank < 0 || rank >= world ||
      (reinterpret_cast<uintptr_t>(mc) & 15u))
    return bdl_fail("bdl_nvls_allreduce_bf16_inplace: bad arguments");
  const long long n8 = n / 8;
  const long long per = (n8 + world - 1) / world;
  const long long first = per * rank;
  long long count = n8 - first;
  if (count > per) count = per;
  if (count <= 0) return BDL_OK;
  bdl::nvls_allreduce_bf16_inplace_kernel<<<bdl::nvls_grid(count), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<__nv_bfloat16*>(mc), first, count);
  return bdl_check_launch("nvls_allreduce_bf16_inplace_kernel");
}

extern "C" int bdl_nvls_allreduce_f32(const void* mc, float* out, long long n, float scale, void* stream) {
  if (mc == nullptr || out == nullptr || n <= 0 || (n % 4) || (reinterpret_cast<uintptr_t>(mc) & 15u) ||
      (reinterpret_cast<uintptr_t>(out) & 15u))
    return bdl_fail("bdl_nvls_allreduce_f32: n must be a positive multiple of 4 and pointers 16-byte aligned");
  bdl::nvls_allreduce_f32_kernel<<<bdl::nvls_grid(n / 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const float*>(mc), out, n / 4, scale);
  return bdl_check_launch("nvls_allreduce_f32_kernel");
}
