// Fused LogSoftMax + ClassNLLCriterion (= CrossEntropyCriterion) forward/backward, plus Top-1/Top-5 counters.
// One CTA per row: a single pass keeps the row in registers/shared memory, so logits are read once in forward and
// once in backward; the [B, C] log-prob matrix is never materialised.
//
// Semantics follow S/nn/ClassNLLCriterion.scala:70-242 and S/nn/CrossEntropyCriterion.scala:35-53:
//   * targets are 1-based class indices; target == padding_value rows contribute nothing,
//   * optional per-class weights, sizeAverage divides by the sum of the weights of the non-padded targets,
//   * loss = -sum_i w[t_i] * logp[i, t_i] / (sizeAverage ? sum_i w[t_i] : 1).
// Top-k counters mirror S/optim/ValidationMethod.scala:173 (Top1Accuracy) and :827 (Top5Accuracy).
#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

template <typename T>
__device__ __forceinline__ float ldf(const T* p, long long i);
template <>
__device__ __forceinline__ float ldf<float>(const float* p, long long i) { return p[i]; }
template <>
__device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p, long long i) { return __bfloat162float(p[i]); }

template <typename T>
__device__ __forceinline__ void stf(T* p, long long i, float v);
template <>
__device__ __forceinline__ void stf<float>(float* p, long long i, float v) { p[i] = v; }
template <>
__device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, long long i, float v) { p[i] = __float2bfloat16_rn(v); }

__device__ __forceinline__ float block_reduce(float v, float* sh, bool is_max) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = (lane < nw) ? sh[lane] : (is_max ? -INFINITY : 0.f);
  r = is_max ? warp_max(r) : warp_sum(r);
  return r;
}

// Forward: per-row (weighted nll, weight, logsumexp). Targets are float (BigDL tensors) or int64, 1-based.
template <typename T, typename TT>
__global__ void xent_fwd_kernel(const T* __restrict__ logits, const TT* __restrict__ target,
                                const float* __restrict__ class_w, float* __restrict__ row_loss,
                                float* __restrict__ row_w, float* __restrict__ row_lse, int C, long long ld,
                                int padding_value) {
  __shared__ float sh[32];
  const long long r = blockIdx.x;
  const T* row = logits + r * ld;
  float mx = -INFINITY;
  for (int c = threadIdx.x; c < C; c += blockDim.x) mx = fmaxf(mx, ldf(row, c));
  mx = block_reduce(mx, sh, true);
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s += __expf(ldf(row, c) - mx);
  s = block_reduce(s, sh, false);
  if (threadIdx.x == 0) {
    const float lse = mx + __logf(s);
    row_lse[r] = lse;
    const int t = static_cast<int>(target[r]);
    float w = 0.f, l = 0.f;
    if (t != padding_value && t >= 1 && t <= C) {
      w = class_w ? class_w[t - 1] : 1.f;
      l = -w * (ldf(row, t - 1) - lse);
    }
    row_loss[r] = l;
    row_w[r] = w;
  }
}

// Deterministic single-block reduction of the per-row terms: out[0] = loss, out[1] = total weight.
__global__ void xent_finalize_kernel(const float* __restrict__ row_loss, const float* __restrict__ row_w,
                                     float* __restrict__ out, long long B, int size_average) {
  __shared__ float sh[32];
  float l = 0.f, w = 0.f;
  for (long long i = threadIdx.x; i < B; i += blockDim.x) {
    l += row_loss[i];
    w += row_w[i];
  }
  l = block_reduce(l, sh, false);
  w = block_reduce(w, sh, false);
  if (threadIdx.x == 0) {
    out[1] = w;
    out[0] = size_average ? (w > 0.f ? l / w : 0.f) : l;
  }
}

// Backward: dlogits[r, c] = gscale * w_r / W * (softmax[r, c] - [c == t_r])
template <typename T, typename TT>
__global__ void xent_bwd_kernel(const T* __restrict__ logits, const TT* __restrict__ target,
                                const float* __restrict__ class_w, const float* __restrict__ row_lse,
                                const float* __restrict__ loss_and_w, T* __restrict__ dlogits, int C, long long ld,
                                long long ldd, int padding_value, int size_average, float gscale) {
  const long long r = blockIdx.x;
  const T* row = logits + r * ld;
  T* drow = dlogits + r * ldd;
  const int t = static_cast<int>(target[r]);
  float w = 0.f;
  if (t != padding_value && t >= 1 && t <= C) w = class_w ? class_w[t - 1] : 1.f;
  const float W = size_average ? loss_and_w[1] : 1.f;
  const float coef = (W > 0.f) ? gscale * w / W : 0.f;
  const float lse = row_lse[r];
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float p = __expf(ldf(row, c) - lse);
    stf(drow, c, coef * (p - (c == t - 1 ? 1.f : 0.f)));
  }
}

// counts[0] += #rows whose target is the arg-max; counts[1] += #rows whose target is within the top 5.
template <typename T, typename TT>
__global__ void topk_correct_kernel(const T* __restrict__ logits, const TT* __restrict__ target,
                                    unsigned int* __restrict__ counts, int C, long long ld) {
  __shared__ float sh[32];
  const long long r = blockIdx.x;
  const T* row = logits + r * ld;
  const int t = static_cast<int>(target[r]) - 1;
  if (t < 0 || t >= C) return;
  const float tv = ldf(row, t);
  float greater = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float v = ldf(row, c);
    // ties broken towards the lower index, like a stable arg-max
    greater += (v > tv || (v == tv && c < t)) ? 1.f : 0.f;
  }
  greater = block_reduce(greater, sh, false);
  if (threadIdx.x == 0) {
    if (greater < 0.5f) atomicAdd(counts, 1u);
    if (greater < 4.5f) atomicAdd(counts + 1, 1u);
  }
}

}  // namespace bdl

using namespace bdl;

#define BDL_XENT_DISPATCH(KERNEL, ...)                                                             \
  do {                                                                                             \
    if (logits_bf16 && target_i64)                                                                 \
      KERNEL<__nv_bfloat16, long long><<<grid, 256, 0, st>>>(__VA_ARGS__);                         \
    else if (logits_bf16)                                                                          \
      KERNEL<__nv_bfloat16, float><<<grid, 256, 0, st>>>(__VA_ARGS__);                             \
    else if (target_i64)                                                                           \
      KERNEL<float, long long><<<grid, 256, 0, st>>>(__VA_ARGS__);                                 \
    else                                                                                           \
      KERNEL<float, float><<<grid, 256, 0, st>>>(__VA_ARGS__);                                     \
  } while (0)

template <typename T>
static const T* cp(const void* p) { return static_cast<const T*>(p); }

// scratch: 3*B floats (row_loss | row_w | row_lse); out: 2 floats (loss, total weight)
extern "C" int bdl_xent_fwd(const void* logits, const void* target, const float* class_w, float* scratch, float* out,
                            long long B, int C, long long ld, int logits_bf16, int target_i64, int padding_value,
                            int size_average, void* stream) {
  if (B <= 0 || C <= 0) return bdl_fail("bdl_xent_fwd: empty input");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = static_cast<int>(B);
  float* rl = scratch;
  float* rw = scratch + B;
  float* lse = scratch + 2 * B;
  if (logits_bf16 && target_i64)
    xent_fwd_kernel<<<grid, 256, 0, st>>>(cp<__nv_bfloat16>(logits), cp<long long>(target), class_w, rl, rw, lse, C, ld, padding_value);
  else if (logits_bf16)
    xent_fwd_kernel<<<grid, 256, 0, st>>>(cp<__nv_bfloat16>(logits), cp<float>(target), class_w, rl, rw, lse, C, ld, padding_value);
  else if (target_i64)
    xent_fwd_kernel<<<grid, 256, 0, st>>>(cp<float>(logits), cp<long long>(target), class_w, rl, rw, lse, C, ld, padding_value);
  else
    xent_fwd_kernel<<<grid, 256, 0, st>>>(cp<float>(logits), cp<float>(target), class_w, rl, rw, lse, C, ld, padding_value);
  int rc = bdl_check_launch("xent_fwd_kernel");
  if (rc) return rc;
  xent_finalize_kernel<<<1, 1024, 0, st>>>(rl, rw, out, B, size_average);
  return bdl_check_launch("xent_finalize_kernel");
}

extern "C" int bdl_xent_bwd(const void* logits, const void* target, const float* class_w, const float* scratch,
                            const float* out, void* dlogits, long long B, int C, long long ld, long long ldd,
                            int logits_bf16, int target_i64, int padding_value, int size_average, float gscale,
                            void* stream) {
  if (B <= 0 || C <= 0) return bdl_fail("bdl_xent_bwd: empty input");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = static_cast<int>(B);
  const float* lse = scratch + 2 * B;
  if (logits_bf16 && target_i64)
    xent_bwd_kernel<<<grid, 256, 0, st>>>(cp<__nv_bfloat16>(logits), cp<long long>(target), class_w, lse, out, static_cast<__nv_bfloat16*>(dlogits), C, ld, ldd, padding_value, size_average, gscale);
  else if (logits_bf16)
    xent_bwd_kernel<<<grid, 256, 0, st>>>(cp<__nv_bfloat16>(logits), cp<float>(target), class_w, lse, out, static_cast<__nv_bfloat16*>(dlogits), C, ld, ldd, padding_value, size_average, gscale);
  else if (target_i64)
    xent_bwd_kernel<<<grid, 256, 0, st>>>(cp<float>(logits), cp<long long>(target), class_w, lse, out, static_cast<float*>(dlogits), C, ld, ldd, padding_value, size_average, gscale);
  else
    xent_bwd_kernel<<<grid, 256, 0, st>>>(cp<float>(logits), cp<float>(target), class_w, lse, out, static_cast<float*>(dlogits), C, ld, ldd, padding_value, size_average, gscale);
  return bdl_check_launch("xent_bwd_kernel");
}

// ------------------------------------------------------------------------------------------------ row softmax
// SoftMax / LogSoftMax over the last dimension of a [rows, C] matrix, one warp per row, fp32 math, bf16 or fp32 I/O
// (reference: S/nn/SoftMax.scala, S/nn/LogSoftMax.scala:37-120, S/nn/mkldnn/SoftMax.scala). Used by the layers, by
// Attention's score normalisation and (log form) wherever the loss cannot take the logits directly.
namespace bdl {
template <typename T> __device__ __forceinline__ float ld_as_f32(const T* p, long long i);
template <> __device__ __forceinline__ float ld_as_f32<float>(const float* p, long long i) { return p[i]; }
template <> __device__ __forceinline__ float ld_as_f32<__nv_bfloat16>(const __nv_bfloat16* p, long long i) {
  return __bfloat162float(p[i]);
}
template <typename T> __device__ __forceinline__ void st_from_f32(T* p, long long i, float v);
template <> __device__ __forceinline__ void st_from_f32<float>(float* p, long long i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void st_from_f32<__nv_bfloat16>(__nv_bfloat16* p, long long i, float v) {
  p[i] = __float2bfloat16_rn(v);
}

template <typename TI, typename TO>
__global__ void __launch_bounds__(256) softmax_fwd_kernel(const TI* __restrict__ x, TO* __restrict__ y, long long rows,
                                                          int C, long long ldx, long long ldy, int log_form) {
  const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const TI* xr = x + row * ldx;
  float m = -INFINITY;
  for (int c = lane; c < C; c += 32) m = fmaxf(m, ld_as_f32(xr, c));
  m = warp_max(m);
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += __expf(ld_as_f32(xr, c) - m);
  s = warp_sum(s);
  TO* yr = y + row * ldy;
  if (log_form) {
    const float ls = m + __logf(s);
    for (int c = lane; c < C; c += 32) st_from_f32(yr, c, ld_as_f32(xr, c) - ls);
  } else {
    const float inv = 1.f / s;
    for (int c = lane; c < C; c += 32) st_from_f32(yr, c, __expf(ld_as_f32(xr, c) - m) * inv);
  }
}

// softmax: dx = y * (dy - sum(dy * y));   log-softmax: dx = dy - exp(y) * sum(dy)
template <typename TY, typename TD>
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const TY* __restrict__ y, const TD* __restrict__ dy,
                                                          TD* __restrict__ dx, long long rows, int C, long long ldy,
                                                          long long ldd, long long ldx, int log_form) {
  const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const TY* yr = y + row * ldy;
  const TD* dr = dy + row * ldd;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += log_form ? ld_as_f32(dr, c) : ld_as_f32(dr, c) * ld_as_f32(yr, c);
  s = warp_sum(s);
  TD* xr = dx + row * ldx;
  for (int c = lane; c < C; c += 32) {
    const float yv = ld_as_f32(yr, c), dv = ld_as_f32(dr, c);
    st_from_f32(xr, c, log_form ? dv - __expf(yv) * s : yv * (dv - s));
  }
}
}  // namespace bdl

// dtype codes: 0 = fp32, 1 = bf16
extern "C" int bdl_softmax_fwd(const void* x, void* y, long long rows, int C, long long ldx, long long ldy, int in_bf16,
                               int out_bf16, int log_form, void* stream_) {
  if (rows <= 0 || C <= 0) return BDL_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  const int blocks = static_cast<int>((rows + 7) / 8);
  if (in_bf16 && out_bf16)
    bdl::softmax_fwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), rows, C, ldx, ldy, log_form);
  else if (in_bf16)
    bdl::softmax_fwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), static_cast<float*>(y), rows, C, ldx, ldy, log_form);
  else if (out_bf16)
    bdl::softmax_fwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const float*>(x), static_cast<__nv_bfloat16*>(y), rows, C, ldx, ldy, log_form);
  else
    bdl::softmax_fwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const float*>(x), static_cast<float*>(y), rows, C, ldx, ldy, log_form);
  return bdl_check_launch("softmax_fwd_kernel");
}

extern "C" int bdl_softmax_bwd(const void* y, const void* dy, void* dx, long long rows, int C, long long ldy, long long ldd,
                               long long ldx, int y_bf16, int d_bf16, int log_form, void* stream_) {
  if (rows <= 0 || C <= 0) return BDL_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  const int blocks = static_cast<int>((rows + 7) / 8);
  if (y_bf16 && d_bf16)
    bdl::softmax_bwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), rows, C, ldy, ldd, ldx, log_form);
  else if (y_bf16)
    bdl::softmax_bwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(y), static_cast<const float*>(dy), static_cast<float*>(dx), rows, C, ldy, ldd, ldx, log_form);
  else if (d_bf16)
    bdl::softmax_bwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const float*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), rows, C, ldy, ldd, ldx, log_form);
  else
    bdl::softmax_bwd_kernel<<<blocks, 256, 0, st>>>(static_cast<const float*>(y), static_cast<const float*>(dy), static_cast<float*>(dx), rows, C, ldy, ldd, ldx, log_form);
  return bdl_check_launch("softmax_bwd_kernel");
}

extern "C" int bdl_topk_correct(const void* logits, const void* target, void* counts, long long B, int C, long long ld,
                                int logits_bf16, int target_i64, void* stream) {
  if (B <= 0 || C <= 0) return BDL_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = static_cast<int>(B);
  unsigned int* cn = static_cast<unsigned int*>(counts);
  if (logits_bf16 && target_i64)
    topk_correct_kernel<<<grid, 256, 0, st>>>(cp<__nv_bfloat16>(logits), cp<long long>(target), cn, C, ld);
  else if (logits_bf16)
    topk_correct_kernel<<<grid, 256, 0, st>>>(cp<__nv_bfloat16>(logits), cp<float>(target), cn, C, ld);
  else if (target_i64)
    topk_correct_kernel<<<grid, 256, 0, st>>>(cp<float>(logits), cp<long long>(target), cn, C, ld);
  else
    topk_correct_kernel<<<grid, 256, 0, st>>>(cp<float>(logits), cp<float>(target), cn, C, ld);
  return bdl_check_launch("topk_correct_kernel");
}

// ------------------------------------------------------------------------------------------------------------------
// Allocation-free forward (no counterpart in the reference).
//
// Some hosts of this library are not allowed to obtain device memory at call time: a serving process whose whole
// HBM budget was handed to one arena at start-up, a real-time loop that cannot absorb the device-wide stall a
// cudaMalloc may cause, or a plug-in running inside another runtime that owns the allocator and gives a plain C caller
// no way to ask it for a buffer. bdl_xent_fwd() needs 3*B floats of scratch from its caller; for those hosts the
// variant below keeps that scratch inside the library image instead, in the fixed array right here. The price is a
// hard cap on the batch (BDL_XENT_NOALLOC_MAX_ROWS rows = 768 KiB, resident on every GPU that loads the library until
// the process exits) and no backward pass (the per-row log-sum-exp stays private to the library).
//
// The array name is only a host-side handle; the device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and cached. That look-up happens inside the first call made on a device,
// so make that first call before starting any stream capture. Nothing in the scratch has to start at zero: every slot
// the reduction reads was written earlier in the same call, so contents left by a previous call are harmless. All
// calls on a device share the one array — two calls that can overlap on the GPU (different streams) race; serialise
// them.
#include <mutex>

#define BDL_XENT_NOALLOC_MAX_ROWS 65536
static __device__ __align__(256) float bdl_xent_static_scratch[3 * BDL_XENT_NOALLOC_MAX_ROWS];

static int xent_static_scratch_ptr(float** out) {
  static std::mutex mu;
  static float* per_device[64] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return bdl_fail_cuda("bdl_xent_fwd_noalloc: cudaGetDevice", e);
  if (dev < 0 || dev >= 64) return bdl_fail("bdl_xent_fwd_noalloc: device ordinal outside [0, 64)");
  std::lock_guard<std::mutex> lock(mu);
  if (per_device[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), bdl_xent_static_scratch);
    if (e != cudaSuccess) return bdl_fail_cuda("bdl_xent_fwd_noalloc: cudaGetSymbolAddress", e);
    if (p == nullptr) return bdl_fail("bdl_xent_fwd_noalloc: cudaGetSymbolAddress returned a null address");
    per_device[dev] = p;
  }
  *out = per_device[dev];
  return BDL_OK;
}

extern "C" int bdl_xent_fwd_noalloc(const void* logits, const void* target, const float* class_w, float* out,
                                    long long B, int C, long long ld, int logits_bf16, int target_i64,
                                    int padding_value, int size_average, void* stream) {
  if (B > BDL_XENT_NOALLOC_MAX_ROWS)
    return bdl_fail("bdl_xent_fwd_noalloc: batch exceeds the built-in scratch (at most 65536 rows); "
                    "use bdl_xent_fwd with caller-provided scratch for larger batches");
  float* scratch = nullptr;
  const int rc = xent_static_scratch_ptr(&scratch);
  if (rc != BDL_OK) return rc;
  return bdl_xent_fwd(logits, target, class_w, scratch, out, B, C, ld, logits_bf16, target_i64, padding_value,
                      size_average, stream);
}
