// Quantisation kernels for fp8 (e4m3) inference: per-tensor / per-row abs-max → scale, bf16|fp32 → e4m3 with saturation,
// and the reference's int8 formulas for compatibility (BigQuant symmetric per-row weights:
// q = round(w / max(|max_row|, |min_row|) * 127), S/nn/quantized/Quantization.scala:35-62).
//
// Reference: core/bigquant/native ops (FindMinMax, quantize, pad+shuffle im2col), S/nn/MklInt8Convertible.scala
// (calcScales: abs-max per tensor, mask 0, and per output channel for weights, mask 1).
#include <cuda_fp8.h>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

constexpr float kE4M3Max = 448.f;

// out[r] = max_k |x[r, k]|  (one warp per row)
template <typename T>
__global__ void row_absmax_kernel(const T* __restrict__ x, float* __restrict__ out, long long rows, long long cols,
                                  long long ld) {
  const long long row = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float m = 0.f;
  for (long long c = lane; c < cols; c += 32) m = fmaxf(m, fabsf(static_cast<float>(x[row * ld + c])));
  m = warp_max(m);
  if (lane == 0) out[row] = m;
}

// q = sat_e4m3(x * inv_scale[row or 0]); inv_scale = 448 / absmax. 8 elements per thread-iteration.
template <typename T>
__global__ void quantize_e4m3_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                     const float* __restrict__ absmax, int per_row, long long rows, long long cols,
                                     long long ldx, long long ldq) {
  const long long total = rows * cols;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = i / cols;
    const long long c = i - r * cols;
    const float am = absmax[per_row ? r : 0];
    const float inv = am > 0.f ? kE4M3Max / am : 0.f;
    const float v = static_cast<float>(x[r * ldx + c]) * inv;
    q[r * ldq + c] = static_cast<uint8_t>(__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3));
  }
}

// Vector form for bf16 rows whose length, pitches and base are multiples of 8 elements: 16-byte loads, 8-byte stores.
__global__ void quantize_e4m3_bf16x8_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q,
                                            const float* __restrict__ absmax, int per_row, uint32_t rows,
                                            uint32_t chunks_per_row, long long ldx, long long ldq) {
  const uint32_t total = rows * chunks_per_row;
  const uint32_t stride = gridDim.x * blockDim.x;
  const float am0 = per_row ? 0.f : __ldg(absmax);
  const float inv0 = am0 > 0.f ? kE4M3Max / am0 : 0.f;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const uint32_t r = i / chunks_per_row;
    const uint32_t c = (i - r * chunks_per_row) * 8u;
    float inv = inv0;
    if (per_row) {
      const float am = __ldg(absmax + r);
      inv = am > 0.f ? kE4M3Max / am : 0.f;
    }
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * ldx + c));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[2] = {0u, 0u};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float lo = __uint_as_float(w[j] << 16) * inv;
      const float hi = __uint_as_float(w[j] & 0xffff0000u) * inv;
      const uint32_t p = __nv_cvt_float2_to_fp8x2(make_float2(lo, hi), __NV_SATFINITE, __NV_E4M3);
      o[j >> 1] |= p << ((j & 1) * 16);
    }
    *reinterpret_cast<uint2*>(q + static_cast<long long>(r) * ldq + c) = make_uint2(o[0], o[1]);
  }
}

// col_scale[n] = act_absmax / 448 * w_absmax[n] / 448  (dequantisation factor of the fp8 GEMM epilogue)
__global__ void fp8_colscale_kernel(const float* __restrict__ act_absmax, const float* __restrict__ w_absmax,
                                    float* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (act_absmax[0] / kE4M3Max) * (w_absmax[i] / kE4M3Max);
}

// BigQuant-compatible symmetric int8 rows: q = round(w / absmax_row * 127)
__global__ void quantize_int8_rows_kernel(const float* __restrict__ w, int8_t* __restrict__ q,
                                          const float* __restrict__ absmax, long long rows, long long cols) {
  const long long total = rows * cols;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = i / cols;
    const float am = absmax[r];
    const float v = am > 0.f ? w[i] / am * 127.f : 0.f;
    q[i] = static_cast<int8_t>(fminf(fmaxf(rintf(v), -127.f), 127.f));
  }
}

static int qgrid(long long total) {
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  long long b = (total + 255) / 256;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}


// ------------------------------------------------------------------------------------------------ MX block scaling
// e4m3 values with one UE8M0 (power-of-two) scale per 32 consecutive K elements of every row — the operand format of
// tcgen05.mma.kind::mxf8f6f4.block_scale. This is the B200 form of BigQuant's "local" quantisation, which keeps one
// scale per im2col row window instead of one per tensor (core/bigquant/native/ops/shuffle/shuffle_im2col.h,
// S/nn/quantized/Quantization.scala:35-62): the scale follows the data block by block.
// Scale bytes are written in the layout the tensor core's scale-factor copy (tcgen05.cp 32x128b.warpx4) expects:
// 512-byte atoms of 128 rows x 4 k blocks, byte offset (r % 32) * 16 + (r % 128) / 32 * 4 + (kb % 4), atoms ordered
// [row block][k group]. One warp per (row, 32-element block).
template <typename T>
__global__ void quantize_mxfp8_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, uint8_t* __restrict__ sf,
                                      long long rows, long long cols, long long ldx, long long ldq, int kgroups) {
  // the warp grid covers the PADDED extent (rows to 128, k blocks to 4): absent rows / blocks get scale 1 (byte 127), so
  // no separate fill of the atom buffer is needed
  const long long kblocks = static_cast<long long>(kgroups) * 4;
  const long long rows_p = (rows + 127) / 128 * 128;
  const long long total = rows_p * kblocks;
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  if (w >= total) return;
  const int lane = threadIdx.x & 31;
  const long long r = w / kblocks;
  const long long kb = w - r * kblocks;
  const long long c = kb * 32 + lane;
  const bool live = r < rows && c < cols;
  const float v = live ? static_cast<float>(x[r * ldx + c]) : 0.f;
  const float am = warp_max(fabsf(v));
  // smallest power of two s with am / s <= 448 (e4m3 max); exponent biased by 127, 0 .. 254
  int e = 127;
  if (am > 0.f) {
    int ex;
    const float fr = frexpf(am / kE4M3Max, &ex);      // am / 448 = fr * 2^ex, fr in [0.5, 1)
    e = (fr == 0.5f ? ex - 1 : ex) + 127;
    e = e < 0 ? 0 : (e > 254 ? 254 : e);
  }
  const float inv = exp2f(static_cast<float>(127 - e));
  if (live) q[r * ldq + c] = static_cast<uint8_t>(__nv_cvt_float_to_fp8(v * inv, __NV_SATFINITE, __NV_E4M3));
  if (lane == 0) {
    const long long atom = (r >> 7) * kgroups + (kb >> 2);
    sf[atom * 512 + (r & 31) * 16 + ((r & 127) >> 5) * 4 + (kb & 3)] = static_cast<uint8_t>(e);
  }
}

__global__ void fill_u8_kernel(uint8_t* p, size_t n, uint8_t v) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
    p[i] = v;
}
}  // namespace bdl

using namespace bdl;

extern "C" int bdl_row_absmax(const void* x, float* out, long long rows, long long cols, long long ld, int is_bf16,
                              void* stream) {
  if (rows <= 0 || cols <= 0) return BDL_OK;
  const int grid = static_cast<int>((rows + 7) / 8);
  if (is_bf16)
    row_absmax_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x), out,
                                                                            rows, cols, ld);
  else
    row_absmax_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const float*>(x), out, rows,
                                                                            cols, ld);
  return bdl_check_launch("row_absmax_kernel");
}

extern "C" int bdl_quantize_e4m3(const void* x, void* q, const float* absmax, int per_row, long long rows,
                                 long long cols, long long ldx, long long ldq, int is_bf16, void* stream) {
  if (rows <= 0 || cols <= 0) return BDL_OK;
  if (is_bf16 && cols % 8 == 0 && ldx % 8 == 0 && ldq % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0 &&
      (reinterpret_cast<uintptr_t>(q) & 7u) == 0 && rows * (cols / 8) < (1ll << 31)) {
    quantize_e4m3_bf16x8_kernel<<<qgrid(rows * (cols / 8)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<uint8_t*>(q), absmax, per_row, static_cast<uint32_t>(rows),
        static_cast<uint32_t>(cols / 8), ldx, ldq);
    return bdl_check_launch("quantize_e4m3_bf16x8_kernel");
  }
  if (is_bf16)
    quantize_e4m3_kernel<<<qgrid(rows * cols), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<uint8_t*>(q), absmax, per_row, rows, cols, ldx, ldq);
  else
    quantize_e4m3_kernel<<<qgrid(rows * cols), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const float*>(x), static_cast<uint8_t*>(q), absmax, per_row, rows, cols, ldx, ldq);
  return bdl_check_launch("quantize_e4m3_kernel");
}

extern "C" int bdl_fp8_colscale(const float* act_absmax, const float* w_absmax, float* out, int n, void* stream) {
  fp8_colscale_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(act_absmax, w_absmax, out, n);
  return bdl_check_launch("fp8_colscale_kernel");
}

extern "C" int bdl_quantize_int8_rows(const float* w, void* q, const float* absmax, long long rows, long long cols,
                                      void* stream) {
  if (rows <= 0 || cols <= 0) return BDL_OK;
  quantize_int8_rows_kernel<<<qgrid(rows * cols), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w, static_cast<int8_t*>(q), absmax, rows, cols);
  return bdl_check_launch("quantize_int8_rows_kernel");
}

// q: [rows, ldq] e4m3 bytes (ldq >= cols, multiple of 16; the pad must already be zero); sf: ceil(rows/128) *
// ceil(cols/128) atoms of 512 bytes (see quantize_mxfp8_kernel), bytes of absent rows / k blocks set to 127 (scale 1).
extern "C" int bdl_quantize_mxfp8(const void* x, void* q, void* sf, long long rows, long long cols, long long ldx,
                                  long long ldq, int is_bf16, void* stream_) {
  if (rows <= 0 || cols <= 0) return BDL_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  const int kgroups = static_cast<int>((cols + 127) / 128);
  const long long warps = ((rows + 127) / 128 * 128) * (static_cast<long long>(kgroups) * 4);
  const long long blocks = (warps * 32 + 255) / 256;
  if (blocks >= (1ll << 31)) return bdl_fail("bdl_quantize_mxfp8: too large");
  if (is_bf16)
    bdl::quantize_mxfp8_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), static_cast<uint8_t*>(q),
                                                                           static_cast<uint8_t*>(sf), rows, cols, ldx, ldq, kgroups);
  else
    bdl::quantize_mxfp8_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(static_cast<const float*>(x), static_cast<uint8_t*>(q),
                                                                           static_cast<uint8_t*>(sf), rows, cols, ldx, ldq, kgroups);
  return bdl_check_launch("quantize_mxfp8_kernel");
}
