// Memory-bound elementwise / reduction kernels on bf16 activations: casts (incl. the reference's truncating
// "fp16" codec), ReLU fwd/bwd (+ fused bias-gradient column sums), residual add, dropout (Philox), table ops.
// All kernels are grid-stride over 16-byte vectors with a scalar tail, sized as a multiple of the SM count.
//
// Reference counterparts: S/nn/ReLU.scala, S/nn/Threshold.scala, S/nn/Dropout.scala:40-75, S/nn/CAddTable.scala,
// S/parameters/FP16CompressedTensor.scala:173-196,271-278 (truncate), :80-98 (expand), :152-171 (add).
#include <cstdlib>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

static inline int ew_grid(size_t nvec, int threads = 256) {
  const long long sms = bdl_num_sms();
  long long blocks = static_cast<long long>((nvec + threads - 1) / threads);
  const long long cap = sms * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

struct alignas(16) bf16x8 {
  __nv_bfloat162 v[4];
};

__device__ __forceinline__ void unpack8(const bf16x8& p, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(p.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ bf16x8 pack8(const float (&f)[8]) {
  bf16x8 p;
#pragma unroll
  for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return p;
}

// x[lane] = y[lane] > 0 ? x[lane] : +0 on packed bf16x2 (HSET2 mask + AND)
__device__ __forceinline__ void relu_mask8(bf16x8& x, const bf16x8& y) {
  const __nv_bfloat162 z = __floats2bfloat162_rn(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t m = __hgt2_mask(y.v[i], z);
    uint32_t b = *reinterpret_cast<uint32_t*>(&x.v[i]) & m;
    x.v[i] = *reinterpret_cast<__nv_bfloat162*>(&b);
  }
}

// ------------------------------------------------------------------ casts
template <bool kTruncate>
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, uint16_t* __restrict__ dst, size_t n) {
  const size_t nvec = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
    uint4 o;
    if (kTruncate) {
      o.x = (__float_as_uint(a.x) >> 16) | (__float_as_uint(a.y) & 0xFFFF0000u);
      o.y = (__float_as_uint(a.z) >> 16) | (__float_as_uint(a.w) & 0xFFFF0000u);
      o.z = (__float_as_uint(b.x) >> 16) | (__float_as_uint(b.y) & 0xFFFF0000u);
      o.w = (__float_as_uint(b.z) >> 16) | (__float_as_uint(b.w) & 0xFFFF0000u);
    } else {
      o.x = pack_bf16x2(a.x, a.y);
      o.y = pack_bf16x2(a.z, a.w);
      o.z = pack_bf16x2(b.x, b.y);
      o.w = pack_bf16x2(b.z, b.w);
    }
    reinterpret_cast<uint4*>(dst)[i] = o;
  }
  for (size_t i = nvec * 8 + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (kTruncate)
      dst[i] = static_cast<uint16_t>(__float_as_uint(src[i]) >> 16);
    else {
      __nv_bfloat16 h = __float2bfloat16_rn(src[i]);
      dst[i] = *reinterpret_cast<uint16_t*>(&h);
    }
  }
}

__global__ void cast_bf16_f32_kernel(const uint16_t* __restrict__ src, float* __restrict__ dst, size_t n) {
  const size_t nvec = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 v = reinterpret_cast<const uint4*>(src)[i];
    float4 a, b;
    a.x = __uint_as_float(v.x << 16);
    a.y = __uint_as_float(v.x & 0xFFFF0000u);
    a.z = __uint_as_float(v.y << 16);
    a.w = __uint_as_float(v.y & 0xFFFF0000u);
    b.x = __uint_as_float(v.z << 16);
    b.y = __uint_as_float(v.z & 0xFFFF0000u);
    b.z = __uint_as_float(v.w << 16);
    b.w = __uint_as_float(v.w & 0xFFFF0000u);
    reinterpret_cast<float4*>(dst)[2 * i] = a;
    reinterpret_cast<float4*>(dst)[2 * i + 1] = b;
  }
  for (size_t i = nvec * 8 + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = bf16_bits_to_float(src[i]);
}

// out = trunc(expand(a) + expand(b)) — the reference's compressed add, used by the bit-exact parity mode.
__global__ void bf16_trunc_add_kernel(const uint16_t* __restrict__ a, const uint16_t* __restrict__ b,
                                      uint16_t* __restrict__ out, size_t n) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float s = bf16_bits_to_float(a[i]) + bf16_bits_to_float(b[i]);
    out[i] = static_cast<uint16_t>(__float_as_uint(s) >> 16);
  }
}

// ------------------------------------------------------------------ unary / binary on bf16
enum EwOp : int {
  EW_RELU = 0,       // y = max(x,0)
  EW_RELU_BWD = 1,   // out = b > 0 ? a : 0          (a = dy, b = y)
  EW_ADD = 2,        // out = a + b
  EW_ADD_RELU = 3,   // out = max(a + b, 0)
  EW_MUL = 4,        // out = a * b
  EW_TANH = 5,
  EW_SIGMOID = 6,
  EW_TANH_BWD = 7,     // out = a * (1 - b^2)         (a = dy, b = y)
  EW_SIGMOID_BWD = 8,  // out = a * b * (1 - b)
  EW_SCALE = 9,        // out = a * alpha
  EW_AXPY = 10,        // out = a * alpha + b
  EW_SUB = 11,
  EW_MAX = 12,
  EW_RELU6 = 13,
};

template <int kOp>
__device__ __forceinline__ float ew_apply(float a, float b, float alpha) {
  switch (kOp) {
    case EW_RELU: return fmaxf(a, 0.f);
    case EW_RELU_BWD: return b > 0.f ? a : 0.f;
    case EW_ADD: return a + b;
    case EW_ADD_RELU: return fmaxf(a + b, 0.f);
    case EW_MUL: return a * b;
    case EW_TANH: return tanhf(a);
    case EW_SIGMOID: return 1.f / (1.f + __expf(-a));
    case EW_TANH_BWD: return a * (1.f - b * b);
    case EW_SIGMOID_BWD: return a * b * (1.f - b);
    case EW_SCALE: return a * alpha;
    case EW_AXPY: return a * alpha + b;
    case EW_SUB: return a - b;
    case EW_MAX: return fmaxf(a, b);
    case EW_RELU6: return fminf(fmaxf(a, 0.f), 6.f);
  }
  return a;
}

template <int kOp, bool kBinary>
__global__ void ew_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                               __nv_bfloat16* __restrict__ out, size_t n, float alpha) {
  const size_t nvec = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float fa[8], fb[8], fo[8];
    unpack8(reinterpret_cast<const bf16x8*>(a)[i], fa);
    if (kBinary) unpack8(reinterpret_cast<const bf16x8*>(b)[i], fb);
#pragma unroll
    for (int j = 0; j < 8; ++j) fo[j] = ew_apply<kOp>(fa[j], kBinary ? fb[j] : 0.f, alpha);
    reinterpret_cast<bf16x8*>(out)[i] = pack8(fo);
  }
  for (size_t i = nvec * 8 + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float fa = __bfloat162float(a[i]);
    const float fb = kBinary ? __bfloat162float(b[i]) : 0.f;
    out[i] = __float2bfloat16_rn(ew_apply<kOp>(fa, fb, alpha));
  }
}

// ------------------------------------------------------------------ column sums (bias gradient)
// x is [M, C] with row pitch ld (elements); out[c] (+)= sum_m f(x[m,c]). Optional fused ReLU backward:
// when y != nullptr, z = (y > 0 ? x : 0) is written to dz and summed instead of x.
// Block = 256 threads arranged as (C/8 vector columns) x (rows in flight); one atomicAdd per channel per block.
template <bool kReluBwd>
__global__ void colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                   __nv_bfloat16* __restrict__ dz, float* __restrict__ out, long long M, int C,
                                   long long ldx, long long ldy, long long ldz, int rows_per_block) {
  extern __shared__ float red[];  // [rows_in_flight][cvec*8]
  const int cvec = C / 8;
  const int vcols = cvec < static_cast<int>(blockDim.x) ? cvec : blockDim.x;  // vector columns handled concurrently
  const int rif = blockDim.x / vcols;                                       // rows in flight
  const int tc = threadIdx.x % vcols;
  const int tr = threadIdx.x / vcols;
  const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > M) r1 = M;
  for (int cv = tc; cv < cvec; cv += vcols) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (tr < rif) {
      long long r = r0 + tr;
      // kU independent rows per iteration: 2*kU 16-byte loads in flight per thread keep HBM busy
      constexpr int kU = 8;
      for (; r + (kU - 1ll) * rif < r1; r += static_cast<long long>(kU) * rif) {
        bf16x8 vx[kU], vy[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) vx[u] = *reinterpret_cast<const bf16x8*>(x + (r + static_cast<long long>(u) * rif) * ldx + cv * 8);
        if (kReluBwd) {
#pragma unroll
          for (int u = 0; u < kU; ++u) vy[u] = *reinterpret_cast<const bf16x8*>(y + (r + static_cast<long long>(u) * rif) * ldy + cv * 8);
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          float fx[8];
          if (kReluBwd) {
            relu_mask8(vx[u], vy[u]);     // packed: x &= (y > 0) lane mask, no unpack / repack of dz
            *reinterpret_cast<bf16x8*>(dz + (r + static_cast<long long>(u) * rif) * ldz + cv * 8) = vx[u];
          }
          unpack8(vx[u], fx);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += fx[j];
        }
      }
      for (; r < r1; r += rif) {
        float fx[8];
        bf16x8 vx1 = *reinterpret_cast<const bf16x8*>(x + r * ldx + cv * 8);
        if (kReluBwd) {
          relu_mask8(vx1, *reinterpret_cast<const bf16x8*>(y + r * ldy + cv * 8));
          *reinterpret_cast<bf16x8*>(dz + r * ldz + cv * 8) = vx1;
        }
        unpack8(vx1, fx);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += fx[j];
      }
    }
    // reduce over tr through shared memory
    __syncthreads();
    if (tr < rif) {
#pragma unroll
      for (int j = 0; j < 8; ++j) red[(tr * vcols + tc) * 8 + j] = acc[j];
    }
    __syncthreads();
    if (tr == 0) {
      float s8[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float s = 0.f;
        for (int t = 0; t < rif; ++t) s += red[(t * vcols + tc) * 8 + j];
        s8[j] = s;
      }
      // two 16-byte reductions instead of eight scalar atomics: 4x fewer L2 atomic transactions per block
      float* dst = out + cv * 8;
      if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(s8[0]), "f"(s8[1]), "f"(s8[2]),
                     "f"(s8[3]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4), "f"(s8[4]), "f"(s8[5]), "f"(s8[6]),
                     "f"(s8[7]) : "memory");
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(dst + j, s8[j]);
      }
    }
  }
}

// scalar fallback for C % 8 != 0 (or unaligned pitches)
template <bool kReluBwd>
__global__ void colsum_bf16_scalar_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                          __nv_bfloat16* __restrict__ dz, float* __restrict__ out, long long M, int C,
                                          long long ldx, long long ldy, long long ldz, int rows_per_block) {
  const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > M) r1 = M;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = 0.f;
    for (long long r = r0; r < r1; ++r) {
      float v = __bfloat162float(x[r * ldx + c]);
      if (kReluBwd) {
        v = __bfloat162float(y[r * ldy + c]) > 0.f ? v : 0.f;
        dz[r * ldz + c] = __float2bfloat16_rn(v);
      }
      acc += v;
    }
    atomicAdd(out + c, acc);
  }
}

// ------------------------------------------------------------------ dropout (Philox4x32-10)
__device__ __forceinline__ uint4 philox4x32(uint2 key, uint4 ctr) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// y = x * keep / (1-p); the keep mask is packed one byte per 8 elements for the backward pass.
__global__ void dropout_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                   uint8_t* __restrict__ mask, size_t n, float p, float scale,
                                   unsigned long long seed, unsigned long long offset,
                                   const unsigned long long* __restrict__ offset_dev) {
  if (offset_dev) offset += *offset_dev;   // device-resident stream position (CUDA-graph replays get fresh masks)
  const size_t nvec = (n + 7) / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const uint32_t thresh = static_cast<uint32_t>(fminf(p, 1.f) * 16777216.f);  // compare on 24 bits
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const unsigned long long c = offset + i;
    const uint4 r0 = philox4x32(make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)),
                                make_uint4(static_cast<uint32_t>(c), static_cast<uint32_t>(c >> 32), 0u, 0u));
    const uint4 r1 = philox4x32(make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)),
                                make_uint4(static_cast<uint32_t>(c), static_cast<uint32_t>(c >> 32), 1u, 0u));
    const uint32_t rnd[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    uint8_t m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) m |= static_cast<uint8_t>(((rnd[j] >> 8) >= thresh) ? (1u << j) : 0u);
    mask[i] = m;
    if (i * 8 + 8 <= n) {
      float f[8];
      unpack8(reinterpret_cast<const bf16x8*>(x)[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (m >> j) & 1 ? f[j] * scale : 0.f;
      reinterpret_cast<bf16x8*>(y)[i] = pack8(f);
    } else {
      for (size_t e = i * 8; e < n; ++e)
        y[e] = __float2bfloat16_rn(((m >> (e - i * 8)) & 1) ? __bfloat162float(x[e]) * scale : 0.f);
    }
  }
}

__global__ void dropout_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ mask,
                                   __nv_bfloat16* __restrict__ dx, size_t n, float scale) {
  const size_t nvec = (n + 7) / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint8_t m = mask[i];
    if (i * 8 + 8 <= n) {
      float f[8];
      unpack8(reinterpret_cast<const bf16x8*>(dy)[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (m >> j) & 1 ? f[j] * scale : 0.f;
      reinterpret_cast<bf16x8*>(dx)[i] = pack8(f);
    } else {
      for (size_t e = i * 8; e < n; ++e)
        dx[e] = __float2bfloat16_rn(((m >> (e - i * 8)) & 1) ? __bfloat162float(dy[e]) * scale : 0.f);
    }
  }
}

// ------------------------------------------------------------------ strided 2-D copy (concat / narrow on NHWC)
// dst[r, 0:C] = src[r, 0:C] with independent row pitches; C % 8 == 0 fast path.
__global__ void copy2d_bf16_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                   long long rows, int C, long long lds, long long ldd, int accumulate) {
  const int cvec = C / 8;
  const size_t total = static_cast<size_t>(rows) * cvec;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = i / cvec;
    const int cv = static_cast<int>(i - r * cvec);
    bf16x8 v = *reinterpret_cast<const bf16x8*>(src + r * lds + cv * 8);
    if (accumulate) {
      float a[8], b[8];
      unpack8(v, a);
      unpack8(*reinterpret_cast<const bf16x8*>(dst + r * ldd + cv * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += b[j];
      v = pack8(a);
    }
    *reinterpret_cast<bf16x8*>(dst + r * ldd + cv * 8) = v;
  }
}

__global__ void counter_add_kernel(unsigned long long* c, unsigned long long v) { *c += v; }

}  // namespace bdl

using namespace bdl;

// *counter += v, stream ordered (device-side step counters for graph-captured training steps)
// Finish of a split-K forward GEMM: y = act(acc + bias) as bf16, and the fp32 accumulator is re-zeroed in the same pass so
// the next launch can red.add into it again without a memset (Linear forward with few rows: the K split is what fills
// the SMs, S/nn/Linear.scala:78-104).
__global__ void bias_act_finish_kernel(float* __restrict__ acc, const float* __restrict__ bias, __nv_bfloat16* __restrict__ y,
                                       long long M, int N, long long ld_acc, long long ldy, int relu) {
  const int nv = N / 4;
  const long long total = M * nv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / nv;
    const int c = static_cast<int>(i - r * nv) * 4;
    float4* ap = reinterpret_cast<float4*>(acc + r * ld_acc + c);
    float4 v = *ap;
    *ap = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias) {
      const float4 b = *reinterpret_cast<const float4*>(bias + c);
      v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
    }
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo);
    o.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(y + r * ldy + c) = o;
  }
}

extern "C" int bdl_bias_act_finish(float* acc, const float* bias, void* y, long long M, int N, long long ld_acc,
                                   long long ldy, int relu, void* stream) {
  if (M <= 0 || N <= 0) return BDL_OK;
  if ((N % 4) || (ld_acc % 4) || (ldy % 4) || (reinterpret_cast<uintptr_t>(acc) & 15u) || (reinterpret_cast<uintptr_t>(y) & 7u) ||
      (bias && (reinterpret_cast<uintptr_t>(bias) & 15u)))
    return bdl_fail("bdl_bias_act_finish: N and the pitches must be multiples of 4, pointers vector-aligned");
  const long long total = M * (N / 4);
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(bdl_num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  bias_act_finish_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      acc, bias, static_cast<__nv_bfloat16*>(y), M, N, ld_acc, ldy, relu);
  return bdl_check_launch("bias_act_finish_kernel");
}

// Bias gradients of the branch tails of a channel Concat: the fused ReLU-mask + column-sum pass leaves ONE fp32 vector
// [Ctot]; this adds each tail's segment into that tail's bias-gradient slot of the flat gradient buffer and clears the
// vector again, so it can be a persistent buffer (no fill kernel per step) and the tails issue no add of their own.
struct ScatterSegs {
  float* dst[8];
  int off[8];
  int len[8];
  int count;
};
__global__ void scatter_acc_clear_kernel(float* __restrict__ src, const ScatterSegs sg) {
  for (int k = 0; k < sg.count; ++k)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < sg.len[k]; i += gridDim.x * blockDim.x) {
      const float v = src[sg.off[k] + i];
      src[sg.off[k] + i] = 0.f;
      sg.dst[k][i] += v;
    }
}
extern "C" int bdl_scatter_acc_clear_f32(float* src, void* const* dst, const int* off, const int* len, int count,
                                         void* stream) {
  if (count < 1 || count > 8) return bdl_fail("bdl_scatter_acc_clear_f32: 1..8 segments");
  ScatterSegs sg{};
  sg.count = count;
  int longest = 0;
  for (int k = 0; k < count; ++k) {
    sg.dst[k] = static_cast<float*>(dst[k]);
    sg.off[k] = off[k];
    sg.len[k] = len[k];
    if (len[k] > longest) longest = len[k];
  }
  const int blocks = longest > 1024 ? 4 : 1;
  scatter_acc_clear_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(src, sg);
  return bdl_check_launch("scatter_acc_clear_kernel");
}

extern "C" int bdl_counter_add(void* counter_dev, unsigned long long v, void* stream) {
  counter_add_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<unsigned long long*>(counter_dev), v);
  return bdl_check_launch("counter_add_kernel");
}

extern "C" int bdl_cast_f32_to_bf16(const float* src, void* dst, size_t n, int truncate, void* stream) {
  if (n == 0) return BDL_OK;
  if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u))
    return bdl_fail("bdl_cast_f32_to_bf16: pointers must be 16-byte aligned");
  const int g = ew_grid(n / 8 + 1);
  if (truncate)
    cast_f32_bf16_kernel<true><<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(src, static_cast<uint16_t*>(dst), n);
  else
    cast_f32_bf16_kernel<false><<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(src, static_cast<uint16_t*>(dst), n);
  return bdl_check_launch("cast_f32_bf16_kernel");
}

extern "C" int bdl_cast_bf16_to_f32(const void* src, float* dst, size_t n, void* stream) {
  if (n == 0) return BDL_OK;
  if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u))
    return bdl_fail("bdl_cast_bf16_to_f32: pointers must be 16-byte aligned");
  cast_bf16_f32_kernel<<<ew_grid(n / 8 + 1), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint16_t*>(src), dst, n);
  return bdl_check_launch("cast_bf16_f32_kernel");
}

extern "C" int bdl_bf16_trunc_add(const void* a, const void* b, void* out, size_t n, void* stream) {
  if (n == 0) return BDL_OK;
  bf16_trunc_add_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint16_t*>(a), static_cast<const uint16_t*>(b), static_cast<uint16_t*>(out), n);
  return bdl_check_launch("bf16_trunc_add_kernel");
}

#define BDL_EW_CASE(OP, BIN)                                                                             \
  case OP:                                                                                               \
    ew_bf16_kernel<OP, BIN><<<g, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(a),                     \
                                               static_cast<const __nv_bfloat16*>(b),                     \
                                               static_cast<__nv_bfloat16*>(out), n, alpha);              \
    break;

extern "C" int bdl_eltwise_bf16(int op, const void* a, const void* b, void* out, size_t n, float alpha, void* stream) {
  if (n == 0) return BDL_OK;
  if ((reinterpret_cast<uintptr_t>(a) & 15u) || (reinterpret_cast<uintptr_t>(out) & 15u) ||
      (b && (reinterpret_cast<uintptr_t>(b) & 15u)))
    return bdl_fail("bdl_eltwise_bf16: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int g = ew_grid(n / 8 + 1);
  switch (op) {
    BDL_EW_CASE(EW_RELU, false)
    BDL_EW_CASE(EW_RELU_BWD, true)
    BDL_EW_CASE(EW_ADD, true)
    BDL_EW_CASE(EW_ADD_RELU, true)
    BDL_EW_CASE(EW_MUL, true)
    BDL_EW_CASE(EW_TANH, false)
    BDL_EW_CASE(EW_SIGMOID, false)
    BDL_EW_CASE(EW_TANH_BWD, true)
    BDL_EW_CASE(EW_SIGMOID_BWD, true)
    BDL_EW_CASE(EW_SCALE, false)
    BDL_EW_CASE(EW_AXPY, true)
    BDL_EW_CASE(EW_SUB, true)
    BDL_EW_CASE(EW_MAX, true)
    BDL_EW_CASE(EW_RELU6, false)
    default:
      return bdl_fail("bdl_eltwise_bf16: unknown op");
  }
  return bdl_check_launch("ew_bf16_kernel");
}

// out = sum of up to 8 dense bf16 tensors (fp32 accumulate): the gradient of an activation that fans out to several
// branches (an Inception module input feeds four) in one pass instead of a chain of binary adds.
struct AddNPtrs {
  const __nv_bfloat16* p[8];
};
template <int kN>
__global__ void addn_bf16_kernel(const AddNPtrs in, __nv_bfloat16* __restrict__ out, size_t n8) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    bf16x8 v[kN];
#pragma unroll
    for (int t = 0; t < kN; ++t) v[t] = reinterpret_cast<const bf16x8*>(in.p[t])[i];
    float acc[8];
    unpack8(v[0], acc);
#pragma unroll
    for (int t = 1; t < kN; ++t) {
      float f[8];
      unpack8(v[t], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
    reinterpret_cast<bf16x8*>(out)[i] = pack8(acc);
  }
}

// Fan-out gradient sum + ReLU backward + bias gradient in one pass: z = (y > 0) ? sum_t in_t : 0 is written to dz and
// column-summed into out. All tensors are dense [M, C] (C % 8 == 0). Same thread layout as colsum_bf16_kernel.
template <int kN>
__global__ void colsum_sumn_relu_kernel(const AddNPtrs in, const __nv_bfloat16* __restrict__ y,
                                        __nv_bfloat16* __restrict__ dz, float* __restrict__ out, long long M, int C,
                                        int rows_per_block) {
  extern __shared__ float red[];
  const int cvec = C / 8;
  const int vcols = cvec < static_cast<int>(blockDim.x) ? cvec : blockDim.x;
  const int rif = blockDim.x / vcols;
  const int tc = threadIdx.x % vcols;
  const int tr = threadIdx.x / vcols;
  const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > M) r1 = M;
  for (int cv = tc; cv < cvec; cv += vcols) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (tr < rif) {
      constexpr int kU = 2;
      long long r = r0 + tr;
      for (; r + (kU - 1ll) * rif < r1; r += static_cast<long long>(kU) * rif) {
        bf16x8 v[kU][kN], vy[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const long long off = (r + static_cast<long long>(u) * rif) * C + cv * 8;
#pragma unroll
          for (int t = 0; t < kN; ++t) v[u][t] = *reinterpret_cast<const bf16x8*>(in.p[t] + off);
          vy[u] = *reinterpret_cast<const bf16x8*>(y + off);
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          float s[8];
          unpack8(v[u][0], s);
#pragma unroll
          for (int t = 1; t < kN; ++t) {
            float f[8];
            unpack8(v[u][t], f);
#pragma unroll
            for (int j = 0; j < 8; ++j) s[j] += f[j];
          }
          bf16x8 z = pack8(s);
          relu_mask8(z, vy[u]);
          *reinterpret_cast<bf16x8*>(dz + (r + static_cast<long long>(u) * rif) * C + cv * 8) = z;
          unpack8(z, s);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += s[j];
        }
      }
      for (; r < r1; r += rif) {
        const long long off = r * C + cv * 8;
        float s[8];
        unpack8(*reinterpret_cast<const bf16x8*>(in.p[0] + off), s);
#pragma unroll
        for (int t = 1; t < kN; ++t) {
          float f[8];
          unpack8(*reinterpret_cast<const bf16x8*>(in.p[t] + off), f);
#pragma unroll
          for (int j = 0; j < 8; ++j) s[j] += f[j];
        }
        bf16x8 z = pack8(s);
        relu_mask8(z, *reinterpret_cast<const bf16x8*>(y + off));
        *reinterpret_cast<bf16x8*>(dz + off) = z;
        unpack8(z, s);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += s[j];
      }
    }
    __syncthreads();
    if (tr < rif) {
#pragma unroll
      for (int j = 0; j < 8; ++j) red[(tr * vcols + tc) * 8 + j] = acc[j];
    }
    __syncthreads();
    if (tr == 0) {
      float s8[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float sum = 0.f;
        for (int t = 0; t < rif; ++t) sum += red[(t * vcols + tc) * 8 + j];
        s8[j] = sum;
      }
      float* dst = out + cv * 8;
      if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(s8[0]), "f"(s8[1]), "f"(s8[2]),
                     "f"(s8[3]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4), "f"(s8[4]), "f"(s8[5]), "f"(s8[6]),
                     "f"(s8[7]) : "memory");
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(dst + j, s8[j]);
      }
    }
  }
}

extern "C" int bdl_colsum_sumn_relu_bf16(const void* const* ptrs, int count, const void* y, void* dz, float* out,
                                         long long M, int C, void* stream) {
  if (count < 2 || count > 4) return bdl_fail("bdl_colsum_sumn_relu_bf16: 2..4 addends");
  if (C % 8) return bdl_fail("bdl_colsum_sumn_relu_bf16: C must be a multiple of 8");
  if (M <= 0) return BDL_OK;
  AddNPtrs a{};
  for (int t = 0; t < count; ++t) {
    if (reinterpret_cast<uintptr_t>(ptrs[t]) & 15u) return bdl_fail("bdl_colsum_sumn_relu_bf16: unaligned addend");
    a.p[t] = static_cast<const __nv_bfloat16*>(ptrs[t]);
  }
  if ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dz)) & 15u)
    return bdl_fail("bdl_colsum_sumn_relu_bf16: unaligned y / dz");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int sms = bdl_num_sms();
  long long rpb = (M + static_cast<long long>(sms) * 4 - 1) / (static_cast<long long>(sms) * 4);
  if (rpb < 16) rpb = 16;
  const int grid = static_cast<int>((M + rpb - 1) / rpb);
  const int cvec = C / 8;
  const int vcols = cvec < 256 ? cvec : 256;
  const int rif = 256 / vcols;
  const size_t smem = static_cast<size_t>(rif) * vcols * 8 * sizeof(float);
  const __nv_bfloat16* yp = static_cast<const __nv_bfloat16*>(y);
  __nv_bfloat16* zp = static_cast<__nv_bfloat16*>(dz);
  const int rp = static_cast<int>(rpb);
  if (count == 2) colsum_sumn_relu_kernel<2><<<grid, 256, smem, st>>>(a, yp, zp, out, M, C, rp);
  else if (count == 3) colsum_sumn_relu_kernel<3><<<grid, 256, smem, st>>>(a, yp, zp, out, M, C, rp);
  else colsum_sumn_relu_kernel<4><<<grid, 256, smem, st>>>(a, yp, zp, out, M, C, rp);
  return bdl_check_launch("colsum_sumn_relu_kernel");
}

extern "C" int bdl_addn_bf16(const void* const* ptrs, int count, void* out, size_t n, void* stream) {
  if (count < 2 || count > 8) return bdl_fail("bdl_addn_bf16: 2..8 inputs");
  if (n % 8) return bdl_fail("bdl_addn_bf16: element count must be a multiple of 8");
  AddNPtrs a{};
  for (int t = 0; t < count; ++t) {
    if (reinterpret_cast<uintptr_t>(ptrs[t]) & 15u) return bdl_fail("bdl_addn_bf16: pointers must be 16-byte aligned");
    a.p[t] = static_cast<const __nv_bfloat16*>(ptrs[t]);
  }
  if (reinterpret_cast<uintptr_t>(out) & 15u) return bdl_fail("bdl_addn_bf16: pointers must be 16-byte aligned");
  if (n == 0) return BDL_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int g = ew_grid(n / 8 + 1);
  __nv_bfloat16* o = static_cast<__nv_bfloat16*>(out);
  switch (count) {
    case 2: addn_bf16_kernel<2><<<g, 256, 0, st>>>(a, o, n / 8); break;
    case 3: addn_bf16_kernel<3><<<g, 256, 0, st>>>(a, o, n / 8); break;
    case 4: addn_bf16_kernel<4><<<g, 256, 0, st>>>(a, o, n / 8); break;
    case 5: addn_bf16_kernel<5><<<g, 256, 0, st>>>(a, o, n / 8); break;
    case 6: addn_bf16_kernel<6><<<g, 256, 0, st>>>(a, o, n / 8); break;
    case 7: addn_bf16_kernel<7><<<g, 256, 0, st>>>(a, o, n / 8); break;
    default: addn_bf16_kernel<8><<<g, 256, 0, st>>>(a, o, n / 8); break;
  }
  return bdl_check_launch("addn_bf16_kernel");
}

extern "C" int bdl_colsum_bf16(const void* x, const void* y, void* dz, float* out, long long M, int C, long long ldx,
                               long long ldy, long long ldz, void* stream) {
  if (M <= 0 || C <= 0) return BDL_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool relu = (y != nullptr);
  const int sms = bdl_num_sms();
  // Few, long-running blocks: every block ends with C atomics on the same C addresses, so block count (not bytes) is
  // what limits the small and medium layers. BDL_COLSUM_BPS overrides blocks-per-SM for tuning.
  static const int bps = [] {
    const char* e = getenv("BDL_COLSUM_BPS");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : 4;
  }();
  long long rpb = (M + static_cast<long long>(sms) * bps - 1) / (static_cast<long long>(sms) * bps);
  if (rpb < 16) rpb = 16;
  const int grid = static_cast<int>((M + rpb - 1) / rpb);
  const bool vec = (C % 8 == 0) && (ldx % 8 == 0) && (!relu || ((ldy % 8 == 0) && (ldz % 8 == 0))) &&
                   !(reinterpret_cast<uintptr_t>(x) & 15u) && (!relu || (!(reinterpret_cast<uintptr_t>(y) & 15u) &&
                                                                         !(reinterpret_cast<uintptr_t>(dz) & 15u)));
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* yp = static_cast<const __nv_bfloat16*>(y);
  __nv_bfloat16* zp = static_cast<__nv_bfloat16*>(dz);
  if (vec) {
    const int cvec = C / 8;
    const int vcols = cvec < 256 ? cvec : 256;
    const int rif = 256 / vcols;
    const size_t smem = static_cast<size_t>(rif) * vcols * 8 * sizeof(float);
    if (relu)
      colsum_bf16_kernel<true><<<grid, 256, smem, st>>>(xp, yp, zp, out, M, C, ldx, ldy, ldz, static_cast<int>(rpb));
    else
      colsum_bf16_kernel<false><<<grid, 256, smem, st>>>(xp, yp, zp, out, M, C, ldx, ldy, ldz, static_cast<int>(rpb));
  } else {
    if (relu)
      colsum_bf16_scalar_kernel<true><<<grid, 256, 0, st>>>(xp, yp, zp, out, M, C, ldx, ldy, ldz,
                                                            static_cast<int>(rpb));
    else
      colsum_bf16_scalar_kernel<false><<<grid, 256, 0, st>>>(xp, yp, zp, out, M, C, ldx, ldy, ldz,
                                                             static_cast<int>(rpb));
  }
  return bdl_check_launch("colsum_bf16_kernel");
}

extern "C" int bdl_dropout_fwd_bf16(const void* x, void* y, void* mask, size_t n, float p, unsigned long long seed,
                                    unsigned long long offset, void* stream) {
  if (n == 0) return BDL_OK;
  const float scale = p < 1.f ? 1.f / (1.f - p) : 0.f;
  dropout_fwd_kernel<<<ew_grid((n + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(mask), n, p, scale,
      seed, offset, nullptr);
  return bdl_check_launch("dropout_fwd_kernel");
}

// Same, but the Philox stream position is read from device memory and advanced afterwards by bdl_counter_add, so a
// captured CUDA graph draws a new mask on every replay.
extern "C" int bdl_dropout_fwd_bf16_ctr(const void* x, void* y, void* mask, size_t n, float p, unsigned long long seed,
                                        void* counter_dev, void* stream) {
  if (n == 0) return BDL_OK;
  const float scale = p < 1.f ? 1.f / (1.f - p) : 0.f;
  dropout_fwd_kernel<<<ew_grid((n + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(mask), n, p, scale,
      seed, 0ull, static_cast<const unsigned long long*>(counter_dev));
  int rc = bdl_check_launch("dropout_fwd_kernel");
  if (rc) return rc;
  return bdl_counter_add(counter_dev, (n + 7) / 8, stream);
}

extern "C" int bdl_dropout_bwd_bf16(const void* dy, const void* mask, void* dx, size_t n, float p, void* stream) {
  if (n == 0) return BDL_OK;
  const float scale = p < 1.f ? 1.f / (1.f - p) : 0.f;
  dropout_bwd_kernel<<<ew_grid((n + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(mask), static_cast<__nv_bfloat16*>(dx), n,
      scale);
  return bdl_check_launch("dropout_bwd_kernel");
}

extern "C" int bdl_copy2d_bf16(const void* src, void* dst, long long rows, int C, long long lds, long long ldd,
                               int accumulate, void* stream) {
  if (rows <= 0 || C <= 0) return BDL_OK;
  if ((C % 8) || (lds % 8) || (ldd % 8) || (reinterpret_cast<uintptr_t>(src) & 15u) ||
      (reinterpret_cast<uintptr_t>(dst) & 15u))
    return bdl_fail("bdl_copy2d_bf16: C and pitches must be multiples of 8 and pointers 16-byte aligned");
  const size_t total = static_cast<size_t>(rows) * (C / 8);
  copy2d_bf16_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(src), static_cast<__nv_bfloat16*>(dst), rows, C, lds, ldd, accumulate);
  return bdl_check_launch("copy2d_bf16_kernel");
}
