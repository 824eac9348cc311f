// NHWC im2col / col2im for the explicit-GEMM convolution path (fprop, wgrad use im2col; dgrad uses col2im).
// K ordering of the column matrix is (kh, kw, c) so that it matches weights stored OHWI (= torch channels_last),
// and the row pitch Kp is padded to a multiple of 8 elements so the GEMM's TMA descriptors stay 16-byte aligned.
//
// Reference counterparts: S/nn/NNPrimitive.scala:24-1719 (im2col/col2im NCHW+NHWC, dilation variants) called from
// S/nn/SpatialConvolution.scala:598-655 (updateOutputFrame), :656 (updateGradInputFrame), :738 (accGradParametersFrame).
#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

struct ConvGeom {
  int N, H, W, C;       // input
  int Ho, Wo;           // output spatial
  int kh, kw, sh, sw, ph, pw, dh, dw;
  long long ldx;        // input pixel pitch (elements) — C for a dense tensor, more for a channel slice
  int Kp;               // column-matrix row pitch
};

// vector path: C % 8 == 0. One thread moves 8 channels (16 B) of one (m, kh, kw) cell.
__global__ void im2col_nhwc_vec_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col,
                                       const ConvGeom g) {
  const int cvec = g.C / 8;
  const long long M = static_cast<long long>(g.N) * g.Ho * g.Wo;
  const long long per_row = static_cast<long long>(g.kh) * g.kw * cvec;
  const long long total = M * per_row;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long m = i / per_row;
    int r = static_cast<int>(i - m * per_row);
    const int cv = r % cvec;
    r /= cvec;
    const int kx = r % g.kw;
    const int ky = r / g.kw;
    const int wo = static_cast<int>(m % g.Wo);
    const long long t = m / g.Wo;
    const int ho = static_cast<int>(t % g.Ho);
    const int n = static_cast<int>(t / g.Ho);
    const int hi = ho * g.sh - g.ph + ky * g.dh;
    const int wi = wo * g.sw - g.pw + kx * g.dw;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
      v = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * g.H + hi) * g.W + wi) * g.ldx + cv * 8);
    *reinterpret_cast<uint4*>(col + m * g.Kp + (static_cast<long long>(ky) * g.kw + kx) * g.C + cv * 8) = v;
  }
}

// 4-channel pixels (the RGB stem padded to 4): 8-byte gathers; the thread with tap == taps zero-fills the K padding.
__global__ void im2col_nhwc_vec4_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col,
                                        const ConvGeom g) {
  const long long M = static_cast<long long>(g.N) * g.Ho * g.Wo;
  const int taps = g.kh * g.kw;
  const int K = taps * 4;
  const int per_row = taps + (g.Kp > K ? 1 : 0);
  const long long total = M * per_row;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long m = i / per_row;
    const int tap = static_cast<int>(i - m * per_row);
    if (tap == taps) {
      for (int k = K; k < g.Kp; k += 4) *reinterpret_cast<uint2*>(col + m * g.Kp + k) = make_uint2(0, 0);
      continue;
    }
    const int kx = tap % g.kw;
    const int ky = tap / g.kw;
    const int wo = static_cast<int>(m % g.Wo);
    const long long t = m / g.Wo;
    const int ho = static_cast<int>(t % g.Ho);
    const int n = static_cast<int>(t / g.Ho);
    const int hi = ho * g.sh - g.ph + ky * g.dh;
    const int wi = wo * g.sw - g.pw + kx * g.dw;
    uint2 v = make_uint2(0, 0);
    if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
      v = *reinterpret_cast<const uint2*>(x + ((static_cast<long long>(n) * g.H + hi) * g.W + wi) * g.ldx);
    *reinterpret_cast<uint2*>(col + m * g.Kp + tap * 4) = v;
  }
}

// scalar path (e.g. the 3-channel stem): one thread per column-matrix element, padding columns written as zero.
__global__ void im2col_nhwc_scalar_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col,
                                          const ConvGeom g) {
  const long long M = static_cast<long long>(g.N) * g.Ho * g.Wo;
  const int K = g.kh * g.kw * g.C;
  const long long total = M * g.Kp;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long m = i / g.Kp;
    const int k = static_cast<int>(i - m * g.Kp);
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (k < K) {
      const int c = k % g.C;
      const int r = k / g.C;
      const int kx = r % g.kw;
      const int ky = r / g.kw;
      const int wo = static_cast<int>(m % g.Wo);
      const long long t = m / g.Wo;
      const int ho = static_cast<int>(t % g.Ho);
      const int n = static_cast<int>(t / g.Ho);
      const int hi = ho * g.sh - g.ph + ky * g.dh;
      const int wi = wo * g.sw - g.pw + kx * g.dw;
      if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
        v = x[((static_cast<long long>(n) * g.H + hi) * g.W + wi) * g.ldx + c];
    }
    col[i] = v;
  }
}

// col2im in gather form (no atomics): dx[n,h,w,c] = sum over filter taps whose window covers (h,w).
template <bool kVec>
__global__ void col2im_nhwc_kernel(const __nv_bfloat16* __restrict__ col, __nv_bfloat16* __restrict__ dx,
                                   const ConvGeom g) {
  const int cvec = kVec ? g.C / 8 : g.C;
  const long long total = static_cast<long long>(g.N) * g.H * g.W * cvec;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    long long t = i / cvec;
    const int w = static_cast<int>(t % g.W);
    t /= g.W;
    const int h = static_cast<int>(t % g.H);
    const int n = static_cast<int>(t / g.H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int ky = 0; ky < g.kh; ++ky) {
      const int hh = h + g.ph - ky * g.dh;
      if (hh < 0 || hh % g.sh) continue;
      const int ho = hh / g.sh;
      if (ho >= g.Ho) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int ww = w + g.pw - kx * g.dw;
        if (ww < 0 || ww % g.sw) continue;
        const int wo = ww / g.sw;
        if (wo >= g.Wo) continue;
        const long long m = (static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo;
        const __nv_bfloat16* src = col + m * g.Kp + (static_cast<long long>(ky) * g.kw + kx) * g.C;
        if (kVec) {
          const uint4 v = *reinterpret_cast<const uint4*>(src + cv * 8);
          const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = __bfloat1622float2(p[j]);
            acc[2 * j] += f.x;
            acc[2 * j + 1] += f.y;
          }
        } else {
          acc[0] += __bfloat162float(src[cv]);
        }
      }
    }
    __nv_bfloat16* dst = dx + ((static_cast<long long>(n) * g.H + h) * g.W + w) * g.ldx;
    if (kVec) {
      uint4 o;
      o.x = pack_bf16x2(acc[0], acc[1]);
      o.y = pack_bf16x2(acc[2], acc[3]);
      o.z = pack_bf16x2(acc[4], acc[5]);
      o.w = pack_bf16x2(acc[6], acc[7]);
      *reinterpret_cast<uint4*>(dst + cv * 8) = o;
    } else {
      dst[cv] = __float2bfloat16_rn(acc[0]);
    }
  }
}

static int grid_for(long long total) {
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  long long b = (total + 255) / 256;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace bdl

using namespace bdl;

static ConvGeom make_geom(int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh, int sw, int ph, int pw,
                          int dh, int dw, long long ldx, int Kp) {
  ConvGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.Ho = Ho; g.Wo = Wo;
  g.kh = kh; g.kw = kw; g.sh = sh; g.sw = sw; g.ph = ph; g.pw = pw; g.dh = dh; g.dw = dw;
  g.ldx = ldx; g.Kp = Kp;
  return g;
}

extern "C" int bdl_im2col_nhwc_bf16(const void* x, void* col, int N, int H, int W, int C, int Ho, int Wo, int kh,
                                    int kw, int sh, int sw, int ph, int pw, int dh, int dw, long long ldx, int Kp,
                                    void* stream) {
  if (Kp < kh * kw * C || (Kp % 8)) return bdl_fail("bdl_im2col_nhwc_bf16: Kp must be >= kh*kw*C and a multiple of 8");
  const ConvGeom g = make_geom(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, dh, dw, ldx, Kp);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long M = static_cast<long long>(N) * Ho * Wo;
  const bool vec = (C % 8 == 0) && (ldx % 8 == 0) && (Kp == kh * kw * C) && !(reinterpret_cast<uintptr_t>(x) & 15u) &&
                   !(reinterpret_cast<uintptr_t>(col) & 15u);
  const bool vec4 = (C == 4) && (ldx % 4 == 0) && (Kp % 4 == 0) && !(reinterpret_cast<uintptr_t>(x) & 7u) &&
                    !(reinterpret_cast<uintptr_t>(col) & 7u);
  if (vec) {
    const long long total = M * kh * kw * (C / 8);
    im2col_nhwc_vec_kernel<<<grid_for(total), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x),
                                                            static_cast<__nv_bfloat16*>(col), g);
  } else if (vec4) {
    im2col_nhwc_vec4_kernel<<<grid_for(M * (kh * kw + 1)), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x),
                                                                         static_cast<__nv_bfloat16*>(col), g);
  } else {
    im2col_nhwc_scalar_kernel<<<grid_for(M * Kp), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x),
                                                                static_cast<__nv_bfloat16*>(col), g);
  }
  return bdl_check_launch("im2col_nhwc_kernel");
}

// Space-to-depth for stride-2 RGB stems, with the horizontal filter taps folded into channels as well:
//   y[n, I, J, j*16 + (a*2+b)*C + c] = xpad[n, c, 2I+a, 2(J+j)+b]      j < tw, J < Wo, channels 4C..15 of each tap zero
// where xpad is x with (pt, pl) zero rows/columns in front. A k x k stride-2 convolution over x then equals a
// ceil(k/2) x 1 stride-1 VALID convolution over y with tw*16 channels (64 for the 7x7 stems), which the TMA-im2col
// implicit GEMM runs with full 128-byte rows: the gather is bound by rows fetched, not bytes, so 4 taps of 128 B beat
// 16 taps of 32 B by ~4x. One thread gathers one 2x2 block and stores it to its tw pixels; x is read through element
// strides so NCHW fp32 (what the data pipeline delivers) and channels-last bf16 both work without a conversion pass.
template <typename T>
__global__ void s2d_stem_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int C, int H, int W,
                                long long sn, long long sc, long long sh, long long sw, int pt, int pl, int R,
                                int Wo, int tw) {
  // One thread gathers one 2x2 pixel block (16 channels) ONCE and stores it to the tw output pixels it belongs to:
  // block column Jb is tap j of output pixel Jb - j. (Gathering per (pixel, tap) read every input element tw times.)
  const int S = Wo + tw - 1;                                   // block columns that feed some output pixel
  const long long total = static_cast<long long>(N) * R * S;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int cout = tw * 16;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int Jb = static_cast<int>(i % S);
    long long t = i / S;
    const int I = static_cast<int>(t % R);
    const int n = static_cast<int>(t / R);
    alignas(16) __nv_bfloat16 v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = __float2bfloat16_rn(0.f);
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int h = 2 * I + a - pt;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int w = 2 * Jb + b - pl;
        if (w < 0 || w >= W) continue;
        const T* px = x + n * sn + h * sh + w * sw;
        for (int c = 0; c < C; ++c) v[(a * 2 + b) * C + c] = __float2bfloat16_rn(static_cast<float>(px[c * sc]));
      }
    }
    const uint4 lo = *reinterpret_cast<const uint4*>(v), hi = *reinterpret_cast<const uint4*>(v + 8);
    __nv_bfloat16* row = y + (static_cast<long long>(n) * R + I) * Wo * cout;
    for (int j = 0; j < tw; ++j) {
      const int J = Jb - j;
      if (J < 0 || J >= Wo) continue;
      uint4* dst = reinterpret_cast<uint4*>(row + static_cast<long long>(J) * cout + j * 16);
      dst[0] = lo;
      dst[1] = hi;
    }
  }
}

extern "C" int bdl_s2d_stem_bf16(const void* x, int x_is_fp32, void* y, int N, int C, int H, int W, long long sn,
                                 long long sc, long long sh, long long sw, int pt, int pl, int R, int Wo, int tw,
                                 void* stream) {
  if (C < 1 || C > 4) return bdl_fail("bdl_s2d_stem_bf16: C must be 1..4");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(N) * R * (Wo + tw - 1);
  if (x_is_fp32)
    s2d_stem_kernel<float><<<grid_for(total), 256, 0, st>>>(static_cast<const float*>(x),
                                                            static_cast<__nv_bfloat16*>(y), N, C, H, W, sn, sc, sh,
                                                            sw, pt, pl, R, Wo, tw);
  else
    s2d_stem_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x),
                                                                    static_cast<__nv_bfloat16*>(y), N, C, H, W, sn,
                                                                    sc, sh, sw, pt, pl, R, Wo, tw);
  return bdl_check_launch("s2d_stem_kernel");
}

extern "C" int bdl_col2im_nhwc_bf16(const void* col, void* dx, int N, int H, int W, int C, int Ho, int Wo, int kh,
                                    int kw, int sh, int sw, int ph, int pw, int dh, int dw, long long ldx, int Kp,
                                    void* stream) {
  const ConvGeom g = make_geom(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, dh, dw, ldx, Kp);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = (C % 8 == 0) && (ldx % 8 == 0) && (Kp % 8 == 0) && !(reinterpret_cast<uintptr_t>(dx) & 15u) &&
                   !(reinterpret_cast<uintptr_t>(col) & 15u);
  const long long pix = static_cast<long long>(N) * H * W;
  if (vec)
    col2im_nhwc_kernel<true><<<grid_for(pix * (C / 8)), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(col),
                                                                      static_cast<__nv_bfloat16*>(dx), g);
  else
    col2im_nhwc_kernel<false><<<grid_for(pix * C), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(col),
                                                                 static_cast<__nv_bfloat16*>(dx), g);
  return bdl_check_launch("col2im_nhwc_kernel");
}
