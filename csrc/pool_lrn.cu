// NHWC bf16 pooling (max with argmax byte, average) and cross-channel LRN, forward and backward.
// All are HBM-bound gather kernels: one thread owns 8 channels (16 B) of one output element, no atomics.
//
// Reference counterparts: S/nn/SpatialMaxPooling.scala (ceil mode via .ceil()), S/nn/SpatialAveragePooling.scala
// (countIncludePad, divide), S/nn/SpatialCrossMapLRN.scala:204-237 (y = x * (k + alpha/size * sum x^2)^-beta),
// and their mkldnn twins S/nn/mkldnn/{MaxPooling,AvgPooling,LRN}.scala.
#include <cstdlib>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

struct PoolGeom {
  int N, H, W, C, Ho, Wo;
  int kh, kw, sh, sw, ph, pw;
  long long ldx, ldy;  // pixel pitches of input / output (channel slices allowed)
  int count_include_pad;
  int divide;
};

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 v = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = __bfloat1622float2(q[j]);
    f[2 * j] = t.x;
    f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]);
  o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]);
  o.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = o;
}

// ---------------------------------------------------------------- max pooling
template <typename I>
__global__ void maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                   uint8_t* __restrict__ idx, const PoolGeom g) {
  const int cvec = g.C / 8;
  const I total = static_cast<I>(g.N) * g.Ho * g.Wo * cvec;
  const I stride = static_cast<I>(gridDim.x) * blockDim.x;
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    I t = i / cvec;
    const int wo = static_cast<int>(t % g.Wo);
    t /= g.Wo;
    const int ho = static_cast<int>(t % g.Ho);
    const int n = static_cast<int>(t / g.Ho);
    // packed bf16x2 compare/max (bf16 order == fp32 order): 3 instructions per channel pair and tap
    uint32_t best[4] = {0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u};   // -inf
    uint32_t bi[4] = {0, 0, 0, 0};                                              // tap index per 16-bit lane
    for (int ky = 0; ky < g.kh; ++ky) {
      const int hi = ho * g.sh - g.ph + ky;
      if (hi < 0 || hi >= g.H) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int wi = wo * g.sw - g.pw + kx;
        if (wi < 0 || wi >= g.W) continue;
        const uint4 v = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * g.H + hi) * g.W + wi) * g.ldx + cv * 8);
        const uint32_t f[4] = {v.x, v.y, v.z, v.w};
        const uint32_t tap = static_cast<uint32_t>(ky * g.kw + kx) * 0x00010001u;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const __nv_bfloat162 fq = *reinterpret_cast<const __nv_bfloat162*>(&f[q]);
          const __nv_bfloat162 bq = *reinterpret_cast<const __nv_bfloat162*>(&best[q]);
          const uint32_t gt = __hgt2_mask(fq, bq);
          const __nv_bfloat162 m = __hmax2(bq, fq);
          best[q] = *reinterpret_cast<const uint32_t*>(&m);
          bi[q] = (bi[q] & ~gt) | (tap & gt);
        }
      }
    }
    const long long opix = (static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo;
    *reinterpret_cast<uint4*>(y + opix * g.ldy + cv * 8) = make_uint4(best[0], best[1], best[2], best[3]);
    if (idx)
      *reinterpret_cast<uint2*>(idx + opix * g.C + cv * 8) =
          make_uint2(__byte_perm(bi[0], bi[1], 0x6420), __byte_perm(bi[2], bi[3], 0x6420));
  }
}

// dx[n,h,w,c] = sum of dy over output cells whose argmax tap lands on (h,w)
template <typename I>
__global__ void maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                   __nv_bfloat16* __restrict__ dx, const PoolGeom g) {
  const int cvec = g.C / 8;
  const I total = static_cast<I>(g.N) * g.H * g.W * cvec;
  const I stride = static_cast<I>(gridDim.x) * blockDim.x;
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    I t = i / cvec;
    const int w = static_cast<int>(t % g.W);
    t /= g.W;
    const int h = static_cast<int>(t % g.H);
    const int n = static_cast<int>(t / g.H);
    // packed bf16x2 accumulation: exact when windows do not overlap (k == stride), a few roundings otherwise
    uint32_t acc[4] = {0, 0, 0, 0};
    // output windows covering (h, w): ho in [ceil((h+ph-kh+1)/sh), floor((h+ph)/sh)] — at most ceil(k/s)^2 of them,
    // found with two divisions per axis instead of testing every tap
    const int hp = h + g.ph, wp = w + g.pw;
    int ho0 = hp - g.kh + 1;
    ho0 = ho0 <= 0 ? 0 : (ho0 + g.sh - 1) / g.sh;
    int ho1 = hp / g.sh;
    if (ho1 >= g.Ho) ho1 = g.Ho - 1;
    int wo0 = wp - g.kw + 1;
    wo0 = wo0 <= 0 ? 0 : (wo0 + g.sw - 1) / g.sw;
    int wo1 = wp / g.sw;
    if (wo1 >= g.Wo) wo1 = g.Wo - 1;
    for (int ho = ho0; ho <= ho1; ++ho) {
      const int ky = hp - ho * g.sh;
      for (int wo = wo0; wo <= wo1; ++wo) {
        const int kx = wp - wo * g.sw;
        const long long opix = (static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo;
        uint2 pk = *reinterpret_cast<const uint2*>(idx + opix * g.C + cv * 8);
        const uint4 d = *reinterpret_cast<const uint4*>(dy + opix * g.ldy + cv * 8);
        // tap byte b -> bf16 lane 0x3F80 | b (a normal number) so HSET2.EQ can compare two channels at once
        pk.x |= 0x80808080u;
        pk.y |= 0x80808080u;
        const uint32_t lanes[4] = {__byte_perm(pk.x, 0x3F3F3F3Fu, 0x4140), __byte_perm(pk.x, 0x3F3F3F3Fu, 0x4342),
                                   __byte_perm(pk.y, 0x3F3F3F3Fu, 0x4140), __byte_perm(pk.y, 0x3F3F3F3Fu, 0x4342)};
        const uint32_t want = (0x3F80u | static_cast<uint32_t>(ky * g.kw + kx)) * 0x00010001u;
        const uint32_t dq[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t m = __heq2_mask(*reinterpret_cast<const __nv_bfloat162*>(&lanes[q]),
                                         *reinterpret_cast<const __nv_bfloat162*>(&want)) & dq[q];
          const __nv_bfloat162 sum = __hadd2(*reinterpret_cast<const __nv_bfloat162*>(&acc[q]),
                                             *reinterpret_cast<const __nv_bfloat162*>(&m));
          acc[q] = *reinterpret_cast<const uint32_t*>(&sum);
        }
      }
    }
    *reinterpret_cast<uint4*>(dx + ((static_cast<long long>(n) * g.H + h) * g.W + w) * g.ldx + cv * 8) =
        make_uint4(acc[0], acc[1], acc[2], acc[3]);
  }
}

// 3x3 windows (every pooling layer of Inception-v1 and ResNet): a thread walks down kRows output rows and keeps the
// per-input-row horizontal maxima of the last rows in registers, so every input row is fetched once per strip instead
// of once per overlapping window (3 loads per output for stride 1 instead of 9, 6 for stride 2). The plain kernel
// above is L2-bound by those re-reads. Tie-breaking (first maximum in (ky, kx) scan order) is unchanged.
// All arithmetic stays on packed bf16x2 registers (HMNMX2 / HSET2 masks): bf16 ordering equals fp32 ordering, so the
// result is bit-identical to the fp32 compare while costing ~6x fewer instructions than unpack + FSETP + SEL per lane.
struct RowMax {
  __nv_bfloat162 v[4];
  uint32_t kx[4];   // per 16-bit lane: column tap of the row maximum
};

__device__ __forceinline__ uint32_t sel_mask(uint32_t old, uint32_t neu, uint32_t m) { return (old & ~m) | (neu & m); }

template <int S>
__device__ __forceinline__ void pool3_row(const __nv_bfloat16* __restrict__ x, const PoolGeom& g, long long img_row0,
                                          int hi, int wi0, int cv, RowMax& r) {
  const uint32_t ninf = 0xFF80FF80u;
  uint4 t[3];
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) t[kx] = make_uint4(ninf, ninf, ninf, ninf);
  if (hi >= 0 && hi < g.H) {
    const __nv_bfloat16* row = x + (img_row0 + hi) * g.W * g.ldx + cv * 8;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int wi = wi0 + kx;
      if (wi >= 0 && wi < g.W) t[kx] = *reinterpret_cast<const uint4*>(row + static_cast<long long>(wi) * g.ldx);
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const __nv_bfloat162 t0 = reinterpret_cast<const __nv_bfloat162*>(&t[0])[q];
    const __nv_bfloat162 t1 = reinterpret_cast<const __nv_bfloat162*>(&t[1])[q];
    const __nv_bfloat162 t2 = reinterpret_cast<const __nv_bfloat162*>(&t[2])[q];
    const uint32_t g1 = __hgt2_mask(t1, t0);
    const __nv_bfloat162 m1 = __hmax2(t0, t1);
    const uint32_t g2 = __hgt2_mask(t2, m1);
    r.v[q] = __hmax2(m1, t2);
    r.kx[q] = sel_mask(g1 & 0x00010001u, 0x00020002u, g2);
  }
}

template <int S, int kRows>
__global__ void maxpool3_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                    uint8_t* __restrict__ idx, const PoolGeom g) {
  const int cvec = g.C / 8;
  const int strips = (g.Ho + kRows - 1) / kRows;
  const long long total = static_cast<long long>(g.N) * strips * g.Wo * cvec;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    long long t = i / cvec;
    const int wo = static_cast<int>(t % g.Wo);
    t /= g.Wo;
    const int strip = static_cast<int>(t % strips);
    const int n = static_cast<int>(t / strips);
    const int ho0 = strip * kRows;
    const int wi0 = wo * S - g.pw;
    const long long img_row0 = static_cast<long long>(n) * g.H;
    int hi = ho0 * S - g.ph;
    RowMax a, b, c;
    pool3_row<S>(x, g, img_row0, hi, wi0, cv, a);
    if (S == 1) pool3_row<S>(x, g, img_row0, hi + 1, wi0, cv, b);
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int ho = ho0 + r;
      if (ho >= g.Ho) break;
      if (S == 1) {
        pool3_row<S>(x, g, img_row0, hi + 2, wi0, cv, c);
      } else {
        pool3_row<S>(x, g, img_row0, hi + 1, wi0, cv, b);
        pool3_row<S>(x, g, img_row0, hi + 2, wi0, cv, c);
      }
      uint4 best;
      uint32_t tap[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t gb = __hgt2_mask(b.v[q], a.v[q]);
        const __nv_bfloat162 m1 = __hmax2(a.v[q], b.v[q]);
        const uint32_t gc = __hgt2_mask(c.v[q], m1);
        const __nv_bfloat162 m2 = __hmax2(m1, c.v[q]);
        reinterpret_cast<__nv_bfloat162*>(&best)[q] = m2;
        tap[q] = sel_mask(sel_mask(a.kx[q], b.kx[q] + 0x00030003u, gb), c.kx[q] + 0x00060006u, gc);
      }
      const long long opix = (static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo;
      *reinterpret_cast<uint4*>(y + opix * g.ldy + cv * 8) = best;
      if (idx)   // low byte of each 16-bit lane -> 8 consecutive bytes
        *reinterpret_cast<uint2*>(idx + opix * g.C + cv * 8) =
            make_uint2(__byte_perm(tap[0], tap[1], 0x6420), __byte_perm(tap[2], tap[3], 0x6420));
      if (S == 1) {
        a = b;
        b = c;
      } else {
        a = c;
      }
      hi += S;
    }
  }
}

// Backward twin: a thread owns kRows input rows of one (n, w, 8-channel) column and walks the output rows whose windows
// touch them; three running accumulators hold the input rows a given output row can still reach.
template <int S, int kRows>
__global__ void maxpool3_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                    __nv_bfloat16* __restrict__ dx, const PoolGeom g) {
  const int cvec = g.C / 8;
  const int strips = (g.H + kRows - 1) / kRows;
  const long long total = static_cast<long long>(g.N) * strips * g.W * cvec;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    long long t = i / cvec;
    const int w = static_cast<int>(t % g.W);
    t /= g.W;
    const int strip = static_cast<int>(t % strips);
    const int n = static_cast<int>(t / strips);
    const int h0 = strip * kRows;
    const int h1 = min(h0 + kRows, g.H);           // exclusive
    // output rows reaching [h0, h1): ho*S - ph <= h1-1 and ho*S - ph + 2 >= h0
    const int lo_num = h0 + g.ph - 2;
    const int ho_first = lo_num <= 0 ? 0 : (lo_num + S - 1) / S;
    const int ho_last = min((h1 - 1 + g.ph) / S, g.Ho - 1);
    const int wp = w + g.pw;
    int wo0 = wp - 2;
    wo0 = wo0 <= 0 ? 0 : (wo0 + S - 1) / S;
    const int wo1 = min(wp / S, g.Wo - 1);
    // accumulators are packed bf16x2: an input element receives one contribution in the common case (exact), and a
    // few adds otherwise round like the final bf16 store does
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0, a2 = a0;
    __nv_bfloat16* dcol = dx + (static_cast<long long>(n) * g.H * g.W + w) * g.ldx + cv * 8;
    const uint4 zero4 = make_uint4(0, 0, 0, 0);
    int next_emit = h0;                             // rows [h0, next_emit) are already written
    for (int ho = ho_first; ho <= ho_last; ++ho) {
      const int hbase = ho * S - g.ph;              // input row of ky = 0
      // rows before hbase that were never reached by any window (only possible at the strip start) are zero
      for (; next_emit < min(hbase, h1); ++next_emit)
        *reinterpret_cast<uint4*>(dcol + static_cast<long long>(next_emit) * g.W * g.ldx) = zero4;
      for (int wo = wo0; wo <= wo1; ++wo) {
        const uint32_t kx = static_cast<uint32_t>(wp - wo * S);
        const long long opix = (static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo;
        uint2 pk = *reinterpret_cast<const uint2*>(idx + opix * g.C + cv * 8);
        const uint4 d = *reinterpret_cast<const uint4*>(dy + opix * g.ldy + cv * 8);
        // tap byte b -> bf16 lane 0x3F80 | b (1 + b/128, a normal number), compared against the three taps of this
        // column that land on rows hbase, hbase+1, hbase+2
        pk.x |= 0x80808080u;
        pk.y |= 0x80808080u;
        const uint32_t lanes[4] = {__byte_perm(pk.x, 0x3F3F3F3Fu, 0x4140), __byte_perm(pk.x, 0x3F3F3F3Fu, 0x4342),
                                   __byte_perm(pk.y, 0x3F3F3F3Fu, 0x4140), __byte_perm(pk.y, 0x3F3F3F3Fu, 0x4342)};
        const uint32_t t0 = (0x3F80u | kx) * 0x00010001u, t1 = t0 + 0x00030003u, t2 = t0 + 0x00060006u;
        const uint32_t dq[4] = {d.x, d.y, d.z, d.w};
        uint32_t* p0 = reinterpret_cast<uint32_t*>(&a0);
        uint32_t* p1 = reinterpret_cast<uint32_t*>(&a1);
        uint32_t* p2 = reinterpret_cast<uint32_t*>(&a2);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const __nv_bfloat162 lq = *reinterpret_cast<const __nv_bfloat162*>(&lanes[q]);
          const uint32_t m0 = __heq2_mask(lq, *reinterpret_cast<const __nv_bfloat162*>(&t0)) & dq[q];
          const uint32_t m1 = __heq2_mask(lq, *reinterpret_cast<const __nv_bfloat162*>(&t1)) & dq[q];
          const uint32_t m2 = __heq2_mask(lq, *reinterpret_cast<const __nv_bfloat162*>(&t2)) & dq[q];
          __nv_bfloat162 s0 = __hadd2(*reinterpret_cast<__nv_bfloat162*>(&p0[q]), *reinterpret_cast<const __nv_bfloat162*>(&m0));
          __nv_bfloat162 s1 = __hadd2(*reinterpret_cast<__nv_bfloat162*>(&p1[q]), *reinterpret_cast<const __nv_bfloat162*>(&m1));
          __nv_bfloat162 s2 = __hadd2(*reinterpret_cast<__nv_bfloat162*>(&p2[q]), *reinterpret_cast<const __nv_bfloat162*>(&m2));
          p0[q] = *reinterpret_cast<uint32_t*>(&s0);
          p1[q] = *reinterpret_cast<uint32_t*>(&s1);
          p2[q] = *reinterpret_cast<uint32_t*>(&s2);
        }
      }
      // rows no later output row can reach are complete: hbase .. hbase+S-1
      if (hbase >= h0 && hbase < h1 && hbase >= next_emit) {
        *reinterpret_cast<uint4*>(dcol + static_cast<long long>(hbase) * g.W * g.ldx) = a0;
        next_emit = hbase + 1;
      }
      if (S == 2 && hbase + 1 >= h0 && hbase + 1 < h1 && hbase + 1 >= next_emit) {
        *reinterpret_cast<uint4*>(dcol + static_cast<long long>(hbase + 1) * g.W * g.ldx) = a1;
        next_emit = hbase + 2;
      }
      if (S == 1) {
        a0 = a1;
        a1 = a2;
      } else {
        a0 = a2;
        a1 = zero4;
      }
      a2 = zero4;
    }
    // tail: rows still held in the accumulators, then rows below every window
    const int hb = (ho_last + 1) * S - g.ph;        // row that a0 now stands for
    for (; next_emit < h1; ++next_emit) {
      uint4 v = zero4;
      if (ho_last >= ho_first && next_emit == hb) v = a0;
      else if (ho_last >= ho_first && S == 1 && next_emit == hb + 1) v = a1;
      *reinterpret_cast<uint4*>(dcol + static_cast<long long>(next_emit) * g.W * g.ldx) = v;
    }
  }
}

// ---------------------------------------------------------------- average pooling
__device__ __forceinline__ float avg_divisor(const PoolGeom& g, int ho, int wo) {
  if (!g.divide) return 1.f;
  int hs = ho * g.sh - g.ph, ws = wo * g.sw - g.pw;
  int he = min(hs + g.kh, g.H + g.ph), we = min(ws + g.kw, g.W + g.pw);
  const int pool_size = (he - hs) * (we - ws);
  if (g.count_include_pad) return static_cast<float>(pool_size);
  hs = max(hs, 0);
  ws = max(ws, 0);
  he = min(he, g.H);
  we = min(we, g.W);
  return static_cast<float>(max((he - hs) * (we - ws), 1));
}

template <typename I>
__global__ void avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                   const PoolGeom g) {
  const int cvec = g.C / 8;
  const I total = static_cast<I>(g.N) * g.Ho * g.Wo * cvec;
  const I stride = static_cast<I>(gridDim.x) * blockDim.x;
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    I t = i / cvec;
    const int wo = static_cast<int>(t % g.Wo);
    t /= g.Wo;
    const int ho = static_cast<int>(t % g.Ho);
    const int n = static_cast<int>(t / g.Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int ky = 0; ky < g.kh; ++ky) {
      const int hi = ho * g.sh - g.ph + ky;
      if (hi < 0 || hi >= g.H) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int wi = wo * g.sw - g.pw + kx;
        if (wi < 0 || wi >= g.W) continue;
        float f[8];
        load8(x + ((static_cast<long long>(n) * g.H + hi) * g.W + wi) * g.ldx + cv * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    const float inv = 1.f / avg_divisor(g, ho, wo);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    store8(y + ((static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo) * g.ldy + cv * 8, acc);
  }
}

template <typename I>
__global__ void avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
                                   const PoolGeom g) {
  const int cvec = g.C / 8;
  const I total = static_cast<I>(g.N) * g.H * g.W * cvec;
  const I stride = static_cast<I>(gridDim.x) * blockDim.x;
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    I t = i / cvec;
    const int w = static_cast<int>(t % g.W);
    t /= g.W;
    const int h = static_cast<int>(t % g.H);
    const int n = static_cast<int>(t / g.H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int ky = 0; ky < g.kh; ++ky) {
      const int hh = h + g.ph - ky;
      if (hh < 0 || hh % g.sh) continue;
      const int ho = hh / g.sh;
      if (ho >= g.Ho) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int ww = w + g.pw - kx;
        if (ww < 0 || ww % g.sw) continue;
        const int wo = ww / g.sw;
        if (wo >= g.Wo) continue;
        float f[8];
        load8(dy + ((static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo) * g.ldy + cv * 8, f);
        const float inv = 1.f / avg_divisor(g, ho, wo);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j] * inv;
      }
    }
    store8(dx + ((static_cast<long long>(n) * g.H + h) * g.W + w) * g.ldx + cv * 8, acc);
  }
}

// ---------------------------------------------------------------- cross-channel LRN
// One thread owns 8 consecutive channels of a pixel. It loads the 24-channel neighbourhood [c0-8, c0+16) with three
// 16-byte loads, builds squared values once, and derives every window sum from registers (window <= 9 wide), so each
// input element is read from L1/L2 a constant three times regardless of the window size.
// scale_c = k + alpha/size * sum_{c' in [c-pre, c+post]} x_c'^2, pre = (size-1)/2, post = size-1-pre.
__device__ __forceinline__ void lrn_load24(const __nv_bfloat16* row, int c0, int C, float (&v)[24]) {
#pragma unroll
  for (int blk = 0; blk < 3; ++blk) {
    const int cb = c0 - 8 + 8 * blk;
    if (cb >= 0 && cb + 8 <= C) {
      float f[8];
      load8(row + cb, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[8 * blk + j] = f[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[8 * blk + j] = 0.f;
    }
  }
}

template <typename I, int kHalf>
__global__ void lrn_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long pixels,
                               int C, long long ldx, long long ldy, int size, float alpha, float beta, float k) {
  const int cvec = C / 8;
  const I total = static_cast<I>(pixels) * cvec;
  const I stride = static_cast<I>(gridDim.x) * blockDim.x;
  // kHalf > 0: symmetric window of 2*kHalf+1 known at compile time (size 5 in Inception/AlexNet) — the predicated
  // 17-tap loops below collapse to exactly the taps used
  const int pre = kHalf ? kHalf : (size - 1) / 2, post = kHalf ? kHalf : size - 1 - pre;
  const float as = alpha / size;
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    const I pix = i / cvec;
    const int c0 = cv * 8;
    float v[24];
    lrn_load24(x + pix * ldx, c0, C, v);
    float sq[24];
#pragma unroll
    for (int j = 0; j < 24; ++j) sq[j] = v[j] * v[j];
    float out[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = 0.f;
#pragma unroll
      for (int d = -8; d <= 8; ++d)
        if (d >= -pre && d <= post) s += sq[8 + j + d];
      out[j] = v[8 + j] * __powf(k + as * s, -beta);
    }
    store8(y + pix * ldy + c0, out);
  }
}

// (kernel templates take <index type, kHalf>)
// dx_c = dy_c * s_c^-beta - 2*alpha*beta/size * x_c * sum_{c' : c in window(c')} dy_c' * x_c' * s_c'^(-beta-1)
// c is inside window(c') iff c - post <= c' <= c + pre; s_c' needs x over [c'-pre, c'+post], so the 24-wide
// neighbourhood covers everything for windows up to 9 (pre, post <= 4).
template <typename I, int kHalf>
__global__ void lrn_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                               __nv_bfloat16* __restrict__ dx, long long pixels, int C, long long ldx, long long ldy,
                               long long lddx, int size, float alpha, float beta, float k) {
  const int cvec = C / 8;
  const I total = static_cast<I>(pixels) * cvec;
  const I stride = static_cast<I>(gridDim.x) * blockDim.x;
  const int pre = kHalf ? kHalf : (size - 1) / 2, post = kHalf ? kHalf : size - 1 - pre;
  const float as = alpha / size;
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = static_cast<int>(i % cvec);
    const I pix = i / cvec;
    const int c0 = cv * 8;
    float v[24], g[24];
    lrn_load24(x + pix * ldx, c0, C, v);
    lrn_load24(dy + pix * ldy, c0, C, g);
    float sq[24];
#pragma unroll
    for (int j = 0; j < 24; ++j) sq[j] = v[j] * v[j];
    // ratio[c'] = dy_c' * x_c' * s_c'^(-beta-1) and pw[c'] = s_c'^-beta for c' in [c0-4, c0+12)  (index t = c' - c0 + 4)
    float ratio[16], pw[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      if (kHalf && (t < 4 - kHalf || t >= 12 + kHalf)) {   // never read by the cross sum below
        ratio[t] = 0.f;
        pw[t] = 0.f;
        continue;
      }
      const int cp = c0 - 4 + t;
      if (cp < 0 || cp >= C) {   // outside the tensor: contributes nothing (and k may be 0)
        ratio[t] = 0.f;
        pw[t] = 0.f;
        continue;
      }
      float s = 0.f;
#pragma unroll
      for (int d = -4; d <= 4; ++d)
        if (d >= -pre && d <= post) s += sq[4 + t + d];
      const float sc = k + as * s;
      const float p = __powf(sc, -beta);
      pw[t] = p;
      ratio[t] = __fdividef(g[4 + t] * v[4 + t] * p, sc);
    }
    float out[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float cross = 0.f;
#pragma unroll
      for (int d = -4; d <= 4; ++d)
        if (d >= -post && d <= pre) cross += ratio[4 + j + d];
      out[j] = g[8 + j] * pw[4 + j] - 2.f * as * beta * v[8 + j] * cross;
    }
    store8(dx + pix * lddx + c0, out);
  }
}

// generic fallback for windows wider than 9 (never used by the model zoo)
__global__ void lrn_fwd_wide_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                    long long pixels, int C, long long ldx, long long ldy, int size, float alpha,
                                    float beta, float k) {
  const long long total = pixels * C;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int pre = (size - 1) / 2, post = size - 1 - pre;
  const float as = alpha / size;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    const long long pix = i / C;
    const __nv_bfloat16* row = x + pix * ldx;
    float s = 0.f;
    for (int cc = max(c - pre, 0); cc <= min(c + post, C - 1); ++cc) {
      const float t = __bfloat162float(row[cc]);
      s += t * t;
    }
    y[pix * ldy + c] = __float2bfloat16_rn(__bfloat162float(row[c]) * __powf(k + as * s, -beta));
  }
}

__global__ void lrn_bwd_wide_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                    __nv_bfloat16* __restrict__ dx, long long pixels, int C, long long ldx,
                                    long long ldy, long long lddx, int size, float alpha, float beta, float k) {
  const long long total = pixels * C;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int pre = (size - 1) / 2, post = size - 1 - pre;
  const float as = alpha / size;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    const long long pix = i / C;
    const __nv_bfloat16* xr = x + pix * ldx;
    const __nv_bfloat16* gr = dy + pix * ldy;
    float cross = 0.f, self_scale = 1.f;
    for (int cp = max(c - post, 0); cp <= min(c + pre, C - 1); ++cp) {
      float s = 0.f;
      for (int cc = max(cp - pre, 0); cc <= min(cp + post, C - 1); ++cc) {
        const float t = __bfloat162float(xr[cc]);
        s += t * t;
      }
      const float sc = k + as * s;
      const float p = __powf(sc, -beta);
      if (cp == c) self_scale = p;
      cross += __bfloat162float(gr[cp]) * __bfloat162float(xr[cp]) * p / sc;
    }
    dx[pix * lddx + c] = __float2bfloat16_rn(__bfloat162float(gr[c]) * self_scale -
                                             2.f * as * beta * __bfloat162float(xr[c]) * cross);
  }
}

static bool pool3_sliding(int kh, int kw, int sh, int sw) {
  static const bool on = [] { const char* e = getenv("BDL_POOL_SLIDE"); return !(e && e[0] == '0'); }();
  return on && kh == 3 && kw == 3 && sh == sw && (sh == 1 || sh == 2);
}

#define BDL_IDX_DISPATCH(KERNEL, TOTAL, ...)                                      \
  do {                                                                           \
    if ((TOTAL) < (1ll << 31) - (1ll << 24))                                     \
      KERNEL<unsigned int><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__);       \
    else                                                                         \
      KERNEL<long long><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__);          \
  } while (0)

#define BDL_LRN_DISPATCH(KERNEL, SIZE, TOTAL, ...)                                          \
  do {                                                                                    \
    const bool small_ = (TOTAL) < (1ll << 31) - (1ll << 24);                              \
    if ((SIZE) == 5) {                                                                    \
      if (small_) KERNEL<unsigned int, 2><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__); \
      else KERNEL<long long, 2><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__);           \
    } else if ((SIZE) == 3) {                                                             \
      if (small_) KERNEL<unsigned int, 1><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__); \
      else KERNEL<long long, 1><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__);           \
    } else {                                                                              \
      if (small_) KERNEL<unsigned int, 0><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__); \
      else KERNEL<long long, 0><<<grid_for(TOTAL), 256, 0, st_>>>(__VA_ARGS__);           \
    }                                                                                     \
  } while (0)

static int grid_for(long long total) {
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  long long b = (total + 255) / 256;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace bdl

using namespace bdl;

static int check_pool(const char* who, int C, long long ldx, long long ldy, const void* a, const void* b) {
  if ((C % 8) || (ldx % 8) || (ldy % 8) || (reinterpret_cast<uintptr_t>(a) & 15u) || (reinterpret_cast<uintptr_t>(b) & 15u))
    return bdl_fail(who);
  return BDL_OK;
}

static PoolGeom make_pool(int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh, int sw, int ph, int pw,
                          long long ldx, long long ldy, int cip, int divide) {
  PoolGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.Ho = Ho; g.Wo = Wo;
  g.kh = kh; g.kw = kw; g.sh = sh; g.sw = sw; g.ph = ph; g.pw = pw;
  g.ldx = ldx; g.ldy = ldy; g.count_include_pad = cip; g.divide = divide;
  return g;
}

extern "C" int bdl_maxpool_fwd_nhwc_bf16(const void* x, void* y, void* idx, int N, int H, int W, int C, int Ho, int Wo,
                                         int kh, int kw, int sh, int sw, int ph, int pw, long long ldx, long long ldy,
                                         void* stream) {
  if (check_pool("bdl_maxpool_fwd_nhwc_bf16: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx,
                 ldy, x, y))
    return BDL_ERR_INVALID;
  if (kh * kw > 127) return bdl_fail("bdl_maxpool_fwd_nhwc_bf16: window larger than 127 taps");
  const PoolGeom g = make_pool(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, ldx, ldy, 0, 0);
  const long long total = static_cast<long long>(N) * Ho * Wo * (C / 8);
  if (pool3_sliding(kh, kw, sh, sw)) {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    const long long threads = static_cast<long long>(N) * ((Ho + 6) / 7) * Wo * (C / 8);
    if (sh == 1)
      maxpool3_fwd_kernel<1, 7><<<grid_for(threads), 256, 0, st_>>>(
          static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx), g);
    else
      maxpool3_fwd_kernel<2, 7><<<grid_for(threads), 256, 0, st_>>>(
          static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx), g);
    return bdl_check_launch("maxpool3_fwd_kernel");
  }
  {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    BDL_IDX_DISPATCH(maxpool_fwd_kernel, static_cast<long long>(total), 
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx), g);
  }
  return bdl_check_launch("maxpool_fwd_kernel");
}

extern "C" int bdl_maxpool_bwd_nhwc_bf16(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int Ho,
                                         int Wo, int kh, int kw, int sh, int sw, int ph, int pw, long long ldx,
                                         long long ldy, void* stream) {
  if (check_pool("bdl_maxpool_bwd_nhwc_bf16: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx,
                 ldy, dx, dy))
    return BDL_ERR_INVALID;
  const PoolGeom g = make_pool(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, ldx, ldy, 0, 0);
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  if (pool3_sliding(kh, kw, sh, sw)) {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    const long long threads = static_cast<long long>(N) * ((H + 6) / 7) * W * (C / 8);
    if (sh == 1)
      maxpool3_bwd_kernel<1, 7><<<grid_for(threads), 256, 0, st_>>>(
          static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), g);
    else
      maxpool3_bwd_kernel<2, 7><<<grid_for(threads), 256, 0, st_>>>(
          static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), g);
    return bdl_check_launch("maxpool3_bwd_kernel");
  }
  {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    BDL_IDX_DISPATCH(maxpool_bwd_kernel, static_cast<long long>(total), 
      static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), g);
  }
  return bdl_check_launch("maxpool_bwd_kernel");
}

extern "C" int bdl_avgpool_fwd_nhwc_bf16(const void* x, void* y, int N, int H, int W, int C, int Ho, int Wo, int kh,
                                         int kw, int sh, int sw, int ph, int pw, long long ldx, long long ldy,
                                         int count_include_pad, int divide, void* stream) {
  if (check_pool("bdl_avgpool_fwd_nhwc_bf16: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx,
                 ldy, x, y))
    return BDL_ERR_INVALID;
  const PoolGeom g = make_pool(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, ldx, ldy, count_include_pad, divide);
  const long long total = static_cast<long long>(N) * Ho * Wo * (C / 8);
  {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    BDL_IDX_DISPATCH(avgpool_fwd_kernel, static_cast<long long>(total), 
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), g);
  }
  return bdl_check_launch("avgpool_fwd_kernel");
}

extern "C" int bdl_avgpool_bwd_nhwc_bf16(const void* dy, void* dx, int N, int H, int W, int C, int Ho, int Wo, int kh,
                                         int kw, int sh, int sw, int ph, int pw, long long ldx, long long ldy,
                                         int count_include_pad, int divide, void* stream) {
  if (check_pool("bdl_avgpool_bwd_nhwc_bf16: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx,
                 ldy, dx, dy))
    return BDL_ERR_INVALID;
  const PoolGeom g = make_pool(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, ldx, ldy, count_include_pad, divide);
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    BDL_IDX_DISPATCH(avgpool_bwd_kernel, static_cast<long long>(total), 
      static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), g);
  }
  return bdl_check_launch("avgpool_bwd_kernel");
}

extern "C" int bdl_lrn_fwd_nhwc_bf16(const void* x, void* y, long long pixels, int C, long long ldx, long long ldy,
                                     int size, float alpha, float beta, float k, void* stream) {
  if (check_pool("bdl_lrn_fwd_nhwc_bf16: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx, ldy, x,
                 y))
    return BDL_ERR_INVALID;
  if (size <= 9)
    {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    BDL_LRN_DISPATCH(lrn_fwd_kernel, size, static_cast<long long>(pixels * (C / 8)), 
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), pixels, C, ldx, ldy, size, alpha, beta,
        k);
  }
  else
    lrn_fwd_wide_kernel<<<grid_for(pixels * C), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), pixels, C, ldx, ldy, size, alpha, beta,
        k);
  return bdl_check_launch("lrn_fwd_kernel");
}

extern "C" int bdl_lrn_bwd_nhwc_bf16(const void* x, const void* dy, void* dx, long long pixels, int C, long long ldx,
                                     long long ldy, long long lddx, int size, float alpha, float beta, float k,
                                     void* stream) {
  if (check_pool("bdl_lrn_bwd_nhwc_bf16: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx, ldy, x,
                 dy))
    return BDL_ERR_INVALID;
  if (size <= 9)
    {
    cudaStream_t st_ = static_cast<cudaStream_t>(stream);
    BDL_LRN_DISPATCH(lrn_bwd_kernel, size, static_cast<long long>(pixels * (C / 8)), 
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx),
        pixels, C, ldx, ldy, lddx, size, alpha, beta, k);
  }
  else
    lrn_bwd_wide_kernel<<<grid_for(pixels * C), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx),
        pixels, C, ldx, ldy, lddx, size, alpha, beta, k);
  return bdl_check_launch("lrn_bwd_kernel");
}
