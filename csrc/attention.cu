// Fused multi-head attention forward: O = softmax(scale * Q K^T + bias) V in ONE kernel, on the tcgen05 tensor cores.
//
// Reference: S/nn/Attention.scala:95-114 (split heads -> scaled QK^T -> + bias -> softmax -> dropout -> PV -> combine
// heads; four separate MM / SoftMax / transpose modules, the [B, heads, Lq, Lk] logits materialised twice).
// Here one CTA owns a (batch, head, 128-query) tile and streams the keys / values in 128-row tiles:
//   warp 4      TMA producer: Q once, then K_j and V_j (3-D tensor maps over the [B, L, hidden] projections, so the
//               "split heads" transpose is just the column coordinate h * D, and rows past L arrive as zeros)
//   warp 5      MMA issuer:   S_j = Q K_j^T (tcgen05.mma, fp32 in TMEM columns [0,128)), then O_j = P_j V_j (TMEM [128,128+D))
//   warps 0-3   softmax: thread r owns query row r (TMEM lane r) — row max / sum need no shuffles; reads S_j with
//               tcgen05.ld, applies scale + bias + masks, writes P_j (bf16, 128B-swizzled K-major) to shared memory for
//               the second MMA, then folds the partial O_j into its register accumulator with the running-max rescale.
// The logits never leave the SM. For D = 64 two CTAs fit per SM, so one CTA's softmax overlaps the other's MMAs.
// Output is written straight in the [B, Lq, heads * D] "combined heads" layout, plus the row log-sum-exp for backward.
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

constexpr int kAttThreads = 192;
constexpr int kTileQ = 128, kTileK = 128;
constexpr uint32_t kSub = 16384u;      // one [128 rows x 64 bf16] swizzled sub-tile

struct AttnArgs {
  const float* bias;                   // additive, fp32, inner (key) stride 1; null = none
  long long bias_sb, bias_sh, bias_sq; // element strides over batch / head / query (0 = broadcast)
  __nv_bfloat16* O;                    // [B, Lq, heads * D] (+ pitches)
  long long o_sb, o_sl;
  float* lse;                          // [B, heads, Lq] natural-log row log-sum-exp, or null
  int B, heads, Lq, Lk;
  float scale;
  int causal;                          // mask key > query + (Lk - Lq)
};

struct __align__(8) AttnBars {
  uint64_t q_full, k_full, k_empty, v_full, v_empty, s_full, p_full, o_full, o_free;
  uint32_t tmem_base;
};

// 4 consecutive bias values of one row (0 where there is no bias or the key is past Lk); one 16-byte load when the row
// allows it — the scalar form issued 256 loads per thread and tile and made a full [Lq, Lk] bias 10x slower than none.
__device__ __forceinline__ float4 attn_bias4(const float* brow, int key, int Lk, bool vec_ok) {
  float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
  if (brow == nullptr) return r;
  if (vec_ok && key + 3 < Lk) return __ldg(reinterpret_cast<const float4*>(brow + key));
  if (key < Lk) r.x = __ldg(brow + key);
  if (key + 1 < Lk) r.y = __ldg(brow + key + 1);
  if (key + 2 < Lk) r.z = __ldg(brow + key + 2);
  if (key + 3 < Lk) r.w = __ldg(brow + key + 3);
  return r;
}

// One softmax / output thread: owns query row q0 + r (TMEM lane r of its tile) for all nkv key tiles. bar_s: S tile ready,
// bar_p: P written (count 128), bar_o: partial O ready, bar_f: partial O consumed (count 128).
template <int kD>
__device__ __forceinline__ void attn_row_worker(const AttnArgs& a, int b, int h, int q0, int nkv, int off, int r,
                                                uint32_t lane_base, uint32_t tS, uint32_t tO, uint32_t sP,
                                                uint32_t bar_s, uint32_t bar_p, uint32_t bar_o, uint32_t bar_f) {
    // ------------------------------------------------------------------ softmax + output accumulation (one row each)
    const int q = q0 + r;
    constexpr float kLog2e = 1.4426950408889634f;
    const float sl2 = a.scale * kLog2e;
    const float* brow = nullptr;
    if (a.bias != nullptr && q < a.Lq) brow = a.bias + b * a.bias_sb + h * a.bias_sh + static_cast<long long>(q) * a.bias_sq;
    const int key_limit = a.causal ? min(a.Lk, q + off + 1) : a.Lk;   // keys [0, key_limit) are visible to this row
    const bool bvec = brow != nullptr && (reinterpret_cast<uintptr_t>(brow) & 15u) == 0;
    float m_run = -INFINITY, l_run = 0.f;
    float o_acc[kD];
#pragma unroll
    for (int c = 0; c < kD; ++c) o_acc[c] = 0.f;

    for (int j = 0; j < nkv; ++j) {
      const uint32_t ph = static_cast<uint32_t>(j) & 1u;
      const int k0 = j * kTileK;
      mbar_wait(bar_s, ph);
      tc_fence_after();
      // Interior tiles without a bias tensor need no masks and no per-element scale in the max pass: the softmax warps
      // are bound by instruction issue (one thread per row), so the common case is kept to ~4 instructions per logit.
      const bool plain = (brow == nullptr) && (k0 + kTileK <= key_limit);
      float m_tile = -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < kTileK; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32(tS + lane_base + static_cast<uint32_t>(c0), v);    // warp-collective: never inside a divergent branch
        tmem_ld_wait();
        if (plain) {
#pragma unroll
          for (int e = 0; e < 32; e += 2)
            m_tile = fmaxf(m_tile, fmaxf(__uint_as_float(v[e]), __uint_as_float(v[e + 1])));
        } else {
#pragma unroll
          for (int e4 = 0; e4 < 32; e4 += 4) {
            const float4 bv = attn_bias4(brow, k0 + c0 + e4, a.Lk, bvec);
            const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int key = k0 + c0 + e4 + u;
              const float s = fmaf(bb[u], kLog2e, __uint_as_float(v[e4 + u]) * sl2);
              if (key < key_limit) m_tile = fmaxf(m_tile, s);
            }
          }
        }
      }
      if (plain) m_tile *= sl2;                              // scale > 0: max commutes with it
      const float m_new = fmaxf(m_run, m_tile);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = ex2_approx(m_run - m_use);        // m_run = -inf -> 0
      // pass 2: p = 2^(s - m), row sum, P tile to shared memory
      float l_tile = 0.f;
#pragma unroll 1
      for (int c0 = 0; c0 < kTileK; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32(tS + lane_base + static_cast<uint32_t>(c0), v);
        tmem_ld_wait();
        float pv[32];
        if (plain) {
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            pv[e] = ex2_approx(fmaf(__uint_as_float(v[e]), sl2, -m_use));
            l_tile += pv[e];
          }
        } else {
#pragma unroll
          for (int e4 = 0; e4 < 32; e4 += 4) {
            const float4 bv = attn_bias4(brow, k0 + c0 + e4, a.Lk, bvec);
            const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int key = k0 + c0 + e4 + u;
              const float s = fmaf(bb[u], kLog2e, __uint_as_float(v[e4 + u]) * sl2);
              pv[e4 + u] = key < key_limit ? ex2_approx(s - m_use) : 0.f;
              l_tile += pv[e4 + u];
            }
          }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int col = c0 + 8 * g;                       // 8 keys = one 16-byte chunk of the swizzled row
          const uint32_t chunk = static_cast<uint32_t>((col & 63) >> 3);
          const uint32_t addr = sP + static_cast<uint32_t>(col >> 6) * kSub + static_cast<uint32_t>(r) * 128u +
                                ((chunk ^ (static_cast<uint32_t>(r) & 7u)) << 4);
          const uint32_t w0 = pack_bf16x2(pv[8 * g + 0], pv[8 * g + 1]), w1 = pack_bf16x2(pv[8 * g + 2], pv[8 * g + 3]);
          const uint32_t w2 = pack_bf16x2(pv[8 * g + 4], pv[8 * g + 5]), w3 = pack_bf16x2(pv[8 * g + 6], pv[8 * g + 7]);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar_p);
      l_run = l_run * alpha + l_tile;
      m_run = m_new;
      // fold the partial output of this tile
      mbar_wait(bar_o, ph);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < kD; c0 += 16) {
        uint32_t v[16];
        tmem_ld_16(tO + lane_base + static_cast<uint32_t>(c0), v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) o_acc[c0 + e] = fmaf(o_acc[c0 + e], alpha, __uint_as_float(v[e]));
      }
      tc_fence_before();
      mbar_arrive(bar_f);
    }

    if (q < a.Lq) {
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
      __nv_bfloat16* orow = a.O + b * a.o_sb + static_cast<long long>(q) * a.o_sl + h * kD;
#pragma unroll
      for (int c0 = 0; c0 < kD; c0 += 8) {
        uint4 w;
        w.x = pack_bf16x2(o_acc[c0 + 0] * inv, o_acc[c0 + 1] * inv);
        w.y = pack_bf16x2(o_acc[c0 + 2] * inv, o_acc[c0 + 3] * inv);
        w.z = pack_bf16x2(o_acc[c0 + 4] * inv, o_acc[c0 + 5] * inv);
        w.w = pack_bf16x2(o_acc[c0 + 6] * inv, o_acc[c0 + 7] * inv);
        *reinterpret_cast<uint4*>(orow + c0) = w;
      }
      if (a.lse != nullptr)
        a.lse[(static_cast<long long>(b) * a.heads + h) * a.Lq + q] =
            l_run > 0.f ? (m_run + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
    }
  }

template <int kD>
__global__ void __launch_bounds__(kAttThreads, kD == 64 ? 2 : 1)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnArgs a) {
  constexpr int kDS = kD / 64;                      // 64-column sub-tiles per operand
  constexpr uint32_t kOpBytes = kDS * kSub;         // Q, K or V tile
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sQ = smem_u32(smem), sK = sQ + kOpBytes, sV = sK + kOpBytes, sP = sV + kOpBytes;
  AttnBars* bars = reinterpret_cast<AttnBars*>(smem + 3 * kOpBytes + 2 * kSub);

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kTileQ, h = blockIdx.y, b = blockIdx.z;
  const int off = a.Lk - a.Lq;
  int nkv = (a.Lk + kTileK - 1) / kTileK;
  if (a.causal) {
    const int last_key = min(a.Lk - 1, q0 + kTileQ - 1 + off);
    nkv = last_key < 0 ? 0 : min(nkv, last_key / kTileK + 1);
  }

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars->q_full), 1);
    mbar_init(smem_u32(&bars->k_full), 1);
    mbar_init(smem_u32(&bars->k_empty), 1);
    mbar_init(smem_u32(&bars->v_full), 1);
    mbar_init(smem_u32(&bars->v_empty), 1);
    mbar_init(smem_u32(&bars->s_full), 1);
    mbar_init(smem_u32(&bars->p_full), 128);
    mbar_init(smem_u32(&bars->o_full), 1);
    mbar_init(smem_u32(&bars->o_free), 128);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc<256>(smem_u32(&bars->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  const uint32_t tS = tmem, tO = tmem + 128u;

  if (warp == 4) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && nkv > 0) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      mbar_expect_tx(smem_u32(&bars->q_full), kOpBytes);
      for (int s = 0; s < kDS; ++s) tma_load_3d(sQ + s * kSub, &tmQ, smem_u32(&bars->q_full), h * kD + 64 * s, q0, b);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = static_cast<uint32_t>(j) & 1u;
        if (j > 0) mbar_wait(smem_u32(&bars->k_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->k_full), kOpBytes);
        for (int s = 0; s < kDS; ++s)
          tma_load_3d(sK + s * kSub, &tmK, smem_u32(&bars->k_full), h * kD + 64 * s, j * kTileK, b);
        if (j > 0) mbar_wait(smem_u32(&bars->v_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->v_full), kOpBytes);
        for (int s = 0; s < kDS; ++s)
          tma_load_3d(sV + s * kSub, &tmV, smem_u32(&bars->v_full), h * kD + 64 * s, j * kTileK, b);
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && nkv > 0) {
      // idesc: c=f32 [4,6) | a=bf16 [7,10) | b=bf16 [10,13) | a_major 15 | b_major 16 | N>>3 [17,23) | M>>4 [24,29)
      const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kTileK >> 3) << 17) |
                               (static_cast<uint32_t>(kTileQ >> 4) << 24);
      const uint32_t idesc_o = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | (static_cast<uint32_t>(kD >> 3) << 17) |
                               (static_cast<uint32_t>(kTileQ >> 4) << 24);
      mbar_wait(smem_u32(&bars->q_full), 0);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = static_cast<uint32_t>(j) & 1u;
        mbar_wait(smem_u32(&bars->k_full), ph);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk) {
          const uint32_t o = static_cast<uint32_t>(kk / 4) * kSub + static_cast<uint32_t>(kk % 4) * 32u;
          umma_f16(tS, make_smem_desc_sw128(sQ + o, 16u, 1024u), make_smem_desc_sw128(sK + o, 16u, 1024u), idesc_s,
                   kk > 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&bars->s_full));
        umma_commit(smem_u32(&bars->k_empty));
        mbar_wait(smem_u32(&bars->p_full), ph);
        mbar_wait(smem_u32(&bars->v_full), ph);
        if (j > 0) mbar_wait(smem_u32(&bars->o_free), ph ^ 1u);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kTileK / 16; ++kk) {
          const uint32_t pa = sP + static_cast<uint32_t>(kk / 4) * kSub + static_cast<uint32_t>(kk % 4) * 32u;
          umma_f16(tO, make_smem_desc_sw128(pa, 16u, 1024u),
                   make_smem_desc_sw128(sV + static_cast<uint32_t>(kk) * 2048u, kSub, 1024u), idesc_o, kk > 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&bars->o_full));
        umma_commit(smem_u32(&bars->v_empty));
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + output accumulation (one row each)
    attn_row_worker<kD>(a, b, h, q0, nkv, off, static_cast<int>(threadIdx.x), (warp * 32u) << 16, tS, tO, sP,
                        smem_u32(&bars->s_full), smem_u32(&bars->p_full), smem_u32(&bars->o_full),
                        smem_u32(&bars->o_free));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc<256>(tmem);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Two query tiles per CTA (256 queries): warps 0-3 own tile 0, warps 4-7 tile 1, warp 8 = TMA, warp 9 = MMA. K_j / V_j
// are fetched once for both tiles, and while one group runs its softmax the tensor core works for the other
// (S1_j during softmax0_j, PV0_j during softmax1_j): inside one CTA instead of across two co-resident CTAs, with 8
// softmax warps per SM either way. TMEM: S0 [0,128) | S1 [128,256) | O0 [256,256+D) | O1 [256+D,256+2D).
constexpr int kAtt2Threads = 320;

struct __align__(8) Attn2Bars {
  uint64_t q_full, k_full, k_empty, v_full, v_empty;
  uint64_t s_full[2], p_full[2], o_full[2], o_free[2];
  uint32_t tmem_base;
};

template <int kD>
__global__ void __launch_bounds__(kAtt2Threads, 1)
attention_fwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const AttnArgs a) {
  constexpr int kDS = kD / 64;
  constexpr uint32_t kOpBytes = kDS * kSub;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sQ0 = smem_u32(smem), sQ1 = sQ0 + kOpBytes, sK = sQ1 + kOpBytes, sV = sK + kOpBytes;
  const uint32_t sP0 = sV + kOpBytes, sP1 = sP0 + 2 * kSub;
  Attn2Bars* bars = reinterpret_cast<Attn2Bars*>(smem + 4 * kOpBytes + 4 * kSub);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 2 * kTileQ, h = blockIdx.y, b = blockIdx.z;
  const int off = a.Lk - a.Lq;
  const int nkv_all = (a.Lk + kTileK - 1) / kTileK;
  int nkv_g[2];
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    const int qg = q0 + g * kTileQ;
    int n = qg < a.Lq ? nkv_all : 0;                 // a second tile past the end of the sequence does nothing
    if (a.causal && n > 0) {
      const int last_key = min(a.Lk - 1, qg + kTileQ - 1 + off);
      n = last_key < 0 ? 0 : min(n, last_key / kTileK + 1);
    }
    nkv_g[g] = n;
  }
  const int nkv = max(nkv_g[0], nkv_g[1]);

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars->q_full), 1);
    mbar_init(smem_u32(&bars->k_full), 1);
    mbar_init(smem_u32(&bars->k_empty), 1);
    mbar_init(smem_u32(&bars->v_full), 1);
    mbar_init(smem_u32(&bars->v_empty), 1);
    for (int g = 0; g < 2; ++g) {
      mbar_init(smem_u32(&bars->s_full[g]), 1);
      mbar_init(smem_u32(&bars->p_full[g]), 128);
      mbar_init(smem_u32(&bars->o_full[g]), 1);
      mbar_init(smem_u32(&bars->o_free[g]), 128);
    }
    fence_mbar_init();
  }
  if (warp == 9) tmem_alloc<512>(smem_u32(&bars->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;

  if (warp == 8) {
    if (lane == 0 && nkv > 0) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      mbar_expect_tx(smem_u32(&bars->q_full), 2 * kOpBytes);
      for (int s = 0; s < kDS; ++s) {
        tma_load_3d(sQ0 + s * kSub, &tmQ, smem_u32(&bars->q_full), h * kD + 64 * s, q0, b);
        tma_load_3d(sQ1 + s * kSub, &tmQ, smem_u32(&bars->q_full), h * kD + 64 * s, q0 + kTileQ, b);
      }
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = static_cast<uint32_t>(j) & 1u;
        if (j > 0) mbar_wait(smem_u32(&bars->k_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->k_full), kOpBytes);
        for (int s = 0; s < kDS; ++s)
          tma_load_3d(sK + s * kSub, &tmK, smem_u32(&bars->k_full), h * kD + 64 * s, j * kTileK, b);
        if (j > 0) mbar_wait(smem_u32(&bars->v_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->v_full), kOpBytes);
        for (int s = 0; s < kDS; ++s)
          tma_load_3d(sV + s * kSub, &tmV, smem_u32(&bars->v_full), h * kD + 64 * s, j * kTileK, b);
      }
    }
  } else if (warp == 9) {
    if (lane == 0 && nkv > 0) {
      const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kTileK >> 3) << 17) |
                               (static_cast<uint32_t>(kTileQ >> 4) << 24);
      const uint32_t idesc_o = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | (static_cast<uint32_t>(kD >> 3) << 17) |
                               (static_cast<uint32_t>(kTileQ >> 4) << 24);
      const uint32_t sQ[2] = {sQ0, sQ1}, sP[2] = {sP0, sP1};
      const uint32_t tS[2] = {tmem, tmem + 128u}, tO[2] = {tmem + 256u, tmem + 256u + kD};
      mbar_wait(smem_u32(&bars->q_full), 0);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = static_cast<uint32_t>(j) & 1u;
        mbar_wait(smem_u32(&bars->k_full), ph);
        tc_fence_after();
        // S_g = Q_g K_j^T for both tiles (the S buffer of tile g is free: its PV of step j-1 was issued after P_g(j-1))
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          if (j >= nkv_g[g]) continue;
#pragma unroll
          for (int kk = 0; kk < kD / 16; ++kk) {
            const uint32_t o = static_cast<uint32_t>(kk / 4) * kSub + static_cast<uint32_t>(kk % 4) * 32u;
            umma_f16(tS[g], make_smem_desc_sw128(sQ[g] + o, 16u, 1024u), make_smem_desc_sw128(sK + o, 16u, 1024u),
                     idesc_s, kk > 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(&bars->s_full[g]));
        }
        umma_commit(smem_u32(&bars->k_empty));
        mbar_wait(smem_u32(&bars->v_full), ph);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          if (j >= nkv_g[g]) continue;
          mbar_wait(smem_u32(&bars->p_full[g]), ph);
          if (j > 0) mbar_wait(smem_u32(&bars->o_free[g]), ph ^ 1u);
          tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < kTileK / 16; ++kk) {
            const uint32_t pa = sP[g] + static_cast<uint32_t>(kk / 4) * kSub + static_cast<uint32_t>(kk % 4) * 32u;
            umma_f16(tO[g], make_smem_desc_sw128(pa, 16u, 1024u),
                     make_smem_desc_sw128(sV + static_cast<uint32_t>(kk) * 2048u, kSub, 1024u), idesc_o,
                     kk > 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(&bars->o_full[g]));
        }
        umma_commit(smem_u32(&bars->v_empty));
      }
    }
  } else {
    const int g = static_cast<int>(warp >> 2);                      // 0 or 1: which query tile
    const int r = static_cast<int>(threadIdx.x) - g * 128;
    attn_row_worker<kD>(a, b, h, q0 + g * kTileQ, nkv_g[g], off, r, ((warp & 3u) * 32u) << 16,
                        tmem + static_cast<uint32_t>(g) * 128u, tmem + 256u + static_cast<uint32_t>(g) * kD,
                        g ? sP1 : sP0, smem_u32(&bars->s_full[g]), smem_u32(&bars->p_full[g]),
                        smem_u32(&bars->o_full[g]), smem_u32(&bars->o_free[g]));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// 1: two query tiles per CTA wherever Lq > 128; 0 / -1 (default): one tile per CTA, two CTAs per SM
static int g_attn_two_mode = [] { const char* e = getenv("BDL_ATTN_TWO"); return e ? atoi(e) : -1; }();

template <int kD>
static int launch_attention(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnArgs& a,
                            cudaStream_t s) {
  const size_t smem = 3 * (kD / 64) * kSub + 2 * kSub + sizeof(AttnBars) + 1024;
  const int two_mode = g_attn_two_mode;
  // two query tiles per CTA when there are at least two and the halved grid still covers the SMs
  const long long ctas2 = static_cast<long long>((a.Lq + 2 * kTileQ - 1) / (2 * kTileQ)) * a.heads * a.B;
  // Measured (profiles/attention_r2.md): the 256-query CTA is 8-20 % SLOWER than two co-resident 128-query CTAs (the two
  // softmax groups wait on one MMA issuer and one K/V stage), so it only runs when asked for.
  (void)ctas2;
  if (two_mode == 1 && a.Lq > kTileQ) {
    const size_t smem2 = 4 * (kD / 64) * kSub + 4 * kSub + sizeof(Attn2Bars) + 1024;
    auto k2 = attention_fwd2_kernel<kD>;
    cudaError_t e2 = cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem2));
    if (e2 != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(attention_fwd2)", e2);
    dim3 grid2((a.Lq + 2 * kTileQ - 1) / (2 * kTileQ), a.heads, a.B);
    k2<<<grid2, kAtt2Threads, smem2, s>>>(tq, tk, tv, a);
    return bdl_check_launch("attention_fwd2_kernel");
  }
  auto kfn = attention_fwd_kernel<kD>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(attention_fwd)", e);
  dim3 grid((a.Lq + kTileQ - 1) / kTileQ, a.heads, a.B);
  kfn<<<grid, kAttThreads, smem, s>>>(tq, tk, tv, a);
  return bdl_check_launch("attention_fwd_kernel");
}

}  // namespace bdl

extern "C" int bdl_attention_set_two_mode(int mode) {
  const int old = bdl::g_attn_two_mode;
  bdl::g_attn_two_mode = mode < 0 ? -1 : (mode > 0 ? 1 : 0);
  return old;
}

// Q [B, Lq, heads*D], K / V [B, Lk, heads*D] bf16 with row pitches q_sl / k_sl / v_sl and batch pitches *_sb (elements);
// O like Q. bias: fp32, strides in elements over (batch, head, query), key stride 1, 0 = broadcast; null = none.
extern "C" int bdl_attention_fwd(const void* Q, long long q_sb, long long q_sl, const void* K, long long k_sb,
                                 long long k_sl, const void* V, long long v_sb, long long v_sl, const float* bias,
                                 long long bias_sb, long long bias_sh, long long bias_sq, void* O, long long o_sb,
                                 long long o_sl, float* lse, int B, int heads, int Lq, int Lk, int D, float scale,
                                 int causal, void* stream) {
  using namespace bdl;
  if (B <= 0 || heads <= 0 || Lq <= 0 || Lk <= 0) return bdl_fail("bdl_attention_fwd: empty problem");
  if (D != 64 && D != 128) return bdl_fail("bdl_attention_fwd: head size must be 64 or 128");
  if ((q_sl % 8) || (k_sl % 8) || (v_sl % 8) || (q_sb % 8) || (k_sb % 8) || (v_sb % 8) || (o_sl % 8) || (o_sb % 8))
    return bdl_fail("bdl_attention_fwd: pitches must be multiples of 8 elements");
  if ((reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(K) | reinterpret_cast<uintptr_t>(V) |
       reinterpret_cast<uintptr_t>(O)) & 15u)
    return bdl_fail("bdl_attention_fwd: operands must be 16-byte aligned");
  if (B > 65535 || heads > 65535) return bdl_fail("bdl_attention_fwd: batch / heads exceed the grid limits");
  CUtensorMap tq, tk, tv;
  const uint64_t width = static_cast<uint64_t>(heads) * D;
  int rc = bdl_make_tmap_3d_bf16(&tq, Q, width, Lq, B, q_sl, B > 1 ? q_sb : q_sl * Lq, 64, 128);
  if (rc) return rc;
  rc = bdl_make_tmap_3d_bf16(&tk, K, width, Lk, B, k_sl, B > 1 ? k_sb : k_sl * Lk, 64, 128);
  if (rc) return rc;
  rc = bdl_make_tmap_3d_bf16(&tv, V, width, Lk, B, v_sl, B > 1 ? v_sb : v_sl * Lk, 64, 128);
  if (rc) return rc;
  AttnArgs a{};
  a.bias = bias; a.bias_sb = bias_sb; a.bias_sh = bias_sh; a.bias_sq = bias_sq;
  a.O = static_cast<__nv_bfloat16*>(O); a.o_sb = o_sb; a.o_sl = o_sl; a.lse = lse;
  a.B = B; a.heads = heads; a.Lq = Lq; a.Lk = Lk; a.scale = scale; a.causal = causal;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return D == 64 ? launch_attention<64>(tq, tk, tv, a, s) : launch_attention<128>(tq, tk, tv, a, s);
}
