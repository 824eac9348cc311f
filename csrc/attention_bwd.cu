// Fused multi-head attention backward on tcgen05: the probabilities are recomputed from the saved row log-sum-exp, so
// neither the [B, heads, Lq, Lk] logits nor their gradient ever reach HBM (S/nn/Attention.scala's backward runs the MM /
// SoftMax / transpose modules in reverse and materialises both).
//   P = exp(scale * Q K^T + bias - lse),  dP = dO V^T,  Delta = rowsum(dO o O),  dS = scale * P o (dP - Delta)
//   dQ = dS K          -> attention_dq_kernel : CTA per (batch, head, 128 queries), streams K_j / V_j; also writes Delta
//   dV = P^T dO, dK = dS^T Q -> attention_dkv_kernel: CTA per (batch, head, 128 keys), streams Q_i / dO_i and works on
//                              the TRANSPOSED tiles S^T = K Q^T, dP^T = V dO^T so that a thread owns one key row.
// Same roles as the forward kernel: warp 4 = TMA, warp 5 = MMA issue, warps 0-3 = one thread per TMEM lane.
#include <cuda_bf16.h>
#include <math.h>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

constexpr int kBwThreads = 192;
constexpr int kBT = 128;                 // tile edge (queries and keys)
constexpr uint32_t kBSub = 16384u;       // [128 rows x 64 bf16] swizzled sub-tile
constexpr float kL2e = 1.4426950408889634f;

struct AttnBwdArgs {
  const float* bias;
  long long bias_sb, bias_sh, bias_sq;
  const float* lse;                      // [B, heads, Lq]
  float* delta;                          // [B, heads, Lq] scratch: written by the dQ kernel, read by the dK/dV kernel
  const __nv_bfloat16* O;                // forward output and its gradient, [B, Lq, heads*D]
  const __nv_bfloat16* dO;
  long long o_sb, o_sl, do_sb, do_sl;
  __nv_bfloat16* dQ; long long dq_sb, dq_sl;
  __nv_bfloat16* dK; long long dk_sb, dk_sl;
  __nv_bfloat16* dV; long long dv_sb, dv_sl;
  int B, heads, Lq, Lk;
  float scale;
  int causal;
};

// 8 consecutive K-elements (16 bytes) of row `r`, columns [col, col+8), of a [128 x 128] bf16 K-major tile made of two
// 128B-swizzled sub-tiles of 64 columns.
__device__ __forceinline__ void store_kmajor8(uint32_t tile, int r, int col, const float* v) {
  const uint32_t chunk = static_cast<uint32_t>((col & 63) >> 3);
  const uint32_t addr = tile + static_cast<uint32_t>(col >> 6) * kBSub + static_cast<uint32_t>(r) * 128u +
                        ((chunk ^ (static_cast<uint32_t>(r) & 7u)) << 4);
  const uint32_t w0 = pack_bf16x2(v[0], v[1]), w1 = pack_bf16x2(v[2], v[3]);
  const uint32_t w2 = pack_bf16x2(v[4], v[5]), w3 = pack_bf16x2(v[6], v[7]);
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
}

__device__ __forceinline__ uint32_t kmajor_step(int kk) {          // byte offset of K step kk (16 elements) in such a tile
  return static_cast<uint32_t>(kk / 4) * kBSub + static_cast<uint32_t>(kk % 4) * 32u;
}

// tcgen05.ld is warp-collective (.sync.aligned): every thread of the warp runs the loads, only the store is predicated.
template <int kD>
__device__ __forceinline__ void store_row_bf16(__nv_bfloat16* dst, uint32_t taddr, float mul, bool pred) {
#pragma unroll
  for (int c0 = 0; c0 < kD; c0 += 16) {
    uint32_t v[16];
    tmem_ld_16(taddr + static_cast<uint32_t>(c0), v);
    tmem_ld_wait();
    uint4 w0, w1;
    w0.x = pack_bf16x2(__uint_as_float(v[0]) * mul, __uint_as_float(v[1]) * mul);
    w0.y = pack_bf16x2(__uint_as_float(v[2]) * mul, __uint_as_float(v[3]) * mul);
    w0.z = pack_bf16x2(__uint_as_float(v[4]) * mul, __uint_as_float(v[5]) * mul);
    w0.w = pack_bf16x2(__uint_as_float(v[6]) * mul, __uint_as_float(v[7]) * mul);
    w1.x = pack_bf16x2(__uint_as_float(v[8]) * mul, __uint_as_float(v[9]) * mul);
    w1.y = pack_bf16x2(__uint_as_float(v[10]) * mul, __uint_as_float(v[11]) * mul);
    w1.z = pack_bf16x2(__uint_as_float(v[12]) * mul, __uint_as_float(v[13]) * mul);
    w1.w = pack_bf16x2(__uint_as_float(v[14]) * mul, __uint_as_float(v[15]) * mul);
    if (pred) {
      *reinterpret_cast<uint4*>(dst + c0) = w0;
      *reinterpret_cast<uint4*>(dst + c0 + 8) = w1;
    }
  }
}

struct __align__(8) BwdBars {
  uint64_t once_full, a_full, a_empty, b_full, b_empty, sd_full, ds_full, acc_full;
  uint32_t tmem_base;
};

// ---------------------------------------------------------------------------------------------------- dQ (+ Delta)
// shared memory: Q | dO | K_j | V_j | dS (K-major [q x key]) ; TMEM: S [0,128) | dP [128,256) | dQ [256, 256+D)
template <int kD>
__global__ void __launch_bounds__(kBwThreads, 1)
attention_dq_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                    const AttnBwdArgs a) {
  constexpr int kDS = kD / 64;
  constexpr uint32_t kOp = kDS * kBSub;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sQ = smem_u32(smem), sdO = sQ + kOp, sK = sdO + kOp, sV = sK + kOp, sDS = sV + kOp;
  BwdBars* bars = reinterpret_cast<BwdBars*>(smem + 4 * kOp + 2 * kBSub);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kBT, h = blockIdx.y, b = blockIdx.z;
  const int off = a.Lk - a.Lq;
  int nkv = (a.Lk + kBT - 1) / kBT;
  if (a.causal) {
    const int last_key = min(a.Lk - 1, q0 + kBT - 1 + off);
    nkv = last_key < 0 ? 0 : min(nkv, last_key / kBT + 1);
  }
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars->once_full), 1);
    mbar_init(smem_u32(&bars->a_full), 1);   // K_j
    mbar_init(smem_u32(&bars->a_empty), 1);
    mbar_init(smem_u32(&bars->b_full), 1);   // V_j
    mbar_init(smem_u32(&bars->b_empty), 1);
    mbar_init(smem_u32(&bars->sd_full), 1);
    mbar_init(smem_u32(&bars->ds_full), 128);
    mbar_init(smem_u32(&bars->acc_full), 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc<512>(smem_u32(&bars->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  const uint32_t tS = tmem, tP = tmem + 128u, tQ = tmem + 256u;

  if (warp == 4) {
    if (lane == 0 && nkv > 0) {
      mbar_expect_tx(smem_u32(&bars->once_full), 2 * kOp);
      for (int s = 0; s < kDS; ++s) {
        tma_load_3d(sQ + s * kBSub, &tmQ, smem_u32(&bars->once_full), h * kD + 64 * s, q0, b);
        tma_load_3d(sdO + s * kBSub, &tmdO, smem_u32(&bars->once_full), h * kD + 64 * s, q0, b);
      }
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = static_cast<uint32_t>(j) & 1u;
        if (j > 0) mbar_wait(smem_u32(&bars->b_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->b_full), kOp);
        for (int s = 0; s < kDS; ++s) tma_load_3d(sV + s * kBSub, &tmV, smem_u32(&bars->b_full), h * kD + 64 * s, j * kBT, b);
        if (j > 0) mbar_wait(smem_u32(&bars->a_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->a_full), kOp);
        for (int s = 0; s < kDS; ++s) tma_load_3d(sK + s * kBSub, &tmK, smem_u32(&bars->a_full), h * kD + 64 * s, j * kBT, b);
      }
    }
  } else if (warp == 5) {
    if (lane == 0 && nkv > 0) {
      const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kBT >> 3) << 17) |
                               (static_cast<uint32_t>(kBT >> 4) << 24);
      const uint32_t idesc_d = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | (static_cast<uint32_t>(kD >> 3) << 17) |
                               (static_cast<uint32_t>(kBT >> 4) << 24);
      mbar_wait(smem_u32(&bars->once_full), 0);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = static_cast<uint32_t>(j) & 1u;
        mbar_wait(smem_u32(&bars->a_full), ph);
        mbar_wait(smem_u32(&bars->b_full), ph);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk)       // S = Q K^T
          umma_f16(tS, make_smem_desc_sw128(sQ + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sK + kmajor_step(kk), 16u, 1024u), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk)       // dP = dO V^T
          umma_f16(tP, make_smem_desc_sw128(sdO + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sV + kmajor_step(kk), 16u, 1024u), idesc_s, kk > 0 ? 1u : 0u);
        umma_commit(smem_u32(&bars->sd_full));
        umma_commit(smem_u32(&bars->b_empty));
        mbar_wait(smem_u32(&bars->ds_full), ph);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kBT / 16; ++kk)      // dQ += dS K   (K_j as an MN-major B operand)
          umma_f16(tQ, make_smem_desc_sw128(sDS + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sK + static_cast<uint32_t>(kk) * 2048u, kBSub, 1024u), idesc_d,
                   (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit(smem_u32(&bars->a_empty));
      }
      umma_commit(smem_u32(&bars->acc_full));
    }
  } else {
    const int r = static_cast<int>(threadIdx.x);
    const int q = q0 + r;
    const uint32_t lane_base = (warp * 32u) << 16;
    const float sl2 = a.scale * kL2e;
    const bool live = q < a.Lq;
    const float* brow = nullptr;
    if (a.bias != nullptr && live) brow = a.bias + b * a.bias_sb + h * a.bias_sh + static_cast<long long>(q) * a.bias_sq;
    const int key_limit = a.causal ? min(a.Lk, q + off + 1) : a.Lk;
    const bool bvec = brow != nullptr && (reinterpret_cast<uintptr_t>(brow) & 15u) == 0;
    const long long stat = (static_cast<long long>(b) * a.heads + h) * a.Lq + q;
    float lse2 = 0.f, delta = 0.f;
    if (live) {
      lse2 = a.lse[stat] * kL2e;
      const __nv_bfloat16* orow = a.O + b * a.o_sb + static_cast<long long>(q) * a.o_sl + h * kD;
      const __nv_bfloat16* grow = a.dO + b * a.do_sb + static_cast<long long>(q) * a.do_sl + h * kD;
#pragma unroll
      for (int c = 0; c < kD; c += 8) {
        const uint4 ov = *reinterpret_cast<const uint4*>(orow + c), gv = *reinterpret_cast<const uint4*>(grow + c);
        const uint32_t ow[4] = {ov.x, ov.y, ov.z, ov.w}, gw[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          delta = fmaf(__uint_as_float(ow[e] << 16), __uint_as_float(gw[e] << 16), delta);
          delta = fmaf(__uint_as_float(ow[e] & 0xFFFF0000u), __uint_as_float(gw[e] & 0xFFFF0000u), delta);
        }
      }
      a.delta[stat] = delta;
    }
    for (int j = 0; j < nkv; ++j) {
      const uint32_t ph = static_cast<uint32_t>(j) & 1u;
      const int k0 = j * kBT;
      mbar_wait(smem_u32(&bars->sd_full), ph);
      tc_fence_after();
      const bool plain = brow == nullptr && live && (k0 + kBT <= key_limit);
      const float nds = -delta * a.scale;
#pragma unroll 1
      for (int c0 = 0; c0 < kBT; c0 += 16) {
        uint32_t sv[16], pv[16];
        tmem_ld_16(tS + lane_base + static_cast<uint32_t>(c0), sv);
        tmem_ld_16(tP + lane_base + static_cast<uint32_t>(c0), pv);
        tmem_ld_wait();
        float ds[16];
        if (plain) {                                   // interior tile, no bias tensor: 4 instructions per logit
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float p = ex2_approx(fmaf(__uint_as_float(sv[e]), sl2, -lse2));
            ds[e] = p * fmaf(__uint_as_float(pv[e]), a.scale, nds);
          }
        } else {
#pragma unroll
          for (int e4 = 0; e4 < 16; e4 += 4) {
            float bb[4] = {0.f, 0.f, 0.f, 0.f};
            const int kb = k0 + c0 + e4;
            if (brow != nullptr) {
              if (bvec && kb + 3 < a.Lk) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(brow + kb));
                bb[0] = t.x; bb[1] = t.y; bb[2] = t.z; bb[3] = t.w;
              } else {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                  if (kb + u < a.Lk) bb[u] = __ldg(brow + kb + u);
              }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = e4 + u;
              const float s = fmaf(bb[u], kL2e, __uint_as_float(sv[e]) * sl2);
              const float p = (live && kb + u < key_limit) ? ex2_approx(s - lse2) : 0.f;
              ds[e] = p * fmaf(__uint_as_float(pv[e]), a.scale, nds);
            }
          }
        }
        store_kmajor8(sDS, r, c0, ds);
        store_kmajor8(sDS, r, c0 + 8, ds + 8);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars->ds_full));
    }
    if (nkv > 0) {
      mbar_wait(smem_u32(&bars->acc_full), 0);
      tc_fence_after();
      store_row_bf16<kD>(a.dQ + b * a.dq_sb + static_cast<long long>(q) * a.dq_sl + h * kD, tQ + lane_base, 1.f, live);
    } else if (live) {
      __nv_bfloat16* d = a.dQ + b * a.dq_sb + static_cast<long long>(q) * a.dq_sl + h * kD;
      for (int c = 0; c < kD; c += 8) *reinterpret_cast<uint4*>(d + c) = make_uint4(0, 0, 0, 0);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// ---------------------------------------------------------------------------------------------------- dK, dV
// shared memory: K | V | Q_i | dO_i | P^T | dS^T (both K-major [key x q]) | lse, delta of the q tile
// TMEM: S^T [0,128) | dP^T [128,256) | dV [256,256+D) | dK [256+D, 256+2D)
template <int kD>
__global__ void __launch_bounds__(kBwThreads, 1)
attention_dkv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                     const AttnBwdArgs a) {
  constexpr int kDS = kD / 64;
  constexpr uint32_t kOp = kDS * kBSub;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sK = smem_u32(smem), sV = sK + kOp, sQ = sV + kOp, sdO = sQ + kOp, sPT = sdO + kOp, sDST = sPT + 2 * kBSub;
  float* s_lse = reinterpret_cast<float*>(smem + 4 * kOp + 4 * kBSub);
  float* s_delta = s_lse + kBT;
  BwdBars* bars = reinterpret_cast<BwdBars*>(s_delta + kBT);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * kBT, h = blockIdx.y, b = blockIdx.z;
  const int off = a.Lk - a.Lq;
  const int nq = (a.Lq + kBT - 1) / kBT;
  int i0 = 0;                                        // causal: query tiles entirely before this key tile see none of it
  if (a.causal) i0 = max(0, (k0 - off) / kBT);
  if (i0 > nq) i0 = nq;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars->once_full), 1);
    mbar_init(smem_u32(&bars->a_full), 1);   // Q_i + dO_i
    mbar_init(smem_u32(&bars->a_empty), 1);
    mbar_init(smem_u32(&bars->b_full), 128); // lse / delta staged by the row threads
    mbar_init(smem_u32(&bars->b_empty), 1);
    mbar_init(smem_u32(&bars->sd_full), 1);
    mbar_init(smem_u32(&bars->ds_full), 128);
    mbar_init(smem_u32(&bars->acc_full), 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc<512>(smem_u32(&bars->tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  const uint32_t tS = tmem, tP = tmem + 128u, tV = tmem + 256u, tK = tmem + 256u + kD;
  const bool any = i0 < nq;

  if (warp == 4) {
    if (lane == 0 && any) {
      mbar_expect_tx(smem_u32(&bars->once_full), 2 * kOp);
      for (int s = 0; s < kDS; ++s) {
        tma_load_3d(sK + s * kBSub, &tmK, smem_u32(&bars->once_full), h * kD + 64 * s, k0, b);
        tma_load_3d(sV + s * kBSub, &tmV, smem_u32(&bars->once_full), h * kD + 64 * s, k0, b);
      }
      for (int i = i0; i < nq; ++i) {
        const uint32_t ph = static_cast<uint32_t>(i - i0) & 1u;
        if (i > i0) mbar_wait(smem_u32(&bars->a_empty), ph ^ 1u);
        mbar_expect_tx(smem_u32(&bars->a_full), 2 * kOp);
        for (int s = 0; s < kDS; ++s) {
          tma_load_3d(sQ + s * kBSub, &tmQ, smem_u32(&bars->a_full), h * kD + 64 * s, i * kBT, b);
          tma_load_3d(sdO + s * kBSub, &tmdO, smem_u32(&bars->a_full), h * kD + 64 * s, i * kBT, b);
        }
      }
    }
  } else if (warp == 5) {
    if (lane == 0 && any) {
      const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kBT >> 3) << 17) |
                               (static_cast<uint32_t>(kBT >> 4) << 24);
      const uint32_t idesc_d = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | (static_cast<uint32_t>(kD >> 3) << 17) |
                               (static_cast<uint32_t>(kBT >> 4) << 24);
      mbar_wait(smem_u32(&bars->once_full), 0);
      for (int i = i0; i < nq; ++i) {
        const uint32_t ph = static_cast<uint32_t>(i - i0) & 1u;
        mbar_wait(smem_u32(&bars->a_full), ph);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk)       // S^T = K Q^T
          umma_f16(tS, make_smem_desc_sw128(sK + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sQ + kmajor_step(kk), 16u, 1024u), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk)       // dP^T = V dO^T
          umma_f16(tP, make_smem_desc_sw128(sV + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sdO + kmajor_step(kk), 16u, 1024u), idesc_s, kk > 0 ? 1u : 0u);
        umma_commit(smem_u32(&bars->sd_full));
        mbar_wait(smem_u32(&bars->ds_full), ph);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kBT / 16; ++kk)      // dV += P^T dO_i   (dO_i as an MN-major B operand)
          umma_f16(tV, make_smem_desc_sw128(sPT + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sdO + static_cast<uint32_t>(kk) * 2048u, kBSub, 1024u), idesc_d,
                   (i > i0 || kk > 0) ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kBT / 16; ++kk)      // dK += dS^T Q_i
          umma_f16(tK, make_smem_desc_sw128(sDST + kmajor_step(kk), 16u, 1024u),
                   make_smem_desc_sw128(sQ + static_cast<uint32_t>(kk) * 2048u, kBSub, 1024u), idesc_d,
                   (i > i0 || kk > 0) ? 1u : 0u);
        umma_commit(smem_u32(&bars->a_empty));
        umma_commit(smem_u32(&bars->b_empty));
      }
      umma_commit(smem_u32(&bars->acc_full));
    }
  } else {
    const int r = static_cast<int>(threadIdx.x);
    const int key = k0 + r;
    const uint32_t lane_base = (warp * 32u) << 16;
    const float sl2 = a.scale * kL2e;
    const bool live = key < a.Lk;
    const float* bcol = nullptr;                       // bias[b, h, q, key]: consecutive threads -> consecutive keys
    if (a.bias != nullptr && live) bcol = a.bias + b * a.bias_sb + h * a.bias_sh + key;
    const long long stat0 = (static_cast<long long>(b) * a.heads + h) * a.Lq;
    for (int i = i0; i < nq; ++i) {
      const uint32_t ph = static_cast<uint32_t>(i - i0) & 1u;
      const int qb = i * kBT;
      // stage lse / delta of this query tile (the previous tile's values are dead once its dS^T was written)
      if (i > i0) mbar_wait(smem_u32(&bars->b_empty), ph ^ 1u);
      {
        const int q = qb + r;
        s_lse[r] = q < a.Lq ? a.lse[stat0 + q] * kL2e : 0.f;
        s_delta[r] = q < a.Lq ? a.delta[stat0 + q] * a.scale : 0.f;      // pre-scaled: dS = P (scale dP - scale Delta)
      }
      mbar_arrive(smem_u32(&bars->b_full));
      mbar_wait(smem_u32(&bars->b_full), ph);
      mbar_wait(smem_u32(&bars->sd_full), ph);
      tc_fence_after();
      // whole query tile visible to this key, inside Lq, no bias tensor -> no masks
      const bool plain = bcol == nullptr && a.bias == nullptr && live && (qb + kBT <= a.Lq) && (!a.causal || key <= qb + off);
#pragma unroll 1
      for (int c0 = 0; c0 < kBT; c0 += 16) {
        uint32_t sv[16], pv[16];
        tmem_ld_16(tS + lane_base + static_cast<uint32_t>(c0), sv);
        tmem_ld_16(tP + lane_base + static_cast<uint32_t>(c0), pv);
        tmem_ld_wait();
        float pt[16], ds[16];
        float ls[16], dl[16];
#pragma unroll
        for (int e = 0; e < 16; e += 4) {
          *reinterpret_cast<float4*>(ls + e) = *reinterpret_cast<const float4*>(s_lse + c0 + e);
          *reinterpret_cast<float4*>(dl + e) = *reinterpret_cast<const float4*>(s_delta + c0 + e);
        }
        if (plain) {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            pt[e] = ex2_approx(fmaf(__uint_as_float(sv[e]), sl2, -ls[e]));
            ds[e] = pt[e] * (fmaf(__uint_as_float(pv[e]), a.scale, -dl[e]));
          }
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int q = qb + c0 + e;
            float s = __uint_as_float(sv[e]) * sl2;
            if (bcol != nullptr && q < a.Lq) s += __ldg(bcol + static_cast<long long>(q) * a.bias_sq) * kL2e;
            const bool vis = live && q < a.Lq && (!a.causal || key <= q + off);
            pt[e] = vis ? ex2_approx(s - ls[e]) : 0.f;
            ds[e] = pt[e] * (fmaf(__uint_as_float(pv[e]), a.scale, -dl[e]));
          }
        }
        store_kmajor8(sPT, r, c0, pt);
        store_kmajor8(sPT, r, c0 + 8, pt + 8);
        store_kmajor8(sDST, r, c0, ds);
        store_kmajor8(sDST, r, c0 + 8, ds + 8);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars->ds_full));
    }
    __nv_bfloat16* dv = a.dV + b * a.dv_sb + static_cast<long long>(key) * a.dv_sl + h * kD;
    __nv_bfloat16* dk = a.dK + b * a.dk_sb + static_cast<long long>(key) * a.dk_sl + h * kD;
    if (any) {
      mbar_wait(smem_u32(&bars->acc_full), 0);
      tc_fence_after();
      store_row_bf16<kD>(dv, tV + lane_base, 1.f, live);
      store_row_bf16<kD>(dk, tK + lane_base, 1.f, live);
    } else if (live) {
      for (int c = 0; c < kD; c += 8) {
        *reinterpret_cast<uint4*>(dv + c) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(dk + c) = make_uint4(0, 0, 0, 0);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

template <int kD>
static int launch_attention_bwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                                const CUtensorMap& tdo, const AttnBwdArgs& a, cudaStream_t s) {
  constexpr uint32_t kOp = (kD / 64) * kBSub;
  const size_t smem_dq = 4 * kOp + 2 * kBSub + sizeof(BwdBars) + 1024;
  const size_t smem_dkv = 4 * kOp + 4 * kBSub + 2 * kBT * sizeof(float) + sizeof(BwdBars) + 1024;
  auto kq = attention_dq_kernel<kD>;
  auto kkv = attention_dkv_kernel<kD>;
  cudaError_t e = cudaFuncSetAttribute(kq, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_dq));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(attention_dq)", e);
  e = cudaFuncSetAttribute(kkv, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_dkv));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(attention_dkv)", e);
  kq<<<dim3((a.Lq + kBT - 1) / kBT, a.heads, a.B), kBwThreads, smem_dq, s>>>(tq, tk, tv, tdo, a);
  int rc = bdl_check_launch("attention_dq_kernel");
  if (rc) return rc;
  kkv<<<dim3((a.Lk + kBT - 1) / kBT, a.heads, a.B), kBwThreads, smem_dkv, s>>>(tq, tk, tv, tdo, a);
  return bdl_check_launch("attention_dkv_kernel");
}

}  // namespace bdl

// Gradients of bdl_attention_fwd w.r.t. Q, K, V (all [B, L, heads*D] bf16 with element pitches). lse is the forward's
// row log-sum-exp, delta a [B, heads, Lq] fp32 scratch. The bias gets no gradient.
extern "C" int bdl_attention_bwd(const void* Q, long long q_sb, long long q_sl, const void* K, long long k_sb,
                                 long long k_sl, const void* V, long long v_sb, long long v_sl, const void* O,
                                 long long o_sb, long long o_sl, const void* dO, long long do_sb, long long do_sl,
                                 const float* lse, float* delta, const float* bias, long long bias_sb, long long bias_sh,
                                 long long bias_sq, void* dQ, long long dq_sb, long long dq_sl, void* dK, long long dk_sb,
                                 long long dk_sl, void* dV, long long dv_sb, long long dv_sl, int B, int heads, int Lq,
                                 int Lk, int D, float scale, int causal, void* stream) {
  using namespace bdl;
  if (B <= 0 || heads <= 0 || Lq <= 0 || Lk <= 0) return bdl_fail("bdl_attention_bwd: empty problem");
  if (D != 64 && D != 128) return bdl_fail("bdl_attention_bwd: head size must be 64 or 128");
  const long long pitches[] = {q_sb, q_sl, k_sb, k_sl, v_sb, v_sl, o_sb, o_sl, do_sb, do_sl, dq_sb, dq_sl, dk_sb, dk_sl,
                               dv_sb, dv_sl};
  for (long long p : pitches)
    if (p % 8) return bdl_fail("bdl_attention_bwd: pitches must be multiples of 8 elements");
  const void* ptrs[] = {Q, K, V, O, dO, dQ, dK, dV};
  for (const void* p : ptrs)
    if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u)) return bdl_fail("bdl_attention_bwd: operands must be 16-byte aligned");
  if (lse == nullptr || delta == nullptr) return bdl_fail("bdl_attention_bwd: lse / delta missing");
  if (B > 65535 || heads > 65535) return bdl_fail("bdl_attention_bwd: batch / heads exceed the grid limits");
  CUtensorMap tq, tk, tv, tdo;
  const uint64_t width = static_cast<uint64_t>(heads) * D;
  int rc = bdl_make_tmap_3d_bf16(&tq, Q, width, Lq, B, q_sl, B > 1 ? q_sb : q_sl * Lq, 64, 128);
  if (rc) return rc;
  rc = bdl_make_tmap_3d_bf16(&tk, K, width, Lk, B, k_sl, B > 1 ? k_sb : k_sl * Lk, 64, 128);
  if (rc) return rc;
  rc = bdl_make_tmap_3d_bf16(&tv, V, width, Lk, B, v_sl, B > 1 ? v_sb : v_sl * Lk, 64, 128);
  if (rc) return rc;
  rc = bdl_make_tmap_3d_bf16(&tdo, dO, width, Lq, B, do_sl, B > 1 ? do_sb : do_sl * Lq, 64, 128);
  if (rc) return rc;
  AttnBwdArgs a{};
  a.bias = bias; a.bias_sb = bias_sb; a.bias_sh = bias_sh; a.bias_sq = bias_sq;
  a.lse = lse; a.delta = delta;
  a.O = static_cast<const __nv_bfloat16*>(O); a.dO = static_cast<const __nv_bfloat16*>(dO);
  a.o_sb = o_sb; a.o_sl = o_sl; a.do_sb = do_sb; a.do_sl = do_sl;
  a.dQ = static_cast<__nv_bfloat16*>(dQ); a.dq_sb = dq_sb; a.dq_sl = dq_sl;
  a.dK = static_cast<__nv_bfloat16*>(dK); a.dk_sb = dk_sb; a.dk_sl = dk_sl;
  a.dV = static_cast<__nv_bfloat16*>(dV); a.dv_sb = dv_sb; a.dv_sl = dv_sl;
  a.B = B; a.heads = heads; a.Lq = Lq; a.Lk = Lk; a.scale = scale; a.causal = causal;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return D == 64 ? launch_attention_bwd<64>(tq, tk, tv, tdo, a, s) : launch_attention_bwd<128>(tq, tk, tv, tdo, a, s);
}
