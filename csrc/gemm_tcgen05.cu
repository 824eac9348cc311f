// bf16 GEMM on 5th-gen tensor cores: D[M,N] = opA(A) * opB(B)^T (+bias)(ReLU), fp32 accumulation in TMEM.
//
// One persistent CTA per SM, 10 warps:
//   warp 0      TMA producer   (cp.async.bulk.tensor.2d -> 128B-swizzled smem ring, mbarrier complete_tx)
//   warp 1      MMA issuer     (one lane issues tcgen05.mma.cta_group::1.kind::f16, tcgen05.commit frees slots)
//   warps 2..9  epilogue       (tcgen05.ld TMEM -> registers -> scale/bias/ReLU -> bf16|fp32|e4m3 global, or fp32
//               red.add). Two warps per TMEM lane quadrant split a tile's columns; each keeps the next 32 columns'
//               tcgen05.ld in flight while it converts and stores the current ones, and the tile's bias/scale columns
//               are staged in shared memory before the accumulator is ready (small-K GEMMs are epilogue-bound).
// Two TMEM accumulator stages (2 x 256 columns) let the epilogue of tile i overlap the MMAs of tile i+1.
//
// This single kernel is the engine behind Linear fwd/bwd, 1x1 conv, im2col conv fprop/dgrad/wgrad and the
// LSTM gate GEMMs. It replaces the reference's MKL sgemm JNI (core/mkl2017-xeon-blas/src/main/c/mkl.c,
// S/tensor/DenseTensorBLAS.scala gemm) and MKL-DNN inner product (S/nn/mkldnn/Linear.scala).
//
// Operand "majors": K-major means the contraction index is contiguous in global memory ([rows, K] row-major);
// MN-major means the non-contraction index is contiguous ([K, rows] row-major). wgrad needs MN-major for both
// operands, dgrad needs an MN-major B; the tensor core consumes both directly (idesc a_major/b_major bits), so no
// transposes are ever materialised.
#include <cuda_fp8.h>

#include <cstdlib>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
// dynamic shared memory available to the operand ring: 227 KB minus alignment slack, barriers and the epilogue columns
constexpr size_t kSmemBudget = 227 * 1024 - 1024 - 512 - 4096;
constexpr int kMaxStages = 8;
constexpr int kGemmThreads = 320;   // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quadrant)
constexpr int kEpiThreads = 256;
constexpr int kTmemCols = 512;
constexpr int kAccStride = 256;  // TMEM columns between the two accumulator stages

// Division by a runtime constant without the ~100-cycle integer divide sequence: the single producer thread decodes
// pixel / tap coordinates for every k block, so its divides sit on the critical path of the TMA issue rate.
struct FastDiv {
  uint32_t mul, shr, div;
  __host__ void init(uint32_t d) {
    div = d;
    if (d <= 1) { mul = 0; shr = 0; return; }
    uint32_t l = 0;
    while ((1ull << l) < d) ++l;
    mul = static_cast<uint32_t>(((1ull << 32) * ((1ull << l) - d)) / d + 1);
    shr = l;
  }
  __device__ __forceinline__ uint32_t quot(uint32_t n) const {
    if (div <= 1) return n;
    const uint32_t t = __umulhi(n, mul);
    return (t + ((n - t) >> 1)) >> (shr - 1);
  }
};

struct GemmParams {
  int M, N, K;
  int block_n;     // UMMA N (multiple of 16, <= 256)
  int b_rows;      // smem rows reserved for B per stage
  int num_stages;
  int splits;      // split-K factor (>=1)
  int tiles_m, tiles_n;
  long long ldd;
  int out_fp32;
  int flags;
  void* D;
  const float* bias;
  const float* col_scale;    // fp8 only: per-output-column dequantisation scale
  const float* out_qscale;   // fp8 only, out_fp32 == 2: device scalar 448 / absmax of the next layer's input
  float* stat_sums;                // optional [2 * N] fp32: per-column sum and sum of squares of the (bf16-rounded) output are
                                   // ADDED here by the epilogue — the batch-norm statistics of a following BN without a pass
                                   // over the output (S/nn/mkldnn/SpatialBatchNormalization.scala needs its own pass)
  const __nv_bfloat16* residual;   // optional [M, N] bf16 tensor (row pitch ldr) added before the ReLU: the shortcut of a
  long long ldr;                   // residual block, so conv + add + ReLU is one pass (reference: Fusion.fusionCAddTable)
  int res_mask;                    // 1: the operand is not added but used as a ReLU mask, out = (residual > 0 ? out : 0) —
                                   // dgrad of a layer whose input is a ReLU output: the ReLU backward of the PRODUCING layer
                                   // happens here instead of in a pass of its own
  // implicit-GEMM convolution (kConv != 0): the A (fprop/dgrad) or B (wgrad) operand is gathered from an NHWC
  // activation tensor by TMA im2col loads instead of being read from a materialised column matrix.
  int cv_Ho, cv_Wo;          // output spatial size of the traversal
  int cv_sh, cv_sw;          // traversal strides
  int cv_ph, cv_pw;          // lower padding (first base position = -pad)
  int cv_dh, cv_dw;          // dilation
  int cv_kh, cv_kw;          // filter size
  int cv_C;                  // channels of the gathered tensor
  int cv_cblocks;            // ceil(C / 64)
  int cv_flip;               // dgrad: B uses the spatially flipped tap
  int cv_num_kb;             // number of 64-wide k blocks (taps * cblocks for fprop/dgrad, ceil(M/64) for wgrad)
  int cv_nimg;               // batch size of the gathered tensor (first out-of-bounds image index)
  int dual_m;                // one CTA owns two stacked 128-row tiles (both TMEM accumulator stages), so the B operand
                             // is fetched from L2 once per 256 output rows (bf16 kernels only)
  int cv_aw;                 // channels per gathered atom: 64 (128B swizzle) or, for C <= 32, 32 / 16 (64B / 32B swizzle)
  // block-scaled fp8 (kFp8 plain GEMM only): UE8M0 scale bytes per (row, 32 k) in 512-byte atoms of 128 rows x 4 k
  // blocks, [row block][k group] (csrc/quantize.cu). Non-null = tcgen05.mma.kind::mxf8f6f4.block_scale.
  const uint8_t* mx_sfa;
  const uint8_t* mx_sfb;
  int group_m;               // tile rasterisation: > 0 walks the output in bands of group_m m-tiles, m fastest inside a
                             // band, so the tiles in flight at one time share few A row-blocks AND few B column-blocks
                             // (L2 working set ~ (group_m + W / group_m) operand panels instead of 1 + tiles_n)
  FastDiv fd_hw, fd_wo, fd_cblocks, fd_kw, fd_c;
};

__device__ __forceinline__ void decode_tile(const GemmParams& p, int rem, int& m_blk, int& n_blk) {
  if (p.group_m <= 0) {
    m_blk = rem / p.tiles_n;
    n_blk = rem - m_blk * p.tiles_n;
    return;
  }
  const int band = p.group_m * p.tiles_n;
  const int g = rem / band;
  const int first_m = g * p.group_m;
  const int gsize = p.tiles_m - first_m < p.group_m ? p.tiles_m - first_m : p.group_m;
  const int in = rem - g * band;
  n_blk = in / gsize;
  m_blk = first_m + (in - n_blk * gsize);
}

// kConv: 0 plain GEMM; 1 A operand via im2col TMA (conv fprop, and dgrad when kBMN); 2 B operand via im2col TMA (wgrad)
// One 16-column chunk of one accumulator row: dequantise / bias / ReLU, then store as fp32 (plain or red.add), bf16 or
// requantised e4m3. ``sb``/``ss`` point at this chunk's bias / scale columns in shared memory (16-byte aligned).
template <bool kFp8>
__device__ __forceinline__ void epilogue_store16(const GemmParams& p, const uint32_t* r, int row, int n0, int nvalid,
                                                 const float* sb, const float* ss, bool has_bias, bool use_scale,
                                                 bool atomic) {
  // n0 = first OUTPUT column of this chunk, nvalid = how many of its 16 columns exist (<= 0: nothing to store)
  if (nvalid <= 0) return;
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
  const bool full = nvalid >= 16;
  if (kFp8 && use_scale) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 s4 = *reinterpret_cast<const float4*>(ss + j);
      v[j] *= s4.x; v[j + 1] *= s4.y; v[j + 2] *= s4.z; v[j + 3] *= s4.w;
    }
  }
  if (has_bias) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 b4 = *reinterpret_cast<const float4*>(sb + j);
      v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
    }
  }
  if (p.residual) {
    const __nv_bfloat16* rp = p.residual + static_cast<long long>(row) * p.ldr + n0;
    if (full && ((reinterpret_cast<uintptr_t>(rp) & 15u) == 0)) {
      const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(rp)), r1 = __ldg(reinterpret_cast<const uint4*>(rp) + 1);
      const uint32_t w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
      if (p.res_mask) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[2 * j] = __uint_as_float(w[j] << 16) > 0.f ? v[2 * j] : 0.f;
          v[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u) > 0.f ? v[2 * j + 1] : 0.f;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[2 * j] += __uint_as_float(w[j] << 16);
          v[2 * j + 1] += __uint_as_float(w[j] & 0xFFFF0000u);
        }
      }
    } else {
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) {
          const float rv = __bfloat162float(rp[j]);
          v[j] = p.res_mask ? (rv > 0.f ? v[j] : 0.f) : v[j] + rv;
        }
    }
  }
  if (p.flags & BDL_GEMM_RELU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (kFp8 && p.out_fp32 == 2) {
    // requantise for the next fp8 layer: q = sat_e4m3(v * out_qscale), 16 bytes per lane-row
    const float qs = __ldg(p.out_qscale);
    uint8_t* dst = reinterpret_cast<uint8_t*>(p.D) + static_cast<long long>(row) * p.ldd + n0;
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * j] * qs, v[4 * j + 1] * qs), __NV_SATFINITE,
                                                   __NV_E4M3);
      const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * j + 2] * qs, v[4 * j + 3] * qs), __NV_SATFINITE,
                                                   __NV_E4M3);
      w[j] = lo | (hi << 16);
    }
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
      *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) dst[j] = static_cast<uint8_t>((w[j >> 2] >> ((j & 3) * 8)) & 0xffu);
    }
  } else if (p.out_fp32) {
    float* dst = reinterpret_cast<float*>(p.D) + static_cast<long long>(row) * p.ldd + n0;
    const bool al = full && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0);
    if (atomic) {
      if (al) {
#pragma unroll
        for (int j = 0; j < 16; j += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(v[j]), "f"(v[j + 1]),
                       "f"(v[j + 2]), "f"(v[j + 3])
                       : "memory");
      } else {
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) atomicAdd(dst + j, v[j]);
      }
    } else if (al) {
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) dst[j] = v[j];
    }
  } else {
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.D) + static_cast<long long>(row) * p.ldd + n0;
    const bool al = full && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0);
    if (al) {
      uint4 q0, q1;
      q0.x = pack_bf16x2(v[0], v[1]);
      q0.y = pack_bf16x2(v[2], v[3]);
      q0.z = pack_bf16x2(v[4], v[5]);
      q0.w = pack_bf16x2(v[6], v[7]);
      q1.x = pack_bf16x2(v[8], v[9]);
      q1.y = pack_bf16x2(v[10], v[11]);
      q1.z = pack_bf16x2(v[12], v[13]);
      q1.w = pack_bf16x2(v[14], v[15]);
      *reinterpret_cast<uint4*>(dst) = q0;
      *reinterpret_cast<uint4*>(dst + 8) = q1;
    } else {
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) dst[j] = __float2bfloat16_rn(v[j]);
    }
  }
}

// Math of one 16-column chunk only (bias, residual, ReLU) -> 16 bf16 packed into two 16-byte vectors, for the
// shared-memory staged TMA-store epilogue. nvalid <= 0 or row >= M: the values are never written (the TMA store clips).
// kFp8: the per-column dequantisation scales ``ss`` are applied first; with ``to_e4m3`` the 16 values are requantised
// (v * out_qscale, saturating) into q0 alone.
template <bool kFp8>
__device__ __forceinline__ void epilogue_pack16(const GemmParams& p, const uint32_t* r, int row, int n0, int nvalid,
                                                const float* sb, const float* ss, bool has_bias, bool use_scale,
                                                bool to_e4m3, float qs, uint4& q0, uint4& q1) {
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
  if (kFp8 && use_scale) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 s4 = *reinterpret_cast<const float4*>(ss + j);
      v[j] *= s4.x; v[j + 1] *= s4.y; v[j + 2] *= s4.z; v[j + 3] *= s4.w;
    }
  }
  if (has_bias) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 b4 = *reinterpret_cast<const float4*>(sb + j);
      v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
    }
  }
  if (p.residual && row < p.M && nvalid > 0) {
    const __nv_bfloat16* rp = p.residual + static_cast<long long>(row) * p.ldr + n0;
    if (nvalid >= 16 && ((reinterpret_cast<uintptr_t>(rp) & 15u) == 0)) {
      const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(rp)), r1 = __ldg(reinterpret_cast<const uint4*>(rp) + 1);
      const uint32_t w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
      if (p.res_mask) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[2 * j] = __uint_as_float(w[j] << 16) > 0.f ? v[2 * j] : 0.f;
          v[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u) > 0.f ? v[2 * j + 1] : 0.f;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[2 * j] += __uint_as_float(w[j] << 16);
          v[2 * j + 1] += __uint_as_float(w[j] & 0xFFFF0000u);
        }
      }
    } else {
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) {
          const float rv = __bfloat162float(rp[j]);
          v[j] = p.res_mask ? (rv > 0.f ? v[j] : 0.f) : v[j] + rv;
        }
    }
  }
  if (p.flags & BDL_GEMM_RELU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (kFp8 && to_e4m3) {
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * j] * qs, v[4 * j + 1] * qs), __NV_SATFINITE,
                                                   __NV_E4M3);
      const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * j + 2] * qs, v[4 * j + 3] * qs), __NV_SATFINITE,
                                                   __NV_E4M3);
      w[j] = lo | (hi << 16);
    }
    q0 = make_uint4(w[0], w[1], w[2], w[3]);
    return;
  }
  q0.x = pack_bf16x2(v[0], v[1]);   q0.y = pack_bf16x2(v[2], v[3]);
  q0.z = pack_bf16x2(v[4], v[5]);   q0.w = pack_bf16x2(v[6], v[7]);
  q1.x = pack_bf16x2(v[8], v[9]);   q1.y = pack_bf16x2(v[10], v[11]);
  q1.z = pack_bf16x2(v[12], v[13]); q1.w = pack_bf16x2(v[14], v[15]);
}

// Flush the per-CTA column accumulators of n tile `nblk` (none if < 0) to global memory and clear them; called by all
// epilogue threads (named barrier 1). Out of line for the same reason as epi_col_stats.
__device__ __noinline__ void epi_stats_flush(float* sums, int N, int block_n, int nblk, uint32_t et, float* stat_sum,
                                             float* stat_sq) {
  asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
  if (nblk >= 0) {
    const int n = nblk * block_n + static_cast<int>(et);
    if (static_cast<int>(et) < block_n && n < N) {
      atomicAdd(sums + n, stat_sum[et]);
      atomicAdd(sums + N + n, stat_sq[et]);
    }
  }
  stat_sum[et] = 0.f;
  stat_sq[et] = 0.f;
  asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
}

// Column sums of one staged 32-row x 64-column bf16 box (this warp's TMA-store staging buffer, 128B-swizzled) into the
// per-CTA accumulators. Kept out of line: inlined it cost the kernels that never take statistics ~1 % (registers / code
// layout of the epilogue warps).
__device__ __noinline__ void epi_col_stats(uint32_t stg, uint32_t lane, int rows_valid, int c, float* stat_sum,
                                           float* stat_sq) {
  // lane l sums columns c + 2l, c + 2l + 1 of the staged rows (bank-conflict free: one 128-byte row per step)
  rows_valid = rows_valid > 32 ? 32 : rows_valid;
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
  // the swizzle repeats every 8 rows: 8 lane-dependent offsets, then constant strides
  uint32_t off8[8];
#pragma unroll
  for (int k = 0; k < 8; ++k)
    off8[k] = stg + static_cast<uint32_t>(k) * 128u + (((lane >> 2) ^ static_cast<uint32_t>(k)) << 4) + ((lane & 3u) << 2);
  if (rows_valid == 32) {
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      uint32_t w;
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(off8[r & 7] + static_cast<uint32_t>(r >> 3) * 1024u));
      const float v0 = __uint_as_float(w << 16), v1 = __uint_as_float(w & 0xFFFF0000u);
      s0 += v0;
      s1 += v1;
      q0 = fmaf(v0, v0, q0);
      q1 = fmaf(v1, v1, q1);
    }
  } else {
    for (int r = 0; r < rows_valid; ++r) {
      uint32_t w;
      const uint32_t a = stg + static_cast<uint32_t>(r) * 128u + (((lane >> 2) ^ (static_cast<uint32_t>(r) & 7u)) << 4) +
                         ((lane & 3u) << 2);
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(a));
      const float v0 = __uint_as_float(w << 16), v1 = __uint_as_float(w & 0xFFFF0000u);
      s0 += v0;
      s1 += v1;
      q0 = fmaf(v0, v0, q0);
      q1 = fmaf(v1, v1, q1);
    }
  }
  const uint32_t col = static_cast<uint32_t>(c) + 2u * lane;
  atomicAdd(&stat_sum[col], s0);
  atomicAdd(&stat_sum[col + 1], s1);
  atomicAdd(&stat_sq[col], q0);
  atomicAdd(&stat_sq[col + 1], q1);
}

// kFp8: operands are e4m3 bytes (128 per swizzle row, UMMA K = 32) multiplied with kind::f8f6f4; the epilogue applies a
//       per-output-column fp32 scale (activation scale x per-channel weight scale) before bias/ReLU. K-major only.
// kTmaStore: bf16 outputs leave through shared memory and TMA stores (one 32-row x 128-byte box per warp and 64-column
//       group) instead of 32-byte per-row stores; the staging area follows the pipeline stages in dynamic smem.
// kTwo: CTA-pair variant (launched as 2-CTA clusters): one tcgen05.mma.cta_group::2 covers 256 output rows; each CTA
//       stages its own 128 rows of A and half of the B tile (no multicast — every CTA fetches its own slice), the leader
//       (cluster rank 0) issues the MMAs and commits to the barriers of both CTAs, each CTA's epilogue drains its own
//       128 TMEM rows. Per SM the B bytes from L2 and through shared memory are halved.
// kAG: all-gather fused into the GEMM (Linear forward under the sharded parameter plane). The weight matrix B [N, K]
//       is partitioned by rows among the ranks; an n tile whose rows another rank owns is loaded straight from THAT
//       rank's memory — plain cp.async.bulk.tensor loads through a tensor map over the owner's buffer, one per CTA and k
//       block, never multicast (TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
//       from another GPU's memory hung the GPU) — and, once landed, the tile is also TMA-stored into the local copy of
//       the weights, so backward's dgrad and any later forward read local memory. The transfer of tile i+1..i+stages
//       overlaps the MMAs of tile i through the ordinary operand ring.
constexpr int kAgMaxRanks = 8;
struct AgMaps {
  CUtensorMap tm[kAgMaxRanks];   // B as seen in rank r's buffer (same shape / box / swizzle as the local map)
  int t0[kAgMaxRanks];           // n tiles [t0[r], t1[r]) are fetched from rank r
  int t1[kAgMaxRanks];
  int nranks, rank;
};

template <bool kAMN, bool kBMN, int kConv, bool kFp8, bool kDual, bool kTmaStore, bool kTwo, bool kAG>
__device__ __forceinline__ void gemm_tcgen05_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD,
                                                  const GemmParams& p, const AgMaps* ag) {
  static_assert(!kAG || (!kAMN && !kBMN && kConv == 0 && !kFp8 && !kDual && !kTwo), "all-gather GEMM: plain K-major form");
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float epi_cols[2][2][256];   // [accumulator stage][bias | scale][column of the tile]

  static_assert(!kTwo || (!kAMN && kConv != 2), "CTA pairs: A must be K-major, no wgrad");
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t crank = kTwo ? cluster_ctarank() : 0u;   // rank inside the CTA pair; rank 0 leads
  const int work0 = kTwo ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int work_step = kTwo ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = BM * BK * 2 * (kDual ? 2 : 1);
  const uint32_t b_bytes = static_cast<uint32_t>(p.b_rows) * BK * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  // block-scaled fp8: 2 x 512-byte scale-factor atoms per stage, behind the operand ring (and the store staging)
  const uint32_t sf_region = tiles + static_cast<uint32_t>(p.num_stages) * stage_bytes +
                             (kTmaStore ? static_cast<uint32_t>(8 * 4096) : 0u);

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&tfull_bar[s]), 1);
      mbar_init(smem_u32(&tempty_bar[s]), kTwo ? 16 : 8);   // pair: the leader also collects the peer's 8 warps
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (kTwo) tmem_alloc_pair<kTmemCols>(smem_u32(&tmem_base_slot));
    else tmem_alloc<kTmemCols>(smem_u32(&tmem_base_slot));
  }
  tc_fence_before();
  if (kTwo) cluster_sync_all();   // the peer's barriers must be initialised before anything arrives on them remotely
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  constexpr int KE = kFp8 ? 128 : 64;   // elements per 128-byte k block
  const bool small_c = kConv != 0 && !kFp8 && p.cv_aw < 64;
  const int num_kb = kConv ? p.cv_num_kb : (p.K + KE - 1) / KE;
  const int tiles_mn = p.tiles_m * p.tiles_n;
  const int total_work = tiles_mn * p.splits;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int work = work0; work < total_work; work += work_step) {
        const int split = work / tiles_mn;
        const int rem = work - split * tiles_mn;
        int m_blk, n_blk;
        decode_tile(p, rem, m_blk, n_blk);
        const int kb0 = static_cast<int>((static_cast<long long>(split) * num_kb) / p.splits);
        const int kb1 = static_cast<int>((static_cast<long long>(split + 1) * num_kb) / p.splits);
        // pair: this CTA's 128 rows of the 256-row tile, and its half of the tile's B rows
        // pair + dual: the pair owns 512 rows as two 256-row MMAs; this CTA stages rows [+128 crank, +128) of each
        const int m0 = kTwo ? (m_blk * (kDual ? 4 : 2) + static_cast<int>(crank)) * BM : m_blk * BM * (kDual ? 2 : 1);
        const int m0b = m0 + (kTwo ? 2 * BM : BM);   // first row of this CTA's second stacked half (kDual)
        const CUtensorMap* ag_src = &tmB;
        if (kAG) {
          for (int r = 0; r < ag->nranks; ++r)
            if (r != ag->rank && n_blk >= ag->t0[r] && n_blk < ag->t1[r]) ag_src = &ag->tm[r];
        }
        const int n0 = n_blk * p.block_n + (kTwo ? static_cast<int>(crank) * (p.block_n >> 1) : 0);
        // implicit-GEMM bookkeeping hoisted out of the k loop
        int cv_img = 0, cv_h = 0, cv_w = 0, cv_r = 0, cv_s = 0, cv_cb = 0;
        int cv_img2 = 0, cv_h2 = 0, cv_w2 = 0;
        int at_c0[4] = {0, 0, 0, 0};
        uint16_t at_ow[4] = {0, 0, 0, 0}, at_oh[4] = {0, 0, 0, 0};
        bool at_ok[4] = {false, false, false, false};
        uint32_t sm_off[16];
        if (kConv == 1) {
          const uint32_t img = p.fd_hw.quot(static_cast<uint32_t>(m0));
          const uint32_t rem = static_cast<uint32_t>(m0) - img * p.fd_hw.div;
          const uint32_t ho = p.fd_wo.quot(rem);
          const uint32_t wo = rem - ho * p.fd_wo.div;
          cv_img = static_cast<int>(img);
          cv_h = static_cast<int>(ho) * p.cv_sh - p.cv_ph;
          cv_w = static_cast<int>(wo) * p.cv_sw - p.cv_pw;
          if (kDual) {   // base pixel of the second 128-row half
            const uint32_t img2 = p.fd_hw.quot(static_cast<uint32_t>(m0b));
            const uint32_t rem2 = static_cast<uint32_t>(m0b) - img2 * p.fd_hw.div;
            const uint32_t ho2 = p.fd_wo.quot(rem2);
            cv_img2 = static_cast<int>(img2);
            cv_h2 = static_cast<int>(ho2) * p.cv_sh - p.cv_ph;
            cv_w2 = static_cast<int>(rem2 - ho2 * p.fd_wo.div) * p.cv_sw - p.cv_pw;
          }
          const uint32_t tap = small_c ? static_cast<uint32_t>(kb0 * (64 / p.cv_aw))
                                       : p.fd_cblocks.quot(static_cast<uint32_t>(kb0));
          cv_cb = small_c ? 0 : kb0 - static_cast<int>(tap) * p.cv_cblocks;
          cv_r = static_cast<int>(p.fd_kw.quot(tap));
          cv_s = static_cast<int>(tap) - cv_r * p.cv_kw;
        } else if (kConv == 2 && small_c) {
          // one atom per filter tap (C <= atom width): atom j of this tile is tap n0 / aw + j
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const uint32_t tap = static_cast<uint32_t>(n0 / p.cv_aw + j);
            const uint32_t r = p.fd_kw.quot(tap);
            sm_off[j] = tap < static_cast<uint32_t>(p.cv_kh * p.cv_kw)
                            ? (((tap - r * p.fd_kw.div) * p.cv_dw) | ((r * p.cv_dh) << 16))
                            : 0xFFFFFFFFu;
          }
        } else if (kConv == 2) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t ncol = static_cast<uint32_t>(n0 + 64 * j);
            const uint32_t tap = p.fd_c.quot(ncol);
            const uint32_t r = p.fd_kw.quot(tap);
            at_c0[j] = static_cast<int>(ncol - tap * p.fd_c.div);
            at_ow[j] = static_cast<uint16_t>((tap - r * p.fd_kw.div) * p.cv_dw);
            at_oh[j] = static_cast<uint16_t>(r * p.cv_dh);
            at_ok[j] = tap < static_cast<uint32_t>(p.cv_kh * p.cv_kw);
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          // pair: both CTAs' bytes are counted on the LEADER's full barrier (it alone waits on it, to issue the MMAs)
          const uint32_t fb = kTwo ? mapa_shared(smem_u32(&full_bar[stage]), 0u) : smem_u32(&full_bar[stage]);
          const uint32_t sa = tiles + stage * stage_bytes;
          const uint32_t sb = sa + a_bytes;
          const int k0 = kb * KE;
          if (kConv == 1 && !kBMN && small_c) {
            // k block = 64 / aw filter taps, each gathered as its own [128 pixels x aw channels] narrow-swizzle region
            const int tpb = 64 / p.cv_aw;
            const int taps = p.cv_kh * p.cv_kw;
            int tap = kb * tpb;
            const int nv = taps - tap < tpb ? taps - tap : tpb;
            const uint32_t a_tap = BM * p.cv_aw * 2, b_tap = static_cast<uint32_t>(p.block_n) * p.cv_aw * 2;
            mbar_expect_tx(fb, nv * (a_tap + b_tap));
            for (int j = 0; j < nv; ++j, ++tap) {
              tma_load_im2col_4d(sa + j * a_tap, &tmA, fb, 0, cv_w, cv_h, cv_img,
                                 static_cast<uint16_t>(cv_s * p.cv_dw), static_cast<uint16_t>(cv_r * p.cv_dh));
              tma_load_2d(sb + j * b_tap, &tmB, fb, tap * p.cv_C, n0);
              if (++cv_s == p.cv_kw) {
                cv_s = 0;
                ++cv_r;
              }
            }
          } else if (kConv == 2 && small_c) {
            mbar_expect_tx(fb, stage_bytes);
            tma_load_2d(sa, &tmA, fb, m0, k0);
            tma_load_2d(sa + 8192u, &tmA, fb, m0 + 64, k0);
            if (kDual) {
              tma_load_2d(sa + 16384u, &tmA, fb, m0 + 128, k0);
              tma_load_2d(sa + 24576u, &tmA, fb, m0 + 192, k0);
            }
            const uint32_t img = p.fd_hw.quot(static_cast<uint32_t>(k0));
            const uint32_t rem = static_cast<uint32_t>(k0) - img * p.fd_hw.div;
            const uint32_t ho = p.fd_wo.quot(rem);
            const uint32_t wo = rem - ho * p.fd_wo.div;
            const int w = static_cast<int>(wo) * p.cv_sw - p.cv_pw, h = static_cast<int>(ho) * p.cv_sh - p.cv_ph;
            const uint32_t b_atom = 64u * p.cv_aw * 2;
            const int natoms = p.b_rows / p.cv_aw;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (j < natoms) {
                if (sm_off[j] != 0xFFFFFFFFu)
                  tma_load_im2col_4d(sb + j * b_atom, &tmB, fb, 0, w, h, static_cast<int>(img),
                                     static_cast<uint16_t>(sm_off[j] & 0xFFFFu), static_cast<uint16_t>(sm_off[j] >> 16));
                else
                  tma_load_im2col_4d(sb + j * b_atom, &tmB, fb, 0, -p.cv_pw, -p.cv_ph, p.cv_nimg, 0, 0);
              }
            }
          } else if (kConv == 1 && kTwo) {
            if (crank == 0) mbar_expect_tx(smem_u32(&full_bar[stage]), 2u * stage_bytes);
            tma_load_im2col_4d_pair(sa, &tmA, fb, cv_cb * KE, cv_w, cv_h, cv_img, static_cast<uint16_t>(cv_s * p.cv_dw),
                                    static_cast<uint16_t>(cv_r * p.cv_dh));
            if (kDual)
              tma_load_im2col_4d_pair(sa + 16384u, &tmA, fb, cv_cb * KE, cv_w2, cv_h2, cv_img2,
                                      static_cast<uint16_t>(cv_s * p.cv_dw), static_cast<uint16_t>(cv_r * p.cv_dh));
            const int tap = cv_r * p.cv_kw + cv_s;
            if (kBMN) {
              const int btap = p.cv_flip ? (p.cv_kh - 1 - cv_r) * p.cv_kw + (p.cv_kw - 1 - cv_s) : tap;
              for (int j = 0; j < p.b_rows / 64; ++j)
                tma_load_2d_pair(sb + j * 8192u, &tmB, fb, btap * p.N + n0 + 64 * j, cv_cb * 64);
            } else {
              tma_load_2d_pair(sb, &tmB, fb, tap * p.cv_C + cv_cb * KE, n0);
            }
            if (++cv_cb == p.cv_cblocks) {
              cv_cb = 0;
              if (++cv_s == p.cv_kw) {
                cv_s = 0;
                ++cv_r;
              }
            }
          } else if (kConv == 1) {
            mbar_expect_tx(fb, stage_bytes);
            // k block = (filter tap, 64-channel block); A tile = 128 consecutive output pixels of that tap
            tma_load_im2col_4d(sa, &tmA, fb, cv_cb * KE, cv_w, cv_h, cv_img, static_cast<uint16_t>(cv_s * p.cv_dw),
                               static_cast<uint16_t>(cv_r * p.cv_dh));
            if (kDual)
              tma_load_im2col_4d(sa + 16384u, &tmA, fb, cv_cb * KE, cv_w2, cv_h2, cv_img2,
                                 static_cast<uint16_t>(cv_s * p.cv_dw), static_cast<uint16_t>(cv_r * p.cv_dh));
            const int tap = cv_r * p.cv_kw + cv_s;
            if (kBMN) {
              // dgrad: B[k = cout block][n = cin] is the (flipped) tap's [Cout, Cin] slice of the OHWI weights
              const int btap = p.cv_flip ? (p.cv_kh - 1 - cv_r) * p.cv_kw + (p.cv_kw - 1 - cv_s) : tap;
              for (int j = 0; j < p.b_rows / 64; ++j)
                tma_load_2d(sb + j * 8192u, &tmB, fb, btap * p.N + n0 + 64 * j, cv_cb * 64);
            } else {
              tma_load_2d(sb, &tmB, fb, tap * p.cv_C + cv_cb * KE, n0);
            }
            if (++cv_cb == p.cv_cblocks) {   // advance (tap, channel block) without dividing
              cv_cb = 0;
              if (++cv_s == p.cv_kw) {
                cv_s = 0;
                ++cv_r;
              }
            }
          } else if (kConv == 2) {
            // wgrad: k block = 64 consecutive output pixels; A = dY^T (MN-major), B atoms = im2col gathers
            mbar_expect_tx(fb, stage_bytes);
            tma_load_2d(sa, &tmA, fb, m0, k0);
            tma_load_2d(sa + 8192u, &tmA, fb, m0 + 64, k0);
            if (kDual) {
              tma_load_2d(sa + 16384u, &tmA, fb, m0 + 128, k0);
              tma_load_2d(sa + 24576u, &tmA, fb, m0 + 192, k0);
            }
            const uint32_t img = p.fd_hw.quot(static_cast<uint32_t>(k0));
            const uint32_t rem = static_cast<uint32_t>(k0) - img * p.fd_hw.div;
            const uint32_t ho = p.fd_wo.quot(rem);
            const uint32_t wo = rem - ho * p.fd_wo.div;
            const int w = static_cast<int>(wo) * p.cv_sw - p.cv_pw, h = static_cast<int>(ho) * p.cv_sh - p.cv_ph;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (j < p.b_rows / 64) {
                if (at_ok[j])
                  tma_load_im2col_4d(sb + j * 8192u, &tmB, fb, at_c0[j], w, h, static_cast<int>(img), at_ow[j], at_oh[j]);
                else  // past the last tap: a box that starts after the last image (zero fill)
                  tma_load_im2col_4d(sb + j * 8192u, &tmB, fb, 0, -p.cv_pw, -p.cv_ph, p.cv_nimg, 0, 0);
              }
            }
          } else if (kTwo) {
            if (crank == 0) mbar_expect_tx(smem_u32(&full_bar[stage]), 2u * stage_bytes);
            tma_load_2d_pair(sa, &tmA, fb, k0, m0);
            if (kDual) tma_load_2d_pair(sa + 16384u, &tmA, fb, k0, m0b);
            if (kBMN) {
              for (int j = 0; j < p.b_rows / 64; ++j) tma_load_2d_pair(sb + j * 8192u, &tmB, fb, n0 + 64 * j, k0);
            } else {
              tma_load_2d_pair(sb, &tmB, fb, k0, n0);
            }
          } else {
          mbar_expect_tx(fb, stage_bytes + ((kFp8 && kConv == 0 && p.mx_sfa) ? 1024u : 0u));
          if (kAMN) {
            // A stored [K, M]: two boxes of {64 m, 64 k}
            tma_load_2d(sa, &tmA, fb, m0, k0);
            tma_load_2d(sa + 8192u, &tmA, fb, m0 + 64, k0);
            if (kDual) {
              tma_load_2d(sa + 16384u, &tmA, fb, m0 + 128, k0);
              tma_load_2d(sa + 24576u, &tmA, fb, m0 + 192, k0);
            }
          } else {
            tma_load_2d(sa, &tmA, fb, k0, m0);  // box {64 k, 128 m}
            if (kDual) tma_load_2d(sa + 16384u, &tmA, fb, k0, m0 + BM);
          }
          if (kBMN) {
            for (int j = 0; j < p.b_rows / 64; ++j) tma_load_2d(sb + j * 8192u, &tmB, fb, n0 + 64 * j, k0);
          } else if (kAG) {
            tma_load_2d(sb, ag_src, fb, k0, n0);   // the owner's copy of this weight tile (local map for my own rows)
          } else {
            tma_load_2d(sb, &tmB, fb, k0, n0);  // box {64 k, block_n}
          }
          if (kFp8 && kConv == 0 && p.mx_sfa) {
            // scale-factor atoms of this k group: A rows [m0, +128) and B rows [n0, +128) (block_n <= 128 here)
            const uint32_t sfs = sf_region + static_cast<uint32_t>(stage) * 1024u;
            bulk_load_1d(sfs, p.mx_sfa + (static_cast<long long>(m0 >> 7) * num_kb + kb) * 512, 512u, fb);
            bulk_load_1d(sfs + 512u, p.mx_sfb + (static_cast<long long>(n0 >> 7) * num_kb + kb) * 512, 512u, fb);
          }
          }
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0 && crank == 0) {
      // instruction descriptor: c=f32 [4,6), a=bf16 [7,10), b=bf16 [10,13), a_major 15, b_major 16, N>>3 [17,23), M>>4 [24,29)
      // fp8: a/b format 0 = E4M3; bf16: format 1
      const uint32_t idesc = (1u << 4) | (kFp8 ? 0u : ((1u << 7) | (1u << 10))) | (static_cast<uint32_t>(kAMN) << 15) |
                             (static_cast<uint32_t>(kBMN) << 16) | (static_cast<uint32_t>(p.block_n >> 3) << 17) |
                             (static_cast<uint32_t>((kTwo ? 2 * BM : BM) >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (int work = work0; work < total_work; work += work_step) {
        const int split = work / tiles_mn;
        const int kb0 = static_cast<int>((static_cast<long long>(split) * num_kb) / p.splits);
        const int kb1 = static_cast<int>((static_cast<long long>(split + 1) * num_kb) / p.splits);
        constexpr bool dual = kDual;
        bool ag_remote = false;
        int ag_n0 = 0;
        if (kAG) {
          int m_blk_, n_blk_;
          decode_tile(p, work - split * tiles_mn, m_blk_, n_blk_);
          ag_n0 = n_blk_ * p.block_n;
          for (int r = 0; r < ag->nranks; ++r)
            if (r != ag->rank && n_blk_ >= ag->t0[r] && n_blk_ < ag->t1[r]) ag_remote = true;
        }
        if (kTwo) mbar_wait_cluster(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
        else mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
        if (dual) {   // dual: acc is always 0, both stages are ours
          if (kTwo) mbar_wait_cluster(smem_u32(&tempty_bar[1]), acc_phase ^ 1u);
          else mbar_wait(smem_u32(&tempty_bar[1]), acc_phase ^ 1u);
        }
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kAccStride;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t sa = tiles + stage * stage_bytes;
          const uint32_t sb = sa + a_bytes;
          if (kConv == 1 && !kBMN && small_c) {
            const int tpb = 64 / p.cv_aw;
            const int left = p.cv_kh * p.cv_kw - kb * tpb;
            const int nv = left < tpb ? left : tpb;
            const uint32_t a_tap = BM * p.cv_aw * 2, b_tap = static_cast<uint32_t>(p.block_n) * p.cv_aw * 2;
            const uint32_t sbo = 8u * p.cv_aw * 2, layout = p.cv_aw == 32 ? 4u : 6u;
            for (int j = 0; j < nv; ++j)
              for (int k = 0; k < p.cv_aw / 16; ++k)
                umma_f16(d_tmem, make_smem_desc(sa + j * a_tap + k * 32u, 16u, sbo, layout),
                         make_smem_desc(sb + j * b_tap + k * 32u, 16u, sbo, layout), idesc,
                         (kb > kb0 || j > 0 || k > 0) ? 1u : 0u);
          } else if (kConv == 2 && small_c) {
            const uint32_t row = p.cv_aw * 2u, layout = p.cv_aw == 32 ? 4u : 6u;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t bdesc = make_smem_desc(sb + k * 16u * row, 64u * row, 8u * row, layout);
              umma_f16(d_tmem, make_smem_desc_sw128(sa + k * 2048u, 8192u, 1024u), bdesc, idesc,
                       (kb > kb0 || k > 0) ? 1u : 0u);
              if (dual)
                umma_f16(d_tmem + kAccStride, make_smem_desc_sw128(sa + 16384u + k * 2048u, 8192u, 1024u), bdesc,
                         idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
          } else {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t adesc = kAMN ? make_smem_desc_sw128(sa + k * 2048u, 8192u, 1024u)
                                        : make_smem_desc_sw128(sa + k * 32u, 16u, 1024u);
            const uint64_t bdesc = kBMN ? make_smem_desc_sw128(sb + k * 2048u, 8192u, 1024u)
                                        : make_smem_desc_sw128(sb + k * 32u, 16u, 1024u);
            if (kFp8 && kConv == 0 && p.mx_sfa) {
              // scale factors of this k group: shared -> TMEM once per stage (tcgen05.cp and tcgen05.mma execute in
              // issue order, so the copy cannot overtake the previous stage's MMAs), then one MMA per 32-element block
              // with the block's scale-byte index in the instruction descriptor
              const uint32_t sfa_t = tmem_base + 128u, sfb_t = tmem_base + 132u;
              if (k == 0) {
                const uint32_t sfs = sf_region + static_cast<uint32_t>(stage) * 1024u;
                utccp_sf(sfa_t, make_smem_desc(sfs, 0u, 128u, 0u));
                utccp_sf(sfb_t, make_smem_desc(sfs + 512u, 0u, 128u, 0u));
              }
              const uint32_t idesc_mx = (static_cast<uint32_t>(k) << 4) | (static_cast<uint32_t>(p.block_n >> 3) << 17) |
                                        (1u << 23) | (static_cast<uint32_t>(BM >> 4) << 24) |
                                        (static_cast<uint32_t>(k) << 29);
              umma_mxf8(d_tmem, adesc, bdesc, idesc_mx, sfa_t, sfb_t, (kb > kb0 || k > 0) ? 1u : 0u);
            } else if (kTwo) {
              if (kFp8) umma_f8_pair(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
              else umma_f16_pair(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            } else if (kFp8)
              umma_f8(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else
              umma_f16(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            if (dual) {
              const uint64_t adesc2 = kAMN ? make_smem_desc_sw128(sa + 16384u + k * 2048u, 8192u, 1024u)
                                           : make_smem_desc_sw128(sa + 16384u + k * 32u, 16u, 1024u);
              if (kTwo) umma_f16_pair(d_tmem + kAccStride, adesc2, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
              else umma_f16(d_tmem + kAccStride, adesc2, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
          }
          }
          if (kAG && ag_remote) {
            // write-through: the remote tile that just landed also goes into the local copy of the weights. The slot is
            // released only after the bulk store has READ shared memory (the MMAs above read it concurrently).
            tma_store_2d(&tmB, sb, kb * KE, ag_n0);
            tma_store_commit();
            tma_store_wait_read0();
          }
          if (kTwo) umma_commit_pair(smem_u32(&empty_bar[stage]));
          else umma_commit(smem_u32(&empty_bar[stage]));
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (kTwo) umma_commit_pair(smem_u32(&tfull_bar[acc]));
        else umma_commit(smem_u32(&tfull_bar[acc]));
        if (dual) {
          if (kTwo) umma_commit_pair(smem_u32(&tfull_bar[1]));
          else umma_commit(smem_u32(&tfull_bar[1]));
          acc_phase ^= 1u;
        } else {
          acc ^= 1u;
          if (acc == 0) acc_phase ^= 1u;
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..9)
    const uint32_t quad = warp & 3u;  // TMEM lane quadrant this warp may read
    const uint32_t et = threadIdx.x - 64u;   // 0..255 within the epilogue warps
    const uint32_t half = (warp - 2u) >> 2;  // warps 2..5 take the low columns of a tile, 6..9 the high ones
    uint32_t acc = 0, acc_phase = 0;
    const bool has_bias = (p.flags & BDL_GEMM_BIAS) != 0;
    const bool atomic = (p.flags & BDL_GEMM_ATOMIC_ADD) != 0;
    const bool use_scale = kFp8 && p.col_scale != nullptr;
    // pair: TMEM drained -> arrive on the LEADER's barrier (remote for the peer CTA)
    const uint32_t tempty_addr[2] = {kTwo ? mapa_shared(smem_u32(&tempty_bar[0]), 0u) : smem_u32(&tempty_bar[0]),
                                     kTwo ? mapa_shared(smem_u32(&tempty_bar[1]), 0u) : smem_u32(&tempty_bar[1])};
    // column statistics (stat_sums): accumulated per CTA in shared memory across the m tiles of one n tile and flushed
    // with one atomic per column when the n tile changes / at the end, so the global atomics stay ~#CTA per column
    const bool do_stats = kTmaStore && !kFp8 && p.stat_sums != nullptr;
    // the dequantisation-scale half of epi_cols is only used by the fp8 kernels: bf16 kernels keep the per-CTA column
    // accumulators there ([0] sums, [1] sums of squares), so the statistics cost no shared memory
    float* const stat_sum = &epi_cols[0][1][0];
    float* const stat_sq = &epi_cols[1][1][0];
    int stat_nblk = -1;
    auto stat_flush = [&]() { epi_stats_flush(p.stat_sums, p.N, p.block_n, stat_nblk, et, stat_sum, stat_sq); };
    for (int work = work0; work < total_work; work += work_step) {
      const int split = work / tiles_mn;
      const int rem = work - split * tiles_mn;
      int m_blk, n_blk;
      decode_tile(p, rem, m_blk, n_blk);
      if (do_stats && n_blk != stat_nblk) {
        stat_flush();
        stat_nblk = n_blk;
      }
      const int nhalves = kDual ? 2 : 1;   // dual: accumulator stage h holds rows [256 m_blk + 128 h, +128)
      for (int hm = 0; hm < nhalves; ++hm) {
      const int tile_row0 = kTwo ? (m_blk * 2 * nhalves + hm * 2 + static_cast<int>(crank)) * BM
                                 : (m_blk * nhalves + hm) * BM;
      const int row = tile_row0 + static_cast<int>(quad * 32u + lane);
      const int nbase = n_blk * p.block_n;
      // Stage this tile's bias / dequantisation columns in shared memory while the MMAs are still running: reading
      // them with __ldg inside the column loop put a long-scoreboard stall on every 16-column chunk.
      float* sb = &epi_cols[acc][0][0];
      float* ss = &epi_cols[acc][1][0];
      const bool tile_bias = has_bias && (!atomic || split == 0);
      if (has_bias || use_scale) {
        for (int c = static_cast<int>(et); c < p.block_n; c += kEpiThreads) {
          const int n = nbase + c;
          sb[c] = (tile_bias && n < p.N) ? __ldg(p.bias + n) : 0.f;
          if (kFp8) ss[c] = (use_scale && n < p.N) ? __ldg(p.col_scale + n) : 1.f;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
      }
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((quad * 32u) << 16) + acc * kAccStride;
      if (kTmaStore) {
        // this warp's staging box: 32 rows x 128 bytes, 16-byte chunks XOR-swizzled by (row & 7) like the tensor map
        const uint32_t stg = tiles + static_cast<uint32_t>(p.num_stages) * stage_bytes + (warp - 2u) * 4096u;
        const int row0 = tile_row0 + static_cast<int>(quad * 32u);
        const int g_split = ((p.block_n + 127) / 128) * 64;
        const int g_begin = half ? g_split : 0;
        const int g_end = half ? ((p.block_n + 63) / 64) * 64 : g_split;
        for (int c = g_begin; c < g_end; c += 64) {
          uint32_t r0[16], r1[16], r2[16], r3[16];
          tmem_ld_16(t_row + static_cast<uint32_t>(c), r0);
          tmem_ld_16(t_row + static_cast<uint32_t>(c + 16), r1);
          tmem_ld_16(t_row + static_cast<uint32_t>(c + 32), r2);
          tmem_ld_16(t_row + static_cast<uint32_t>(c + 48), r3);
          if (lane == 0) tma_store_wait_read0();      // the previous box has left the staging buffer
          __syncwarp();
          tmem_ld_wait();
          uint4 q[8];
          const bool e4 = kFp8 && p.out_fp32 == 2;
          const float qs = e4 ? __ldg(p.out_qscale) : 1.f;
          epilogue_pack16<kFp8>(p, r0, row, nbase + c, p.N - (nbase + c), sb + c, ss + c, has_bias, use_scale, e4, qs,
                                q[0], q[1]);
          epilogue_pack16<kFp8>(p, r1, row, nbase + c + 16, p.N - (nbase + c + 16), sb + c + 16, ss + c + 16, has_bias,
                                use_scale, e4, qs, q[2], q[3]);
          epilogue_pack16<kFp8>(p, r2, row, nbase + c + 32, p.N - (nbase + c + 32), sb + c + 32, ss + c + 32, has_bias,
                                use_scale, e4, qs, q[4], q[5]);
          epilogue_pack16<kFp8>(p, r3, row, nbase + c + 48, p.N - (nbase + c + 48), sb + c + 48, ss + c + 48, has_bias,
                                use_scale, e4, qs, q[6], q[7]);
          if (e4) {
            // e4m3 output: 64 columns = 64-byte rows, 64B-swizzled (chunk ^ ((row >> 1) & 3)) like its tensor map
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t a = stg + lane * 64u + ((static_cast<uint32_t>(j) ^ ((lane >> 1) & 3u)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(q[2 * j].x), "r"(q[2 * j].y),
                           "r"(q[2 * j].z), "r"(q[2 * j].w) : "memory");
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t a = stg + lane * 128u + ((static_cast<uint32_t>(j) ^ (lane & 7u)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(q[j].x), "r"(q[j].y), "r"(q[j].z),
                           "r"(q[j].w) : "memory");
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmD, stg, nbase + c, row0);
            tma_store_commit();
          }
          if (do_stats) epi_col_stats(stg, lane, p.M - row0, c, stat_sum, stat_sq);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (kTwo) mbar_arrive_cluster(tempty_addr[acc]);
          else mbar_arrive(smem_u32(&tempty_bar[acc]));
        }
        acc ^= 1u;
        if (acc == 0) acc_phase ^= 1u;
        continue;
      }
      const bool row_ok = row < p.M;
      // the two warps of a TMEM lane quadrant split the tile's columns
      const int c_split = ((p.block_n / 16 + 1) / 2) * 16;
      const int c_begin = half ? c_split : 0;
      const int c_end = half ? p.block_n : c_split;
      // Column mapping of a 16-column chunk starting at tile column cc. Plain GEMM / fprop / dgrad: output column =
      // nbase + cc. wgrad (kConv == 2): the GEMM N axis is padded per filter tap to a multiple of 64 channels
      // (tap * Cpad + ch) so that the im2col atoms tile it; the real dW column is tap * C + ch and channels >= C
      // (zero-filled by TMA) are dropped here.
#define EPI_STORE(REGS, CC)                                                                                       \
  do {                                                                                                            \
    int n_out = nbase + (CC), n_valid;                                                                            \
    if (kConv == 2) {                                                                                             \
      const uint32_t tap_ = p.fd_c.quot(static_cast<uint32_t>(n_out));                                            \
      const int rem_ = n_out - static_cast<int>(tap_ * p.fd_c.div);                                               \
      n_valid = (n_out < p.N) ? (p.cv_C - rem_) : 0;                                                              \
      n_out = static_cast<int>(tap_) * p.cv_C + rem_;                                                             \
    } else {                                                                                                      \
      n_valid = p.N - n_out;                                                                                      \
    }                                                                                                             \
    epilogue_store16<kFp8>(p, REGS, row, n_out, n_valid, sb + (CC), ss + (CC), has_bias, use_scale, atomic);       \
  } while (0)
      // Two register sets: the TMEM loads of the next 32 columns are in flight while this set is converted and stored.
      uint32_t a0[16], a1[16], b0[16], b1[16];
      int c = c_begin;
      if (c < c_end) {
        tmem_ld_16(t_row + static_cast<uint32_t>(c), a0);
        if (c + 16 < c_end) tmem_ld_16(t_row + static_cast<uint32_t>(c + 16), a1);
      }
      while (c < c_end) {
        tmem_ld_wait();
        if (c + 32 < c_end) {
          tmem_ld_16(t_row + static_cast<uint32_t>(c + 32), b0);
          if (c + 48 < c_end) tmem_ld_16(t_row + static_cast<uint32_t>(c + 48), b1);
        }
        if (row_ok) {
          EPI_STORE(a0, c);
          if (c + 16 < c_end) EPI_STORE(a1, c + 16);
        }
        c += 32;
        if (c >= c_end) break;
        tmem_ld_wait();
        if (c + 32 < c_end) {
          tmem_ld_16(t_row + static_cast<uint32_t>(c + 32), a0);
          if (c + 48 < c_end) tmem_ld_16(t_row + static_cast<uint32_t>(c + 48), a1);
        }
        if (row_ok) {
          EPI_STORE(b0, c);
          if (c + 16 < c_end) EPI_STORE(b1, c + 16);
        }
        c += 32;
      }
#undef EPI_STORE
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kTwo) mbar_arrive_cluster(tempty_addr[acc]);
        else mbar_arrive(smem_u32(&tempty_bar[acc]));
      }
      acc ^= 1u;
      if (acc == 0) acc_phase ^= 1u;
      }
    }
    if (do_stats) stat_flush();
  }

  if (kTmaStore && warp >= 2 && lane == 0) tma_store_wait0();   // shared memory must outlive the bulk stores
  tc_fence_before();
  if (kTwo) cluster_sync_all();   // neither CTA may leave while the other can still signal it or read its operands
  else __syncthreads();
  if (kAG && warp == 1 && lane == 0) tma_store_wait0();   // write-through stores complete before the CTA retires
  if (warp == 1) {
    if (kTwo) tmem_dealloc_pair<kTmemCols>(tmem_base);
    else tmem_dealloc<kTmemCols>(tmem_base);
  }
}

template <bool kAMN, bool kBMN, int kConv, bool kFp8 = false, bool kDual = false, bool kTmaStore = false,
          bool kTwo = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmD, const GemmParams p) {
  gemm_tcgen05_body<kAMN, kBMN, kConv, kFp8, kDual, kTmaStore, kTwo, false>(tmA, tmB, tmD, p, nullptr);
}

// Linear forward with the all-gather of its weight rows fused in (see kAG above).
template <bool kTmaStore>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_allgather_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                           const __grid_constant__ CUtensorMap tmD, const GemmParams p,
                           const __grid_constant__ AgMaps ag) {
  gemm_tcgen05_body<false, false, 0, false, false, kTmaStore, false, true>(tmA, tmB, tmD, p, &ag);
}

// ------------------------------------------------------------------------------------------- host
// The residual operand of the NEXT launch from this host thread (bdl_gemm_set_residual): consumed and cleared by
// bdl_gemm_bf16 / bdl_conv_igemm_bf16 / bdl_gemm_fp8, so the launch signatures stay as they are.
static thread_local const void* g_next_residual = nullptr;
static thread_local long long g_next_ldr = 0;
static thread_local int g_next_res_mask = 0;          // bdl_gemm_set_residual_mask: the armed residual is a ReLU mask
static thread_local float* g_next_stats = nullptr;   // bdl_gemm_set_stats: column statistics of the next launch's output
static thread_local int g_stats_done = 0;            // did the last launch on this thread produce them?
static void take_residual(GemmParams& p) {
  p.residual = static_cast<const __nv_bfloat16*>(g_next_residual);
  p.ldr = g_next_ldr;
  g_next_residual = nullptr;
  g_next_ldr = 0;
  p.res_mask = g_next_res_mask;
  g_next_res_mask = 0;
  p.stat_sums = g_next_stats;
  g_next_stats = nullptr;
  g_stats_done = 0;
}

// bf16, non-accumulating outputs can leave through shared memory + TMA stores (kTmaStore): needs a 16-byte aligned D
// whose row pitch is a multiple of 8 elements. max_kb: only epilogue-heavy problems (short reductions) take it.
constexpr size_t kStageOutBytes = 8 * 4096;
static bool want_tma_store(const void* D, long long ldd, int out_fp32, int flags, bool a_mn, int num_kb) {
  static const int max_kb = [] { const char* e = getenv("BDL_GEMM_TMA_STORE_KB"); return e ? atoi(e) : 1 << 30; }();
  return max_kb > 0 && num_kb <= max_kb && !a_mn && !out_fp32 && !(flags & BDL_GEMM_ATOMIC_ADD) && (ldd % 8) == 0 &&
         (reinterpret_cast<uintptr_t>(D) & 15u) == 0;
}

// Two stacked 128-row tiles per CTA (see GemmParams::dual_m): only when the halved tile count still fills the SMs,
// directly or through the K split.
static bool want_dual_m(int M, long long tiles_n, int num_kb, int num_sms, bool atomic) {
  static const bool on = [] { const char* e = getenv("BDL_GEMM_DUALM"); return !(e && e[0] == '0'); }();
  if (!on || !atomic || M <= BM) return false;
  const long long tiles = static_cast<long long>((M + 2 * BM - 1) / (2 * BM)) * tiles_n;
  const long long max_splits = num_kb / 4 > 0 ? num_kb / 4 : 1;
  return tiles * max_splits >= num_sms;
}

// CTA-pair tiles (256 rows x block_n per 2-SM cluster) whenever they need no more waves than 128-row tiles do: the MMA
// rate per SM is the same, the B operand bytes per SM are halved. BDL_GEMM_TWO=0 disables, =1 forces (when legal).
static int g_pair_mode = [] { const char* e = getenv("BDL_GEMM_TWO"); return e ? atoi(e) : -1; }();
// Measured (profiles/gemm_pair_r2.md): pairs pay off where the B panel is re-fetched by many m tiles and the reduction is
// long — big Linear-style GEMMs (8192^3: 1.18 -> 1.53 PFLOP/s) and 512-channel 3x3 convolutions (+5..9 %); on short
// reductions and narrow outputs (the early, bandwidth-bound conv layers) the 1-CTA kernel with stacked tiles and the
// TMA-store epilogue is faster, so the default thresholds are conservative. `conv`: implicit-GEMM call.
static bool want_two(int M, int block_n, int tiles_n, bool b_mn, int num_sms, int num_kb, bool conv = false) {
  const int mode = g_pair_mode;
  if (mode == 0 || M <= BM) return false;
  if (b_mn ? (block_n % 128) != 0 : (block_n % 32) != 0) return false;
  if (mode == 1) return true;
  static const int min_kb = [] { const char* e = getenv("BDL_GEMM_TWO_MIN_KB"); return e ? atoi(e) : 16; }();
  static const int min_kb_conv = [] { const char* e = getenv("BDL_CONV_TWO_MIN_KB"); return e ? atoi(e) : 64; }();
  if (num_kb < (conv ? min_kb_conv : min_kb)) return false;
  if (conv ? block_n < 256 : tiles_n < 2) return false;
  const long long t1 = static_cast<long long>((M + BM - 1) / BM) * tiles_n;
  const long long t2 = static_cast<long long>((M + 2 * BM - 1) / (2 * BM)) * tiles_n;
  const long long w1 = (t1 + num_sms - 1) / num_sms, w2 = (t2 + num_sms / 2 - 1) / (num_sms / 2);
  // wave efficiency of the pair tiling must not fall below the 1-CTA tiling's
  return static_cast<double>(t2) / (w2 * (num_sms / 2)) + 1e-9 >= static_cast<double>(t1) / (w1 * num_sms) - 0.05;
}

// 512-row pair tiles (two stacked 256-row MMAs, all 512 TMEM columns): fewest operand bytes per MAC from L2 (the fabric,
// not the tensor pipe, bounds large GEMMs), at the price of an epilogue that no longer overlaps the next main loop —
// so only for long reductions with enough tiles to keep every pair busy.
static bool want_pair_dual(int M, int tiles_n, int num_kb, int num_sms) {
  static const int min_kb = [] { const char* e = getenv("BDL_GEMM_PAIR_DUAL_KB"); return e ? atoi(e) : 32; }();
  if (min_kb <= 0 || num_kb < min_kb) return false;
  const long long t4 = static_cast<long long>((M + 4 * BM - 1) / (4 * BM)) * tiles_n;
  const long long slots = num_sms / 2, w4 = (t4 + slots - 1) / slots;
  return t4 >= slots && static_cast<double>(t4) / (w4 * slots) >= 0.85;   // full waves only: the tile is 512 x block_n
}

template <bool AMN, bool BMN, int CV, bool FP8, bool TMAOUT, bool DUAL = false>
static auto pair_kernel() {
  if constexpr (!AMN && CV != 2)
    return &gemm_bf16_tcgen05_kernel<AMN, BMN, CV, FP8, DUAL, TMAOUT, true>;
  else
    return &gemm_bf16_tcgen05_kernel<AMN, BMN, CV, FP8, false, TMAOUT, false>;
}

// One launch path for both forms: plain grid, or 2-CTA clusters for the pair kernels.
using GemmKernel = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const GemmParams);
static int launch_gemm_kernel(GemmKernel kfn, long long total_work, int num_sms, size_t smem, cudaStream_t stream,
                              bool two, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD,
                              const GemmParams& p, const char* what) {
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBudget + 1024));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(gemm)", e);
  cudaLaunchConfig_t cfg{};
  const long long slots = two ? num_sms / 2 : num_sms;
  const int units = static_cast<int>(total_work < slots ? total_work : slots);
  cfg.gridDim = dim3(two ? 2 * units : units);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  unsigned na = 0;
  if (two) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (na) {
    cfg.attrs = attr;
    cfg.numAttrs = na;
  }
  e = cudaLaunchKernelEx(&cfg, kfn, tmA, tmB, tmD, p);
  if (e != cudaSuccess) return bdl_fail_cuda(what, e);
  return bdl_check_launch(what);
}

static int pick_block_n(int M, int N, int num_sms, bool split_k = false) {
  if (N <= 256) return ((N + 15) / 16) * 16;
  // Prefer the widest tile that still gives every SM work; otherwise fall back to narrower tiles.
  const int tiles_m = (M + BM - 1) / BM;
  const int cands[3] = {256, 192, 128};
  if (split_k && static_cast<long long>(tiles_m) * ((N + 127) / 128) < num_sms) {
    // Split-K problems (weight gradients: small M x N, huge K) fill the SMs through the K split, and every n tile
    // re-reads the whole A operand from L2, which is what bounds them: take the fewest n tiles, narrowest on a tie.
    static const bool on = [] { const char* e = getenv("BDL_WGRAD_WIDE"); return !(e && e[0] == '0'); }();
    if (on) {
      int best = 128, best_tn = (N + 127) / 128;
      for (int i = 1; i >= 0; --i) {
        const int tn = (N + cands[i] - 1) / cands[i];
        if (tn < best_tn) {
          best_tn = tn;
          best = cands[i];
        }
      }
      return best;
    }
  }
  int best = 128;
  double best_cost = 1e30;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    const int tn = (N + bn - 1) / bn;
    const long long work = static_cast<long long>(tiles_m) * tn;
    const long long waves = (work + num_sms - 1) / num_sms;
    // cost ~ waves * tile time (proportional to bn), penalise padding implicitly through tn*bn
    const double cost = static_cast<double>(waves) * bn;
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

}  // namespace bdl

using namespace bdl;

extern "C" int bdl_gemm_bf16(const void* A, const void* B, void* D, const float* bias, int M, int N, int K,
                             long long lda, long long ldb, long long ldd, int a_mn, int b_mn, int out_fp32, int flags,
                             int splits, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  GemmParams p{};
  take_residual(p);   // first: a call that fails validation must not leave the residual armed for the next GEMM
  if (M <= 0 || N <= 0 || K <= 0) return bdl_fail("bdl_gemm_bf16: empty problem");
  if ((flags & BDL_GEMM_ATOMIC_ADD) && !out_fp32) return bdl_fail("bdl_gemm_bf16: ATOMIC_ADD needs fp32 output");
  if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u) || (lda % 8) || (ldb % 8))
    return bdl_fail("bdl_gemm_bf16: A/B must be 16-byte aligned with leading dimensions that are multiples of 8");
  const int num_sms = bdl_num_sms();
  if (p.residual && ((flags & BDL_GEMM_ATOMIC_ADD) || out_fp32)) return bdl_fail("bdl_gemm_bf16: residual needs a bf16 output");
  p.M = M;
  p.N = N;
  p.K = K;
  p.block_n = pick_block_n(M, N, num_sms, (flags & BDL_GEMM_ATOMIC_ADD) != 0 && splits <= 0);
  p.b_rows = b_mn ? ((p.block_n + 63) / 64) * 64 : p.block_n;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + p.block_n - 1) / p.block_n;
  const int num_kb = (K + BK - 1) / BK;
  if (a_mn && want_dual_m(M, p.tiles_n, num_kb, num_sms, (flags & BDL_GEMM_ATOMIC_ADD) != 0)) {
    p.dual_m = 1;
    p.tiles_m = (M + 2 * BM - 1) / (2 * BM);
  }
  const bool two = !a_mn && !p.dual_m && !(flags & BDL_GEMM_ATOMIC_ADD) && splits <= 1 &&
                   want_two(M, p.block_n, p.tiles_n, b_mn != 0, num_sms, num_kb);
  if (two) {
    p.tiles_m = (M + 2 * BM - 1) / (2 * BM);
    p.b_rows = p.block_n / 2;      // each CTA of the pair stages half of the tile's B rows
    if (want_pair_dual(M, p.tiles_n, num_kb, num_sms)) {
      p.dual_m = 1;                // 512 x block_n per pair: A 32 KB + B 16 KB per k block for twice the MACs
      p.tiles_m = (M + 4 * BM - 1) / (4 * BM);
    }
  }
  if (p.tiles_n > 1 && p.tiles_m > 1) p.group_m = 8;
  const long long tiles_mn = static_cast<long long>(p.tiles_m) * p.tiles_n;
  if (splits <= 0) {
    splits = 1;
    if ((flags & BDL_GEMM_ATOMIC_ADD) && tiles_mn < num_sms) {
      splits = static_cast<int>(num_sms / tiles_mn);
      const int max_by_k = num_kb / 4 > 0 ? num_kb / 4 : 1;  // keep >= 4 k-blocks per split
      if (splits > max_by_k) splits = max_by_k;
    }
  }
  if (splits > 1 && !(flags & BDL_GEMM_ATOMIC_ADD)) return bdl_fail("bdl_gemm_bf16: split-K needs ATOMIC_ADD");
  if (splits > num_kb) splits = num_kb;
  p.splits = splits;
  p.ldd = ldd;
  p.out_fp32 = out_fp32;
  p.flags = flags;
  p.D = D;
  p.bias = bias;
  const size_t stage_bytes = static_cast<size_t>(BM) * BK * 2 * (p.dual_m ? 2 : 1) + static_cast<size_t>(p.b_rows) * BK * 2;
  // 512-row pair tiles spend the staging area on a fourth pipeline stage instead (their epilogue is a small share)
  const bool tma_out = !(two && p.dual_m) && want_tma_store(D, ldd, out_fp32, flags, a_mn != 0, num_kb);
  if (p.stat_sums != nullptr && tma_out && !(flags & BDL_GEMM_RELU)) g_stats_done = 1;   // the staged-output epilogue sums the columns
  else p.stat_sums = nullptr;
  const size_t budget = kSmemBudget - (tma_out ? kStageOutBytes : 0);
  int stages = static_cast<int>(budget / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return bdl_fail("bdl_gemm_bf16: tile does not fit in shared memory");
  p.num_stages = stages;
  const size_t smem = stages * stage_bytes + 1024 + (tma_out ? kStageOutBytes : 0);
  CUtensorMap tmD{};
  if (tma_out) {
    const int rcd = bdl_make_tmap_2d_bf16(&tmD, D, static_cast<uint64_t>(N), static_cast<uint64_t>(M), ldd, 64, 32);
    if (rcd) return rcd;
  }

  CUtensorMap tmA, tmB;
  int rc;
  if (a_mn)
    rc = bdl_make_tmap_2d_bf16(&tmA, A, static_cast<uint64_t>(M), static_cast<uint64_t>(K), lda, 64, 64);
  else
    rc = bdl_make_tmap_2d_bf16(&tmA, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M), lda, 64, BM);
  if (rc) return rc;
  if (b_mn)
    rc = bdl_make_tmap_2d_bf16(&tmB, B, static_cast<uint64_t>(N), static_cast<uint64_t>(K), ldb, 64, 64);
  else
    rc = bdl_make_tmap_2d_bf16(&tmB, B, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 64,
                               two ? p.block_n / 2 : p.block_n);
  if (rc) return rc;

  const long long total = tiles_mn * splits;
  GemmKernel kfn;
#define BDL_PICK_GEMM(AMN, BMN)                                                                                   \
  kfn = two ? (p.dual_m ? (tma_out ? pair_kernel<AMN, BMN, 0, false, !(AMN), true>()                              \
                                   : pair_kernel<AMN, BMN, 0, false, false, true>())                               \
                        : (tma_out ? pair_kernel<AMN, BMN, 0, false, !(AMN)>() : pair_kernel<AMN, BMN, 0, false, false>())) \
        : p.dual_m ? gemm_bf16_tcgen05_kernel<AMN, BMN, 0, false, true, false>                                    \
        : (tma_out && !(AMN)) ? gemm_bf16_tcgen05_kernel<AMN, BMN, 0, false, false, !(AMN)>                       \
                              : gemm_bf16_tcgen05_kernel<AMN, BMN, 0, false, false, false>
  if (a_mn && b_mn)
    BDL_PICK_GEMM(true, true);
  else if (a_mn)
    BDL_PICK_GEMM(true, false);
  else if (b_mn)
    BDL_PICK_GEMM(false, true);
  else
    BDL_PICK_GEMM(false, false);
#undef BDL_PICK_GEMM
  return launch_gemm_kernel(kfn, total, num_sms, smem, stream, two, tmA, tmB, tmD, p, "gemm_bf16_tcgen05_kernel");
}


// ------------------------------------------------------------------------------------------------ all-gather + GEMM
// D[M,N] = A[M,K] * B[N,K]^T (+bias)(ReLU) where the rows of B are sharded among `nranks` GPUs: n tiles
// [t0[r], t1[r]) (tiles of `block_n` rows) are read from peer_B[r] — rank r's copy of the same matrix, a peer-mapped
// pointer — and written through to the local B; every other tile is read from the local B. The caller guarantees (by
// acquiring the owners' ready flags, bdl_pull_gather) that the owners' rows are final before this launch.
// Reference: AllReduceParameter.getWeights (S/parameters/AllReduceParameter.scala:201-228) — the blocking all-gather at
// the top of every iteration, S/optim/DistriOptimizer.scala:212 — fused into S/nn/Linear.scala:78-104 (updateOutput).
extern "C" int bdl_gemm_bf16_gathered(const void* A, void* B_local, void* D, const float* bias, int M, int N, int K,
                                      long long lda, long long ldb, long long ldd, int out_fp32, int flags, int splits,
                                      int block_n, const void* const* peer_B, const int* t0, const int* t1, int nranks,
                                      int rank, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (M <= 0 || N <= 0 || K <= 0) return bdl_fail("bdl_gemm_bf16_gathered: empty problem");
  if (nranks < 1 || nranks > kAgMaxRanks || rank < 0 || rank >= nranks)
    return bdl_fail("bdl_gemm_bf16_gathered: 1..8 ranks");
  if (block_n < 16 || block_n > 256 || (block_n % 16)) return bdl_fail("bdl_gemm_bf16_gathered: block_n must be a multiple of 16 in [16,256]");
  if ((flags & BDL_GEMM_ATOMIC_ADD) && !out_fp32) return bdl_fail("bdl_gemm_bf16_gathered: ATOMIC_ADD needs fp32 output");
  if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B_local) & 15u) || (lda % 8) || (ldb % 8))
    return bdl_fail("bdl_gemm_bf16_gathered: A/B must be 16-byte aligned with leading dimensions that are multiples of 8");
  const int num_sms = bdl_num_sms();
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.block_n = block_n;
  p.b_rows = block_n;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + block_n - 1) / block_n;
  const int num_kb = (K + BK - 1) / BK;
  const long long tiles_mn = static_cast<long long>(p.tiles_m) * p.tiles_n;
  if (splits <= 0) {
    splits = 1;
    if ((flags & BDL_GEMM_ATOMIC_ADD) && tiles_mn < num_sms) {
      splits = static_cast<int>(num_sms / tiles_mn);
      const int max_by_k = num_kb / 4 > 0 ? num_kb / 4 : 1;
      if (splits > max_by_k) splits = max_by_k;
    }
  }
  if (splits > 1 && !(flags & BDL_GEMM_ATOMIC_ADD)) return bdl_fail("bdl_gemm_bf16_gathered: split-K needs ATOMIC_ADD");
  if (splits > num_kb) splits = num_kb;
  p.splits = splits;
  p.ldd = ldd; p.out_fp32 = out_fp32; p.flags = flags; p.D = D; p.bias = bias;
  const size_t stage_bytes = static_cast<size_t>(BM) * BK * 2 + static_cast<size_t>(p.b_rows) * BK * 2;
  const bool tma_out = want_tma_store(D, ldd, out_fp32, flags, false, num_kb);
  const size_t budget = kSmemBudget - (tma_out ? kStageOutBytes : 0);
  int stages = static_cast<int>(budget / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  p.num_stages = stages;
  const size_t smem = stages * stage_bytes + 1024 + (tma_out ? kStageOutBytes : 0);
  CUtensorMap tmA, tmB, tmD{};
  int rc = bdl_make_tmap_2d_bf16(&tmA, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M), lda, 64, BM);
  if (rc) return rc;
  rc = bdl_make_tmap_2d_bf16(&tmB, B_local, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 64, block_n);
  if (rc) return rc;
  if (tma_out) {
    rc = bdl_make_tmap_2d_bf16(&tmD, D, static_cast<uint64_t>(N), static_cast<uint64_t>(M), ldd, 64, 32);
    if (rc) return rc;
  }
  AgMaps ag{};
  ag.nranks = nranks;
  ag.rank = rank;
  for (int r = 0; r < nranks; ++r) {
    ag.t0[r] = t0[r];
    ag.t1[r] = t1[r];
    if (r == rank || t0[r] >= t1[r]) {
      ag.t0[r] = ag.t1[r] = 0;
      ag.tm[r] = tmB;
      continue;
    }
    if (t0[r] < 0 || t1[r] > p.tiles_n) return bdl_fail("bdl_gemm_bf16_gathered: tile range outside the matrix");
    if (!peer_B[r] || (reinterpret_cast<uintptr_t>(peer_B[r]) & 15u)) return bdl_fail("bdl_gemm_bf16_gathered: bad peer pointer");
    // an ordinary tiled map over the owner's buffer; the kernel only ever issues plain (non-multicast) loads through it
    rc = bdl_make_tmap_2d_bf16(&ag.tm[r], peer_B[r], static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 64, block_n);
    if (rc) return rc;
  }
  auto kfn = tma_out ? gemm_bf16_allgather_kernel<true> : gemm_bf16_allgather_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBudget + 1024));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(allgather gemm)", e);
  const long long total = tiles_mn * splits;
  const int grid = static_cast<int>(total < num_sms ? total : num_sms);
  kfn<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tmD, p, ag);
  return bdl_check_launch("gemm_bf16_allgather_kernel");
}

// ------------------------------------------------------------------------------------------------ implicit-GEMM conv
// mode 0 (fprop):  D[N*Ho*Wo, O]      = im2col(act)[., (tap, c)] * Wt[O, (tap, c)]^T          (+bias)(ReLU)
// mode 1 (dgrad):  D[N*Ho*Wo, O]      = im2col(act)[., (tap, c)] * Wt[c, (flip(tap), o)]       act = dY, O = Cin; the caller
//                  passes the transposed-convolution geometry (Ho/Wo = input size, pad' = dil*(k-1) - pad, stride 1)
// mode 2 (wgrad):  D[O, (tap, c)]    += other[N*Ho*Wo, O]^T * im2col(act)[., (tap, c)]          other = dY, fp32 red.add
// act is an NHWC tensor [N, H, W, C] with pixel pitch ld_act; `other` is the weight matrix (modes 0/1, row pitch
// ld_other) or dY (mode 2). Reference: S/nn/SpatialConvolution.scala:253-521, S/nn/mkldnn/SpatialConvolution.scala.
extern "C" int bdl_conv_igemm_bf16(int mode, const void* act, int N, int H, int W, int C, long long ld_act,
                                   const void* other, long long ld_other, void* D, long long ldd, const float* bias,
                                   int O, int Ho, int Wo, int kh, int kw, int sh, int sw, int ph, int pw, int dh,
                                   int dw, int out_fp32, int flags, int splits, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  GemmParams p{};
  take_residual(p);
  if (mode < 0 || mode > 2) return bdl_fail("bdl_conv_igemm_bf16: mode must be 0 (fprop), 1 (dgrad) or 2 (wgrad)");
  if ((reinterpret_cast<uintptr_t>(act) & 15u) || (reinterpret_cast<uintptr_t>(other) & 15u) || (ld_act % 8) ||
      (ld_other % 8))
    return bdl_fail("bdl_conv_igemm_bf16: operands must be 16-byte aligned with pitches that are multiples of 8");
  if (mode == 2 && (C % 8)) return bdl_fail("bdl_conv_igemm_bf16: wgrad needs C % 8 == 0");
  if (mode == 2 && !(flags & BDL_GEMM_ATOMIC_ADD)) return bdl_fail("bdl_conv_igemm_bf16: wgrad accumulates (ATOMIC_ADD)");
  const int num_sms = bdl_num_sms();
  const long long Mpix = static_cast<long long>(N) * Ho * Wo;
  if (Mpix >= (1ll << 31)) return bdl_fail("bdl_conv_igemm_bf16: too many output pixels");
  const int taps = kh * kw;
  if (p.residual && (mode == 2 || out_fp32)) return bdl_fail("bdl_conv_igemm_bf16: the residual operand needs a bf16 fprop / dgrad output");
  p.cv_Ho = Ho; p.cv_Wo = Wo; p.cv_sh = sh; p.cv_sw = sw; p.cv_ph = ph; p.cv_pw = pw; p.cv_dh = dh; p.cv_dw = dw;
  p.cv_kh = kh; p.cv_kw = kw; p.cv_C = C; p.cv_cblocks = (C + 63) / 64; p.cv_flip = (mode == 1); p.cv_nimg = N;
  p.fd_hw.init(static_cast<uint32_t>(Ho) * Wo);
  p.fd_wo.init(static_cast<uint32_t>(Wo));
  p.fd_cblocks.init(static_cast<uint32_t>(p.cv_cblocks));
  p.fd_kw.init(static_cast<uint32_t>(kw));
  // Few-channel inputs (C <= 32: Inception's 5x5 reductions, the space-to-depth RGB stems) are gathered with 32- or
  // 16-channel boxes into 64B / 32B-swizzled atoms, so a 64-wide k block (fprop) or N atom (wgrad) holds 2 or 4 filter
  // taps instead of one tap plus zero padding.
  static const bool small_ok = [] { const char* e = getenv("BDL_IGEMM_SMALLC"); return !(e && e[0] == '0'); }();
  const int aw = (small_ok && mode != 1 && C <= 32) ? (C <= 16 ? 16 : 32) : 64;
  p.cv_aw = aw;
  // wgrad pads every tap's channel range to whole atoms on the GEMM N axis (see EPI_STORE in the kernel)
  const int c_pad = aw < 64 ? aw : p.cv_cblocks * 64;
  p.fd_c.init(static_cast<uint32_t>(mode == 2 ? c_pad : C));
  int gm, gn;
  if (mode == 2) {
    gm = O;
    gn = taps * c_pad;
    p.K = static_cast<int>(Mpix);
    p.cv_num_kb = static_cast<int>((Mpix + 63) / 64);
  } else {
    gm = static_cast<int>(Mpix);
    gn = O;
    p.K = taps * p.cv_cblocks * 64;
    p.cv_num_kb = aw < 64 ? (taps * aw + 63) / 64 : taps * p.cv_cblocks;
  }
  p.M = gm;
  p.N = gn;
  p.block_n = pick_block_n(gm, gn, num_sms, mode == 2 && splits <= 0);
  const bool b_mn = (mode != 0);
  p.b_rows = b_mn ? ((p.block_n + 63) / 64) * 64 : p.block_n;
  if (mode == 2 && aw < 64) p.b_rows = ((p.block_n + aw - 1) / aw) * aw;
  p.tiles_m = (gm + BM - 1) / BM;
  p.tiles_n = (gn + p.block_n - 1) / p.block_n;
  if (mode == 2 && want_dual_m(gm, p.tiles_n, p.cv_num_kb, num_sms, true)) {
    p.dual_m = 1;
    p.tiles_m = (gm + 2 * BM - 1) / (2 * BM);
  }
  if (mode != 2 && aw == 64) {
    // fprop / dgrad with a long reduction (K >= 1024) and several waves of tiles: 256 pixels per CTA halve the L2
    // reads of the weight operand per MAC (these GEMMs are bound by L2 -> SM bandwidth, not by the tensor pipe); the
    // price is an epilogue that no longer overlaps the next tile's main loop, hence the K threshold.
    static const int min_kb = [] { const char* e = getenv("BDL_GEMM_DUALM_FPROP_KB"); return e ? atoi(e) : 8; }();
    const long long pairs = static_cast<long long>((gm + 2 * BM - 1) / (2 * BM)) * p.tiles_n;
    if (min_kb > 0 && p.cv_num_kb >= min_kb && pairs >= 2ll * num_sms) {
      p.dual_m = 1;
      p.tiles_m = (gm + 2 * BM - 1) / (2 * BM);
    }
  }
  // fprop / dgrad on CTA pairs (256 pixels x block_n per 2-SM cluster, half of the weight tile per CTA) replace the
  // stacked-tile form wherever the pair tiling is legal
  const bool two = mode != 2 && aw == 64 && !(flags & BDL_GEMM_ATOMIC_ADD) && splits <= 1 &&
                   want_two(gm, p.block_n, p.tiles_n, b_mn, num_sms, p.cv_num_kb, true);
  if (two) {
    p.dual_m = 0;
    p.tiles_m = (gm + 2 * BM - 1) / (2 * BM);
    p.b_rows = p.block_n / 2;
    if (want_pair_dual(gm, p.tiles_n, p.cv_num_kb, num_sms)) {
      p.dual_m = 1;
      p.tiles_m = (gm + 4 * BM - 1) / (4 * BM);
    }
  }
  if (p.tiles_n > 1 && p.tiles_m > 1) p.group_m = 8;
  const long long tiles_mn = static_cast<long long>(p.tiles_m) * p.tiles_n;
  if (splits <= 0) {
    splits = 1;
    if ((flags & BDL_GEMM_ATOMIC_ADD) && tiles_mn < num_sms) {
      splits = static_cast<int>(num_sms / tiles_mn);
      const int max_by_k = p.cv_num_kb / 4 > 0 ? p.cv_num_kb / 4 : 1;
      if (splits > max_by_k) splits = max_by_k;
    }
  }
  if (splits > 1 && !(flags & BDL_GEMM_ATOMIC_ADD)) return bdl_fail("bdl_conv_igemm_bf16: split-K needs ATOMIC_ADD");
  if (splits > p.cv_num_kb) splits = p.cv_num_kb;
  p.splits = splits;
  p.ldd = ldd;
  p.out_fp32 = out_fp32;
  p.flags = flags;
  p.D = D;
  p.bias = bias;
  const size_t stage_bytes = static_cast<size_t>(BM) * BK * 2 * (p.dual_m ? 2 : 1) + static_cast<size_t>(p.b_rows) * BK * 2;
  const bool tma_out = !(two && p.dual_m) && want_tma_store(D, ldd, out_fp32, flags, mode == 2, p.cv_num_kb);
  if (p.stat_sums != nullptr && tma_out && mode != 2 && !(flags & BDL_GEMM_RELU)) g_stats_done = 1;
  else p.stat_sums = nullptr;
  const size_t budget = kSmemBudget - (tma_out ? kStageOutBytes : 0);
  int stages = static_cast<int>(budget / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return bdl_fail("bdl_conv_igemm_bf16: tile does not fit in shared memory");
  p.num_stages = stages;
  const size_t smem = stages * stage_bytes + 1024 + (tma_out ? kStageOutBytes : 0);
  CUtensorMap tmD{};
  if (tma_out) {
    const int rcd = bdl_make_tmap_2d_bf16(&tmD, D, static_cast<uint64_t>(gn), static_cast<uint64_t>(gm), ldd, 64, 32);
    if (rcd) return rcd;
  }

  // bounding box of base pixel positions: [-pad, dim - 1 + pad - dil*(k-1)] (inclusive), traversal stride = conv stride
  const int up_w = pw - dw * (kw - 1), up_h = ph - dh * (kh - 1);
  CUtensorMap tmA, tmB;
  int rc;
  if (mode == 2) {
    rc = bdl_make_tmap_2d_bf16(&tmA, other, static_cast<uint64_t>(O), static_cast<uint64_t>(Mpix), ld_other, 64, 64);
    if (rc) return rc;
    rc = bdl_make_tmap_im2col_bf16_sw(&tmB, act, C, W, H, N, ld_act, -pw, -ph, up_w, up_h, aw, 64, sw, sh, aw * 2);
    if (rc) return rc;
  } else {
    rc = bdl_make_tmap_im2col_bf16_sw(&tmA, act, C, W, H, N, ld_act, -pw, -ph, up_w, up_h, aw, BM, sw, sh, aw * 2);
    if (rc) return rc;
    if (mode == 0)
      rc = bdl_make_tmap_2d_bf16_sw(&tmB, other, static_cast<uint64_t>(taps) * C, static_cast<uint64_t>(O), ld_other,
                                    aw, two ? p.block_n / 2 : p.block_n, aw * 2);
    else  // weights [Cact rows, taps*O cols]
      rc = bdl_make_tmap_2d_bf16(&tmB, other, static_cast<uint64_t>(taps) * O, static_cast<uint64_t>(C), ld_other, 64,
                                 64);
    if (rc) return rc;
  }
  const long long total = tiles_mn * splits;
  GemmKernel kfn;
#define BDL_PICK_CONV(AMN, BMN, CV)                                                                               \
  kfn = two ? (p.dual_m ? (tma_out ? pair_kernel<AMN, BMN, CV, false, !(AMN), true>()                             \
                                   : pair_kernel<AMN, BMN, CV, false, false, true>())                              \
                        : (tma_out ? pair_kernel<AMN, BMN, CV, false, !(AMN)>() : pair_kernel<AMN, BMN, CV, false, false>())) \
        : p.dual_m ? (tma_out && !(AMN) ? gemm_bf16_tcgen05_kernel<AMN, BMN, CV, false, true, !(AMN)>             \
                                        : gemm_bf16_tcgen05_kernel<AMN, BMN, CV, false, true, false>)             \
                   : (tma_out && !(AMN) ? gemm_bf16_tcgen05_kernel<AMN, BMN, CV, false, false, !(AMN)>            \
                                        : gemm_bf16_tcgen05_kernel<AMN, BMN, CV, false, false, false>)
  if (mode == 0)
    BDL_PICK_CONV(false, false, 1);
  else if (mode == 1)
    BDL_PICK_CONV(false, true, 1);
  else
    BDL_PICK_CONV(true, true, 2);
#undef BDL_PICK_CONV
  return launch_gemm_kernel(kfn, total, num_sms, smem, stream, two, tmA, tmB, tmD, p, "gemm_bf16_tcgen05_kernel<conv>");
}


// ------------------------------------------------------------------------------------------------ fp8 (e4m3) inference
// D[M,N] = (A_q[M,K] * B_q[N,K]^T) * col_scale[n] (+bias)(ReLU); A_q/B_q are e4m3 bytes, K-major, pitches in bytes
// (multiples of 16). With conv != 0 the A operand is gathered from an NHWC e4m3 tensor by TMA im2col (fprop only).
// This is the B200 form of the reference's int8 paths: BigQuant MixPrecisionGEMM (core/bigquant, S/nn/quantized/*) and
// MKL-DNN x8s8s32x convolution with output scales (S/nn/mkldnn/SpatialConvolution.scala:237-247,325-372).
extern "C" int bdl_gemm_fp8(const void* A, const void* B, void* D, const float* bias, const float* col_scale,
                            const float* out_qscale, const void* conv_geom, int M, int N, int K, long long lda, long long ldb, long long ldd,
                            int out_fp32, int flags, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  GemmParams p{};
  take_residual(p);
  if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u) || (lda % 16) || (ldb % 16))
    return bdl_fail("bdl_gemm_fp8: operands must be 16-byte aligned with pitches that are multiples of 16 bytes");
  if (flags & BDL_GEMM_ATOMIC_ADD) return bdl_fail("bdl_gemm_fp8: inference kernel, no accumulate mode");
  const int num_sms = bdl_num_sms();
  const int* cg = static_cast<const int*>(conv_geom);   // {N,H,W,C,Ho,Wo,kh,kw,sh,sw,ph,pw,dh,dw} or NULL
  int gm = M;
  if (cg) {
    p.cv_Ho = cg[4]; p.cv_Wo = cg[5]; p.cv_kh = cg[6]; p.cv_kw = cg[7]; p.cv_sh = cg[8]; p.cv_sw = cg[9];
    p.cv_ph = cg[10]; p.cv_pw = cg[11]; p.cv_dh = cg[12]; p.cv_dw = cg[13]; p.cv_C = cg[3]; p.cv_nimg = cg[0];
    p.cv_aw = 64;
    p.cv_cblocks = (cg[3] + 127) / 128;
    p.cv_num_kb = cg[6] * cg[7] * p.cv_cblocks;
    p.fd_hw.init(static_cast<uint32_t>(cg[4]) * cg[5]);
    p.fd_wo.init(static_cast<uint32_t>(cg[5]));
    p.fd_cblocks.init(static_cast<uint32_t>(p.cv_cblocks));
    p.fd_kw.init(static_cast<uint32_t>(cg[7]));
    p.fd_c.init(static_cast<uint32_t>(cg[3]));
    gm = cg[0] * cg[4] * cg[5];
    K = cg[6] * cg[7] * cg[3];
  }
  p.M = gm;
  p.N = N;
  p.K = K;
  p.block_n = pick_block_n(gm, N, num_sms);
  p.b_rows = p.block_n;
  p.tiles_m = (gm + BM - 1) / BM;
  p.tiles_n = (N + p.block_n - 1) / p.block_n;
  p.splits = 1;
  p.ldd = ldd;
  p.out_fp32 = out_fp32;
  p.flags = flags;
  p.D = D;
  p.bias = bias;
  p.col_scale = col_scale;
  p.out_qscale = out_qscale;
  if (out_fp32 == 2 && !out_qscale) return bdl_fail("bdl_gemm_fp8: e4m3 output needs out_qscale");
  const size_t stage_bytes = static_cast<size_t>(BM) * 128 + static_cast<size_t>(p.b_rows) * 128;
  // bf16 (kind 0) and e4m3 (kind 2) outputs can go through the staged TMA-store epilogue
  const long long d_esize = out_fp32 == 2 ? 1 : 2;
  const bool tma_out = (out_fp32 == 0 || out_fp32 == 2) && (reinterpret_cast<uintptr_t>(D) & 15u) == 0 &&
                       (ldd * d_esize) % 16 == 0 && want_tma_store(D, 8, 0, flags, false, 0);
  int stages = static_cast<int>((kSmemBudget - (tma_out ? kStageOutBytes : 0)) / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  p.num_stages = stages;
  const size_t smem = stages * stage_bytes + 1024 + (tma_out ? kStageOutBytes : 0);
  CUtensorMap tmD{};
  if (tma_out) {
    const int rcd = out_fp32 == 2
                        ? bdl_make_tmap_2d_u8_sw(&tmD, D, static_cast<uint64_t>(N), static_cast<uint64_t>(gm), ldd, 64, 32, 64)
                        : bdl_make_tmap_2d_bf16(&tmD, D, static_cast<uint64_t>(N), static_cast<uint64_t>(gm), ldd, 64, 32);
    if (rcd) return rcd;
  }
  CUtensorMap tmA, tmB;
  int rc;
  if (cg)
    rc = bdl_make_tmap_im2col_u8(&tmA, A, cg[3], cg[2], cg[1], cg[0], lda, -cg[11], -cg[10],
                                 cg[11] - cg[13] * (cg[7] - 1), cg[10] - cg[12] * (cg[6] - 1), 128, BM, cg[9], cg[8]);
  else
    rc = bdl_make_tmap_2d_u8(&tmA, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M), lda, 128, BM);
  if (rc) return rc;
  rc = bdl_make_tmap_2d_u8(&tmB, B, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 128, p.block_n);
  if (rc) return rc;
  const long long total = static_cast<long long>(p.tiles_m) * p.tiles_n;
  const int grid = static_cast<int>(total < num_sms ? total : num_sms);
  if (cg) {
    auto kfn = tma_out ? gemm_bf16_tcgen05_kernel<false, false, 1, true, false, true>
                       : gemm_bf16_tcgen05_kernel<false, false, 1, true, false, false>;
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBudget + 1024));
    if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(fp8 conv)", e);
    kfn<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tma_out ? tmD : tmA, p);
  } else {
    auto kfn = tma_out ? gemm_bf16_tcgen05_kernel<false, false, 0, true, false, true>
                       : gemm_bf16_tcgen05_kernel<false, false, 0, true, false, false>;
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBudget + 1024));
    if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(fp8 gemm)", e);
    kfn<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tma_out ? tmD : tmA, p);
  }
  return bdl_check_launch("gemm_tcgen05_kernel<fp8>");
}

// ------------------------------------------------------------------------------------------------ block-scaled fp8
// D[M,N] = sum over 32-wide k blocks of (A_q 2^(sfa-127)) (B_q 2^(sfb-127))^T (+bias)(ReLU): e4m3 operands with one
// UE8M0 scale per (row, 32 k), multiplied by tcgen05.mma.kind::mxf8f6f4.block_scale with the scale bytes staged in TMEM.
// A_q [M, lda] / B_q [N, ldb]: e4m3 bytes, K-major, pitches multiples of 16 with zeroed padding; sfa / sfb: scale atoms
// from bdl_quantize_mxfp8. Reference: BigQuant's per-window ("local") scales, core/bigquant/native/ops/shuffle/
// shuffle_im2col.h and S/nn/quantized/Quantization.scala:35-62; S/nn/quantized/Linear.scala.
extern "C" int bdl_gemm_mxfp8(const void* A, const void* B, const void* sfa, const void* sfb, void* D, const float* bias,
                              int M, int N, int K, long long lda, long long ldb, long long ldd, int out_fp32, int flags,
                              void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (M <= 0 || N <= 0 || K <= 0) return bdl_fail("bdl_gemm_mxfp8: empty problem");
  if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u) || (lda % 16) || (ldb % 16))
    return bdl_fail("bdl_gemm_mxfp8: operands must be 16-byte aligned with pitches that are multiples of 16 bytes");
  if (!sfa || !sfb || (reinterpret_cast<uintptr_t>(sfa) & 15u) || (reinterpret_cast<uintptr_t>(sfb) & 15u))
    return bdl_fail("bdl_gemm_mxfp8: scale-factor atoms must be 16-byte aligned");
  if (flags & BDL_GEMM_ATOMIC_ADD) return bdl_fail("bdl_gemm_mxfp8: inference kernel, no accumulate mode");
  if (out_fp32 == 2) return bdl_fail("bdl_gemm_mxfp8: bf16 or fp32 output");
  const int num_sms = bdl_num_sms();
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  // one scale atom covers 128 rows, and the TMEM columns beside the two 128-column accumulators hold the scale bytes
  p.block_n = N <= 128 ? ((N + 15) / 16) * 16 : 128;
  p.b_rows = p.block_n;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + p.block_n - 1) / p.block_n;
  p.splits = 1;
  p.ldd = ldd; p.out_fp32 = out_fp32; p.flags = flags; p.D = D; p.bias = bias;
  p.mx_sfa = static_cast<const uint8_t*>(sfa);
  p.mx_sfb = static_cast<const uint8_t*>(sfb);
  if (p.tiles_n > 1 && p.tiles_m > 1) p.group_m = 8;
  const size_t stage_bytes = static_cast<size_t>(BM) * 128 + static_cast<size_t>(p.b_rows) * 128;
  const bool tma_out = out_fp32 == 0 && (reinterpret_cast<uintptr_t>(D) & 15u) == 0 && (ldd % 8) == 0;
  int stages = static_cast<int>((kSmemBudget - (tma_out ? kStageOutBytes : 0)) / (stage_bytes + 1024));
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return bdl_fail("bdl_gemm_mxfp8: tile does not fit in shared memory");
  p.num_stages = stages;
  const size_t smem = stages * (stage_bytes + 1024) + 1024 + (tma_out ? kStageOutBytes : 0);
  CUtensorMap tmA, tmB, tmD{};
  int rc = bdl_make_tmap_2d_u8(&tmA, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M), lda, 128, BM);
  if (rc) return rc;
  rc = bdl_make_tmap_2d_u8(&tmB, B, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 128, p.block_n);
  if (rc) return rc;
  if (tma_out) {
    rc = bdl_make_tmap_2d_bf16(&tmD, D, static_cast<uint64_t>(N), static_cast<uint64_t>(M), ldd, 64, 32);
    if (rc) return rc;
  }
  GemmKernel kfn = tma_out ? gemm_bf16_tcgen05_kernel<false, false, 0, true, false, true>
                           : gemm_bf16_tcgen05_kernel<false, false, 0, true, false, false>;
  return launch_gemm_kernel(kfn, static_cast<long long>(p.tiles_m) * p.tiles_n, num_sms, smem, stream, false, tmA, tmB,
                            tma_out ? tmD : tmA, p, "gemm_tcgen05_kernel<mxfp8>");
}

// -1: CTA-pair tiles where the heuristic wants them (default), 0: never, 1: wherever they are legal. Returns the old mode.
extern "C" int bdl_gemm_set_pair_mode(int mode) {
  const int old = g_pair_mode;
  g_pair_mode = mode < 0 ? -1 : (mode > 0 ? 1 : 0);
  return old;
}

// Arms per-column statistics for the NEXT bdl_gemm_bf16 / bdl_conv_igemm_bf16 on this thread: sums [2 * N] fp32 receives
// += (sum, sum of squares) of the bf16 output columns. Only kernels that stage the output for TMA stores can do it;
// bdl_gemm_stats_done() tells whether the launch did (otherwise run bdl_bn_stats on the output).
extern "C" int bdl_gemm_set_stats(float* sums) {
  g_next_stats = sums;
  return BDL_OK;
}
extern "C" int bdl_gemm_stats_done() { return g_stats_done; }

// The residual armed for the next launch is a ReLU mask source: out = residual > 0 ? out : 0 (see GemmParams::res_mask).
extern "C" int bdl_gemm_set_residual_mask(int on) {
  g_next_res_mask = on ? 1 : 0;
  return BDL_OK;
}

extern "C" int bdl_gemm_set_residual(const void* residual_bf16, long long ld) {
  if (residual_bf16 && ld <= 0) return bdl_fail("bdl_gemm_set_residual: row pitch must be positive");
  g_next_residual = residual_bf16;
  g_next_ldr = ld;
  return BDL_OK;
}
