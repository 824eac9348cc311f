// bf16 GEMM on 5th-gen tensor cores: D[M,N] = opA(A) * opB(B)^T (+bias)(ReLU), fp32 accumulation in TMEM.
//
// One persistent CTA per SM, 6 warps:
//   warp 0      TMA producer   (cp.async.bulk.tensor.2d -> 128B-swizzled smem ring, mbarrier complete_tx)
//   warp 1      MMA issuer     (one lane issues tcgen05.mma.cta_group::1.kind::f16, tcgen05.commit frees slots)
//   warps 2..5  epilogue       (tcgen05.ld TMEM -> registers -> bias/ReLU -> bf16|fp32 global, or fp32 red.add)
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
#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int kMaxStages = 8;
constexpr int kGemmThreads = 192;
constexpr int kTmemCols = 512;
constexpr int kAccStride = 256;  // TMEM columns between the two accumulator stages

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
};

template <bool kAMN, bool kBMN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = BM * BK * 2;
  const uint32_t b_bytes = static_cast<uint32_t>(p.b_rows) * BK * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&tfull_bar[s]), 1);
      mbar_init(smem_u32(&tempty_bar[s]), 4);
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc<kTmemCols>(smem_u32(&tmem_base_slot));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  const int num_kb = (p.K + BK - 1) / BK;
  const int tiles_mn = p.tiles_m * p.tiles_n;
  const int total_work = tiles_mn * p.splits;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int work = blockIdx.x; work < total_work; work += gridDim.x) {
        const int split = work / tiles_mn;
        const int rem = work - split * tiles_mn;
        const int m_blk = rem / p.tiles_n;
        const int n_blk = rem - m_blk * p.tiles_n;
        const int kb0 = static_cast<int>((static_cast<long long>(split) * num_kb) / p.splits);
        const int kb1 = static_cast<int>((static_cast<long long>(split + 1) * num_kb) / p.splits);
        const int m0 = m_blk * BM;
        const int n0 = n_blk * p.block_n;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          const uint32_t sa = tiles + stage * stage_bytes;
          const uint32_t sb = sa + a_bytes;
          mbar_expect_tx(fb, stage_bytes);
          const int k0 = kb * BK;
          if (kAMN) {
            // A stored [K, M]: two boxes of {64 m, 64 k}
            tma_load_2d(sa, &tmA, fb, m0, k0);
            tma_load_2d(sa + 8192u, &tmA, fb, m0 + 64, k0);
          } else {
            tma_load_2d(sa, &tmA, fb, k0, m0);  // box {64 k, 128 m}
          }
          if (kBMN) {
            for (int j = 0; j < p.b_rows / 64; ++j) tma_load_2d(sb + j * 8192u, &tmB, fb, n0 + 64 * j, k0);
          } else {
            tma_load_2d(sb, &tmB, fb, k0, n0);  // box {64 k, block_n}
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
    if (lane == 0) {
      // instruction descriptor: c=f32 [4,6), a=bf16 [7,10), b=bf16 [10,13), a_major 15, b_major 16, N>>3 [17,23), M>>4 [24,29)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kAMN) << 15) |
                             (static_cast<uint32_t>(kBMN) << 16) | (static_cast<uint32_t>(p.block_n >> 3) << 17) |
                             (static_cast<uint32_t>(BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (int work = blockIdx.x; work < total_work; work += gridDim.x) {
        const int split = work / tiles_mn;
        const int kb0 = static_cast<int>((static_cast<long long>(split) * num_kb) / p.splits);
        const int kb1 = static_cast<int>((static_cast<long long>(split + 1) * num_kb) / p.splits);
        mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kAccStride;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t sa = tiles + stage * stage_bytes;
          const uint32_t sb = sa + a_bytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t adesc = kAMN ? make_smem_desc_sw128(sa + k * 2048u, 8192u, 1024u)
                                        : make_smem_desc_sw128(sa + k * 32u, 16u, 1024u);
            const uint64_t bdesc = kBMN ? make_smem_desc_sw128(sb + k * 2048u, 8192u, 1024u)
                                        : make_smem_desc_sw128(sb + k * 32u, 16u, 1024u);
            umma_f16(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(smem_u32(&empty_bar[stage]));
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(smem_u32(&tfull_bar[acc]));
        acc ^= 1u;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..5)
    const uint32_t quad = warp & 3u;  // TMEM lane quadrant this warp may read
    uint32_t acc = 0, acc_phase = 0;
    const bool has_bias = (p.flags & BDL_GEMM_BIAS) != 0;
    const bool relu = (p.flags & BDL_GEMM_RELU) != 0;
    const bool atomic = (p.flags & BDL_GEMM_ATOMIC_ADD) != 0;
    for (int work = blockIdx.x; work < total_work; work += gridDim.x) {
      const int split = work / tiles_mn;
      const int rem = work - split * tiles_mn;
      const int m_blk = rem / p.tiles_n;
      const int n_blk = rem - m_blk * p.tiles_n;
      const int row = m_blk * BM + static_cast<int>(quad * 32u + lane);
      const int nbase = n_blk * p.block_n;
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((quad * 32u) << 16) + acc * kAccStride;
      const bool row_ok = row < p.M;
      for (int c = 0; c < p.block_n; c += 16) {
        uint32_t r[16];
        tmem_ld_16(t_row + static_cast<uint32_t>(c), r);
        tmem_ld_wait();
        const int n0 = nbase + c;
        if (!row_ok || n0 >= p.N) continue;
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
        const bool full = (n0 + 16 <= p.N);
        if (has_bias && (!atomic || split == 0)) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (full || n0 + j < p.N) v[j] += __ldg(p.bias + n0 + j);
        }
        if (relu) {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (p.out_fp32) {
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
                if (n0 + j < p.N) atomicAdd(dst + j, v[j]);
            }
          } else if (al) {
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
            for (int j = 0; j < 16; ++j)
              if (n0 + j < p.N) dst[j] = v[j];
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
              if (n0 + j < p.N) dst[j] = __float2bfloat16_rn(v[j]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
      acc ^= 1u;
      if (acc == 0) acc_phase ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<kTmemCols>(tmem_base);
}

// ------------------------------------------------------------------------------------------- host
static int pick_block_n(int M, int N, int num_sms) {
  if (N <= 256) return ((N + 15) / 16) * 16;
  // Prefer the widest tile that still gives every SM work; otherwise fall back to narrower tiles.
  const int tiles_m = (M + BM - 1) / BM;
  const int cands[3] = {256, 192, 128};
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
  if (M <= 0 || N <= 0 || K <= 0) return bdl_fail("bdl_gemm_bf16: empty problem");
  if ((flags & BDL_GEMM_ATOMIC_ADD) && !out_fp32) return bdl_fail("bdl_gemm_bf16: ATOMIC_ADD needs fp32 output");
  if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u) || (lda % 8) || (ldb % 8))
    return bdl_fail("bdl_gemm_bf16: A/B must be 16-byte aligned with leading dimensions that are multiples of 8");
  const int num_sms = bdl_num_sms();
  GemmParams p{};
  p.M = M;
  p.N = N;
  p.K = K;
  p.block_n = pick_block_n(M, N, num_sms);
  p.b_rows = b_mn ? ((p.block_n + 63) / 64) * 64 : p.block_n;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + p.block_n - 1) / p.block_n;
  const int num_kb = (K + BK - 1) / BK;
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
  const size_t stage_bytes = static_cast<size_t>(BM) * BK * 2 + static_cast<size_t>(p.b_rows) * BK * 2;
  const size_t budget = 227 * 1024 - 1024 - 256;
  int stages = static_cast<int>(budget / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return bdl_fail("bdl_gemm_bf16: tile does not fit in shared memory");
  p.num_stages = stages;
  const size_t smem = stages * stage_bytes + 1024;

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
    rc = bdl_make_tmap_2d_bf16(&tmB, B, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 64, p.block_n);
  if (rc) return rc;

  const long long total = tiles_mn * splits;
  const int grid = static_cast<int>(total < num_sms ? total : num_sms);
#define BDL_LAUNCH_GEMM(AMN, BMN)                                                                                 \
  do {                                                                                                            \
    auto kfn = gemm_bf16_tcgen05_kernel<AMN, BMN>;                                                                \
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 256);     \
    if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(gemm)", e);                                  \
    kfn<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, p);                                                       \
  } while (0)
  if (a_mn && b_mn)
    BDL_LAUNCH_GEMM(true, true);
  else if (a_mn)
    BDL_LAUNCH_GEMM(true, false);
  else if (b_mn)
    BDL_LAUNCH_GEMM(false, true);
  else
    BDL_LAUNCH_GEMM(false, false);
#undef BDL_LAUNCH_GEMM
  return bdl_check_launch("gemm_bf16_tcgen05_kernel");
}
