// Batch normalisation for NHWC bf16 activations (rows = N*H*W pixels, columns = channels), training and inference,
// with optional fused ReLU in both directions. All kernels are HBM-bound: every thread owns 8 channels (one 16-byte
// vector) and walks rows; per-channel partial sums are combined through shared memory and one fp32 atomic per
// channel per block. The statistics buffer is separate from the apply step so that the [2, C] sums can be
// all-reduced across ranks in between (sync-BN, the cross-replica variant of S/nn/BatchNormalization.scala:65-76,
// 224-235 built on S/utils/ParameterSynchronizer.scala).
//
// Semantics follow S/nn/BatchNormalization.scala:40-110 and SpatialBatchNormalization.scala:485-494:
//   y = gamma * (x - mean) / sqrt(var_biased + eps) + beta
//   running = (1 - momentum) * running + momentum * batch, with the UNBIASED batch variance for runningVar.
// mkldnn twin: S/nn/mkldnn/SpatialBatchNormalization.scala (fused ReLU :44-83).
#include <cstdlib>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

__device__ __forceinline__ void bn_load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 v = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = __bfloat1622float2(q[j]);
    f[2 * j] = t.x;
    f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ void bn_store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]);
  o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]);
  o.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = o;
}

// Thread layout shared by the reductions: vcols = min(C/8, 256) vector columns across, rif = 256/vcols rows in flight.
// acc0/acc1 are two per-channel accumulators (sum & sum of squares, or dbeta & dgamma).
// kRemask (backward only): the layer had a fused ReLU and its output is NOT read back for the mask — it is recomputed
// from x with exactly the forward's arithmetic (y > 0  <=>  fma(x, invstd*gamma, beta - mean*invstd*gamma) > 0), which
// saves one of the three tensor reads of this pass.
template <int kMode, bool kRemask = false>  // 0: forward stats (x)   1: backward sums (dy, x̂, optional ReLU mask)
__global__ void bn_reduce_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ x,
                                 const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean,
                                 const float* __restrict__ invstd, float* __restrict__ out, long long M, int C,
                                 long long lda, long long ldx, long long ldy, int rows_per_block,
                                 const float* __restrict__ gamma = nullptr, const float* __restrict__ beta = nullptr) {
  extern __shared__ float red[];  // [rif][vcols*16]
  const int cvec = C / 8;
  const int vcols = cvec < static_cast<int>(blockDim.x) ? cvec : blockDim.x;
  const int rif = blockDim.x / vcols;
  const int tc = threadIdx.x % vcols;
  const int tr = threadIdx.x / vcols;
  const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > M) r1 = M;
  for (int cv = tc; cv < cvec; cv += vcols) {
    float s0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float mu[8], is[8], ma[8], mb[8];
    if (kMode == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        mu[j] = mean[cv * 8 + j];
        is[j] = invstd[cv * 8 + j];
        if (kRemask) {
          ma[j] = is[j] * (gamma ? gamma[cv * 8 + j] : 1.f);
          mb[j] = (beta ? beta[cv * 8 + j] : 0.f) - mu[j] * ma[j];
        }
      }
    }
    if (tr < rif) {
      long long r = r0 + tr;
      if (kMode == 0) {
        // forward statistics: 8 independent 16-byte loads in flight per thread
        for (; r + 7ll * rif < r1; r += 8ll * rif) {
          uint4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const uint4*>(a + (r + u * rif) * lda + cv * 8);
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 t = __bfloat1622float2(q[j]);
              s0[2 * j] += t.x;
              s0[2 * j + 1] += t.y;
              s1[2 * j] = fmaf(t.x, t.x, s1[2 * j]);
              s1[2 * j + 1] = fmaf(t.y, t.y, s1[2 * j + 1]);
            }
          }
        }
      }
      if (kMode == 1) {
        // backward sums: four rows (up to twelve 16-byte loads) in flight per thread
        for (; r + 3ll * rif < r1; r += 4ll * rif) {
          uint4 va[4], vx[4], vy[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            va[u] = *reinterpret_cast<const uint4*>(a + (r + u * rif) * lda + cv * 8);
            vx[u] = *reinterpret_cast<const uint4*>(x + (r + u * rif) * ldx + cv * 8);
            if (!kRemask && y) vy[u] = *reinterpret_cast<const uint4*>(y + (r + u * rif) * ldy + cv * 8);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const __nv_bfloat162* qa = reinterpret_cast<const __nv_bfloat162*>(&va[u]);
            const __nv_bfloat162* qx = reinterpret_cast<const __nv_bfloat162*>(&vx[u]);
            const __nv_bfloat162* qy = reinterpret_cast<const __nv_bfloat162*>(&vy[u]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float2 g2 = __bfloat1622float2(qa[j]);
              const float2 x2 = __bfloat1622float2(qx[j]);
              if (kRemask) {
                g2.x = fmaf(x2.x, ma[2 * j], mb[2 * j]) > 0.f ? g2.x : 0.f;
                g2.y = fmaf(x2.y, ma[2 * j + 1], mb[2 * j + 1]) > 0.f ? g2.y : 0.f;
              } else if (y) {
                const float2 y2 = __bfloat1622float2(qy[j]);
                g2.x = y2.x > 0.f ? g2.x : 0.f;
                g2.y = y2.y > 0.f ? g2.y : 0.f;
              }
              s0[2 * j] += g2.x;
              s0[2 * j + 1] += g2.y;
              s1[2 * j] = fmaf(g2.x, (x2.x - mu[2 * j]) * is[2 * j], s1[2 * j]);
              s1[2 * j + 1] = fmaf(g2.y, (x2.y - mu[2 * j + 1]) * is[2 * j + 1], s1[2 * j + 1]);
            }
          }
        }
      }
      for (; r < r1; r += rif) {
        float fa[8];
        bn_load8(a + r * lda + cv * 8, fa);
        if (kMode == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s0[j] += fa[j];
            s1[j] += fa[j] * fa[j];
          }
        } else {
          float fx[8];
          bn_load8(x + r * ldx + cv * 8, fx);
          if (kRemask) {
#pragma unroll
            for (int j = 0; j < 8; ++j) fa[j] = fmaf(fx[j], ma[j], mb[j]) > 0.f ? fa[j] : 0.f;
          } else if (y) {
            float fy[8];
            bn_load8(y + r * ldy + cv * 8, fy);
#pragma unroll
            for (int j = 0; j < 8; ++j) fa[j] = fy[j] > 0.f ? fa[j] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s0[j] += fa[j];
            s1[j] += fa[j] * (fx[j] - mu[j]) * is[j];
          }
        }
      }
    }
    __syncthreads();
    if (tr < rif) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        red[(tr * vcols + tc) * 16 + j] = s0[j];
        red[(tr * vcols + tc) * 16 + 8 + j] = s1[j];
      }
    }
    __syncthreads();
    if (tr == 0) {
      float t0[8], t1[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float a0 = 0.f, a1 = 0.f;
        for (int t = 0; t < rif; ++t) {
          a0 += red[(t * vcols + tc) * 16 + j];
          a1 += red[(t * vcols + tc) * 16 + 8 + j];
        }
        t0[j] = a0;
        t1[j] = a1;
      }
      // four 16-byte reductions instead of sixteen scalar atomics: every block ends here on the same 2C addresses, and
      // the L2 atomic transaction count (not the bytes streamed) was what bounded this kernel
      float* d0 = out + cv * 8;
      float* d1 = out + C + cv * 8;
      if (((reinterpret_cast<uintptr_t>(d0) | reinterpret_cast<uintptr_t>(d1)) & 15u) == 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d0 + 4 * h), "f"(t0[4 * h]),
                       "f"(t0[4 * h + 1]), "f"(t0[4 * h + 2]), "f"(t0[4 * h + 3]) : "memory");
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d1 + 4 * h), "f"(t1[4 * h]),
                       "f"(t1[4 * h + 1]), "f"(t1[4 * h + 2]), "f"(t1[4 * h + 3]) : "memory");
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          atomicAdd(d0 + j, t0[j]);
          atomicAdd(d1 + j, t1[j]);
        }
      }
    }
  }
}

// mean / invstd from the (possibly all-reduced) sums; running statistics update. One thread per channel.
__global__ void bn_finalize_kernel(const float* __restrict__ sums, float count, int C, float eps, float momentum,
                                   float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                   float* __restrict__ running_mean, float* __restrict__ running_var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = sums[c] / count;
  float var = sums[C + c] / count - mean * mean;
  var = fmaxf(var, 0.f);
  save_mean[c] = mean;
  save_invstd[c] = rsqrtf(var + eps);
  if (running_mean) {
    const float unbiased = count > 1.f ? var * count / (count - 1.f) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// y = (x - mean) * invstd * gamma + beta (+ReLU). For inference pass mean = running mean and invstd computed from
// the running variance (use_var = 1 makes the kernel derive invstd = rsqrt(var + eps) itself).
// Thread layout: blockDim.x = vcols * rif; a thread owns ONE 8-channel vector column (its scale/shift live in
// registers) and streams rows with 4 independent 16-byte loads in flight — no per-element index arithmetic.
template <bool kRes>
__global__ void bn_apply_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                const float* __restrict__ mean, const float* __restrict__ invstd_or_var,
                                const float* __restrict__ gamma, const float* __restrict__ beta, long long M, int C,
                                long long ldx, long long ldy, float eps, int use_var, int relu, int rows_per_block,
                                const __nv_bfloat16* __restrict__ res = nullptr, long long ldr = 0) {
  const int cvec = C / 8;
  const int vcols = cvec < static_cast<int>(blockDim.x) ? cvec : blockDim.x;
  const int rif = blockDim.x / vcols;
  const int tc = threadIdx.x % vcols;
  const int tr = threadIdx.x / vcols;
  if (tr >= rif) return;
  const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > M) r1 = M;
  for (int cv = tc; cv < cvec; cv += vcols) {
    float sa[8], sb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cv * 8 + j;
      const float is = use_var ? rsqrtf(invstd_or_var[c] + eps) : invstd_or_var[c];
      sa[j] = is * (gamma ? gamma[c] : 1.f);
      sb[j] = (beta ? beta[c] : 0.f) - mean[c] * sa[j];
    }
    const __nv_bfloat16* xp = x + cv * 8;
    const __nv_bfloat16* rp = kRes ? res + cv * 8 : nullptr;
    __nv_bfloat16* yp = y + cv * 8;
    long long r = r0 + tr;
    for (; r + 3ll * rif < r1; r += 4ll * rif) {
      uint4 v[4], w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v[u] = *reinterpret_cast<const uint4*>(xp + (r + u * rif) * ldx);
        if (kRes) w[u] = *reinterpret_cast<const uint4*>(rp + (r + u * rif) * ldr);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
        const __nv_bfloat162* qr = reinterpret_cast<const __nv_bfloat162*>(&w[u]);
        float f[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 t = __bfloat1622float2(q[j]);
          f[2 * j] = fmaf(t.x, sa[2 * j], sb[2 * j]);
          f[2 * j + 1] = fmaf(t.y, sa[2 * j + 1], sb[2 * j + 1]);
          if (kRes) {                     // residual block: y = relu(bn(x) + shortcut) in the same pass
            const float2 t2 = __bfloat1622float2(qr[j]);
            f[2 * j] += t2.x;
            f[2 * j + 1] += t2.y;
          }
        }
        if (relu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        bn_store8(yp + (r + u * rif) * ldy, f);
      }
    }
    for (; r < r1; r += rif) {
      float f[8], fr[8];
      bn_load8(xp + r * ldx, f);
      if (kRes) bn_load8(rp + r * ldr, fr);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        f[j] = fmaf(f[j], sa[j], sb[j]);
        if (kRes) f[j] += fr[j];
        if (relu) f[j] = fmaxf(f[j], 0.f);
      }
      bn_store8(yp + r * ldy, f);
    }
  }
}

// dx = gamma * invstd * (dy' - dbeta/count - x̂ * dgamma/count), dy' = dy masked by (y > 0) when ReLU was fused.
// Rewritten per channel as dx = ka * dy' + kb * x + kc with
//   ka = gamma*invstd, kb = -gamma*invstd^2*dgamma/count, kc = -ka*dbeta/count - kb*mean.
template <bool kMask, bool kRemask = false>
__global__ void bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                    const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean,
                                    const float* __restrict__ invstd, const float* __restrict__ gamma,
                                    const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx, long long M,
                                    float count, int C, long long lddy, long long ldx, long long ldy,
                                    long long lddx, int rows_per_block, __nv_bfloat16* __restrict__ dmask = nullptr,
                                    long long lddm = 0, const float* __restrict__ beta = nullptr) {
  const int cvec = C / 8;
  const int vcols = cvec < static_cast<int>(blockDim.x) ? cvec : blockDim.x;
  const int rif = blockDim.x / vcols;
  const int tc = threadIdx.x % vcols;
  const int tr = threadIdx.x / vcols;
  if (tr >= rif) return;
  const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > M) r1 = M;
  const float inv_count = 1.f / count;
  for (int cv = tc; cv < cvec; cv += vcols) {
    float ka[8], kb[8], kc[8], ma[8], mb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cv * 8 + j;
      const float is = invstd[c];
      const float gm = gamma ? gamma[c] : 1.f;
      if (kRemask) {                       // the forward's scale / shift, bit for bit (bn_apply_kernel)
        ma[j] = is * gm;
        mb[j] = (beta ? beta[c] : 0.f) - mean[c] * ma[j];
      }
      ka[j] = gm * is;
      kb[j] = -gm * is * is * sums[C + c] * inv_count;
      kc[j] = -ka[j] * sums[c] * inv_count - kb[j] * mean[c];
    }
    long long r = r0 + tr;
    for (; r + 1ll * rif < r1; r += 2ll * rif) {
      uint4 vg[2], vx[2], vy[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        vg[u] = *reinterpret_cast<const uint4*>(dy + (r + u * rif) * lddy + cv * 8);
        vx[u] = *reinterpret_cast<const uint4*>(x + (r + u * rif) * ldx + cv * 8);
        if (!kRemask && y) vy[u] = *reinterpret_cast<const uint4*>(y + (r + u * rif) * ldy + cv * 8);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const __nv_bfloat162* qg = reinterpret_cast<const __nv_bfloat162*>(&vg[u]);
        const __nv_bfloat162* qx = reinterpret_cast<const __nv_bfloat162*>(&vx[u]);
        const __nv_bfloat162* qy = reinterpret_cast<const __nv_bfloat162*>(&vy[u]);
        float f[8], fm[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 g2 = __bfloat1622float2(qg[j]);
          const float2 x2 = __bfloat1622float2(qx[j]);
          if (kRemask) {
            g2.x = fmaf(x2.x, ma[2 * j], mb[2 * j]) > 0.f ? g2.x : 0.f;
            g2.y = fmaf(x2.y, ma[2 * j + 1], mb[2 * j + 1]) > 0.f ? g2.y : 0.f;
          } else if (y) {
            const float2 y2 = __bfloat1622float2(qy[j]);
            g2.x = y2.x > 0.f ? g2.x : 0.f;
            g2.y = y2.y > 0.f ? g2.y : 0.f;
          }
          f[2 * j] = fmaf(ka[2 * j], g2.x, fmaf(kb[2 * j], x2.x, kc[2 * j]));
          f[2 * j + 1] = fmaf(ka[2 * j + 1], g2.y, fmaf(kb[2 * j + 1], x2.y, kc[2 * j + 1]));
          if (kMask) {                    // the ReLU-masked gradient is also the shortcut's gradient: emit it here
            fm[2 * j] = g2.x;
            fm[2 * j + 1] = g2.y;
          }
        }
        bn_store8(dx + (r + u * rif) * lddx + cv * 8, f);
        if (kMask) bn_store8(dmask + (r + u * rif) * lddm + cv * 8, fm);
      }
    }
    for (; r < r1; r += rif) {
      float g[8], fx[8];
      bn_load8(dy + r * lddy + cv * 8, g);
      bn_load8(x + r * ldx + cv * 8, fx);
      if (kRemask) {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = fmaf(fx[j], ma[j], mb[j]) > 0.f ? g[j] : 0.f;
      } else if (y) {
        float fy[8];
        bn_load8(y + r * ldy + cv * 8, fy);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = fy[j] > 0.f ? g[j] : 0.f;
      }
      if (kMask) bn_store8(dmask + r * lddm + cv * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = fmaf(ka[j], g[j], fmaf(kb[j], fx[j], kc[j]));
      bn_store8(dx + r * lddx + cv * 8, g);
    }
  }
}

// gradWeight += scale_w * dgamma ; gradBias += scale_b * dbeta
__global__ void bn_param_grad_kernel(const float* __restrict__ sums, float* __restrict__ gw, float* __restrict__ gb,
                                     int C, float scale_w, float scale_b) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (gb) gb[c] += scale_b * sums[c];
  if (gw) gw[c] += scale_w * sums[C + c];
}

static int bn_grid(long long total) {
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  long long b = (total + 255) / 256;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

static int bn_check(const char* who, int C, long long l0, long long l1, const void* p0, const void* p1) {
  if ((C % 8) || (l0 % 8) || (l1 % 8) || (reinterpret_cast<uintptr_t>(p0) & 15u) ||
      (reinterpret_cast<uintptr_t>(p1) & 15u))
    return bdl_fail(who);
  return BDL_OK;
}

}  // namespace bdl

using namespace bdl;

// streaming kernels: ~16 CTAs per SM so that every SM keeps 2048 threads x 4 loads in flight
static void stream_geometry(long long M, int C, int* grid, int* rpb, size_t* smem) {
  const int sms = bdl_num_sms();
  const int cvec = C / 8;
  const int vcols = cvec < 256 ? cvec : 256;
  const int rif = 256 / vcols;
  long long r = (M + sms * 16 - 1) / (sms * 16);
  const long long min_rows = 4ll * rif;
  if (r < min_rows) r = min_rows;
  *rpb = static_cast<int>(r);
  *grid = static_cast<int>((M + r - 1) / r);
  *smem = 0;
}

// Reduction kernels end with 2*C red.add on the same addresses in every block, so the block count is kept at one or a
// few per SM and memory-level parallelism comes from 8-12 independent loads in flight per thread.
// (Sweep on B200, benchmarks/eltwise_bench.py --bn: backward is fastest with one block per SM everywhere; the forward
// statistics want more blocks only for the largest activations.) BDL_BN_BPS / BDL_BN_THREADS override for tuning.
static int bn_env(const char* name) {
  const char* e = getenv(name);
  return e ? atoi(e) : 0;
}
static void reduce_geometry(long long M, int C, int mode, int* grid, int* rpb, int* threads, size_t* smem) {
  static const int bps_override = bn_env("BDL_BN_BPS");
  static const int thr_override = bn_env("BDL_BN_THREADS");
  const int sms = bdl_num_sms();
  const int cvec = C / 8;
  const int nthr = thr_override > 0 ? thr_override : 256;   // the backward variant needs >128 registers: 256 threads max
  const int vcols = cvec < nthr ? cvec : nthr;
  const int rif = nthr / vcols;
  const long long elems = M * C;
  int bps = mode == 1 ? 1 : (elems >= (40ll << 20) ? 4 : (elems >= (8ll << 20) ? 2 : 1));
  if (bps_override > 0) bps = bps_override;
  const long long blocks_target = static_cast<long long>(sms) * bps;
  long long r = (M + blocks_target - 1) / blocks_target;
  if (r < 4ll * rif) r = 4ll * rif;
  *rpb = static_cast<int>(r);
  *grid = static_cast<int>((M + r - 1) / r);
  *threads = nthr;
  *smem = static_cast<size_t>(rif) * vcols * 16 * sizeof(float);
}

// sums: [2*C] fp32, must be zero on entry (sum, sum of squares)
extern "C" int bdl_bn_stats(const void* x, float* sums, long long M, int C, long long ldx, void* stream) {
  if (bn_check("bdl_bn_stats: C and pitch must be multiples of 8, pointer 16-byte aligned", C, ldx, ldx, x, x))
    return BDL_ERR_INVALID;
  int grid, rpb, threads;
  size_t smem;
  reduce_geometry(M, C, 0, &grid, &rpb, &threads, &smem);
  bn_reduce_kernel<0><<<grid, threads, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), nullptr, nullptr, nullptr, nullptr, sums, M, C, ldx, 0, 0, rpb);
  return bdl_check_launch("bn_reduce_kernel<0>");
}

extern "C" int bdl_bn_finalize(const float* sums, float count, int C, float eps, float momentum, float* save_mean,
                               float* save_invstd, float* running_mean, float* running_var, void* stream) {
  bn_finalize_kernel<<<(C + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      sums, count, C, eps, momentum, save_mean, save_invstd, running_mean, running_var);
  return bdl_check_launch("bn_finalize_kernel");
}

extern "C" int bdl_bn_apply(const void* x, void* y, const float* mean, const float* invstd_or_var, const float* gamma,
                            const float* beta, long long M, int C, long long ldx, long long ldy, float eps,
                            int use_var, int relu, void* stream) {
  if (bn_check("bdl_bn_apply: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx, ldy, x, y))
    return BDL_ERR_INVALID;
  int grid, rpb;
  size_t smem_unused;
  stream_geometry(M, C, &grid, &rpb, &smem_unused);
  bn_apply_kernel<false><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), mean, invstd_or_var, gamma, beta, M, C,
      ldx, ldy, eps, use_var, relu, rpb);
  return bdl_check_launch("bn_apply_kernel");
}

// y = relu?(bn(x) + res): the BN that ends a residual branch, the shortcut add and the ReLU in one pass
// (S/nn/mkldnn/Fusion.scala fuses conv + sum at inference; this is the training-time counterpart on the BN side).
extern "C" int bdl_bn_apply_res(const void* x, const void* res, void* y, const float* mean, const float* invstd_or_var,
                                const float* gamma, const float* beta, long long M, int C, long long ldx, long long ldr,
                                long long ldy, float eps, int use_var, int relu, void* stream) {
  if (bn_check("bdl_bn_apply_res: C and pitches must be multiples of 8, pointers 16-byte aligned", C, ldx, ldy, x, y))
    return BDL_ERR_INVALID;
  if (res == nullptr || (ldr % 8) || (reinterpret_cast<uintptr_t>(res) & 15u))
    return bdl_fail("bdl_bn_apply_res: residual must be 16-byte aligned with a pitch that is a multiple of 8");
  int grid, rpb;
  size_t smem_unused;
  stream_geometry(M, C, &grid, &rpb, &smem_unused);
  bn_apply_kernel<true><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), mean, invstd_or_var, gamma, beta, M, C,
      ldx, ldy, eps, use_var, relu, rpb, static_cast<const __nv_bfloat16*>(res), ldr);
  return bdl_check_launch("bn_apply_kernel(res)");
}

// bdl_bn_bwd_apply that also writes dy masked by (y > 0) to dmask [M, C] (pitch lddm): the gradient of the shortcut
extern "C" int bdl_bn_bwd_apply_mask(const void* dy, const void* x, const void* y, const float* mean,
                                     const float* invstd, const float* gamma, const float* sums, void* dx, void* dmask,
                                     long long M, float count, int C, long long lddy, long long ldx, long long ldy,
                                     long long lddx, long long lddm, void* stream) {
  if (bn_check("bdl_bn_bwd_apply_mask: C and pitches must be multiples of 8, pointers 16-byte aligned", C, lddy, lddx,
               dy, dx))
    return BDL_ERR_INVALID;
  if (dmask == nullptr || (lddm % 8) || (reinterpret_cast<uintptr_t>(dmask) & 15u))
    return bdl_fail("bdl_bn_bwd_apply_mask: dmask must be 16-byte aligned with a pitch that is a multiple of 8");
  int grid, rpb;
  size_t smem_unused;
  stream_geometry(M, C, &grid, &rpb, &smem_unused);
  bn_bwd_apply_kernel<true><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
      static_cast<const __nv_bfloat16*>(y), mean, invstd, gamma, sums, static_cast<__nv_bfloat16*>(dx), M, count, C,
      lddy, ldx, ldy, lddx, rpb, static_cast<__nv_bfloat16*>(dmask), lddm);
  return bdl_check_launch("bn_bwd_apply_kernel(mask)");
}

// sums: [2*C] fp32, zero on entry; receives (dbeta, dgamma)
extern "C" int bdl_bn_bwd_reduce(const void* dy, const void* x, const void* y, const float* mean, const float* invstd,
                                 float* sums, long long M, int C, long long lddy, long long ldx, long long ldy,
                                 void* stream) {
  if (bn_check("bdl_bn_bwd_reduce: C and pitches must be multiples of 8, pointers 16-byte aligned", C, lddy, ldx, dy,
               x))
    return BDL_ERR_INVALID;
  int grid, rpb, threads;
  size_t smem;
  reduce_geometry(M, C, 1, &grid, &rpb, &threads, &smem);
  bn_reduce_kernel<1><<<grid, threads, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
      static_cast<const __nv_bfloat16*>(y), mean, invstd, sums, M, C, lddy, ldx, ldy, rpb);
  return bdl_check_launch("bn_reduce_kernel<1>");
}

// Backward of BN + fused ReLU without reading the layer's output: the mask is recomputed from x (see kRemask).
extern "C" int bdl_bn_bwd_reduce_remask(const void* dy, const void* x, const float* mean, const float* invstd,
                                        const float* gamma, const float* beta, float* sums, long long M, int C,
                                        long long lddy, long long ldx, void* stream) {
  if (bn_check("bdl_bn_bwd_reduce_remask: C and pitches must be multiples of 8, pointers 16-byte aligned", C, lddy, ldx,
               dy, x))
    return BDL_ERR_INVALID;
  int grid, rpb, threads;
  size_t smem;
  reduce_geometry(M, C, 1, &grid, &rpb, &threads, &smem);
  bn_reduce_kernel<1, true><<<grid, threads, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x), nullptr, mean, invstd, sums, M, C,
      lddy, ldx, 0, rpb, gamma, beta);
  return bdl_check_launch("bn_reduce_kernel<1,remask>");
}

extern "C" int bdl_bn_bwd_apply_remask(const void* dy, const void* x, const float* mean, const float* invstd,
                                       const float* gamma, const float* beta, const float* sums, void* dx, long long M,
                                       float count, int C, long long lddy, long long ldx, long long lddx,
                                       void* stream) {
  if (bn_check("bdl_bn_bwd_apply_remask: C and pitches must be multiples of 8, pointers 16-byte aligned", C, lddy, lddx,
               dy, dx))
    return BDL_ERR_INVALID;
  int grid, rpb;
  size_t smem_unused;
  stream_geometry(M, C, &grid, &rpb, &smem_unused);
  bn_bwd_apply_kernel<false, true><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x), nullptr, mean, invstd, gamma, sums,
      static_cast<__nv_bfloat16*>(dx), M, count, C, lddy, ldx, 0, lddx, rpb, nullptr, 0, beta);
  return bdl_check_launch("bn_bwd_apply_kernel(remask)");
}

extern "C" int bdl_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* invstd,
                                const float* gamma, const float* sums, void* dx, long long M, float count, int C,
                                long long lddy, long long ldx, long long ldy, long long lddx, void* stream) {
  if (bn_check("bdl_bn_bwd_apply: C and pitches must be multiples of 8, pointers 16-byte aligned", C, lddy, lddx, dy,
               dx))
    return BDL_ERR_INVALID;
  int grid, rpb;
  size_t smem_unused;
  stream_geometry(M, C, &grid, &rpb, &smem_unused);
  bn_bwd_apply_kernel<false><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
      static_cast<const __nv_bfloat16*>(y), mean, invstd, gamma, sums, static_cast<__nv_bfloat16*>(dx), M, count, C,
      lddy, ldx, ldy, lddx, rpb);
  return bdl_check_launch("bn_bwd_apply_kernel");
}

extern "C" int bdl_bn_param_grad(const float* sums, float* gw, float* gb, int C, float scale_w, float scale_b,
                                 void* stream) {
  bn_param_grad_kernel<<<(C + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(sums, gw, gb, C, scale_w,
                                                                                        scale_b);
  return bdl_check_launch("bn_param_grad_kernel");
}
