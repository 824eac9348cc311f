// NVLink peer-memory parameter synchronisation: ONE kernel per step does
//   reduce-scatter of bf16-compressed gradients (P2P loads from every peer's gradient buffer)
//   -> 1/N averaging, clipping and the OptimMethod update on this rank's fp32 master slice (+ state)
//   -> bf16 compress of the updated slice and all-gather by P2P stores into every peer's compute-weight buffer,
// bracketed by flag barriers in peer memory (st.release.sys / ld.acquire.sys), so no host round trip and no
// separate compress / decompress / collective passes.
//
// This is the B200 form of the reference's AllReduceParameter iteration
// (S/parameters/AllReduceParameter.scala: putGradients :287-309, aggregateGradientPartition :236-278,
// sendWeightPartition :315-328, getWeights :201-228; FP16CompressedTensor.scala:152-196 for the codec) plus the
// sharded optimizer step of S/optim/DistriOptimizer.scala:358-396. `parity` mode reproduces the reference's
// arithmetic bit for bit: truncating codec, pairwise adds in partition order with re-truncation after every add.
//
// One-GPU testing: with barriers disabled the same kernel can be launched once per *virtual* rank over buffers that
// all live on one device (launch order provides the ordering the flags provide across GPUs).
#include "bigdl_b200.h"
#include "host_common.h"
#include "optim_update.cuh"
#include "ptx.cuh"
#include <stdlib.h>
#include <string.h>

namespace bdl {

constexpr int kMaxPeers = 16;

struct PeerTable {
  const uint16_t* grad16[kMaxPeers];  // each peer's compressed gradient vector (full length P)
  uint16_t* w16[kMaxPeers];           // each peer's bf16 compute-weight vector (full length P)
  uint32_t* flags[kMaxPeers];         // each peer's flag array: [0,N) begin slots, [N,2N) end slots
};

enum CommMode : int {
  CM_REDUCE = 1,         // sum the N compressed gradient slices (else read fp32 grad_part)
  CM_WRITE_GRAD = 2,     // store the averaged fp32 gradient slice to grad_part
  CM_UPDATE = 4,         // apply the OptimMethod to the master slice
  CM_PUBLISH = 8,        // push bf16(master) to every peer
  CM_BARRIER_BEGIN = 16,
  CM_BARRIER_END = 32,
  CM_PARITY = 64,        // reference-exact arithmetic (truncate after every add, truncating publish)
  CM_SUMSQ = 128,        // accumulate sum of squares of the averaged gradient slice into sumsq[0]
  CM_DIVIDE = 256,       // g = g / hp.gscale (the reference divides by the replica count) instead of g * hp.gscale
  // bucketed form (bucket_rs_update_kernel)
  CM_COMPRESS = 512,       // first compress MY fp32 gradients of the whole bucket into my bf16 exchange buffer (and zero them)
  CM_PUBLISH_LOCAL = 1024, // bf16(master) goes to MY compute-weight buffer only; peers pull it (pull_gather / fused GEMM)
  CM_SIGNAL_READY = 2048,  // when every block is done: store the generation into every peer's "ready" slot, do not wait
  CM_CLIP_DEV = 4096,      // multiply the averaged gradient by *clip_scale (device scalar: global-norm clipping)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_v4(void* p, const uint4& v) {
  asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ float trunc_bf16(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFF0000u); }

// local_sync: [0] = "begin barrier passed" epoch (block 0 publishes), [1] = finished-block counter
__global__ void __launch_bounds__(512)
fused_rs_update_kernel(const PeerTable peers, int nranks, int rank, long long start, long long length,
                       float* __restrict__ master, float* __restrict__ grad_part, float* __restrict__ s0,
                       float* __restrict__ s1, int kind, OptimHP hp, const OptimHP* __restrict__ hp_dev, int mode,
                       uint32_t epoch, uint32_t* __restrict__ local_sync, float* __restrict__ sumsq,
                       const unsigned long long* __restrict__ epoch_dev) {
  if (hp_dev) hp = *hp_dev;
  if (epoch_dev) epoch += static_cast<uint32_t>(*epoch_dev);   // graph replays: the step counter lives on the device
  const bool parity = (mode & CM_PARITY) != 0;

  // ---------------- begin barrier: every rank's compressed gradients are complete and visible
  if (mode & CM_BARRIER_BEGIN) {
    if (blockIdx.x == 0) {
      if (threadIdx.x < nranks) {
        __threadfence_system();
        st_release_sys(peers.flags[threadIdx.x] + rank, epoch);
        const uint32_t* mine = peers.flags[rank] + threadIdx.x;
        while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(local_sync), "r"(epoch) : "memory");
      }
    } else {
      if (threadIdx.x == 0) {
        while (static_cast<int32_t>(ld_acquire_gpu(local_sync) - epoch) < 0) {
        }
      }
      __syncthreads();
    }
  }

  // ---------------- main pass: 8 elements (one 16-byte bf16 vector) per thread-iteration
  const long long nvec = length / 8;
  const long long gstride = static_cast<long long>(gridDim.x) * blockDim.x;
  float sq = 0.f;
  for (long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += gstride) {
    const long long off = v * 8;       // offset inside my slice
    const long long goff = start + off;  // offset inside the full vector
    float g[8];
    if (mode & CM_REDUCE) {
      if (parity) {
        // reference order: ((g_0 + g_1)trunc + g_2)trunc ... in partition-id order
        uint4 raw = ld_peer_v4(peers.grad16[0] + goff);
        const uint32_t* r = reinterpret_cast<const uint32_t*>(&raw);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          g[2 * j] = __uint_as_float(r[j] << 16);
          g[2 * j + 1] = __uint_as_float(r[j] & 0xFFFF0000u);
        }
        for (int p = 1; p < nranks; ++p) {
          raw = ld_peer_v4(peers.grad16[p] + goff);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            g[2 * j] = trunc_bf16(g[2 * j] + __uint_as_float(r[j] << 16));
            g[2 * j + 1] = trunc_bf16(g[2 * j + 1] + __uint_as_float(r[j] & 0xFFFF0000u));
          }
        }
      } else {
        // issue all peer loads first (one NVLink round trip), rotate the start peer to spread load over links
        uint4 raw[kMaxPeers];
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
          if (p < nranks) {
            int src = rank + p;
            if (src >= nranks) src -= nranks;
            raw[p] = ld_peer_v4(peers.grad16[src] + goff);
          }
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
          if (p < nranks) {
            const uint32_t* r = reinterpret_cast<const uint32_t*>(&raw[p]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              g[2 * j] += __uint_as_float(r[j] << 16);
              g[2 * j + 1] += __uint_as_float(r[j] & 0xFFFF0000u);
            }
          }
      }
      if (mode & CM_DIVIDE) {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = __fdiv_rn(g[j], hp.gscale);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] *= hp.gscale;
      }
    } else {
      const float4 a = *reinterpret_cast<const float4*>(grad_part + off);
      const float4 b = *reinterpret_cast<const float4*>(grad_part + off + 4);
      g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w;
      g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w;
    }
    if (mode & CM_SUMSQ) {
#pragma unroll
      for (int j = 0; j < 8; ++j) sq += g[j] * g[j];
    }
    if (mode & CM_WRITE_GRAD) {
      *reinterpret_cast<float4*>(grad_part + off) = make_float4(g[0], g[1], g[2], g[3]);
      *reinterpret_cast<float4*>(grad_part + off + 4) = make_float4(g[4], g[5], g[6], g[7]);
    }
    float w[8];
    if (mode & (CM_UPDATE | CM_PUBLISH)) {
      const float4 a = *reinterpret_cast<const float4*>(master + off);
      const float4 b = *reinterpret_cast<const float4*>(master + off + 4);
      w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
      w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    }
    if (mode & CM_UPDATE) {
      OptimHP h = hp;
      h.gscale = (mode & CM_REDUCE) ? 1.f : hp.gscale;  // already applied above when reducing
      float a0[8], a1[8];
      if (s0) {
        const float4 x = *reinterpret_cast<const float4*>(s0 + off);
        const float4 y = *reinterpret_cast<const float4*>(s0 + off + 4);
        a0[0] = x.x; a0[1] = x.y; a0[2] = x.z; a0[3] = x.w; a0[4] = y.x; a0[5] = y.y; a0[6] = y.z; a0[7] = y.w;
      }
      if (s1) {
        const float4 x = *reinterpret_cast<const float4*>(s1 + off);
        const float4 y = *reinterpret_cast<const float4*>(s1 + off + 4);
        a1[0] = x.x; a1[1] = x.y; a1[2] = x.z; a1[3] = x.w; a1[4] = y.x; a1[5] = y.y; a1[6] = y.z; a1[7] = y.w;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float t0 = s0 ? a0[j] : 0.f, t1 = s1 ? a1[j] : 0.f;
        w[j] = optim_dispatch(kind, w[j], g[j], &t0, &t1, h);
        a0[j] = t0;
        a1[j] = t1;
      }
      *reinterpret_cast<float4*>(master + off) = make_float4(w[0], w[1], w[2], w[3]);
      *reinterpret_cast<float4*>(master + off + 4) = make_float4(w[4], w[5], w[6], w[7]);
      if (s0) {
        *reinterpret_cast<float4*>(s0 + off) = make_float4(a0[0], a0[1], a0[2], a0[3]);
        *reinterpret_cast<float4*>(s0 + off + 4) = make_float4(a0[4], a0[5], a0[6], a0[7]);
      }
      if (s1) {
        *reinterpret_cast<float4*>(s1 + off) = make_float4(a1[0], a1[1], a1[2], a1[3]);
        *reinterpret_cast<float4*>(s1 + off + 4) = make_float4(a1[4], a1[5], a1[6], a1[7]);
      }
    }
    if (mode & CM_PUBLISH) {
      uint4 o;
      uint32_t* q = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        q[j] = static_cast<uint32_t>(f32_to_bf16_bits(w[2 * j], parity)) |
               (static_cast<uint32_t>(f32_to_bf16_bits(w[2 * j + 1], parity)) << 16);
#pragma unroll
      for (int p = 0; p < kMaxPeers; ++p)
        if (p < nranks) {
          int dst = rank + p;
          if (dst >= nranks) dst -= nranks;
          st_peer_v4(peers.w16[dst] + goff, o);
        }
    }
  }
  if (mode & CM_SUMSQ) {
    sq = warp_sum(sq);
    if ((threadIdx.x & 31) == 0 && sq != 0.f) atomicAdd(sumsq, sq);
  }

  // ---------------- end barrier: my pushes are visible everywhere and everyone's pushes into my buffer are done
  if (mode & CM_BARRIER_END) {
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
      __threadfence_system();
      const uint32_t done = atomicAdd(local_sync + 1, 1u);
      last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
      if (threadIdx.x == 0) local_sync[1] = 0;  // re-arm for the next launch (stream order protects it)
      if (threadIdx.x < nranks) {
        __threadfence_system();
        st_release_sys(peers.flags[threadIdx.x] + nranks + rank, epoch);
        const uint32_t* mine = peers.flags[rank] + nranks + threadIdx.x;
        while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
        }
      }
    }
  }
}


// ------------------------------------------------------------------------------------------------ bucketed form
// One launch per gradient bucket, issued from inside backward on a communication stream the moment the bucket's
// gradients are final, so the NVLink traffic and the optimizer math overlap the rest of backward
// (reference: S/optim/ParallelOptimizer.scala:481-513 + AbstractModule.asyncGradient, applied to DistriOptimizer's
// sharded update, S/optim/DistriOptimizer.scala:358-396). A bucket [b_start, b_start + b_len) of the flat vector is
// partitioned among the ranks; this rank owns [start, start + length) of it.
//   phase 0  compress MY fp32 gradients of the whole bucket into my bf16 exchange buffer (and zero the fp32 source)
//   barrier  every rank's compressed bucket is visible (generation = epoch_dev * gen_mul + gen_add, unique per launch)
//   phase 1  P2P reduce of my sub-slice over all peers -> average (-> clip) -> OptimMethod on the fp32 master slice ->
//            bf16 weights into MY compute buffer only
//   signal   "ready" generation into every peer's flag area; nobody waits here. Consumers (pull_gather_kernel, or the
//            all-gather-fused GEMM) acquire that flag right before they read the slice from this GPU.
struct BucketArgs {
  PeerTable peers;
  int nranks, rank;
  long long b_start, b_len;
  long long start, length;
  float* grads32;
  float* master;
  float* grad_part;
  float* s0;
  float* s1;
  int kind;
  const OptimHP* hp_dev;
  OptimHP hp;
  int mode;
  uint32_t gen_mul, gen_add;
  const unsigned long long* epoch_dev;
  long long flag_off;      // this bucket's 3N words inside every flag array: [0,N) begin, [N,2N) end, [2N,3N) ready
  uint32_t* local_sync;    // [0] begin-barrier generation, [1] finished blocks, [2] blocks done compressing
  float* sumsq;
  const float* clip_scale;
  long long hole_start, hole_len;   // CM_PUBLISH: elements in this range of the full vector stay local (a fused GEMM
                                    // gathers them from the owner instead)
};

__device__ __forceinline__ void load8(const float* p, float* v) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store8(float* p, const float* v) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// kP: compile-time bound on the number of peers (8 on one NVSwitch box). With 8, the kernel fits 48 registers, so a
// 512-thread CTA is co-resident with one of backward's GEMM CTAs (320 threads x 128 registers) on the same SM instead of
// taking the SM away from the compute stream while it runs or waits for the slowest peer.
template <int kP>
__global__ void __launch_bounds__(512, kP <= 8 ? 2 : 1)
bucket_rs_update_kernel(const BucketArgs a) {
  OptimHP hp = a.hp_dev ? *a.hp_dev : a.hp;
  const uint32_t gen = static_cast<uint32_t>(a.epoch_dev ? *a.epoch_dev : 0ull) * a.gen_mul + a.gen_add;
  const int mode = a.mode;
  const bool parity = (mode & CM_PARITY) != 0;
  const int nranks = a.nranks, rank = a.rank;
  const long long gstride = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;

  if (mode & CM_COMPRESS) {
    uint16_t* mine = const_cast<uint16_t*>(a.peers.grad16[rank]) + a.b_start;
    float* src = a.grads32 + a.b_start;
    for (long long v = tid; v < a.b_len / 8; v += gstride) {
      float f[8];
      load8(src + v * 8, f);
      uint4 o;
      uint32_t* q = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        q[j] = static_cast<uint32_t>(f32_to_bf16_bits(f[2 * j], parity)) |
               (static_cast<uint32_t>(f32_to_bf16_bits(f[2 * j + 1], parity)) << 16);
      reinterpret_cast<uint4*>(mine)[v] = o;
      const float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      store8(src + v * 8, z);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(a.local_sync + 2, 1u);
    }
  }

  if (mode & CM_BARRIER_BEGIN) {
    if (blockIdx.x == 0) {
      if ((mode & CM_COMPRESS) && threadIdx.x == 0) {
        while (ld_acquire_gpu(a.local_sync + 2) < gridDim.x) {
        }
        a.local_sync[2] = 0;     // re-armed for the next launch (launches that share local_sync are stream-ordered)
      }
      __syncthreads();
      if (threadIdx.x < nranks) {
        __threadfence_system();
        st_release_sys(a.peers.flags[threadIdx.x] + a.flag_off + rank, gen);
        const uint32_t* mine = a.peers.flags[rank] + a.flag_off + threadIdx.x;
        while (static_cast<int32_t>(ld_acquire_sys(mine) - gen) < 0) {
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a.local_sync), "r"(gen) : "memory");
      }
    } else {
      if (threadIdx.x == 0) {
        while (static_cast<int32_t>(ld_acquire_gpu(a.local_sync) - gen) < 0) {
        }
      }
      __syncthreads();
    }
  }

  const float clip = (mode & CM_CLIP_DEV) ? *a.clip_scale : 1.f;
  float sq = 0.f;
  for (long long v = tid; v < a.length / 8; v += gstride) {
    const long long off = v * 8;
    const long long goff = a.start + off;
    float g[8];
    if (mode & CM_REDUCE) {
      if (parity) {
        uint4 raw = ld_peer_v4(a.peers.grad16[0] + goff);
        const uint32_t* r = reinterpret_cast<const uint32_t*>(&raw);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          g[2 * j] = __uint_as_float(r[j] << 16);
          g[2 * j + 1] = __uint_as_float(r[j] & 0xFFFF0000u);
        }
        for (int p = 1; p < nranks; ++p) {
          raw = ld_peer_v4(a.peers.grad16[p] + goff);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            g[2 * j] = trunc_bf16(g[2 * j] + __uint_as_float(r[j] << 16));
            g[2 * j + 1] = trunc_bf16(g[2 * j + 1] + __uint_as_float(r[j] & 0xFFFF0000u));
          }
        }
      } else {
        uint4 raw[kP];
#pragma unroll
        for (int p = 0; p < kP; ++p)
          if (p < nranks) {
            int src = rank + p;
            if (src >= nranks) src -= nranks;
            raw[p] = ld_peer_v4(a.peers.grad16[src] + goff);
          }
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
        for (int p = 0; p < kP; ++p)
          if (p < nranks) {
            const uint32_t* r = reinterpret_cast<const uint32_t*>(&raw[p]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              g[2 * j] += __uint_as_float(r[j] << 16);
              g[2 * j + 1] += __uint_as_float(r[j] & 0xFFFF0000u);
            }
          }
      }
      if (mode & CM_DIVIDE) {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = __fdiv_rn(g[j], hp.gscale);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] *= hp.gscale;
      }
    } else {
      load8(a.grad_part + off, g);
    }
    if (mode & CM_SUMSQ) {
#pragma unroll
      for (int j = 0; j < 8; ++j) sq += g[j] * g[j];
    }
    if (mode & CM_WRITE_GRAD) store8(a.grad_part + off, g);
    if (mode & (CM_UPDATE | CM_PUBLISH | CM_PUBLISH_LOCAL)) {
      float w[8];
      load8(a.master + off, w);
      if (mode & CM_UPDATE) {
        OptimHP h = hp;
        // phase 2 of global-norm clipping reads a gradient that phase 1 already averaged: only the clip factor is left
        h.gscale = ((mode & (CM_REDUCE | CM_CLIP_DEV)) ? 1.f : hp.gscale) * clip;
        float a0[8], a1[8];
        if (a.s0) load8(a.s0 + off, a0);
        if (a.s1) load8(a.s1 + off, a1);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float t0 = a.s0 ? a0[j] : 0.f, t1 = a.s1 ? a1[j] : 0.f;
          w[j] = optim_dispatch(a.kind, w[j], g[j], &t0, &t1, h);
          a0[j] = t0;
          a1[j] = t1;
        }
        store8(a.master + off, w);
        if (a.s0) store8(a.s0 + off, a0);
        if (a.s1) store8(a.s1 + off, a1);
      }
      uint4 o;
      uint32_t* q = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        q[j] = static_cast<uint32_t>(f32_to_bf16_bits(w[2 * j], parity)) |
               (static_cast<uint32_t>(f32_to_bf16_bits(w[2 * j + 1], parity)) << 16);
      if ((mode & CM_PUBLISH_LOCAL) || ((mode & CM_PUBLISH) && goff >= a.hole_start && goff < a.hole_start + a.hole_len)) {
        *reinterpret_cast<uint4*>(a.peers.w16[rank] + goff) = o;
      } else if (mode & CM_PUBLISH) {
#pragma unroll
        for (int p = 0; p < kP; ++p)
          if (p < nranks) {
            int dst = rank + p;
            if (dst >= nranks) dst -= nranks;
            st_peer_v4(a.peers.w16[dst] + goff, o);
          }
      }
    }
  }
  if (mode & CM_SUMSQ) {
    sq = warp_sum(sq);
    if ((threadIdx.x & 31) == 0 && sq != 0.f) atomicAdd(a.sumsq, sq);
  }

  if (mode & (CM_BARRIER_END | CM_SIGNAL_READY)) {
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
      __threadfence_system();
      const uint32_t done = atomicAdd(a.local_sync + 1, 1u);
      last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
      if (threadIdx.x == 0) a.local_sync[1] = 0;
      if (threadIdx.x < nranks) {
        __threadfence_system();
        if (mode & CM_SIGNAL_READY) st_release_sys(a.peers.flags[threadIdx.x] + a.flag_off + 2 * nranks + rank, gen);
        if (mode & CM_BARRIER_END) {
          st_release_sys(a.peers.flags[threadIdx.x] + a.flag_off + nranks + rank, gen);
          const uint32_t* mine = a.peers.flags[rank] + a.flag_off + nranks + threadIdx.x;
          while (static_cast<int32_t>(ld_acquire_sys(mine) - gen) < 0) {
          }
        }
      }
    }
  }
}

// Pull side of the all-gather: copy the slices of a bucket that OTHER ranks own from their compute-weight buffers into
// mine, each after acquiring that owner's "ready" generation (reference: AllReduceParameter.getWeights,
// S/parameters/AllReduceParameter.scala:201-228 — one fetch per remote partition). Runs on the communication stream at
// the start of the next forward, bucket by bucket in forward order; layers wait on the bucket's event, so the transfer
// of late buckets hides behind the early layers' compute.
constexpr int kMaxSegs = 2 * kMaxPeers;
struct PullArgs {
  uint16_t* w16[kMaxPeers];
  const uint32_t* my_flags;      // MY flag array
  int nranks, rank, nsegs;
  int seg_peer[kMaxSegs];
  long long seg_start[kMaxSegs], seg_len[kMaxSegs];   // element ranges of the full vector, multiples of 8
  uint32_t gen_mul, gen_add;
  int gen_prev_steps;            // the generation waited for belongs to epoch_dev - gen_prev_steps
  const unsigned long long* epoch_dev;
  long long flag_off;
};

__global__ void __launch_bounds__(512) pull_gather_kernel(const PullArgs a) {
  const unsigned long long ep = a.epoch_dev ? *a.epoch_dev : 0ull;
  const bool wait = ep >= static_cast<unsigned long long>(a.gen_prev_steps);   // nothing was published before step 0
  const uint32_t gen = static_cast<uint32_t>(ep - a.gen_prev_steps) * a.gen_mul + a.gen_add;
  const long long gstride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (int s = 0; s < a.nsegs; ++s) {
    const int peer = a.seg_peer[s];
    if (wait) {
      if (threadIdx.x == 0) {
        const uint32_t* f = a.my_flags + a.flag_off + 2 * a.nranks + peer;
        while (static_cast<int32_t>(ld_acquire_sys(f) - gen) < 0) {
        }
      }
      __syncthreads();
    }
    const uint16_t* src = a.w16[peer] + a.seg_start[s];
    uint16_t* dst = a.w16[a.rank] + a.seg_start[s];
    const long long nvec = a.seg_len[s] / 8;
    long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    // four independent 16-byte peer loads in flight per thread (NVLink latency is ~2 us)
    for (; v + 3 * gstride < nvec; v += 4 * gstride) {
      const uint4 x0 = ld_peer_v4(src + v * 8), x1 = ld_peer_v4(src + (v + gstride) * 8),
                  x2 = ld_peer_v4(src + (v + 2 * gstride) * 8), x3 = ld_peer_v4(src + (v + 3 * gstride) * 8);
      reinterpret_cast<uint4*>(dst)[v] = x0;
      reinterpret_cast<uint4*>(dst)[v + gstride] = x1;
      reinterpret_cast<uint4*>(dst)[v + 2 * gstride] = x2;
      reinterpret_cast<uint4*>(dst)[v + 3 * gstride] = x3;
    }
    for (; v < nvec; v += gstride) reinterpret_cast<uint4*>(dst)[v] = ld_peer_v4(src + v * 8);
  }
}

// Global-norm clipping without the host: every rank holds the sum of squares of ITS slices of the averaged gradient;
// one block per rank stores it into every peer's slot, waits for all slots of this generation and writes
// scale = min(1, max_norm / sqrt(sum)) to device memory, where the update kernels read it (CM_CLIP_DEV).
// Reference: L2NormClippingProcessor, S/parameters/ParameterOperations.scala:94-130.
struct ClipArgs {
  float* slots[kMaxPeers];        // per peer: nranks floats
  uint32_t* flags[kMaxPeers];
  int nranks, rank;
  uint32_t gen_mul, gen_add;
  const unsigned long long* epoch_dev;
  long long flag_off;             // nranks words used
  float* sumsq;                   // my partial sum (reset to 0 here for the next step)
  float max_norm;
  float* scale_out;
  float* norm_out;                // optional: the global norm, for logging
};

__global__ void peer_clip_scale_kernel(const ClipArgs a) {
  const uint32_t gen = static_cast<uint32_t>(a.epoch_dev ? *a.epoch_dev : 0ull) * a.gen_mul + a.gen_add;
  __shared__ float part[kMaxPeers];
  const int t = threadIdx.x;
  if (t < a.nranks) {
    const float mine = *a.sumsq;
    // value first, then the flag that publishes it (release); slots are double-buffered by generation parity so a
    // fast rank's next step cannot overwrite a value a slow rank has not read yet
    float* slot = a.slots[t] + (gen & 1u) * a.nranks + a.rank;
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(slot), "f"(mine) : "memory");
    __threadfence_system();
    st_release_sys(a.flags[t] + a.flag_off + a.rank, gen);
    const uint32_t* f = a.flags[a.rank] + a.flag_off + t;
    while (static_cast<int32_t>(ld_acquire_sys(f) - gen) < 0) {
    }
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(a.slots[a.rank] + (gen & 1u) * a.nranks + t) : "memory");
    part[t] = v;
  }
  __syncthreads();
  if (t == 0) {
    float sum = 0.f;
    for (int i = 0; i < a.nranks; ++i) sum += part[i];   // same order on every rank: identical scale everywhere
    const float nrm = sqrtf(sum);
    *a.scale_out = nrm > a.max_norm ? a.max_norm / nrm : 1.f;
    if (a.norm_out) *a.norm_out = nrm;
    *a.sumsq = 0.f;
  }
}

// fp32 gradients -> bf16 (RN or truncate) into the symmetric buffer, optionally zeroing the fp32 source
// (K1 of SURVEY.md §2.M fused with zeroGradParameters).
__global__ void compress_grads_kernel(float* __restrict__ g, uint16_t* __restrict__ out, size_t n, int truncate,
                                      int zero_src) {
  const size_t nvec = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const float4 a = reinterpret_cast<const float4*>(g)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(g)[2 * i + 1];
    const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint4 o;
    uint32_t* q = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      q[j] = static_cast<uint32_t>(f32_to_bf16_bits(f[2 * j], truncate)) |
             (static_cast<uint32_t>(f32_to_bf16_bits(f[2 * j + 1], truncate)) << 16);
    reinterpret_cast<uint4*>(out)[i] = o;
    if (zero_src) {
      reinterpret_cast<float4*>(g)[2 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
      reinterpret_cast<float4*>(g)[2 * i + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

}  // namespace bdl

using namespace bdl;

extern "C" int bdl_compress_grads(float* g, void* out16, size_t n, int truncate, int zero_src, void* stream) {
  if (n == 0) return BDL_OK;
  if (n % 8) return bdl_fail("bdl_compress_grads: length must be a multiple of 8");
  const long long cap = static_cast<long long>(bdl_num_sms()) * 8;
  long long blocks = static_cast<long long>((n / 8 + 255) / 256);
  if (blocks > cap) blocks = cap;
  compress_grads_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      g, static_cast<uint16_t*>(out16), n, truncate, zero_src);
  return bdl_check_launch("compress_grads_kernel");
}

// grad16_ptrs / w16_ptrs / flag_ptrs: host arrays of `nranks` device pointers (peer-mapped).
extern "C" int bdl_fused_rs_update(const void* const* grad16_ptrs, void* const* w16_ptrs, void* const* flag_ptrs,
                                   int nranks, int rank, long long start, long long length, float* master,
                                   float* grad_part, float* s0, float* s1, int kind, const void* hp_host,
                                   const void* hp_dev, int mode, unsigned int epoch, void* local_sync, float* sumsq,
                                   int max_blocks, const void* epoch_dev, void* stream) {
  if (nranks < 1 || nranks > kMaxPeers) return bdl_fail("bdl_fused_rs_update: nranks must be in [1,16]");
  if ((start % 8) || (length % 8)) return bdl_fail("bdl_fused_rs_update: slice must be 8-element aligned");
  if (length == 0 && !(mode & (CM_BARRIER_BEGIN | CM_BARRIER_END))) return BDL_OK;
  PeerTable pt{};
  for (int i = 0; i < nranks; ++i) {
    pt.grad16[i] = grad16_ptrs ? static_cast<const uint16_t*>(grad16_ptrs[i]) : nullptr;
    pt.w16[i] = w16_ptrs ? static_cast<uint16_t*>(w16_ptrs[i]) : nullptr;
    pt.flags[i] = flag_ptrs ? static_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
  }
  if ((mode & (CM_BARRIER_BEGIN | CM_BARRIER_END)) && (!flag_ptrs || !local_sync))
    return bdl_fail("bdl_fused_rs_update: barriers need flag pointers and a local sync word pair");
  OptimHP hp{};
  if (hp_host) hp = *static_cast<const OptimHP*>(hp_host);
  const int threads = 512;
  long long blocks = (length / 8 + threads - 1) / threads;
  int cap = bdl_num_sms();  // co-residency: at most one CTA per SM so the in-grid spin cannot starve
  if (max_blocks > 0 && max_blocks < cap) cap = max_blocks;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  fused_rs_update_kernel<<<static_cast<int>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      pt, nranks, rank, start, length, master, grad_part, s0, s1, kind, hp, static_cast<const OptimHP*>(hp_dev), mode,
      epoch, static_cast<uint32_t*>(local_sync), sumsq, static_cast<const unsigned long long*>(epoch_dev));
  return bdl_check_launch("fused_rs_update_kernel");
}


// ---------------------------------------------------------------- bucketed entry points
// ptr tables: host arrays of `nranks` peer-mapped device pointers (full-vector bases). master/grad_part/s0/s1 point at
// THIS bucket's sub-slice inside the rank's partition arrays. See BucketArgs for the protocol.
extern "C" int bdl_bucket_rs_update(const void* const* grad16_ptrs, void* const* w16_ptrs, void* const* flag_ptrs,
                                    int nranks, int rank, long long b_start, long long b_len, long long start,
                                    long long length, float* grads32, float* master, float* grad_part, float* s0,
                                    float* s1, int kind, const void* hp_host, const void* hp_dev, int mode,
                                    unsigned int gen_mul, unsigned int gen_add, const void* epoch_dev,
                                    long long flag_off, void* local_sync, float* sumsq, const float* clip_scale,
                                    long long hole_start, long long hole_len, int max_blocks, void* stream) {
  if (nranks < 1 || nranks > kMaxPeers) return bdl_fail("bdl_bucket_rs_update: nranks must be in [1,16]");
  if ((b_start % 8) || (b_len % 8) || (start % 8) || (length % 8))
    return bdl_fail("bdl_bucket_rs_update: ranges must be 8-element aligned");
  if (start < b_start || start + length > b_start + b_len) return bdl_fail("bdl_bucket_rs_update: slice outside its bucket");
  if (!local_sync) return bdl_fail("bdl_bucket_rs_update: local_sync (3 words) is required");
  if ((mode & (CM_BARRIER_BEGIN | CM_BARRIER_END | CM_SIGNAL_READY)) && !flag_ptrs)
    return bdl_fail("bdl_bucket_rs_update: barriers / ready signal need flag pointers");
  if ((mode & CM_COMPRESS) && (!grads32 || !grad16_ptrs)) return bdl_fail("bdl_bucket_rs_update: COMPRESS needs grads32 and grad16");
  if ((mode & CM_CLIP_DEV) && !clip_scale) return bdl_fail("bdl_bucket_rs_update: CLIP_DEV needs clip_scale");
  BucketArgs a{};
  for (int i = 0; i < nranks; ++i) {
    a.peers.grad16[i] = grad16_ptrs ? static_cast<const uint16_t*>(grad16_ptrs[i]) : nullptr;
    a.peers.w16[i] = w16_ptrs ? static_cast<uint16_t*>(w16_ptrs[i]) : nullptr;
    a.peers.flags[i] = flag_ptrs ? static_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
  }
  a.nranks = nranks; a.rank = rank; a.b_start = b_start; a.b_len = b_len; a.start = start; a.length = length;
  a.grads32 = grads32; a.master = master; a.grad_part = grad_part; a.s0 = s0; a.s1 = s1; a.kind = kind;
  a.hp_dev = static_cast<const OptimHP*>(hp_dev);
  if (hp_host) a.hp = *static_cast<const OptimHP*>(hp_host);
  a.mode = mode; a.gen_mul = gen_mul; a.gen_add = gen_add;
  a.epoch_dev = static_cast<const unsigned long long*>(epoch_dev);
  a.flag_off = flag_off; a.local_sync = static_cast<uint32_t*>(local_sync); a.sumsq = sumsq; a.clip_scale = clip_scale;
  a.hole_start = hole_start; a.hole_len = hole_len;
  // CTA shape (measured, profiles/comm_r2.md): 512-thread CTAs on a limited number of SMs beat many small CTAs spread
  // over all SMs — the kernel is bandwidth-bound and finishes sooner, and backward's GEMMs lose fewer SM-milliseconds.
  static const int threads = [] { const char* e = getenv("BDL_COMM_THREADS"); const int t = e ? atoi(e) : 512;
                                  return t >= 32 && t <= 512 ? t / 32 * 32 : 512; }();
  const long long work = (mode & CM_COMPRESS) ? (b_len > length ? b_len : length) : length;
  long long blocks = (work / 8 + threads - 1) / threads;
  // the compress phase makes block 0 wait for every other block, so all blocks must be co-resident next to whatever
  // backward kernels are running
  int cap = bdl_num_sms();       // max_blocks <= 0: nothing else is running (end of the step) — take every SM
  if (max_blocks > 0) cap = max_blocks < bdl_num_sms() ? max_blocks : bdl_num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (nranks <= 8)
    bucket_rs_update_kernel<8><<<static_cast<int>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(a);
  else
    bucket_rs_update_kernel<kMaxPeers><<<static_cast<int>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(a);
  return bdl_check_launch("bucket_rs_update_kernel");
}

// segs: nsegs triples (peer, start, len) as long long[3 * nsegs]; generation awaited = (epoch_dev - prev_steps) * gen_mul
// + gen_add of the owner's bdl_bucket_rs_update launch (no wait while epoch_dev < prev_steps).
extern "C" int bdl_pull_gather(void* const* w16_ptrs, const void* my_flags, int nranks, int rank, const long long* segs,
                               int nsegs, unsigned int gen_mul, unsigned int gen_add, int prev_steps,
                               const void* epoch_dev, long long flag_off, int max_blocks, void* stream) {
  if (nranks < 1 || nranks > kMaxPeers) return bdl_fail("bdl_pull_gather: nranks must be in [1,16]");
  if (nsegs < 0 || nsegs > kMaxSegs) return bdl_fail("bdl_pull_gather: at most 32 segments per launch");
  if (nsegs == 0) return BDL_OK;
  PullArgs a{};
  for (int i = 0; i < nranks; ++i) a.w16[i] = static_cast<uint16_t*>(w16_ptrs[i]);
  a.my_flags = static_cast<const uint32_t*>(my_flags);
  a.nranks = nranks; a.rank = rank; a.nsegs = nsegs;
  long long total = 0;
  for (int i = 0; i < nsegs; ++i) {
    a.seg_peer[i] = static_cast<int>(segs[3 * i]);
    a.seg_start[i] = segs[3 * i + 1];
    a.seg_len[i] = segs[3 * i + 2];
    if (a.seg_peer[i] < 0 || a.seg_peer[i] >= nranks || (a.seg_start[i] % 8) || (a.seg_len[i] % 8) || a.seg_len[i] < 0)
      return bdl_fail("bdl_pull_gather: bad segment");
    if (a.seg_len[i] > total) total = a.seg_len[i];
  }
  a.gen_mul = gen_mul; a.gen_add = gen_add; a.gen_prev_steps = prev_steps;
  a.epoch_dev = static_cast<const unsigned long long*>(epoch_dev);
  a.flag_off = flag_off;
  const int threads = total == 0 ? 32 : 512;     // all segments empty: a pure wait on the owners' ready flags
  long long blocks = (total / 8 + threads * 4 - 1) / (threads * 4);
  int cap = max_blocks > 0 ? max_blocks : 32;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  pull_gather_kernel<<<static_cast<int>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(a);
  return bdl_check_launch("pull_gather_kernel");
}

extern "C" int bdl_peer_clip_scale(void* const* slot_ptrs, void* const* flag_ptrs, int nranks, int rank,
                                   unsigned int gen_mul, unsigned int gen_add, const void* epoch_dev,
                                   long long flag_off, float* sumsq, float max_norm, float* scale_out, float* norm_out,
                                   void* stream) {
  if (nranks < 1 || nranks > kMaxPeers) return bdl_fail("bdl_peer_clip_scale: nranks must be in [1,16]");
  ClipArgs a{};
  for (int i = 0; i < nranks; ++i) {
    a.slots[i] = static_cast<float*>(slot_ptrs[i]);
    a.flags[i] = static_cast<uint32_t*>(flag_ptrs[i]);
  }
  a.nranks = nranks; a.rank = rank; a.gen_mul = gen_mul; a.gen_add = gen_add;
  a.epoch_dev = static_cast<const unsigned long long*>(epoch_dev);
  a.flag_off = flag_off; a.sumsq = sumsq; a.max_norm = max_norm; a.scale_out = scale_out; a.norm_out = norm_out;
  peer_clip_scale_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a);
  return bdl_check_launch("peer_clip_scale_kernel");
}

// ---------------------------------------------------------------- CUDA IPC helpers (one process per GPU)
extern "C" int bdl_ipc_alloc(size_t bytes, void** ptr, void* handle64) {
  cudaError_t e = cudaMalloc(ptr, bytes);
  if (e != cudaSuccess) return bdl_fail_cuda("cudaMalloc(ipc)", e);
  e = cudaMemset(*ptr, 0, bytes);
  if (e != cudaSuccess) return bdl_fail_cuda("cudaMemset(ipc)", e);
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, *ptr);
  if (e != cudaSuccess) return bdl_fail_cuda("cudaIpcGetMemHandle", e);
  static_assert(sizeof(h) == 64, "ipc handle size");
  memcpy(handle64, &h, sizeof(h));
  return BDL_OK;
}
extern "C" int bdl_ipc_open(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return bdl_fail_cuda("cudaIpcOpenMemHandle", e);
  return BDL_OK;
}
extern "C" int bdl_ipc_close(void* ptr) {
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  if (e != cudaSuccess) return bdl_fail_cuda("cudaIpcCloseMemHandle", e);
  return BDL_OK;
}
extern "C" int bdl_ipc_free(void* ptr) {
  cudaError_t e = cudaFree(ptr);
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFree(ipc)", e);
  return BDL_OK;
}
