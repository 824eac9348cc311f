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
