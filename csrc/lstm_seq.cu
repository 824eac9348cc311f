// Persistent LSTM over a whole sequence: ONE launch runs all T time steps of a layer.
//
// The input projection x_t W_ih^T + b of every step is hoisted into one big tcgen05 GEMM by the caller (xw [B,T,4H]).
// What is left is the strictly sequential part, h_{t-1} W_hh^T plus the gate math, which as one launch per step is bound
// by launch latency and by re-reading W_hh from L2 every step. Here W_hh stays in shared memory for the whole sequence,
// split across the SMs by hidden unit: CTA k owns hidden units [j0, j0 + HJ) — the 4 x HJ gate rows of W_hh that produce
// them — so it can finish the cell update of its units without talking to anybody; per step it only needs the full
// h_{t-1} [B, H], which every CTA publishes its slice of, separated by one grid-wide generation barrier
// (cooperative launch: all CTAs are co-resident).
//   per step:   wait for h_{t-1}  ->  stage it in smem  ->  gates^T[4HJ, B] = W_slice[4HJ, H] . h_{t-1}^T  (mma.sync
//   m16n8k16 bf16, fp32 accumulate, K split over the 8 warps and reduced in smem — the GEMM is 44 x 20 x 1500 per CTA,
//   far below a tcgen05 tile)  ->  + xw_t  ->  sigmoid/tanh  ->  c_t (fp32, kept in smem), h_t (bf16, published)
// The backward kernel walks the steps in reverse with a 2-D split of W_hh (gate-row block x hidden-column block), see
// lstm_seq_bwd_kernel.
// Reference: S/nn/Recurrent.scala:243-527 (the time loop), S/nn/LSTM.scala:44-175, and the fused multi-step RNN
// primitive of S/nn/mkldnn/RNN.scala:31-100. Gate order along 4H: [input, candidate, forget, output].
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdlib.h>
#include <mma.h>

#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

__device__ __forceinline__ float sigm_(float v) { return 1.f / (1.f + __expf(-v)); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& a0, uint32_t& a1, uint32_t& a2, uint32_t& a3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t addr, uint32_t& b0, uint32_t& b1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t ld_acq_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Stage `rows` rows of `H` bf16 (row r at src + r * stride elements) into shared memory rows of pitch Hp, zero padded,
// reading through L2 (the data was written by other SMs during this launch). vw = elements per load: 8, 4, 2 or 1,
// whatever the row starts are aligned to (PTB's 650 and 1500 are multiples of 2 and 4 only).
__device__ __forceinline__ void stage_rows_cg(__nv_bfloat16* dst, int Hp, const __nv_bfloat16* src, long long stride,
                                              int rows, int rows_pad, int H, int vw, int tid, int nthreads) {
  const int per_row = Hp / vw;
  const int total = rows_pad * per_row;
  constexpr int U = 8;       // loads in flight per thread: the copy is bound by L2 latency, not bandwidth
  for (int base = tid; base < total; base += nthreads * U) {
    uint32_t v[U][4];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads;
      v[u][0] = v[u][1] = v[u][2] = v[u][3] = 0u;
      if (idx < total) {
        const int r = idx / per_row, c = (idx - r * per_row) * vw;
        if (r < rows && c < H && src != nullptr) {
          const __nv_bfloat16* p = src + static_cast<long long>(r) * stride + c;
          if (vw == 8) {
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]) : "l"(p));
          } else if (vw == 4) {
            asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(v[u][0]), "=r"(v[u][1]) : "l"(p));
          } else if (vw == 2) {
            asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v[u][0]) : "l"(p));
          } else {
            uint16_t h16;
            asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(h16) : "l"(p));
            v[u][0] = h16;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads;
      if (idx < total) {
        const int r = idx / per_row, c = (idx - r * per_row) * vw;
        __nv_bfloat16* d = dst + static_cast<size_t>(r) * Hp + c;
        if (vw == 8) *reinterpret_cast<uint4*>(d) = make_uint4(v[u][0], v[u][1], v[u][2], v[u][3]);
        else if (vw == 4) *reinterpret_cast<uint2*>(d) = make_uint2(v[u][0], v[u][1]);
        else if (vw == 2) *reinterpret_cast<uint32_t*>(d) = v[u][0];
        else *reinterpret_cast<uint16_t*>(d) = static_cast<uint16_t>(v[u][0]);
      }
    }
  }
}
__device__ __forceinline__ int vec_width(int H, long long stride, const void* base) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(base);
  if (H % 8 == 0 && stride % 8 == 0 && (a & 15u) == 0) return 8;
  if (H % 4 == 0 && stride % 4 == 0 && (a & 7u) == 0) return 4;
  if (H % 2 == 0 && stride % 2 == 0 && (a & 3u) == 0) return 2;
  return 1;
}

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define LSTM_MARK(i) do { if (a.dbg && blockIdx.x == 0 && tid == 0 && step == 5) a.dbg[i] = gtime(); } while (0)

constexpr int kSeqThreads = 256;
constexpr int kSeqWarps = 8;
constexpr int kMaxMT = 4;     // gate rows per CTA <= 64
constexpr int kMaxNT = 4;     // batch chunk <= 32

struct LstmSeqArgs {
  const __nv_bfloat16* xw;   // [B, T, 4H] pre-projected inputs (+ bias), strides in elements
  long long xw_sb, xw_st;
  const __nv_bfloat16* whh;  // [4H, H] row-major
  const __nv_bfloat16* h0;   // [B, H] or null (zeros)
  const float* c0;           // [B, H] or null
  __nv_bfloat16* hs;         // [B, T, H] out
  long long hs_sb, hs_st;
  int hs_hread;              // columns of an hs row that may be READ (>= H; the pad columns are zero): lets rows of a
                             // padded buffer be staged with 16-byte loads when H itself is not a multiple of 8
  float* cs;                 // [T, B, H] out (cell states, for backward); may be null
  __nv_bfloat16* gates;      // [T, B, 4H] out (activated gates, for backward); may be null
  float* c_last;             // [B, H] out
  int B, T, H, HJ, reverse;
  int chunk;                 // batch rows staged per pass: 32, or 16 when the W slice leaves no room for more
  int scratch_alias;         // the per-warp partial tiles reuse the h staging buffer (it is dead after the MMAs)
  uint32_t* counter;         // zeroed before the launch
  unsigned long long* dbg;   // optional: %globaltimer of CTA 0 at the phase boundaries of step 5 (tuning aid)
};

// shared memory: W slice [R][Hp] bf16 | h [chunk][Hp] bf16 | c [HJ][32] fp32 | (per-warp partial tiles, unless aliased)
template <int kDummy>
__global__ void __launch_bounds__(kSeqThreads, 1) lstm_seq_fwd_kernel(const LstmSeqArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int H = a.H, B = a.B, HJ = a.HJ;
  const int Hp = (H + 63) / 64 * 64 + 8;                 // row pitch (elements): 16-byte rows, odd multiple of 16 B mod 128
  const int R = (4 * HJ + 15) / 16 * 16;
  const int MT = R / 16;
  __nv_bfloat16* Ws = reinterpret_cast<__nv_bfloat16*>(smem);
  __nv_bfloat16* hsm = Ws + static_cast<size_t>(R) * Hp;
  const int CH = a.chunk;
  float* csm = reinterpret_cast<float*>(hsm + static_cast<size_t>(CH) * Hp);
  // K is split over the 8 warps; each warp parks its partial [R, CH] tile here and the gate math sums the 8 of them
  // (shared-memory float atomics cost ~2 cycles per lane: 9k of them per step were 10 us of a 16 us step)
  float* scratch = a.scratch_alias ? reinterpret_cast<float*>(hsm) : csm + HJ * 32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int j0 = blockIdx.x * HJ;
  const int nj = min(HJ, H - j0);                        // hidden units this CTA really owns
  const int G = gridDim.x;

  // ---- one-time: W slice into smem (local row g*HJ + j <- global row g*H + j0 + j), zero padding
  for (int idx = tid; idx < R * (Hp / 8); idx += kSeqThreads)
    *reinterpret_cast<uint4*>(Ws + static_cast<size_t>(idx) * 8) = make_uint4(0, 0, 0, 0);
  __syncthreads();
  for (int g = 0; g < 4; ++g) {
    const __nv_bfloat16* src = a.whh + static_cast<long long>(g * H + j0) * H;
    stage_rows_cg(Ws + static_cast<size_t>(g * HJ) * Hp, Hp, src, H, nj, nj, H, vec_width(H, H, src), tid, kSeqThreads);
  }
  for (int idx = tid; idx < CH * (Hp / 8); idx += kSeqThreads)
    *reinterpret_cast<uint4*>(hsm + static_cast<size_t>(idx) * 8) = make_uint4(0, 0, 0, 0);
  __syncthreads();

  const int ksteps = (H + 15) / 16;
  for (int b0 = 0; b0 < B; b0 += CH) {                   // batch chunks (training batches beyond one chunk loop here)
    const int nb = min(CH, B - b0);
    const int NT = (nb + 7) / 8;
    // initial cell state of this chunk
    for (int idx = tid; idx < nj * nb; idx += kSeqThreads) {
      const int j = idx / nb, b = idx - j * nb;
      csm[j * 32 + b] = a.c0 ? a.c0[static_cast<long long>(b0 + b) * H + j0 + j] : 0.f;
    }
    for (int step = 0; step < a.T; ++step) {
      const int t = a.reverse ? a.T - 1 - step : step;
      const int tp = a.reverse ? t + 1 : t - 1;          // the step whose h feeds this one
      const uint32_t gen = static_cast<uint32_t>((b0 / CH) * a.T + step);   // steps completed by everyone before this one
      LSTM_MARK(0);
      if (gen > 0) {
        if (tid == 0) {
          const uint32_t want = gen * static_cast<uint32_t>(G);
          while (ld_acq_gpu(a.counter) < want) {
          }
        }
        __syncthreads();
      }
      LSTM_MARK(1);
      // ---- stage h_{prev} [nb, H] and clear the pre-activation tile
      {
        const __nv_bfloat16* src = nullptr;
        long long stride = 0;
        if (step == 0) {
          if (a.h0) { src = a.h0 + static_cast<long long>(b0) * H; stride = H; }
        } else {
          src = a.hs + static_cast<long long>(b0) * a.hs_sb + static_cast<long long>(tp) * a.hs_st;
          stride = a.hs_sb;
        }
        const int Hread = (step == 0) ? H : a.hs_hread;
        const int vw = src ? vec_width(Hread, stride, src) : 8;
        stage_rows_cg(hsm, Hp, src, stride, nb, nb, Hread, vw, tid, kSeqThreads);
      }
      __syncthreads();
      LSTM_MARK(2);

      // ---- gates^T[R, nb] = W_slice[R, H] . h^T : K split over the warps, fp32 partial tiles summed afterwards
      float acc[kMaxMT][kMaxNT][4];
#pragma unroll
      for (int m = 0; m < kMaxMT; ++m)
#pragma unroll
        for (int n = 0; n < kMaxNT; ++n)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[m][n][e] = 0.f;
      for (int ks = warp; ks < ksteps; ks += kSeqWarps) {
        const int k0 = ks * 16;
        uint32_t af[kMaxMT][4], bfr[kMaxNT][2];
#pragma unroll
        for (int m = 0; m < kMaxMT; ++m)
          if (m < MT) {
            const __nv_bfloat16* p = Ws + static_cast<size_t>(m * 16 + (lane & 15)) * Hp + k0 + ((lane >> 4) << 3);
            ldmatrix_x4(smem_u32(p), af[m][0], af[m][1], af[m][2], af[m][3]);
          }
#pragma unroll
        for (int n = 0; n < kMaxNT; ++n)
          if (n < NT) {
            const __nv_bfloat16* p = hsm + static_cast<size_t>(n * 8 + (lane & 7)) * Hp + k0 + (((lane >> 3) & 1) << 3);
            ldmatrix_x2(smem_u32(p), bfr[n][0], bfr[n][1]);
          }
#pragma unroll
        for (int m = 0; m < kMaxMT; ++m)
          if (m < MT)
#pragma unroll
            for (int n = 0; n < kMaxNT; ++n)
              if (n < NT) mma_bf16_16816(acc[m][n], af[m][0], af[m][1], af[m][2], af[m][3], bfr[n][0], bfr[n][1]);
      }
      if (a.scratch_alias) __syncthreads();                 // every warp is done reading h before it is overwritten
      {
        float* mine = scratch + static_cast<size_t>(warp) * R * CH;
#pragma unroll
        for (int m = 0; m < kMaxMT; ++m)
          if (m < MT)
#pragma unroll
            for (int n = 0; n < kMaxNT; ++n)
              if (n < NT) {
                const int r = m * 16 + (lane >> 2), c = n * 8 + (lane & 3) * 2;
                *reinterpret_cast<float2*>(mine + r * CH + c) = make_float2(acc[m][n][0], acc[m][n][1]);
                *reinterpret_cast<float2*>(mine + (r + 8) * CH + c) = make_float2(acc[m][n][2], acc[m][n][3]);
              }
      }
      __syncthreads();
      LSTM_MARK(3);

      // ---- gate math for my hidden units, publish h_t
      for (int idx = tid; idx < nj * nb; idx += kSeqThreads) {
        const int j = idx / nb, b = idx - j * nb;
        const __nv_bfloat16* xr = a.xw + static_cast<long long>(b0 + b) * a.xw_sb + static_cast<long long>(t) * a.xw_st + j0 + j;
        float p4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int w = 0; w < kSeqWarps; ++w) {
          const float* part = scratch + static_cast<size_t>(w) * R * CH + b;
#pragma unroll
          for (int g = 0; g < 4; ++g) p4[g] += part[(g * HJ + j) * CH];
        }
        const float ig = sigm_(p4[0] + __bfloat162float(xr[0]));
        const float zg = tanhf(p4[1] + __bfloat162float(xr[H]));
        const float fg = sigm_(p4[2] + __bfloat162float(xr[2 * H]));
        const float og = sigm_(p4[3] + __bfloat162float(xr[3 * H]));
        const float cn = ig * zg + fg * csm[j * 32 + b];
        csm[j * 32 + b] = cn;
        const long long bh = static_cast<long long>(b0 + b);
        a.hs[bh * a.hs_sb + static_cast<long long>(t) * a.hs_st + j0 + j] = __float2bfloat16_rn(tanhf(cn) * og);
        if (a.cs) a.cs[(static_cast<long long>(t) * B + bh) * H + j0 + j] = cn;
        if (a.gates) {
          __nv_bfloat16* gr = a.gates + (static_cast<long long>(t) * B + bh) * 4 * H + j0 + j;
          gr[0] = __float2bfloat16_rn(ig);
          gr[H] = __float2bfloat16_rn(zg);
          gr[2 * H] = __float2bfloat16_rn(fg);
          gr[3 * H] = __float2bfloat16_rn(og);
        }
        if (step == a.T - 1 && a.c_last) a.c_last[bh * H + j0 + j] = cn;
      }
      __syncthreads();
      LSTM_MARK(4);
      if (tid == 0) {
        __threadfence();
        atomicAdd(a.counter, 1u);
      }
      LSTM_MARK(5);
    }
  }
}


// ------------------------------------------------------------------------------------------------ backward (BPTT)
// Per step (walking the forward order backwards):
//   (1) gate gradients of step t from dh_t (upstream + recurrent part) and the carried dc  — done by the CTAs of gate
//       block 0 for their hidden-column block, written to dG[t] (bf16; also the operand of the hoisted dW / dx GEMMs)
//   (2) dh_rec[b, k] = sum_r dG[t][b, r] W_hh[r, k]: W_hh is split 2-D, CTA (p, q) keeps rows of gate p x columns of
//       block q in shared memory (stored k-major, so the same mma.sync code as forward applies), multiplies the gate-p
//       quarter of dG[t] and adds its partial into dh_rec with fp32 red.add (4 partials per element)
// with a grid-wide generation barrier after each phase.
struct LstmSeqBwdArgs {
  const __nv_bfloat16* dhs;   // [B, T, H] upstream gradient of every h_t (may be null: only the last state has one)
  long long dhs_sb, dhs_st;
  const __nv_bfloat16* whh;   // [4H, H]
  const __nv_bfloat16* gates; // [T, B, 4H] from forward
  const float* cs;            // [T, B, H] from forward
  const float* c0;            // [B, H] or null
  __nv_bfloat16* dG;          // [T, B, 4H] out: gradient w.r.t. the gate pre-activations
  float* dh_rec;              // [2, B, H] fp32 scratch, zeroed before the launch; holds dh0 at the end (slot T & 1)
  float* dc;                  // [B, H] fp32, zeroed (or the upstream dc_T) before the launch; holds dc0 at the end
  int B, T, H, HC, Q, reverse;
  int chunk;                  // batch rows per pass of phase 2 (16 or 24)
  int dg_hp;                  // pitch of one gate block inside a dG row (>= H, multiple of 8: 16-byte aligned blocks whose
                              // pad columns stay zero), dG is [T, B, 4, dg_hp]
  int scratch_alias;          // per-warp partial tiles reuse the dG staging buffer
  uint32_t* counter;
};

constexpr int kBwdMaxNT = 3;

template <int kDummy>
__global__ void __launch_bounds__(kSeqThreads, 1) lstm_seq_bwd_kernel(const LstmSeqBwdArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int H = a.H, B = a.B, HC = a.HC;
  const int Hp = (H + 63) / 64 * 64 + 8;
  const int R = (HC + 15) / 16 * 16;
  const int MT = R / 16;
  __nv_bfloat16* Wt = reinterpret_cast<__nv_bfloat16*>(smem);               // [R][Hp]: Wt[k][r] = W_hh[p*H + r][k0 + k]
  const int kBwdChunk = a.chunk;
  const long long GH = a.dg_hp;                                             // gate-block pitch inside a dG row
  __nv_bfloat16* gsm = Wt + static_cast<size_t>(R) * Hp;                    // [chunk][Hp]: dG[t][b][p][r]
  float* part = a.scratch_alias ? reinterpret_cast<float*>(gsm)
                                : reinterpret_cast<float*>(gsm + static_cast<size_t>(kBwdChunk) * Hp);   // [8][R][chunk]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int p = blockIdx.x / a.Q, q = blockIdx.x - p * a.Q;
  const int k0 = q * HC;
  const int nk = min(HC, H - k0);
  const uint32_t G = gridDim.x;

  // W sub-block, transposed into smem (one time): coalesce over r (rows of W are H apart, so read along k per row)
  for (int idx = tid; idx < R * (Hp / 8); idx += kSeqThreads)
    *reinterpret_cast<uint4*>(Wt + static_cast<size_t>(idx) * 8) = make_uint4(0, 0, 0, 0);
  __syncthreads();
  for (int idx = tid; idx < H * nk; idx += kSeqThreads) {
    const int r = idx / nk, k = idx - r * nk;
    Wt[static_cast<size_t>(k) * Hp + r] = a.whh[static_cast<long long>(p * H + r) * H + k0 + k];
  }
  for (int idx = tid; idx < kBwdChunk * (Hp / 8); idx += kSeqThreads)
    *reinterpret_cast<uint4*>(gsm + static_cast<size_t>(idx) * 8) = make_uint4(0, 0, 0, 0);
  __syncthreads();

  const int ksteps = (H + 15) / 16;
  uint32_t gen = 0;
  auto grid_barrier = [&]() {
    __syncthreads();
    ++gen;
    if (tid == 0) {
      __threadfence();
      atomicAdd(a.counter, 1u);
      while (ld_acq_gpu(a.counter) < gen * G) {
      }
    }
    __syncthreads();
  };

  for (int step = 0; step < a.T; ++step) {
    const int t = a.reverse ? step : a.T - 1 - step;          // forward visited t in the opposite order
    const int tprev = a.reverse ? t + 1 : t - 1;               // the step whose state fed step t in forward
    const bool first_fwd = a.reverse ? (t == a.T - 1) : (t == 0);
    float* dh_in = a.dh_rec + static_cast<long long>(step & 1) * B * H;          // written by the previous bwd step
    float* dh_out = a.dh_rec + static_cast<long long>((step + 1) & 1) * B * H;   // this step's dh for h_{tprev}
    // ---- (1) gate gradients: gate block 0 owns hidden units [k0, k0 + nk)
    if (p == 0) {
      for (int idx = tid; idx < B * nk; idx += kSeqThreads) {
        const int b = idx / nk, j = k0 + (idx - b * nk);
        const long long bh = static_cast<long long>(b) * H + j;
        float dh;
        asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(dh) : "l"(dh_in + bh));
        dh_in[bh] = 0.f;                                        // consumed: the slot is accumulated into again at step + 2
        if (a.dhs) dh += __bfloat162float(a.dhs[static_cast<long long>(b) * a.dhs_sb + static_cast<long long>(t) * a.dhs_st + j]);
        const __nv_bfloat16* gr = a.gates + (static_cast<long long>(t) * B + b) * 4 * H + j;
        const float ig = __bfloat162float(gr[0]), zg = __bfloat162float(gr[H]), fg = __bfloat162float(gr[2 * H]),
                    og = __bfloat162float(gr[3 * H]);
        const float ct = a.cs[(static_cast<long long>(t) * B + b) * H + j];
        const float cp = first_fwd ? (a.c0 ? a.c0[bh] : 0.f) : a.cs[(static_cast<long long>(tprev) * B + b) * H + j];
        const float th = tanhf(ct);
        const float dct = a.dc[bh] + dh * og * (1.f - th * th);
        a.dc[bh] = dct * fg;
        __nv_bfloat16* dg = a.dG + (static_cast<long long>(t) * B + b) * 4 * GH + j;
        dg[0] = __float2bfloat16_rn(dct * zg * ig * (1.f - ig));
        dg[GH] = __float2bfloat16_rn(dct * ig * (1.f - zg * zg));
        dg[2 * GH] = __float2bfloat16_rn(dct * cp * fg * (1.f - fg));
        dg[3 * GH] = __float2bfloat16_rn(dh * th * og * (1.f - og));
      }
    }
    grid_barrier();
    // ---- (2) partial dh_rec for my (gate, column) block, batch in chunks of 16
    for (int b0 = 0; b0 < B; b0 += kBwdChunk) {
      const int nb = min(kBwdChunk, B - b0);
      const int NT = (nb + 7) / 8;
      {
        const __nv_bfloat16* src = a.dG + (static_cast<long long>(t) * B + b0) * 4 * GH + static_cast<long long>(p) * GH;
        const int Hread = static_cast<int>(GH) <= (H + 63) / 64 * 64 ? static_cast<int>(GH) : H;
        stage_rows_cg(gsm, Hp, src, 4 * GH, nb, kBwdChunk, Hread, vec_width(Hread, 4 * GH, src), tid, kSeqThreads);
      }
      __syncthreads();
      float acc[kMaxMT][kBwdMaxNT][4];
#pragma unroll
      for (int m = 0; m < kMaxMT; ++m)
#pragma unroll
        for (int n = 0; n < kBwdMaxNT; ++n)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[m][n][e] = 0.f;
      for (int ks = warp; ks < ksteps; ks += kSeqWarps) {
        const int kk = ks * 16;
        uint32_t af[kMaxMT][4], bfr[kBwdMaxNT][2];
#pragma unroll
        for (int m = 0; m < kMaxMT; ++m)
          if (m < MT) {
            const __nv_bfloat16* pa = Wt + static_cast<size_t>(m * 16 + (lane & 15)) * Hp + kk + ((lane >> 4) << 3);
            ldmatrix_x4(smem_u32(pa), af[m][0], af[m][1], af[m][2], af[m][3]);
          }
#pragma unroll
        for (int n = 0; n < kBwdMaxNT; ++n)
          if (n < NT) {
            const __nv_bfloat16* pb = gsm + static_cast<size_t>(n * 8 + (lane & 7)) * Hp + kk + (((lane >> 3) & 1) << 3);
            ldmatrix_x2(smem_u32(pb), bfr[n][0], bfr[n][1]);
          }
#pragma unroll
        for (int m = 0; m < kMaxMT; ++m)
          if (m < MT)
#pragma unroll
            for (int n = 0; n < kBwdMaxNT; ++n)
              if (n < NT) mma_bf16_16816(acc[m][n], af[m][0], af[m][1], af[m][2], af[m][3], bfr[n][0], bfr[n][1]);
      }
      if (a.scratch_alias) __syncthreads();
      {
        float* mine = part + static_cast<size_t>(warp) * R * kBwdChunk;
#pragma unroll
        for (int m = 0; m < kMaxMT; ++m)
          if (m < MT)
#pragma unroll
            for (int n = 0; n < kBwdMaxNT; ++n)
              if (n < NT) {
                const int r = m * 16 + (lane >> 2), c = n * 8 + (lane & 3) * 2;
                *reinterpret_cast<float2*>(mine + r * kBwdChunk + c) = make_float2(acc[m][n][0], acc[m][n][1]);
                *reinterpret_cast<float2*>(mine + (r + 8) * kBwdChunk + c) = make_float2(acc[m][n][2], acc[m][n][3]);
              }
      }
      __syncthreads();
      for (int idx = tid; idx < nk * nb; idx += kSeqThreads) {
        const int k = idx / nb, b = idx - k * nb;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kSeqWarps; ++w) v += part[(static_cast<size_t>(w) * R + k) * kBwdChunk + b];
        atomicAdd(dh_out + static_cast<long long>(b0 + b) * H + k0 + k, v);
      }
      __syncthreads();
    }
    grid_barrier();
  }
}

}  // namespace bdl

using namespace bdl;

// the 8 per-warp partial tiles ([R, chunk] fp32 each) alias the staging buffer when it is large enough
static bool lstm_alias(size_t Hp, size_t R) { return 2 * Hp >= 4 * kSeqWarps * R; }
static size_t lstm_fwd_smem(int H, int HJ, int chunk) {
  const size_t Hp = (H + 63) / 64 * 64 + 8;
  const size_t R = (4 * HJ + 15) / 16 * 16;
  const size_t scratch = lstm_alias(Hp, R) ? 0 : static_cast<size_t>(kSeqWarps) * R * chunk * 4;
  return R * Hp * 2 + static_cast<size_t>(chunk) * Hp * 2 + static_cast<size_t>(HJ) * 32 * 4 + scratch;
}
static size_t lstm_bwd_smem(int H, int HC, int chunk) {
  const size_t Hp = (H + 63) / 64 * 64 + 8;
  const size_t R = (HC + 15) / 16 * 16;
  const size_t scratch = lstm_alias(Hp, R) ? 0 : static_cast<size_t>(kSeqWarps) * R * chunk * 4;
  return R * Hp * 2 + static_cast<size_t>(chunk) * Hp * 2 + scratch;
}

// All T steps of one LSTM layer in one cooperative launch. xw = x W_ih^T + b for every step ([B,T,4H] bf16, element
// strides xw_sb / xw_st), whh [4H,H] bf16, h0 (bf16) / c0 (fp32) optional. Outputs: hs [B,T,H] bf16 (strides hs_sb /
// hs_st), cs [T,B,H] fp32 and gates [T,B,4H] bf16 (for backward; may be null for inference), c_last [B,H] fp32.
// counter: one zeroed uint32 of device memory. reverse: walk t from T-1 down (BiRecurrent's second direction).
// Returns BDL_ERR-style non-zero when the shape does not fit one CTA per SM (the caller then takes the per-step path).
extern "C" int bdl_lstm_seq_fwd(const void* xw, long long xw_sb, long long xw_st, const void* whh, const void* h0,
                                const float* c0, void* hs, long long hs_sb, long long hs_st, int hs_hread, float* cs,
                                void* gates, float* c_last, int B, int T, int H, int reverse, void* counter,
                                void* stream_) {
  if (B <= 0 || T <= 0 || H <= 0) return bdl_fail("bdl_lstm_seq_fwd: empty problem");
  const int sms = bdl_num_sms();
  int HJ = (H + sms - 1) / sms;
  if (HJ < 4) HJ = 4;
  if (4 * HJ > 16 * kMaxMT) return bdl_fail("bdl_lstm_seq_fwd: hidden size too large for the persistent kernel");
  const int grid = (H + HJ - 1) / HJ;
  // fewest passes over the sequence: the largest batch chunk (32 / 24 / 16 rows) whose staging buffer still fits
  int chunk = 16;
  for (int c : {24, 32})
    if (lstm_fwd_smem(H, HJ, c) <= 227 * 1024 && (B + c - 1) / c < (B + chunk - 1) / chunk) chunk = c;
  const size_t smem = lstm_fwd_smem(H, HJ, chunk);
  if (smem > 227 * 1024) return bdl_fail("bdl_lstm_seq_fwd: W_hh slice does not fit in shared memory");
  LstmSeqArgs a{};
  a.xw = static_cast<const __nv_bfloat16*>(xw); a.xw_sb = xw_sb; a.xw_st = xw_st;
  a.whh = static_cast<const __nv_bfloat16*>(whh);
  a.h0 = static_cast<const __nv_bfloat16*>(h0); a.c0 = c0;
  a.hs = static_cast<__nv_bfloat16*>(hs); a.hs_sb = hs_sb; a.hs_st = hs_st;
  a.hs_hread = (hs_hread >= H && hs_hread <= hs_st && hs_hread <= (H + 63) / 64 * 64) ? hs_hread : H;
  a.cs = cs; a.gates = static_cast<__nv_bfloat16*>(gates); a.c_last = c_last;
  a.B = B; a.T = T; a.H = H; a.HJ = HJ; a.reverse = reverse; a.chunk = chunk;
  a.scratch_alias = lstm_alias((H + 63) / 64 * 64 + 8, (4 * HJ + 15) / 16 * 16) ? 1 : 0;
  a.counter = static_cast<uint32_t*>(counter);
  {
    static unsigned long long* dbg = [] {
      unsigned long long* p = nullptr;
      if (getenv("BDL_LSTM_DBG")) cudaMallocManaged(&p, 64);
      return p;
    }();
    a.dbg = dbg;
    if (dbg && getenv("BDL_LSTM_DBG_PRINT")) {
      cudaDeviceSynchronize();
      fprintf(stderr, "lstm fwd phases (ns): wait %llu stage %llu mma %llu gates %llu signal %llu\n", dbg[1] - dbg[0],
              dbg[2] - dbg[1], dbg[3] - dbg[2], dbg[4] - dbg[3], dbg[5] - dbg[4]);
    }
  }
  auto kfn = lstm_seq_fwd_kernel<0>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(lstm_seq_fwd)", e);
  void* params[] = {&a};
  e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kfn), dim3(grid), dim3(kSeqThreads), params, smem,
                                  static_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaLaunchCooperativeKernel(lstm_seq_fwd)", e);
  return bdl_check_launch("lstm_seq_fwd_kernel");
}

// BPTT of the same layer: dG [T,B,4H] bf16 = gradient w.r.t. the gate pre-activations of every step (the operand of
// the hoisted dW_ih / dW_hh / dx GEMMs). dh_rec: [2,B,H] fp32 zeroed (slot 0 may hold an upstream dh_T); dc: [B,H] fp32
// zeroed or holding the upstream dc_T. At return dh_rec slot (T & 1) is dh0 and dc is dc0.
extern "C" int bdl_lstm_seq_bwd(const void* dhs, long long dhs_sb, long long dhs_st, const void* whh, const void* gates,
                                const float* cs, const float* c0, void* dG, int dg_hp, float* dh_rec, float* dc, int B,
                                int T, int H, int reverse, void* counter, void* stream_) {
  if (B <= 0 || T <= 0 || H <= 0) return bdl_fail("bdl_lstm_seq_bwd: empty problem");
  const int sms = bdl_num_sms();
  const int Qmax = sms / 4;
  int HC = (H + Qmax - 1) / Qmax;
  if (HC < 8) HC = 8;
  if (HC > 16 * kMaxMT) return bdl_fail("bdl_lstm_seq_bwd: hidden size too large for the persistent kernel");
  const int Q = (H + HC - 1) / HC;
  int chunk = 16;
  if (B > 16 && lstm_bwd_smem(H, HC, 24) <= 227 * 1024) chunk = 24;
  const size_t smem = lstm_bwd_smem(H, HC, chunk);
  if (smem > 227 * 1024) return bdl_fail("bdl_lstm_seq_bwd: W_hh block does not fit in shared memory");
  if (dg_hp < H || (dg_hp % 2)) return bdl_fail("bdl_lstm_seq_bwd: bad dG gate pitch");
  LstmSeqBwdArgs a{};
  a.dhs = static_cast<const __nv_bfloat16*>(dhs); a.dhs_sb = dhs_sb; a.dhs_st = dhs_st;
  a.whh = static_cast<const __nv_bfloat16*>(whh);
  a.gates = static_cast<const __nv_bfloat16*>(gates); a.cs = cs; a.c0 = c0;
  a.dG = static_cast<__nv_bfloat16*>(dG); a.dh_rec = dh_rec; a.dc = dc;
  a.B = B; a.T = T; a.H = H; a.HC = HC; a.Q = Q; a.reverse = reverse; a.chunk = chunk; a.dg_hp = dg_hp;
  a.scratch_alias = lstm_alias((H + 63) / 64 * 64 + 8, (HC + 15) / 16 * 16) ? 1 : 0;
  a.counter = static_cast<uint32_t*>(counter);
  auto kfn = lstm_seq_bwd_kernel<0>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaFuncSetAttribute(lstm_seq_bwd)", e);
  void* params[] = {&a};
  e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kfn), dim3(4 * Q), dim3(kSeqThreads), params, smem,
                                  static_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) return bdl_fail_cuda("cudaLaunchCooperativeKernel(lstm_seq_bwd)", e);
  return bdl_check_launch("lstm_seq_bwd_kernel");
}
