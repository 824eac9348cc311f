// fp32 / fp64 BLAS and vector-math entry points for the Tensor library on the GPU.
//
// The reference's TensorNumeric funnels every DenseTensor math call into one JNI library (core/mkl2017-xeon-blas/src/
// main/c/mkl.c: vs/vdAdd, Sub, Mul, Div, Powx, Ln, Exp, Sqrt, Tanh, Abs, Log1p, axpy, dot, scal, ger, gemv, gemm —
// S/tensor/TensorNumeric.scala:253-420 are the call sites). This file is the same surface as plain C entry points over
// device pointers, column-major like BLAS, with offsets folded into the pointers by the caller. The training hot path
// is bf16 on the tcgen05 kernels (gemm_tcgen05.cu); this is the exact-precision path of Tensor[Float] / Tensor[Double]
// (optimizer state math, validation statistics, user tensor code), so it runs on the CUDA cores in the tensor's own
// precision: register-tiled SIMT GEMM, warp-per-column GEMV, grid-stride vector ops.
#include <cuda_runtime.h>
#include <stdint.h>

#include "bigdl_b200.h"
#include "host_common.h"

namespace bdl {

enum VmlOp { kAdd = 0, kSub, kMul, kDiv, kPowx, kLn, kExp, kSqrt, kTanh, kAbs, kLog1p };

template <typename T> __device__ __forceinline__ T vml_apply(int op, T a, T b) {
  switch (op) {
    case kAdd: return a + b;
    case kSub: return a - b;
    case kMul: return a * b;
    case kDiv: return a / b;
    case kPowx: return pow(a, b);
    case kLn: return log(a);
    case kExp: return exp(a);
    case kSqrt: return sqrt(a);
    case kTanh: return tanh(a);
    case kAbs: return fabs(a);
    default: return log1p(a);
  }
}

// y[i] = op(a[i], b[i] | scalar). 16-byte accesses when all three pointers allow it.
template <typename T, int kOp, bool kBinary>
__global__ void __launch_bounds__(256) vml_kernel(long long n, const T* __restrict__ a, const T* __restrict__ b, T scalar,
                                                  T* __restrict__ y, int vec_ok) {
  constexpr int V = 16 / sizeof(T);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (vec_ok) {
    const long long nv = n / V;
    for (long long v = i; v < nv; v += stride) {
      T ra[V], rb[V], ry[V];
      *reinterpret_cast<uint4*>(ra) = *reinterpret_cast<const uint4*>(a + v * V);
      if (kBinary) *reinterpret_cast<uint4*>(rb) = *reinterpret_cast<const uint4*>(b + v * V);
#pragma unroll
      for (int e = 0; e < V; ++e) ry[e] = vml_apply<T>(kOp, ra[e], kBinary ? rb[e] : scalar);
      *reinterpret_cast<uint4*>(y + v * V) = *reinterpret_cast<const uint4*>(ry);
    }
    for (long long t = nv * V + i; t < n; t += stride) y[t] = vml_apply<T>(kOp, a[t], kBinary ? b[t] : scalar);
  } else {
    for (; i < n; i += stride) y[i] = vml_apply<T>(kOp, a[i], kBinary ? b[i] : scalar);
  }
}

template <typename T, int kOp, bool kBinary>
int launch_vml(long long n, const T* a, const T* b, T scalar, T* y, cudaStream_t s) {
  if (n <= 0) return 0;
  const uintptr_t bits = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(y) |
                         (kBinary ? reinterpret_cast<uintptr_t>(b) : 0);
  const int vec_ok = (bits & 15u) == 0;
  long long blocks = (n / (vec_ok ? 16 / sizeof(T) : 1) + 255) / 256;
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  vml_kernel<T, kOp, kBinary><<<static_cast<unsigned>(blocks), 256, 0, s>>>(n, a, b, scalar, y, vec_ok);
  return bdl_check_launch("vml_kernel");
}

template <typename T> int vml_dispatch(int op, long long n, const T* a, const T* b, T scalar, T* y, cudaStream_t s) {
  if (a == nullptr || y == nullptr) return bdl_fail("bdl_vml: null pointer");
  switch (op) {
    case kAdd: return b ? launch_vml<T, kAdd, true>(n, a, b, scalar, y, s) : bdl_fail("bdl_vml: Add needs b");
    case kSub: return b ? launch_vml<T, kSub, true>(n, a, b, scalar, y, s) : bdl_fail("bdl_vml: Sub needs b");
    case kMul: return b ? launch_vml<T, kMul, true>(n, a, b, scalar, y, s) : bdl_fail("bdl_vml: Mul needs b");
    case kDiv: return b ? launch_vml<T, kDiv, true>(n, a, b, scalar, y, s) : bdl_fail("bdl_vml: Div needs b");
    case kPowx: return launch_vml<T, kPowx, false>(n, a, b, scalar, y, s);
    case kLn: return launch_vml<T, kLn, false>(n, a, b, scalar, y, s);
    case kExp: return launch_vml<T, kExp, false>(n, a, b, scalar, y, s);
    case kSqrt: return launch_vml<T, kSqrt, false>(n, a, b, scalar, y, s);
    case kTanh: return launch_vml<T, kTanh, false>(n, a, b, scalar, y, s);
    case kAbs: return launch_vml<T, kAbs, false>(n, a, b, scalar, y, s);
    case kLog1p: return launch_vml<T, kLog1p, false>(n, a, b, scalar, y, s);
  }
  return bdl_fail("bdl_vml: unknown op");
}

// ------------------------------------------------------------------------------------------------ level 1
template <typename T>
__global__ void __launch_bounds__(256) axpy_kernel(long long n, T alpha, const T* __restrict__ x, long long incx,
                                                   T* __restrict__ y, long long incy) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    y[i * incy] += alpha * x[i * incx];
}
template <typename T>
__global__ void __launch_bounds__(256) scal_kernel(long long n, T alpha, T* __restrict__ x, long long incx) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    x[i * incx] *= alpha;
}

// dot: per-block partial sums in a fixed order, the last block to finish adds them up in index order, so the result
// does not depend on the scheduling (no floating-point atomics).
constexpr int kDotBlocks = 592;
constexpr int kDotSlots = 8;       // independent scratch sets, handed out round-robin so dots on different streams may overlap
__device__ double g_dot_partial_all[kDotSlots][kDotBlocks];
__device__ unsigned int g_dot_done_all[kDotSlots];

template <typename T>
__global__ void __launch_bounds__(256) dot_kernel(long long n, const T* __restrict__ x, long long incx,
                                                  const T* __restrict__ y, long long incy, T* __restrict__ out,
                                                  int slot) {
  double* g_dot_partial = g_dot_partial_all[slot];
  unsigned int& g_dot_done = g_dot_done_all[slot];
  __shared__ double red[8];
  __shared__ bool last;
  double acc = 0.0;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  if (sizeof(T) == 4) {
    float f = 0.f;   // fp32 accumulation per thread like sdot, widened only across threads
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
      f = fmaf(static_cast<float>(x[i * incx]), static_cast<float>(y[i * incy]), f);
    acc = f;
  } else {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
      acc = fma(static_cast<double>(x[i * incx]), static_cast<double>(y[i * incy]), acc);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; ++w) s += red[w];
    g_dot_partial[blockIdx.x] = s;
    __threadfence();
    last = atomicAdd(&g_dot_done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double s = 0.0;
    for (unsigned b = 0; b < gridDim.x; ++b) s += reinterpret_cast<volatile double*>(g_dot_partial)[b];
    *out = static_cast<T>(s);
    g_dot_done = 0;
  }
}

template <typename T> int launch_axpy(long long n, T alpha, const T* x, long long incx, T* y, long long incy,
                                      cudaStream_t s) {
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  axpy_kernel<T><<<static_cast<unsigned>(blocks > cap ? cap : blocks), 256, 0, s>>>(n, alpha, x, incx, y, incy);
  return bdl_check_launch("axpy_kernel");
}
template <typename T> int launch_scal(long long n, T alpha, T* x, long long incx, cudaStream_t s) {
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  scal_kernel<T><<<static_cast<unsigned>(blocks > cap ? cap : blocks), 256, 0, s>>>(n, alpha, x, incx);
  return bdl_check_launch("scal_kernel");
}
template <typename T> int launch_dot(long long n, const T* x, long long incx, const T* y, long long incy, T* out,
                                     cudaStream_t s) {
  if (out == nullptr) return bdl_fail("bdl_blas_dot: null result pointer");
  long long blocks = (n + 256 * 8 - 1) / (256 * 8);
  if (blocks < 1) blocks = 1;
  if (blocks > kDotBlocks) blocks = kDotBlocks;
  static unsigned next_slot = 0;
  const int slot = static_cast<int>(__atomic_fetch_add(&next_slot, 1u, __ATOMIC_RELAXED) % kDotSlots);
  dot_kernel<T><<<static_cast<unsigned>(blocks), 256, 0, s>>>(n, x, incx, y, incy, out, slot);
  return bdl_check_launch("dot_kernel");
}

// ------------------------------------------------------------------------------------------------ level 2
// A (m x n, column-major, lda) += alpha * x y^T
template <typename T>
__global__ void __launch_bounds__(256) ger_kernel(int m, int n, T alpha, const T* __restrict__ x, long long incx,
                                                  const T* __restrict__ y, long long incy, T* __restrict__ A,
                                                  long long lda) {
  const int i = blockIdx.x * 64 + (threadIdx.x & 63);
  if (i >= m) return;
  const T xi = alpha * x[i * incx];
  for (int j = blockIdx.y * 16 + (threadIdx.x >> 6); j < n; j += gridDim.y * 16) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int jj = j + u * 4;
      if (jj < n) A[i + jj * lda] += xi * y[jj * incy];
    }
    // (threadIdx.x >> 6) in [0,4): together with u the block covers 16 consecutive columns
  }
}

// y = alpha * A x + beta * y, A not transposed: threads run down the rows (coalesced), the 8 y-slices of a block split
// the columns and are added up in shared memory.
template <typename T>
__global__ void __launch_bounds__(256) gemv_n_kernel(int m, int n, T alpha, const T* __restrict__ A, long long lda,
                                                     const T* __restrict__ x, long long incx, T beta,
                                                     T* __restrict__ y, long long incy) {
  __shared__ T part[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  T acc = 0;
  if (i < m)
    for (int j = slice; j < n; j += 8) acc += A[i + j * lda] * x[j * incx];
  part[slice][lane] = acc;
  __syncthreads();
  if (slice == 0 && i < m) {
    T s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += part[k][lane];
    y[i * incy] = alpha * s + (beta == T(0) ? T(0) : beta * y[i * incy]);
  }
}
// y = alpha * A^T x + beta * y: one warp per column of A (a contiguous run), shuffle reduce.
template <typename T>
__global__ void __launch_bounds__(256) gemv_t_kernel(int m, int n, T alpha, const T* __restrict__ A, long long lda,
                                                     const T* __restrict__ x, long long incx, T beta,
                                                     T* __restrict__ y, long long incy) {
  const int lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (j >= n) return;
  T acc = 0;
  const T* col = A + j * lda;
  for (int i = lane; i < m; i += 32) acc += col[i] * x[i * incx];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) y[j * incy] = alpha * acc + (beta == T(0) ? T(0) : beta * y[j * incy]);
}

// ------------------------------------------------------------------------------------------------ level 3
// C (m x n, ldc) = alpha * op(A) op(B) + beta * C, column-major. 256 threads own a BM x BN tile of C, TM x TN values
// each, K in steps of 16 through shared memory with the next slab prefetched into registers during the FMAs.
template <typename T, int BM, int BN, int TM, int TN, bool TA, bool TB>
__global__ void __launch_bounds__(256) gemm_simt_kernel(int m, int n, int k, T alpha, const T* __restrict__ A,
                                                        long long lda, const T* __restrict__ B, long long ldb, T beta,
                                                        T* __restrict__ C, long long ldc) {
  constexpr int BK = 16;
  static_assert((BM / TM) * (BN / TN) == 256, "tile / thread shape mismatch");
  constexpr int LA = BM * BK / 256, LB = BN * BK / 256;
  __shared__ T As[BK][BM + 4];
  __shared__ T Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid % (BM / TM), ty = tid / (BM / TM);
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  T acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0;
  T ra[LA], rb[LB];

  auto fetch = [&](int k0) {
#pragma unroll
    for (int l = 0; l < LA; ++l) {
      const int idx = tid + l * 256;
      const int i = TA ? idx / BK : idx % BM, kk = TA ? idx % BK : idx / BM;
      const int gi = m0 + i, gk = k0 + kk;
      ra[l] = (gi < m && gk < k) ? (TA ? A[gk + gi * lda] : A[gi + gk * lda]) : T(0);
    }
#pragma unroll
    for (int l = 0; l < LB; ++l) {
      const int idx = tid + l * 256;
      const int kk = TB ? idx / BN : idx % BK, j = TB ? idx % BN : idx / BK;
      const int gj = n0 + j, gk = k0 + kk;
      rb[l] = (gj < n && gk < k) ? (TB ? B[gj + gk * ldb] : B[gk + gj * ldb]) : T(0);
    }
  };
  auto stash = [&]() {
#pragma unroll
    for (int l = 0; l < LA; ++l) {
      const int idx = tid + l * 256;
      const int i = TA ? idx / BK : idx % BM, kk = TA ? idx % BK : idx / BM;
      As[kk][i] = ra[l];
    }
#pragma unroll
    for (int l = 0; l < LB; ++l) {
      const int idx = tid + l * 256;
      const int kk = TB ? idx / BN : idx % BK, j = TB ? idx % BN : idx / BK;
      Bs[kk][j] = rb[l];
    }
  };

  fetch(0);
  for (int k0 = 0; k0 < k; k0 += BK) {
    stash();
    __syncthreads();
    if (k0 + BK < k) fetch(k0 + BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      T a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[kk][tx + i * (BM / TM)];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[kk][ty + j * (BN / TN)];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < TN; ++j) {
    const int gj = n0 + ty + j * (BN / TN);
    if (gj >= n) continue;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int gi = m0 + tx + i * (BM / TM);   // consecutive threads -> consecutive rows: coalesced column-major store
      if (gi < m) {
        T* c = C + gi + gj * ldc;
        *c = alpha * acc[i][j] + (beta == T(0) ? T(0) : beta * *c);
      }
    }
  }
}

template <typename T, int BM, int BN, int TM, int TN>
int launch_gemm_simt(int ta, int tb, int m, int n, int k, T alpha, const T* A, long long lda, const T* B, long long ldb,
                     T beta, T* C, long long ldc, cudaStream_t s) {
  dim3 grid((m + BM - 1) / BM, (n + BN - 1) / BN);
  if (!ta && !tb) gemm_simt_kernel<T, BM, BN, TM, TN, false, false><<<grid, 256, 0, s>>>(m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
  else if (ta && !tb) gemm_simt_kernel<T, BM, BN, TM, TN, true, false><<<grid, 256, 0, s>>>(m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
  else if (!ta && tb) gemm_simt_kernel<T, BM, BN, TM, TN, false, true><<<grid, 256, 0, s>>>(m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
  else gemm_simt_kernel<T, BM, BN, TM, TN, true, true><<<grid, 256, 0, s>>>(m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
  return bdl_check_launch("gemm_simt_kernel");
}

template <typename T>
int blas_gemm(int ta, int tb, int m, int n, int k, T alpha, const T* A, long long lda, const T* B, long long ldb, T beta,
              T* C, long long ldc, cudaStream_t s) {
  if (m < 0 || n < 0 || k < 0) return bdl_fail("bdl_blas_gemm: negative dimension");
  if (m == 0 || n == 0) return 0;
  if (!A || !B || !C) return bdl_fail("bdl_blas_gemm: null pointer");
  if (lda < (ta ? k : m) || ldb < (tb ? n : k) || ldc < m) return bdl_fail("bdl_blas_gemm: leading dimension too small");
  const long long tiles_big = static_cast<long long>((m + 127) / 128) * ((n + 127) / 128);
  if (sizeof(T) == 4 && tiles_big >= bdl_num_sms())
    return launch_gemm_simt<T, 128, 128, 8, 8>(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s);
  return launch_gemm_simt<T, 64, 64, 4, 4>(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s);
}

template <typename T>
int blas_gemv(int trans, int m, int n, T alpha, const T* A, long long lda, const T* x, long long incx, T beta, T* y,
              long long incy, cudaStream_t s) {
  if (m < 0 || n < 0) return bdl_fail("bdl_blas_gemv: negative dimension");
  if (m == 0 || n == 0) return 0;
  if (!A || !x || !y) return bdl_fail("bdl_blas_gemv: null pointer");
  if (lda < m) return bdl_fail("bdl_blas_gemv: leading dimension too small");
  if (trans) gemv_t_kernel<T><<<(n + 7) / 8, 256, 0, s>>>(m, n, alpha, A, lda, x, incx, beta, y, incy);
  else gemv_n_kernel<T><<<(m + 31) / 32, 256, 0, s>>>(m, n, alpha, A, lda, x, incx, beta, y, incy);
  return bdl_check_launch("gemv_kernel");
}

template <typename T>
int blas_ger(int m, int n, T alpha, const T* x, long long incx, const T* y, long long incy, T* A, long long lda,
             cudaStream_t s) {
  if (m <= 0 || n <= 0) return 0;
  if (!A || !x || !y) return bdl_fail("bdl_blas_ger: null pointer");
  int gy = (n + 15) / 16;
  if (gy > 1024) gy = 1024;
  ger_kernel<T><<<dim3((m + 63) / 64, gy), 256, 0, s>>>(m, n, alpha, x, incx, y, incy, A, lda);
  return bdl_check_launch("ger_kernel");
}

}  // namespace bdl

#define BDL_BLAS_ENTRY(SUF, T)                                                                                         \
  extern "C" int bdl_vml_##SUF(int op, long long n, const T* a, const T* b, T scalar, T* y, void* stream) {             \
    return bdl::vml_dispatch<T>(op, n, a, b, scalar, y, static_cast<cudaStream_t>(stream));                             \
  }                                                                                                                    \
  extern "C" int bdl_blas_axpy_##SUF(long long n, T alpha, const T* x, long long incx, T* y, long long incy,            \
                                     void* stream) {                                                                   \
    return bdl::launch_axpy<T>(n, alpha, x, incx, y, incy, static_cast<cudaStream_t>(stream));                          \
  }                                                                                                                    \
  extern "C" int bdl_blas_scal_##SUF(long long n, T alpha, T* x, long long incx, void* stream) {                        \
    return bdl::launch_scal<T>(n, alpha, x, incx, static_cast<cudaStream_t>(stream));                                   \
  }                                                                                                                    \
  extern "C" int bdl_blas_dot_##SUF(long long n, const T* x, long long incx, const T* y, long long incy, T* result,     \
                                    void* stream) {                                                                    \
    return bdl::launch_dot<T>(n, x, incx, y, incy, result, static_cast<cudaStream_t>(stream));                          \
  }                                                                                                                    \
  extern "C" int bdl_blas_ger_##SUF(int m, int n, T alpha, const T* x, long long incx, const T* y, long long incy,      \
                                    T* A, long long lda, void* stream) {                                               \
    return bdl::blas_ger<T>(m, n, alpha, x, incx, y, incy, A, lda, static_cast<cudaStream_t>(stream));                  \
  }                                                                                                                    \
  extern "C" int bdl_blas_gemv_##SUF(int trans, int m, int n, T alpha, const T* A, long long lda, const T* x,           \
                                     long long incx, T beta, T* y, long long incy, void* stream) {                     \
    return bdl::blas_gemv<T>(trans, m, n, alpha, A, lda, x, incx, beta, y, incy, static_cast<cudaStream_t>(stream));    \
  }                                                                                                                    \
  extern "C" int bdl_blas_gemm_##SUF(int transa, int transb, int m, int n, int k, T alpha, const T* A, long long lda,   \
                                     const T* B, long long ldb, T beta, T* C, long long ldc, void* stream) {           \
    return bdl::blas_gemm<T>(transa, transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc,                              \
                             static_cast<cudaStream_t>(stream));                                                       \
  }

BDL_BLAS_ENTRY(f32, float)
BDL_BLAS_ENTRY(f64, double)
