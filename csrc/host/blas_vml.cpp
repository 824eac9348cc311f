// Host (CPU) fp32 / fp64 BLAS + vector math: the same entry points as csrc/blas_vml.cu for tensors that live in host
// memory (the reference's mkl.c JNI surface, core/mkl2017-xeon-blas/src/main/c/mkl.c). Column-major, raw pointers.
// Parallelised over the runtime's worker pool; the inner loops are written so g++ -O3 -march=x86-64-v3 vectorises them.
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <vector>

#include "thread_pool.h"

namespace {

enum VmlOp { kAdd = 0, kSub, kMul, kDiv, kPowx, kLn, kExp, kSqrt, kTanh, kAbs, kLog1p };

template <typename T, typename F> void map1(size_t n, const T* a, T* y, F f) {
  bdlh::ThreadPool::instance().parallel_for(n, 1 << 15, [&](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; ++i) y[i] = f(a[i]);
  });
}
template <typename T, typename F> void map2(size_t n, const T* a, const T* b, T* y, F f) {
  bdlh::ThreadPool::instance().parallel_for(n, 1 << 15, [&](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; ++i) y[i] = f(a[i], b[i]);
  });
}

template <typename T> int vml(int op, long long n_, const T* a, const T* b, T s, T* y) {
  if (n_ <= 0) return 0;
  if (!a || !y) return 1;
  const size_t n = static_cast<size_t>(n_);
  switch (op) {
    case kAdd: if (!b) return 1; map2<T>(n, a, b, y, [](T u, T v) { return u + v; }); return 0;
    case kSub: if (!b) return 1; map2<T>(n, a, b, y, [](T u, T v) { return u - v; }); return 0;
    case kMul: if (!b) return 1; map2<T>(n, a, b, y, [](T u, T v) { return u * v; }); return 0;
    case kDiv: if (!b) return 1; map2<T>(n, a, b, y, [](T u, T v) { return u / v; }); return 0;
    case kPowx: map1<T>(n, a, y, [s](T u) { return std::pow(u, s); }); return 0;
    case kLn: map1<T>(n, a, y, [](T u) { return std::log(u); }); return 0;
    case kExp: map1<T>(n, a, y, [](T u) { return std::exp(u); }); return 0;
    case kSqrt: map1<T>(n, a, y, [](T u) { return std::sqrt(u); }); return 0;
    case kTanh: map1<T>(n, a, y, [](T u) { return std::tanh(u); }); return 0;
    case kAbs: map1<T>(n, a, y, [](T u) { return std::fabs(u); }); return 0;
    case kLog1p: map1<T>(n, a, y, [](T u) { return std::log1p(u); }); return 0;
  }
  return 1;
}

template <typename T> void axpy(long long n, T alpha, const T* x, long long incx, T* y, long long incy) {
  if (n <= 0) return;
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(n), 1 << 15, [&](size_t lo, size_t hi) {
    if (incx == 1 && incy == 1) {
      for (size_t i = lo; i < hi; ++i) y[i] += alpha * x[i];
    } else {
      for (size_t i = lo; i < hi; ++i) y[i * incy] += alpha * x[i * incx];
    }
  });
}

template <typename T> T dot(long long n, const T* x, long long incx, const T* y, long long incy) {
  // fixed 4096-element blocks summed in index order: the value does not depend on the number of worker threads
  if (n <= 0) return T(0);
  const size_t blk = 4096, nb = (static_cast<size_t>(n) + blk - 1) / blk;
  std::vector<double> part(nb, 0.0);
  bdlh::ThreadPool::instance().parallel_for(nb, 8, [&](size_t lo, size_t hi) {
    for (size_t b = lo; b < hi; ++b) {
      const size_t e = std::min(static_cast<size_t>(n), (b + 1) * blk);
      T acc = 0;
      for (size_t i = b * blk; i < e; ++i) acc += x[i * incx] * y[i * incy];
      part[b] = acc;
    }
  });
  double s = 0.0;
  for (double p : part) s += p;
  return static_cast<T>(s);
}

template <typename T> void scal(long long n, T alpha, T* x, long long incx) {
  for (long long i = 0; i < n; ++i) x[i * incx] *= alpha;
}

template <typename T>
void ger(int m, int n, T alpha, const T* x, long long incx, const T* y, long long incy, T* A, long long lda) {
  if (m <= 0 || n <= 0) return;
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(n), 8, [&](size_t lo, size_t hi) {
    for (size_t j = lo; j < hi; ++j) {
      const T yj = alpha * y[j * incy];
      T* col = A + j * lda;
      if (incx == 1) for (int i = 0; i < m; ++i) col[i] += x[i] * yj;
      else for (int i = 0; i < m; ++i) col[i] += x[i * incx] * yj;
    }
  });
}

template <typename T>
void gemv(int trans, int m, int n, T alpha, const T* A, long long lda, const T* x, long long incx, T beta, T* y,
          long long incy) {
  if (m <= 0 || n <= 0) return;
  if (trans) {
    bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(n), 16, [&](size_t lo, size_t hi) {
      for (size_t j = lo; j < hi; ++j) {
        const T* col = A + j * lda;
        T acc = 0;
        for (int i = 0; i < m; ++i) acc += col[i] * x[i * incx];
        y[j * incy] = alpha * acc + (beta == T(0) ? T(0) : beta * y[j * incy]);
      }
    });
  } else {
    bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(m), 256, [&](size_t lo, size_t hi) {
      std::vector<T> acc(hi - lo, T(0));
      for (int j = 0; j < n; ++j) {
        const T xj = x[j * incx];
        const T* col = A + j * lda + lo;
        for (size_t i = 0; i < hi - lo; ++i) acc[i] += col[i] * xj;
      }
      for (size_t i = lo; i < hi; ++i) y[i * incy] = alpha * acc[i - lo] + (beta == T(0) ? T(0) : beta * y[i * incy]);
    });
  }
}

// C = alpha op(A) op(B) + beta C. Columns of C are split across the workers; each worker packs op(A) panels so the
// innermost loop is a unit-stride axpy over a column of C (vectorised).
template <typename T>
void gemm(int ta, int tb, int m, int n, int k, T alpha, const T* A, long long lda, const T* B, long long ldb, T beta,
          T* C, long long ldc) {
  if (m <= 0 || n <= 0) return;
  std::vector<T> At;                     // op(A) as column-major m x k when A is given transposed
  const T* Ap = A;
  long long lap = lda;
  if (ta) {
    At.resize(static_cast<size_t>(m) * k);
    for (int i = 0; i < m; ++i)
      for (int p = 0; p < k; ++p) At[i + static_cast<size_t>(p) * m] = A[p + i * lda];
    Ap = At.data();
    lap = m;
  }
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(n), 4, [&](size_t lo, size_t hi) {
    constexpr int KB = 256;
    for (size_t j = lo; j < hi; ++j) {
      T* c = C + j * ldc;
      if (beta == T(0)) for (int i = 0; i < m; ++i) c[i] = 0;
      else if (beta != T(1)) for (int i = 0; i < m; ++i) c[i] *= beta;
    }
    for (int p0 = 0; p0 < k; p0 += KB) {
      const int p1 = p0 + KB < k ? p0 + KB : k;
      for (size_t j = lo; j < hi; ++j) {
        T* c = C + j * ldc;
        for (int p = p0; p < p1; ++p) {
          const T b = alpha * (tb ? B[j + static_cast<size_t>(p) * ldb] : B[p + j * ldb]);
          const T* a = Ap + static_cast<size_t>(p) * lap;
          for (int i = 0; i < m; ++i) c[i] += a[i] * b;
        }
      }
    }
  });
}

}  // namespace

#define BDLH_BLAS_ENTRY(SUF, T)                                                                                        \
  extern "C" int bdlh_vml_##SUF(int op, long long n, const T* a, const T* b, T scalar, T* y) {                          \
    return vml<T>(op, n, a, b, scalar, y);                                                                             \
  }                                                                                                                    \
  extern "C" void bdlh_blas_axpy_##SUF(long long n, T alpha, const T* x, long long incx, T* y, long long incy) {        \
    axpy<T>(n, alpha, x, incx, y, incy);                                                                               \
  }                                                                                                                    \
  extern "C" void bdlh_blas_scal_##SUF(long long n, T alpha, T* x, long long incx) { scal<T>(n, alpha, x, incx); }      \
  extern "C" T bdlh_blas_dot_##SUF(long long n, const T* x, long long incx, const T* y, long long incy) {               \
    return dot<T>(n, x, incx, y, incy);                                                                                \
  }                                                                                                                    \
  extern "C" void bdlh_blas_ger_##SUF(int m, int n, T alpha, const T* x, long long incx, const T* y, long long incy,    \
                                      T* A, long long lda) {                                                           \
    ger<T>(m, n, alpha, x, incx, y, incy, A, lda);                                                                     \
  }                                                                                                                    \
  extern "C" void bdlh_blas_gemv_##SUF(int trans, int m, int n, T alpha, const T* A, long long lda, const T* x,         \
                                       long long incx, T beta, T* y, long long incy) {                                 \
    gemv<T>(trans, m, n, alpha, A, lda, x, incx, beta, y, incy);                                                       \
  }                                                                                                                    \
  extern "C" void bdlh_blas_gemm_##SUF(int transa, int transb, int m, int n, int k, T alpha, const T* A, long long lda, \
                                       const T* B, long long ldb, T beta, T* C, long long ldc) {                       \
    gemm<T>(transa, transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);                                             \
  }

BDLH_BLAS_ENTRY(f32, float)
BDLH_BLAS_ENTRY(f64, double)
