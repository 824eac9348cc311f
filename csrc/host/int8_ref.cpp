// CPU restatement of BigQuant's mixed-precision arithmetic (core/bigquant native MixPrecisionGEMM, wrapped by
// S/nn/quantized/Linear.scala and SpatialConvolution.scala; quantisation rule S/nn/quantized/Quantization.scala:35-62):
// symmetric per-row int8 (q = round(x / absmax * 127)), int32 accumulation, dequantise with both row scales.
// It is the oracle the fp8 GPU path is compared against in the accuracy tests, and the timed CPU baseline.
#include <cmath>

#include "bigdl_b200_host.h"
#include "thread_pool.h"

extern "C" void bdlh_quantize_rows_int8(const float* src, int8_t* dst, float* absmax, int rows, int cols) {
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(rows), 8, [=](size_t b, size_t e) {
    for (size_t r = b; r < e; ++r) {
      const float* s = src + r * cols;
      float am = 0.f;
      for (int c = 0; c < cols; ++c) am = std::fmax(am, std::fabs(s[c]));
      absmax[r] = am;
      const float inv = am > 0.f ? 127.f / am : 0.f;
      for (int c = 0; c < cols; ++c) {
        float v = std::nearbyint(s[c] * inv);
        v = v > 127.f ? 127.f : (v < -127.f ? -127.f : v);
        dst[r * cols + c] = static_cast<int8_t>(v);
      }
    }
  });
}

extern "C" void bdlh_int8_gemm(const int8_t* A, const float* a_absmax, const int8_t* B, const float* b_absmax,
                               const float* bias, float* C, int M, int N, int K) {
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(M), 1, [=](size_t mb, size_t me) {
    for (size_t m = mb; m < me; ++m) {
      const int8_t* a = A + m * K;
      for (int n = 0; n < N; ++n) {
        const int8_t* b = B + static_cast<size_t>(n) * K;
        int32_t acc = 0;
        for (int k = 0; k < K; ++k) acc += static_cast<int32_t>(a[k]) * static_cast<int32_t>(b[k]);
        C[m * N + n] = static_cast<float>(acc) * (a_absmax[m] / 127.f) * (b_absmax[n] / 127.f) + (bias ? bias[n] : 0.f);
      }
    }
  });
}
