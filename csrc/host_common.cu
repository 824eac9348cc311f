#include "host_common.h"

#include <mutex>
#include <stdio.h>
#include <string.h>

#include "bigdl_b200.h"

static thread_local char g_err[512] = "";

int bdl_fail(const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return BDL_ERR_INVALID;
}
int bdl_fail_cuda(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return BDL_ERR_CUDA;
}
int bdl_check_launch(const char* kernel_name) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return bdl_fail_cuda(kernel_name, e);
  return BDL_OK;
}
extern "C" const char* bdl_last_error() { return g_err; }
extern "C" int bdl_version() { return 100; }

int bdl_num_sms() {
  static int cache[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cache[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev] = n;
  }
  return cache[dev];
}
extern "C" int bdl_sm_count() { return bdl_num_sms(); }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// cuTensorMapEncode* are driver calls: a thread that has not yet made a runtime call that binds the primary context
// (autograd's backward worker threads, served from the caching allocator) gets CUDA_ERROR_INVALID_CONTEXT. Binding is
// what cudaSetDevice does since CUDA 12.
static bool bind_primary_context() {
  int dev = 0;
  return cudaGetDevice(&dev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess;
}

static CUtensorMapSwizzle swizzle_enum(int bytes) {
  return bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
}

static int make_tmap_2d(CUtensorMap* out, CUtensorMapDataType dt, int esize, const void* base, uint64_t dim0,
                        uint64_t dim1, long long ld, uint32_t box0, uint32_t box1, int swizzle_bytes = 128) {
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) return bdl_fail("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
  cuuint64_t dims[2] = {dim0, dim1};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * esize};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = fn(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           swizzle_enum(swizzle_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT || !bind_primary_context()) break;
  }
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof(buf),
             "cuTensorMapEncodeTiled failed (CUresult %d) dims=(%llu,%llu) ld=%lld box=(%u,%u) base=%p", (int)r,
             (unsigned long long)dim0, (unsigned long long)dim1, ld, box0, box1, base);
    return bdl_fail(buf);
  }
  return BDL_OK;
}

int bdl_make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, uint64_t dim2, long long ld1,
                          long long ld2, uint32_t box0, uint32_t box1) {
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) return bdl_fail("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
  cuuint64_t dims[3] = {dim0, dim1, dim2};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld1) * 2, static_cast<cuuint64_t>(ld2) * 2};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT || !bind_primary_context()) break;
  }
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled(3d) failed (CUresult %d) dims=(%llu,%llu,%llu) ld=(%lld,%lld)",
             (int)r, (unsigned long long)dim0, (unsigned long long)dim1, (unsigned long long)dim2, ld1, ld2);
    return bdl_fail(buf);
  }
  return BDL_OK;
}

int bdl_make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld, uint32_t box0,
                          uint32_t box1) {
  return make_tmap_2d(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dim0, dim1, ld, box0, box1);
}
int bdl_make_tmap_2d_bf16_sw(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld,
                             uint32_t box0, uint32_t box1, int swizzle_bytes) {
  return make_tmap_2d(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dim0, dim1, ld, box0, box1, swizzle_bytes);
}
int bdl_make_tmap_2d_u8_sw(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld, uint32_t box0,
                           uint32_t box1, int swizzle_bytes) {
  return make_tmap_2d(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base, dim0, dim1, ld, box0, box1, swizzle_bytes);
}
int bdl_make_tmap_2d_u8(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld, uint32_t box0,
                        uint32_t box1) {
  return make_tmap_2d(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base, dim0, dim1, ld, box0, box1);
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn get_encode_im2col() {
  static EncodeIm2colFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeIm2colFn>(p);
  });
  return fn;
}

static int make_tmap_im2col(CUtensorMap* out, CUtensorMapDataType dt, int esize, const void* base, int C, int W, int H,
                           int N, long long ld_pixel, int lower_w, int lower_h, int upper_w, int upper_h,
                           int channels_per_pixel, int pixels_per_column, int stride_w, int stride_h,
                           int swizzle_bytes = 128);

int bdl_make_tmap_im2col_bf16(CUtensorMap* out, const void* base, int C, int W, int H, int N, long long ld_pixel,
                              int lower_w, int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                              int pixels_per_column, int stride_w, int stride_h) {
  return make_tmap_im2col(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, C, W, H, N, ld_pixel, lower_w, lower_h,
                          upper_w, upper_h, channels_per_pixel, pixels_per_column, stride_w, stride_h);
}
int bdl_make_tmap_im2col_bf16_sw(CUtensorMap* out, const void* base, int C, int W, int H, int N, long long ld_pixel,
                                 int lower_w, int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                                 int pixels_per_column, int stride_w, int stride_h, int swizzle_bytes) {
  return make_tmap_im2col(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, C, W, H, N, ld_pixel, lower_w, lower_h,
                          upper_w, upper_h, channels_per_pixel, pixels_per_column, stride_w, stride_h, swizzle_bytes);
}
int bdl_make_tmap_im2col_u8(CUtensorMap* out, const void* base, int C, int W, int H, int N, long long ld_pixel,
                            int lower_w, int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                            int pixels_per_column, int stride_w, int stride_h) {
  return make_tmap_im2col(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base, C, W, H, N, ld_pixel, lower_w, lower_h, upper_w,
                          upper_h, channels_per_pixel, pixels_per_column, stride_w, stride_h);
}

static int make_tmap_im2col(CUtensorMap* out, CUtensorMapDataType dt, int esize, const void* base, int C, int W, int H,
                           int N, long long ld_pixel, int lower_w, int lower_h, int upper_w, int upper_h,
                           int channels_per_pixel, int pixels_per_column, int stride_w, int stride_h,
                           int swizzle_bytes) {
  EncodeIm2colFn fn = get_encode_im2col();
  if (!fn) return bdl_fail("cuTensorMapEncodeIm2col entry point not available (no CUDA driver?)");
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(N)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld_pixel) * esize, static_cast<cuuint64_t>(ld_pixel) * esize * W,
                           static_cast<cuuint64_t>(ld_pixel) * esize * W * H};
  int lower[2] = {lower_w, lower_h};
  int upper[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(stride_w), static_cast<cuuint32_t>(stride_h), 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = fn(out, dt, 4, const_cast<void*>(base), dims, strides, lower, upper,
           static_cast<cuuint32_t>(channels_per_pixel), static_cast<cuuint32_t>(pixels_per_column), estr,
           CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_enum(swizzle_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT || !bind_primary_context()) break;
  }
  if (r != CUDA_SUCCESS) {
    char buf[320];
    snprintf(buf, sizeof(buf),
             "cuTensorMapEncodeIm2col failed (CUresult %d) dims=(C%d,W%d,H%d,N%d) ld=%lld lower=(%d,%d) upper=(%d,%d) "
             "cpp=%d ppc=%d stride=(%d,%d)",
             (int)r, C, W, H, N, ld_pixel, lower_w, lower_h, upper_w, upper_h, channels_per_pixel, pixels_per_column,
             stride_w, stride_h);
    return bdl_fail(buf);
  }
  return BDL_OK;
}

// 0 = not capturing, 1 = capture active, 2 = capture invalidated (an illegal call happened), <0 = query failed.
// Diagnostic used when a graph capture of a training step fails: poll it after each op to find the offender.
extern "C" int bdl_stream_capture_status(void* stream) {
  cudaStreamCaptureStatus st;
  cudaError_t e = cudaStreamIsCapturing(static_cast<cudaStream_t>(stream), &st);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return -static_cast<int>(e);
  }
  return static_cast<int>(st);
}
