// Host-side helpers shared by every .cu translation unit: error reporting, device properties,
// TMA tensor-map construction through the driver entry point (no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

int bdl_fail(const char* msg);                       // records msg, returns BDL_ERR_INVALID
int bdl_fail_cuda(const char* what, cudaError_t e);  // records "what: cudaGetErrorString(e)", returns BDL_ERR_CUDA
int bdl_check_launch(const char* kernel_name);       // cudaGetLastError() after a launch
int bdl_num_sms();                                   // SM count of the current device (cached per device)

// 2-D bf16 tensor map with 128-byte swizzle. dim0 is the contiguous dimension; ld is the row pitch in elements.
int bdl_make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld, uint32_t box0,
                          uint32_t box1);
// 3-D bf16 tensor map (dims {dim0, dim1, dim2}, pitches ld1 / ld2 in elements), box {box0, box1, 1}, 128-byte swizzle,
// zero fill outside the tensor: a [batch, length, width] activation read one (length x width) tile at a time.
int bdl_make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, uint64_t dim2, long long ld1,
                          long long ld2, uint32_t box0, uint32_t box1);
// Same for 1-byte elements (fp8): box0 is in elements (= bytes), at most 128.
int bdl_make_tmap_2d_bf16_sw(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld,
                             uint32_t box0, uint32_t box1, int swizzle_bytes);
int bdl_make_tmap_im2col_bf16_sw(CUtensorMap* out, const void* base, int C, int W, int H, int N, long long ld_pixel,
                                 int lower_w, int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                                 int pixels_per_column, int stride_w, int stride_h, int swizzle_bytes);
int bdl_make_tmap_2d_u8_sw(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld, uint32_t box0,
                           uint32_t box1, int swizzle_bytes);
int bdl_make_tmap_2d_u8(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, long long ld, uint32_t box0,
                        uint32_t box1);

// 4-D NHWC im2col tensor map (dims {C, W, H, N}, pixel pitch ld elements): bounding box corners for W/H, number of
// channels and pixels fetched per instruction, traversal strides. 128-byte swizzle, zero fill outside the tensor.
int bdl_make_tmap_im2col_bf16(CUtensorMap* out, const void* base, int C, int W, int H, int N, long long ld_pixel,
                              int lower_w, int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                              int pixels_per_column, int stride_w, int stride_h);
int bdl_make_tmap_im2col_u8(CUtensorMap* out, const void* base, int C, int W, int H, int N, long long ld_pixel,
                            int lower_w, int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                            int pixels_per_column, int stride_w, int stride_h);
