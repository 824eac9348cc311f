// GPU side of the image input pipeline: raw uint8 HWC frames (what the decoder produces and what crosses PCIe — 4x fewer
// bytes than fp32) -> per-image random crop + horizontal flip + (x - mean) / std -> bf16 NHWC batch, in one pass.
// Reference counterparts run on CPU threads over OpenCV mats: S/dataset/image/{BGRImgCropper,HFlip,BGRImgNormalizer,
// BGRImgToBatch}.scala and S/transform/vision/image/augmentation/{RandomCropper,HFlip,ChannelNormalize}.scala.
#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

struct Norm3 {
  float mean[4];
  float inv_std[4];
};

// one thread per output pixel; C in 1..4. crop_xy[n] = (x0, y0) of the crop window, flip[n] != 0 mirrors it.
template <typename Out>
__global__ void image_batch_u8_kernel(const uint8_t* __restrict__ src, Out* __restrict__ dst,
                                      const int* __restrict__ crop_xy, const uint8_t* __restrict__ flip, int N, int H,
                                      int W, int C, int Hc, int Wc, Norm3 nm, int out_nchw) {
  const long long total = static_cast<long long>(N) * Hc * Wc;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int x = static_cast<int>(i % Wc);
    long long t = i / Wc;
    const int y = static_cast<int>(t % Hc);
    const int n = static_cast<int>(t / Hc);
    const int x0 = crop_xy ? crop_xy[2 * n] : (W - Wc) / 2;
    const int y0 = crop_xy ? crop_xy[2 * n + 1] : (H - Hc) / 2;
    const int sx = x0 + ((flip && flip[n]) ? (Wc - 1 - x) : x);
    const int sy = y0 + y;
    const bool in = sx >= 0 && sx < W && sy >= 0 && sy < H;
    const uint8_t* p = src + ((static_cast<long long>(n) * H + sy) * W + sx) * C;
    for (int c = 0; c < C; ++c) {
      const float v = in ? (static_cast<float>(p[c]) - nm.mean[c]) * nm.inv_std[c] : 0.f;
      const long long o = out_nchw ? ((static_cast<long long>(n) * C + c) * Hc + y) * Wc + x
                                   : (static_cast<long long>(i) * C + c);
      dst[o] = static_cast<Out>(v);
    }
  }
}

}  // namespace bdl
using namespace bdl;

extern "C" int bdl_image_batch_u8(const void* src_u8, void* dst, const int* crop_xy, const void* flip, int N, int H,
                                  int W, int C, int Hc, int Wc, const float* mean, const float* std_, int out_fp32,
                                  int out_nchw, void* stream) {
  if (C < 1 || C > 4) return bdl_fail("bdl_image_batch_u8: 1..4 channels");
  if (Hc <= 0 || Wc <= 0 || N <= 0) return BDL_OK;
  Norm3 nm{};
  for (int c = 0; c < C; ++c) {
    nm.mean[c] = mean ? mean[c] : 0.f;
    nm.inv_std[c] = (std_ && std_[c] != 0.f) ? 1.f / std_[c] : 1.f;
  }
  const long long total = static_cast<long long>(N) * Hc * Wc;
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(bdl_num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_fp32)
    image_batch_u8_kernel<float><<<static_cast<int>(blocks), 256, 0, st>>>(
        static_cast<const uint8_t*>(src_u8), static_cast<float*>(dst), crop_xy, static_cast<const uint8_t*>(flip), N, H,
        W, C, Hc, Wc, nm, out_nchw);
  else
    image_batch_u8_kernel<__nv_bfloat16><<<static_cast<int>(blocks), 256, 0, st>>>(
        static_cast<const uint8_t*>(src_u8), static_cast<__nv_bfloat16*>(dst), crop_xy,
        static_cast<const uint8_t*>(flip), N, H, W, C, Hc, Wc, nm, out_nchw);
  return bdl_check_launch("image_batch_u8_kernel");
}
