// Multi-threaded mini-batch assembly: crop / mirror / normalise / layout-convert uint8 images straight into the
// (pinned) batch buffer that is then copied to the GPU. One pass per pixel instead of the reference's chain of
// per-image transformers feeding MTLabeledBGRImgToBatch (S/dataset/image/MTLabeledBGRImgToBatch.scala:60-130,
// BGRImgCropper.scala, HFlip.scala, BGRImgNormalizer.scala).
#include <cstring>

#include "bigdl_b200_host.h"
#include "thread_pool.h"

namespace {
inline uint64_t splitmix(uint64_t& s) {
  uint64_t z = (s += 0x9e3779b97f4a7c15ull);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
inline uint16_t to_bf16(float f) {   // round to nearest even
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40u);
  u += 0x7fffu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(u >> 16);
}
}  // namespace

extern "C" int bdlh_batch_images(const uint8_t* const* images, const int* h, const int* w, int n, int crop_h, int crop_w,
                                 int random_crop, int random_flip, const uint64_t* seeds, const float* mean,
                                 const float* stdv, float scale, int nhwc, int out_bf16, void* out) {
  for (int i = 0; i < n; ++i)
    if (h[i] < crop_h || w[i] < crop_w) return -1;
  const float inv_scale = 1.0f / scale;
  float mul[3], sub[3];
  for (int c = 0; c < 3; ++c) {
    mul[c] = inv_scale / stdv[c];
    sub[c] = mean[c] / stdv[c];
  }
  const size_t per = static_cast<size_t>(3) * crop_h * crop_w;
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(n), 1, [&](size_t b, size_t e) {
    for (size_t i = b; i < e; ++i) {
      uint64_t s = seeds ? seeds[i] : 0;
      int oy = (h[i] - crop_h) / 2, ox = (w[i] - crop_w) / 2;
      if (random_crop) {
        oy = static_cast<int>(splitmix(s) % static_cast<uint64_t>(h[i] - crop_h + 1));
        ox = static_cast<int>(splitmix(s) % static_cast<uint64_t>(w[i] - crop_w + 1));
      }
      const bool flip = random_flip && (splitmix(s) & 1ull);
      const uint8_t* img = images[i];
      float* of = out_bf16 ? nullptr : static_cast<float*>(out) + i * per;
      uint16_t* oh = out_bf16 ? static_cast<uint16_t*>(out) + i * per : nullptr;
      for (int y = 0; y < crop_h; ++y) {
        const uint8_t* row = img + (static_cast<size_t>(oy + y) * w[i] + ox) * 3;
        for (int x = 0; x < crop_w; ++x) {
          const uint8_t* px = row + static_cast<size_t>(flip ? crop_w - 1 - x : x) * 3;
          for (int c = 0; c < 3; ++c) {
            const float v = static_cast<float>(px[c]) * mul[c] - sub[c];
            const size_t idx = nhwc ? (static_cast<size_t>(y) * crop_w + x) * 3 + c
                                    : (static_cast<size_t>(c) * crop_h + y) * crop_w + x;
            if (out_bf16) oh[idx] = to_bf16(v); else of[idx] = v;
          }
        }
      }
    }
  });
  return 0;
}

extern "C" void bdlh_gather_rows(const void* const* rows, size_t row_bytes, int n, void* out) {
  uint8_t* o = static_cast<uint8_t*>(out);
  const size_t grain = row_bytes >= (1u << 16) ? 1 : ((1u << 16) / (row_bytes ? row_bytes : 1) + 1);
  bdlh::ThreadPool::instance().parallel_for(static_cast<size_t>(n), grain, [=](size_t b, size_t e) {
    for (size_t i = b; i < e; ++i) std::memcpy(o + i * row_bytes, rows[i], row_bytes);
  });
}
