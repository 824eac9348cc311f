// Truncated-bf16 gradient codec on the host: the wire format of the reference's parameter all-reduce
// (S/parameters/FP16CompressedTensor.scala:150-270 keeps the upper two bytes of every float and sums compressed
// buffers by decompress-add-truncate). Used by the gloo/CPU backend of AllReduceParameter and by checkpoint tooling.
#include <cstring>

#include "bigdl_b200_host.h"
#include "thread_pool.h"

namespace {
inline float up(uint16_t h) {
  const uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline uint16_t down(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  return static_cast<uint16_t>(u >> 16);
}
constexpr size_t kGrain = 1 << 16;
}  // namespace

extern "C" void bdlh_bf16_compress(const float* src, uint16_t* dst, size_t n) {
  bdlh::ThreadPool::instance().parallel_for(n, kGrain, [=](size_t b, size_t e) {
    for (size_t i = b; i < e; ++i) dst[i] = down(src[i]);
  });
}

extern "C" void bdlh_bf16_decompress(const uint16_t* src, float* dst, size_t n) {
  bdlh::ThreadPool::instance().parallel_for(n, kGrain, [=](size_t b, size_t e) {
    for (size_t i = b; i < e; ++i) dst[i] = up(src[i]);
  });
}

extern "C" void bdlh_bf16_add(uint16_t* acc, const uint16_t* other, size_t n) {
  bdlh::ThreadPool::instance().parallel_for(n, kGrain, [=](size_t b, size_t e) {
    for (size_t i = b; i < e; ++i) acc[i] = down(up(acc[i]) + up(other[i]));
  });
}
