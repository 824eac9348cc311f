// CRC32C (Castagnoli, used by TFRecord / TensorBoard event files) and CRC32 (zlib polynomial, S/utils/Crc32.scala).
// SSE4.2 crc32 instructions when the CPU has them, slice-by-8 tables otherwise.
#include <cstring>

#include "bigdl_b200_host.h"

#if defined(__x86_64__)
#include <cpuid.h>
#include <nmmintrin.h>
#endif

namespace {

struct Tables {
  uint32_t c[8][256];   // Castagnoli 0x82F63B78
  uint32_t z[8][256];   // zlib 0xEDB88320
  Tables() {
    build(c, 0x82F63B78u);
    build(z, 0xEDB88320u);
  }
  static void build(uint32_t t[8][256], uint32_t poly) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t r = i;
      for (int k = 0; k < 8; ++k) r = (r >> 1) ^ ((r & 1u) ? poly : 0u);
      t[0][i] = r;
    }
    for (uint32_t i = 0; i < 256; ++i)
      for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xffu];
  }
};
const Tables& tables() {
  static Tables t;
  return t;
}

uint32_t slice8(const uint32_t t[8][256], const uint8_t* p, size_t n, uint32_t crc) {
  while (n >= 8) {
    uint32_t lo, hi;
    std::memcpy(&lo, p, 4);
    std::memcpy(&hi, p + 4, 4);
    lo ^= crc;
    crc = t[7][lo & 0xff] ^ t[6][(lo >> 8) & 0xff] ^ t[5][(lo >> 16) & 0xff] ^ t[4][lo >> 24] ^ t[3][hi & 0xff] ^
          t[2][(hi >> 8) & 0xff] ^ t[1][(hi >> 16) & 0xff] ^ t[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) crc = (crc >> 8) ^ t[0][(crc ^ *p++) & 0xffu];
  return crc;
}

#if defined(__x86_64__)
bool has_sse42() {
  static const bool v = [] {
    unsigned a, b, c, d;
    return __get_cpuid(1, &a, &b, &c, &d) && (c & bit_SSE4_2);
  }();
  return v;
}
__attribute__((target("sse4.2"))) uint32_t crc32c_hw(const uint8_t* p, size_t n, uint32_t crc) {
  uint64_t c = crc;
  while (n >= 8) {
    uint64_t v;
    std::memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8;
    n -= 8;
  }
  uint32_t c32 = static_cast<uint32_t>(c);
  while (n--) c32 = _mm_crc32_u8(c32, *p++);
  return c32;
}
#endif

}  // namespace

extern "C" uint32_t bdlh_crc32c(const void* data, size_t n, uint32_t seed) {
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t crc = ~seed;
#if defined(__x86_64__)
  if (has_sse42()) return ~crc32c_hw(p, n, crc);
#endif
  return ~slice8(tables().c, p, n, crc);
}

extern "C" uint32_t bdlh_crc32c_masked(const void* data, size_t n) {
  const uint32_t crc = bdlh_crc32c(data, n, 0);
  return ((crc >> 15) | (crc << 17)) + 0xa282ead8u;
}

extern "C" uint32_t bdlh_crc32(const void* data, size_t n, uint32_t seed) {
  return ~slice8(tables().z, static_cast<const uint8_t*>(data), n, ~seed);
}

// length(u64 LE) | masked_crc(length) | payload | masked_crc(payload)
extern "C" size_t bdlh_tfrecord_frame(const void* payload, size_t n, void* out) {
  uint8_t* o = static_cast<uint8_t*>(out);
  const uint64_t len = n;
  std::memcpy(o, &len, 8);
  const uint32_t lc = bdlh_crc32c_masked(o, 8);
  std::memcpy(o + 8, &lc, 4);
  std::memcpy(o + 12, payload, n);
  const uint32_t pc = bdlh_crc32c_masked(payload, n);
  std::memcpy(o + 12 + n, &pc, 4);
  return n + 16;
}

extern "C" long long bdlh_tfrecord_next(const void* buf, size_t buf_len, size_t pos, size_t* offset, int verify) {
  const uint8_t* b = static_cast<const uint8_t*>(buf);
  if (pos == buf_len) return -1;
  if (pos + 12 > buf_len) return -2;
  uint64_t len;
  uint32_t lc;
  std::memcpy(&len, b + pos, 8);
  std::memcpy(&lc, b + pos + 8, 4);
  if (verify && lc != bdlh_crc32c_masked(b + pos, 8)) return -2;
  if (pos + 12 + len + 4 > buf_len) return -2;
  if (verify) {
    uint32_t pc;
    std::memcpy(&pc, b + pos + 12 + len, 4);
    if (pc != bdlh_crc32c_masked(b + pos + 12, len)) return -2;
  }
  *offset = pos + 12;
  return static_cast<long long>(len);
}
