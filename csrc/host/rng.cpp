// MT19937 stream with the sampling conventions of Torch7's THRandom, which the reference re-implements in
// S/utils/RandomGenerator.scala:60-260: uniform = u32 * 2^-32 scaled to [a,b); normal = Box-Muller on two uniforms
// (cached second value NOT kept: Torch draws a fresh pair per call in this variant).
#include <cmath>
#include <cstring>
#include <random>

#include "bigdl_b200_host.h"

namespace {
struct Rng {
  std::mt19937 eng;
  explicit Rng(uint64_t seed) : eng(static_cast<uint32_t>(seed)) {}
  double uni() { return static_cast<double>(eng()) * (1.0 / 4294967296.0); }
};
}  // namespace

extern "C" void* bdlh_rng_new(uint64_t seed) { return new Rng(seed); }
extern "C" void bdlh_rng_free(void* r) { delete static_cast<Rng*>(r); }
extern "C" void bdlh_rng_seed(void* r, uint64_t seed) { static_cast<Rng*>(r)->eng.seed(static_cast<uint32_t>(seed)); }
extern "C" uint32_t bdlh_rng_u32(void* r) { return static_cast<uint32_t>(static_cast<Rng*>(r)->eng()); }
extern "C" double bdlh_rng_uniform(void* r, double a, double b) { return static_cast<Rng*>(r)->uni() * (b - a) + a; }
extern "C" double bdlh_rng_normal(void* r, double mean, double stdv) {
  Rng* g = static_cast<Rng*>(r);
  const double u1 = g->uni();
  const double u2 = g->uni();
  const double rad = std::sqrt(-2.0 * std::log(1.0 - u2));
  return rad * std::cos(2.0 * M_PI * u1) * stdv + mean;
}
extern "C" void bdlh_rng_fill_uniform(void* r, float* dst, size_t n, double a, double b) {
  for (size_t i = 0; i < n; ++i) dst[i] = static_cast<float>(bdlh_rng_uniform(r, a, b));
}
extern "C" void bdlh_rng_fill_normal(void* r, float* dst, size_t n, double mean, double stdv) {
  for (size_t i = 0; i < n; ++i) dst[i] = static_cast<float>(bdlh_rng_normal(r, mean, stdv));
}
extern "C" void bdlh_rng_permutation(void* r, int64_t* dst, size_t n) {
  Rng* g = static_cast<Rng*>(r);
  for (size_t i = 0; i < n; ++i) dst[i] = static_cast<int64_t>(i);
  for (size_t i = n; i > 1; --i) {
    const size_t j = static_cast<size_t>(g->uni() * static_cast<double>(i));
    const int64_t t = dst[i - 1];
    dst[i - 1] = dst[j];
    dst[j] = t;
  }
}
