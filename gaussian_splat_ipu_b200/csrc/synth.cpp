// synth.cpp — seeded synthetic scenes, host only.
// xoshiro256** seeded by splitmix64; uniforms from the top 24 bits; normals by Box-Muller in f64.
// Statistics follow the shipped PLYs: position sigma (1.4, 0.8, 1.0) clipped to +-6, per-axis
// log-scale N(-5, 1) clipped to [-13, -0.5] (anisotropic), random unit quaternion, rgb U[0,1),
// opacity U[0.05, 1).  Output records are the reference's Gaussian3D
// (include/splat/ipu_geometry.hpp:305-311) with the app's recentre + z-negate applied
// (src/main/splat.cpp:93-100) and gid = i + 1 (:161).
#include "../../include/gsplat_b200.h"
#include <cmath>
#include <cstdint>

namespace {
struct Rng {
  uint64_t s[4];
  explicit Rng(uint64_t seed) {
    for (auto &v : s) {   // splitmix64
      seed += 0x9E3779B97F4A7C15ull;
      uint64_t z = seed;
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
      z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
      v = z ^ (z >> 31);
    }
  }
  static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  uint64_t next() {   // xoshiro256**
    const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
    s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  float uniform() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }   // [0,1), 24 bits
  bool have = false;
  double spare = 0;
  float normal() {
    if (have) { have = false; return (float)spare; }
    double u1 = ((double)(next() >> 40) + 1.0) * (1.0 / 16777216.0);   // (0,1]
    double u2 = (double)(next() >> 40) * (1.0 / 16777216.0);
    double r = std::sqrt(-2.0 * std::log(u1)), a = 6.283185307179586476925 * u2;
    spare = r * std::sin(a); have = true;
    return (float)(r * std::cos(a));
  }
};
inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
}  // namespace

extern "C" int gsplat_synth_scene(uint64_t seed, size_t n, float *g) {
  if (!g && n) return GSPLAT_E_ARG;
  Rng rng(seed);
  float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
  const float sig[3] = {1.4f, 0.8f, 1.0f};
  for (size_t i = 0; i < n; ++i) {
    float *r = g + i * 16;
    for (int k = 0; k < 3; ++k) {
      r[k] = clampf(rng.normal() * sig[k], -6.f, 6.f);
      mn[k] = r[k] < mn[k] ? r[k] : mn[k];
      mx[k] = r[k] > mx[k] ? r[k] : mx[k];
    }
    r[3] = 1.0f;
    for (int k = 0; k < 3; ++k) r[4 + k] = rng.uniform();
    r[7] = 0.05f + 0.95f * rng.uniform();
    float q[4], len2 = 0.f;
    for (int k = 0; k < 4; ++k) { q[k] = rng.normal(); len2 += q[k] * q[k]; }
    const float inv = len2 > 0.f ? 1.0f / std::sqrt(len2) : 0.f;
    for (int k = 0; k < 4; ++k) r[8 + k] = len2 > 0.f ? q[k] * inv : (k == 0 ? 1.f : 0.f);
    for (int k = 0; k < 3; ++k) r[12 + k] = clampf(-5.0f + rng.normal(), -13.f, -0.5f);
    r[15] = (float)(i + 1);
  }
  // src/main/splat.cpp:93-100: translate the bbox centroid to the origin, then negate z
  float c[3];
  for (int k = 0; k < 3; ++k) c[k] = (mx[k] + mn[k]) * 0.5f;
  for (size_t i = 0; i < n; ++i) {
    float *r = g + i * 16;
    r[0] -= c[0]; r[1] -= c[1]; r[2] -= c[2];
    r[2] = -r[2];
  }
  return GSPLAT_OK;
}
