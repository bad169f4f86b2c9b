// phantom.hpp — deterministic synthetic T1+C-like brain slices.
//
// The reference processes the TCIA Brain-Tumor-Progression set (README.md:98-100), which cannot be
// fetched here; every benchmark and test uses these phantoms instead ("data": "synthetic").
// 12-bit range; background N(40,15) clipped at 0; head ellipse (axes .42w, .48h) tissue N(1000,60);
// skull ring 600; 1-3 enhancing blobs N(1600,120) near the centre, i.e. inside the region-growing raw
// window 1200..2050.  kind = lowcontrast: tissue N(1500,150) over the whole ellipse, so the grown
// region covers more than half of the slice (BASELINE config 5).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace nm03 {

struct Pcg32 {
  uint64_t state, inc;
  explicit Pcg32(uint64_t seed, uint64_t seq = 0x4E4D3033u) : state(0), inc((seq << 1) | 1u) {
    next();
    state += seed;
    next();
  }
  uint32_t next() {
    const uint64_t old = state;
    state = old * 6364136223846793005ULL + inc;
    const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((32u - rot) & 31u));
  }
  float uniform() { return (next() >> 8) * (1.0f / 16777216.0f); }
  float normal(float mean, float sd) {  // Box-Muller, one sample per call
    const float u1 = uniform() + 1e-7f, u2 = uniform();
    return mean + sd * std::sqrt(-2.0f * std::log(u1)) * std::cos(6.2831853f * u2);
  }
};

enum PhantomKind { PHANTOM_DEFAULT = 0, PHANTOM_LOWCONTRAST = 1 };

inline std::vector<int16_t> phantom(int w, int h, int patient, int slice, PhantomKind kind = PHANTOM_DEFAULT) {
  Pcg32 rng(0x4E4D3033u ^ ((uint64_t)patient << 16) ^ (uint64_t)slice);
  std::vector<int16_t> img((size_t)w * h);
  const float cx = w / 2.0f, cy = h / 2.0f;
  const int nblob = 1 + (patient + slice) % 3;
  float bx[3], by[3], br[3];
  for (int b = 0; b < nblob; ++b) {
    bx[b] = cx + (b - (nblob - 1) / 2.0f) * 0.14f * w + (rng.uniform() * 0.04f - 0.02f) * w;
    by[b] = cy + (rng.uniform() * 0.12f - 0.06f) * h;
    br[b] = (0.06f + 0.02f * ((slice % 8) / 8.0f)) * (w < h ? w : h) * (1.0f + 0.3f * b);
  }
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const float ex = (x - cx) / (0.42f * w), ey = (y - cy) / (0.48f * h);
      const float e = ex * ex + ey * ey;
      float v = rng.normal(40.0f, 15.0f);
      if (v < 0) v = 0;
      if (e <= 1.0f) {
        if (kind == PHANTOM_LOWCONTRAST) {
          v = rng.normal(1500.0f, 150.0f);
        } else {
          v = e > 0.88f ? rng.normal(600.0f, 30.0f) : rng.normal(1000.0f, 60.0f);
          for (int b = 0; b < nblob; ++b)
            if ((x - bx[b]) * (x - bx[b]) + (y - by[b]) * (y - by[b]) <= br[b] * br[b]) v = rng.normal(1600.0f, 120.0f);
        }
      }
      const float r = std::nearbyint(v);
      img[(size_t)y * w + x] = (int16_t)(r < 0 ? 0 : (r > 4095 ? 4095 : r));
    }
  return img;
}

}  // namespace nm03
