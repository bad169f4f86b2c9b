// jpeg.hpp — baseline sequential JPEG writer (ITU-T T.81) for 8-bit greyscale or RGB images.
//
// Host-side replacement for FAST's ImageFileExporter("*.jpg") (a Qt image writer inside FAST) that the
// reference calls twice per slice: src/sequential/main_sequential.cpp:61-64,70-73,
// src/parallel/main_parallel.cpp:192-195,207-210, src/test/test_pipeline.cpp:22-25.  It stays on the
// host and is part of the end-to-end figure only (BASELINE.json north_star).  Huffman and
// quantisation tables are the example tables of T.81 Annex K; 4:4:4, no restart markers.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

namespace nm03 {
namespace jpeg {

namespace detail {

static const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                                    15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
                                    62, 63};
static const uint8_t kQLuma[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
                                   69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55,
                                   64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100,
                                   103, 99};
static const uint8_t kQChroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                                     99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                     99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
// Annex K.3 Huffman specifications: BITS (16 counts) + HUFFVAL
static const uint8_t kDcLumaBits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t kDcChromaBits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t kDcVals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t kAcLumaBits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t kAcLumaVals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32,
    0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16,
    0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45,
    0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
    0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94,
    0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6,
    0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8,
    0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};
static const uint8_t kAcChromaBits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t kAcChromaVals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81,
    0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34,
    0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44,
    0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
    0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92,
    0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4,
    0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6,
    0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};

struct Huff {
  uint16_t code[256];
  uint8_t len[256];
  void build(const uint8_t* bits, const uint8_t* vals) {
    for (int i = 0; i < 256; ++i) { code[i] = 0; len[i] = 0; }
    int k = 0;
    uint16_t c = 0;
    for (int l = 1; l <= 16; ++l) {
      for (int i = 0; i < bits[l - 1]; ++i) { code[vals[k]] = c++; len[vals[k]] = (uint8_t)l; ++k; }
      c <<= 1;
    }
  }
};

struct BitWriter {
  std::vector<uint8_t>& out;
  uint32_t acc = 0;
  int nbits = 0;
  explicit BitWriter(std::vector<uint8_t>& o) : out(o) {}
  void put(uint32_t code, int len) {
    acc = (acc << len) | (code & ((1u << len) - 1u));
    nbits += len;
    while (nbits >= 8) {
      const uint8_t b = (uint8_t)(acc >> (nbits - 8));
      out.push_back(b);
      if (b == 0xFF) out.push_back(0);  // byte stuffing
      nbits -= 8;
    }
  }
  void flush() { if (nbits > 0) put(0x7F, 8 - nbits); }
};

// 8-point forward DCT after Arai, Agui and Nakajima: 5 multiplications and 29 additions per
// transform; the outputs are scaled by kAan[u] (times 8 for the 2-D transform), which the caller folds
// into its reciprocal quantisation table.
static const float kAan[8] = {1.0f, 1.387039845f, 1.306562965f, 1.175875602f, 1.0f, 0.785694958f, 0.541196100f, 0.275899379f};

inline void aan8(float* d, int stride) {
  const float t0 = d[0] + d[7 * stride], t7 = d[0] - d[7 * stride];
  const float t1 = d[stride] + d[6 * stride], t6 = d[stride] - d[6 * stride];
  const float t2 = d[2 * stride] + d[5 * stride], t5 = d[2 * stride] - d[5 * stride];
  const float t3 = d[3 * stride] + d[4 * stride], t4 = d[3 * stride] - d[4 * stride];
  const float e0 = t0 + t3, e3 = t0 - t3, e1 = t1 + t2, e2 = t1 - t2;
  d[0] = e0 + e1;
  d[4 * stride] = e0 - e1;
  const float z1 = (e2 + e3) * 0.707106781f;
  d[2 * stride] = e3 + z1;
  d[6 * stride] = e3 - z1;
  const float o0 = t4 + t5, o1 = t5 + t6, o2 = t6 + t7;
  const float z5 = (o0 - o2) * 0.382683433f;
  const float z2 = 0.541196100f * o0 + z5, z4 = 1.306562965f * o2 + z5, z3 = o1 * 0.707106781f;
  const float z11 = t7 + z3, z13 = t7 - z3;
  d[5 * stride] = z13 + z2;
  d[3 * stride] = z13 - z2;
  d[stride] = z11 + z4;
  d[7 * stride] = z11 - z4;
}

inline void fdct8x8_scaled(float* d) {
  for (int y = 0; y < 8; ++y) aan8(d + 8 * y, 1);
  for (int x = 0; x < 8; ++x) aan8(d + x, 8);
}

inline void marker(std::vector<uint8_t>& o, uint8_t m, const std::vector<uint8_t>& payload) {
  o.push_back(0xFF); o.push_back(m);
  const size_t n = payload.size() + 2;
  o.push_back((uint8_t)(n >> 8)); o.push_back((uint8_t)n);
  o.insert(o.end(), payload.begin(), payload.end());
}

}  // namespace detail

// pixels: height*width*channels bytes, channels = 1 (grey) or 3 (RGB interleaved); quality 1..100.
inline std::vector<uint8_t> encode(const uint8_t* pixels, int width, int height, int channels, int quality = 90) {
  using namespace detail;
  if (width <= 0 || height <= 0 || (channels != 1 && channels != 3)) throw std::runtime_error("jpeg: bad image");
  quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  uint8_t q[2][64];
  for (int i = 0; i < 64; ++i) {
    int a = (kQLuma[i] * scale + 50) / 100, b = (kQChroma[i] * scale + 50) / 100;
    q[0][i] = (uint8_t)(a < 1 ? 1 : (a > 255 ? 255 : a));
    q[1][i] = (uint8_t)(b < 1 ? 1 : (b > 255 ? 255 : b));
  }
  float rq[2][64];  // reciprocal quantiser including the DCT's output scaling
  for (int i = 0; i < 64; ++i) {
    const float sc = 1.0f / (kAan[i >> 3] * kAan[i & 7] * 8.0f);
    rq[0][i] = sc / q[0][i];
    rq[1][i] = sc / q[1][i];
  }
  static Huff hdc[2], hac[2];
  static bool built = false;
  if (!built) {
    hdc[0].build(kDcLumaBits, kDcVals); hdc[1].build(kDcChromaBits, kDcVals);
    hac[0].build(kAcLumaBits, kAcLumaVals); hac[1].build(kAcChromaBits, kAcChromaVals);
    built = true;
  }
  std::vector<uint8_t> o;
  o.reserve((size_t)width * height / 4 + 1024);
  o.push_back(0xFF); o.push_back(0xD8);                                       // SOI
  marker(o, 0xE0, {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0});          // APP0
  for (int t = 0; t < (channels == 3 ? 2 : 1); ++t) {                          // DQT
    std::vector<uint8_t> p(65);
    p[0] = (uint8_t)t;
    for (int i = 0; i < 64; ++i) p[1 + i] = q[t][kZigzag[i]];
    marker(o, 0xDB, p);
  }
  {                                                                            // SOF0
    std::vector<uint8_t> p = {8, (uint8_t)(height >> 8), (uint8_t)height, (uint8_t)(width >> 8), (uint8_t)width, (uint8_t)channels};
    for (int c = 0; c < channels; ++c) { p.push_back((uint8_t)(c + 1)); p.push_back(0x11); p.push_back((uint8_t)(c ? 1 : 0)); }
    marker(o, 0xC0, p);
  }
  auto dht = [&](int cls, int id, const uint8_t* bits, const uint8_t* vals, int nvals) {
    std::vector<uint8_t> p;
    p.push_back((uint8_t)((cls << 4) | id));
    p.insert(p.end(), bits, bits + 16);
    p.insert(p.end(), vals, vals + nvals);
    marker(o, 0xC4, p);
  };
  dht(0, 0, kDcLumaBits, kDcVals, 12);
  dht(1, 0, kAcLumaBits, kAcLumaVals, 162);
  if (channels == 3) { dht(0, 1, kDcChromaBits, kDcVals, 12); dht(1, 1, kAcChromaBits, kAcChromaVals, 162); }
  {                                                                            // SOS
    std::vector<uint8_t> p = {(uint8_t)channels};
    for (int c = 0; c < channels; ++c) { p.push_back((uint8_t)(c + 1)); p.push_back((uint8_t)(c ? 0x11 : 0x00)); }
    p.push_back(0); p.push_back(63); p.push_back(0);
    marker(o, 0xDA, p);
  }
  BitWriter bw(o);
  int pred[3] = {0, 0, 0};
  float blk[64];
  for (int by = 0; by < height; by += 8)
    for (int bx = 0; bx < width; bx += 8)
      for (int c = 0; c < channels; ++c) {
        if (channels == 1 && by + 8 <= height && bx + 8 <= width) {
          for (int y = 0; y < 8; ++y) {
            const uint8_t* px = pixels + (size_t)(by + y) * width + bx;
            for (int x = 0; x < 8; ++x) blk[y * 8 + x] = (float)px[x] - 128.0f;
          }
        } else
        for (int y = 0; y < 8; ++y)
          for (int x = 0; x < 8; ++x) {
            const int sy = by + y < height ? by + y : height - 1, sx = bx + x < width ? bx + x : width - 1;
            const uint8_t* px = pixels + ((size_t)sy * width + sx) * channels;
            float v;
            if (channels == 1) v = px[0];
            else {
              const float r = px[0], g = px[1], b = px[2];
              v = c == 0 ? 0.299f * r + 0.587f * g + 0.114f * b
                         : c == 1 ? -0.168736f * r - 0.331264f * g + 0.5f * b + 128.0f
                                  : 0.5f * r - 0.418688f * g - 0.081312f * b + 128.0f;
            }
            blk[y * 8 + x] = v - 128.0f;
          }
        const int t = c ? 1 : 0;
        int zz[64];
        bool flat = true;  // constant block (most of a rendered mask): only the DC term survives
        for (int i = 1; i < 64 && flat; ++i) flat = blk[i] == blk[0];
        if (flat) {
          const float qv = blk[0] * 64.0f * rq[t][0];   // the scaled transform's DC output is the plain sum
          zz[0] = (int)(qv + (qv >= 0.0f ? 0.5f : -0.5f));
          for (int i = 1; i < 64; ++i) zz[i] = 0;
        } else {
          fdct8x8_scaled(blk);
          for (int i = 0; i < 64; ++i) {
            const float qv = blk[kZigzag[i]] * rq[t][kZigzag[i]];
            zz[i] = (int)(qv + (qv >= 0.0f ? 0.5f : -0.5f));
          }
        }
        auto emit = [&](const Huff& h, int run_size, int value, int size) {
          bw.put(h.code[run_size], h.len[run_size]);
          if (size) bw.put((uint32_t)(value < 0 ? value - 1 : value), size);
        };
        auto bitsize = [](int v) { int a = v < 0 ? -v : v, s = 0; while (a) { ++s; a >>= 1; } return s; };
        const int diff = zz[0] - pred[c];
        pred[c] = zz[0];
        emit(hdc[t], bitsize(diff), diff, bitsize(diff));
        int run = 0;
        for (int i = 1; i < 64; ++i) {
          if (zz[i] == 0) { ++run; continue; }
          while (run > 15) { bw.put(hac[t].code[0xF0], hac[t].len[0xF0]); run -= 16; }
          const int s = bitsize(zz[i]);
          emit(hac[t], (run << 4) | s, zz[i], s);
          run = 0;
        }
        if (run) bw.put(hac[t].code[0x00], hac[t].len[0x00]);  // EOB
      }
  bw.flush();
  o.push_back(0xFF); o.push_back(0xD9);  // EOI
  return o;
}

inline void save(const std::string& path, const uint8_t* pixels, int width, int height, int channels, int quality = 90) {
  const std::vector<uint8_t> b = encode(pixels, width, height, channels, quality);
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) throw std::runtime_error("cannot write " + path);
  const size_t n = std::fwrite(b.data(), 1, b.size(), f);
  std::fclose(f);
  if (n != b.size()) throw std::runtime_error("short write " + path);
}

}  // namespace jpeg
}  // namespace nm03
