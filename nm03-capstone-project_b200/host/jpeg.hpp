// jpeg.hpp — baseline sequential JPEG writer (ITU-T T.81) for 8-bit greyscale or RGB images.
//
// Host-side replacement for FAST's ImageFileExporter("*.jpg") (a Qt image writer inside FAST) that the
// reference calls twice per slice: src/sequential/main_sequential.cpp:61-64,70-73,
// src/parallel/main_parallel.cpp:192-195,207-210, src/test/test_pipeline.cpp:22-25.  It stays on the
// host and is part of the end-to-end figure only (BASELINE.json north_star).  Huffman and
// quantisation tables are the example tables of T.81 Annex K; 4:4:4, no restart markers.
#pragma once
#include <cmath>
#if defined(__SSE2__)
#include <emmintrin.h>
#include <xmmintrin.h>
#endif
#include <cstdint>
#include <cstdio>
#include <cstring>
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
static const uint8_t kUnzigzag[64] = {0, 1, 5, 6, 14, 15, 27, 28, 2, 4, 7, 13, 16, 26, 29, 42, 3, 8, 12, 17, 25, 30, 41, 43, 9, 11, 18, 24, 31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60, 21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};  // natural index -> zigzag position
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

// Entropy-coded segment writer: bits collect in a 64-bit accumulator and leave eight bytes at a time, when it is full;
// eight bytes without any 0xFF among them (nearly all) are stored whole, the others byte by byte with the 0x00 stuffing
// byte of T.81 B.1.1.5.  The caller keeps at least 32 bytes of room at `p`.
struct BitWriter {
  uint8_t* p;
  uint64_t acc = 0;
  int free_bits = 64;   // unused low-order room in acc
  explicit BitWriter(uint8_t* dst) : p(dst) {}
  inline void emit8(uint64_t v) {   // eight bytes, most significant first
    if (((v & ~(v + 0x0101010101010101ull)) & 0x8080808080808080ull) == 0) {   // no byte of v is 0xFF
      const uint64_t be = __builtin_bswap64(v);
      std::memcpy(p, &be, 8);
      p += 8;
    } else {
      for (int sh = 56; sh >= 0; sh -= 8) {
        const uint8_t b = (uint8_t)(v >> sh);
        *p++ = b;
        if (b == 0xFF) *p++ = 0;
      }
    }
  }
  inline void put(uint32_t bits, int len) {  // bits < 2^len, 0 < len <= 32
    if (len <= free_bits) {
      acc = (acc << len) | bits;   // (a shift by 64 cannot occur: len <= 32)
      free_bits -= len;
    } else {                       // fill the accumulator with the leading bits, store it, keep the rest
      const int rest = len - free_bits;
      emit8(free_bits ? ((acc << free_bits) | (bits >> rest)) : acc);
      acc = bits;                  // only the low `rest` bits count: the ones above them leave at the top before a store
      free_bits = 64 - rest;
    }
  }
  inline void flush() {  // whole bytes, then the last one padded with 1-bits
    int nbits = 64 - free_bits;
    const uint64_t v = nbits ? (acc << free_bits) : 0;   // left-aligned
    int sh = 56;
    for (; nbits >= 8; nbits -= 8, sh -= 8) {
      const uint8_t b = (uint8_t)(v >> sh);
      *p++ = b;
      if (b == 0xFF) *p++ = 0;
    }
    if (nbits > 0) {
      const uint8_t b = (uint8_t)((v >> sh) | (0xFFu >> nbits));
      *p++ = b;
      if (b == 0xFF) *p++ = 0;
    }
    acc = 0;
    free_bits = 64;
  }
};

// 8-point forward DCT after Arai, Agui and Nakajima: 5 multiplications and 29 additions per
// transform; the outputs are scaled by kAan[u] (times 8 for the 2-D transform), which the caller folds
// into its reciprocal quantisation table.  One template serves the scalar path (T = float) and the
// SSE2 path (T = four transforms side by side); both perform the same operations in the same order,
// so their results are identical.
static const float kAan[8] = {1.0f, 1.387039845f, 1.306562965f, 1.175875602f, 1.0f, 0.785694958f, 0.541196100f, 0.275899379f};

template <typename T, typename C>
inline void aan8(T& d0, T& d1, T& d2, T& d3, T& d4, T& d5, T& d6, T& d7, C c) {
  const T t0 = d0 + d7, t7 = d0 - d7, t1 = d1 + d6, t6 = d1 - d6;
  const T t2 = d2 + d5, t5 = d2 - d5, t3 = d3 + d4, t4 = d3 - d4;
  const T e0 = t0 + t3, e3 = t0 - t3, e1 = t1 + t2, e2 = t1 - t2;
  d0 = e0 + e1;
  d4 = e0 - e1;
  const T z1 = (e2 + e3) * c(0.707106781f);
  d2 = e3 + z1;
  d6 = e3 - z1;
  const T o0 = t4 + t5, o1 = t5 + t6, o2 = t6 + t7;
  const T z5 = (o0 - o2) * c(0.382683433f);
  const T z2 = c(0.541196100f) * o0 + z5, z4 = c(1.306562965f) * o2 + z5, z3 = o1 * c(0.707106781f);
  const T z11 = t7 + z3, z13 = t7 - z3;
  d5 = z13 + z2;
  d3 = z13 - z2;
  d1 = z11 + z4;
  d7 = z11 - z4;
}

inline void fdct8x8_scaled(float* d) {  // rows, then columns
  auto c = [](float x) { return x; };
  for (int y = 0; y < 8; ++y) { float* r = d + 8 * y; aan8(r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], c); }
  for (int x = 0; x < 8; ++x) { float* q = d + x; aan8(q[0], q[8], q[16], q[24], q[32], q[40], q[48], q[56], c); }
}

// level shift, forward DCT and quantisation of one block -> 64 integers in natural (row-major) order
inline void quantise_scalar(float* blk, const float* rq, int* out) {
  fdct8x8_scaled(blk);
  for (int i = 0; i < 64; ++i) {
    const float qv = blk[i] * rq[i];
    out[i] = (int)(qv + (qv >= 0.0f ? 0.5f : -0.5f));
  }
}

#if defined(__SSE2__)
struct V4 {
  __m128 v;
  V4() = default;
  V4(__m128 x) : v(x) {}
  friend inline V4 operator+(V4 a, V4 b) { return _mm_add_ps(a.v, b.v); }
  friend inline V4 operator-(V4 a, V4 b) { return _mm_sub_ps(a.v, b.v); }
  friend inline V4 operator*(V4 a, V4 b) { return _mm_mul_ps(a.v, b.v); }
};

// Same block transform with four 1-D transforms per instruction.  src: 8 rows of 8 bytes (stride in
// bytes).  The row pass runs on transposed 4x4 tiles (lane = row), the column pass on the tiles
// transposed back (lane = column), so the result lands in natural order.
inline void quantise_sse2(const uint8_t* src, size_t stride, const float* rq, int* out) {
  auto c = [](float x) { return V4(_mm_set1_ps(x)); };
  const __m128i zero = _mm_setzero_si128();
  const __m128 bias = _mm_set1_ps(128.0f);
  __m128 r[8][2];
  for (int y = 0; y < 8; ++y) {
    const __m128i b8 = _mm_loadl_epi64(reinterpret_cast<const __m128i*>(src + y * stride));
    const __m128i w16 = _mm_unpacklo_epi8(b8, zero);
    r[y][0] = _mm_sub_ps(_mm_cvtepi32_ps(_mm_unpacklo_epi16(w16, zero)), bias);
    r[y][1] = _mm_sub_ps(_mm_cvtepi32_ps(_mm_unpackhi_epi16(w16, zero)), bias);
  }
  V4 col[2][8];  // col[g][x]: column x, lanes = rows 4g..4g+3
  for (int g = 0; g < 2; ++g)
    for (int xh = 0; xh < 2; ++xh) {
      __m128 a = r[4 * g][xh], b = r[4 * g + 1][xh], cc = r[4 * g + 2][xh], d = r[4 * g + 3][xh];
      _MM_TRANSPOSE4_PS(a, b, cc, d);
      col[g][4 * xh] = a; col[g][4 * xh + 1] = b; col[g][4 * xh + 2] = cc; col[g][4 * xh + 3] = d;
    }
  for (int g = 0; g < 2; ++g)
    aan8(col[g][0], col[g][1], col[g][2], col[g][3], col[g][4], col[g][5], col[g][6], col[g][7], c);
  V4 row[2][8];  // row[uh][y]: row y, lanes = horizontal frequencies 4uh..4uh+3
  for (int g = 0; g < 2; ++g)
    for (int uh = 0; uh < 2; ++uh) {
      __m128 a = col[g][4 * uh].v, b = col[g][4 * uh + 1].v, cc = col[g][4 * uh + 2].v, d = col[g][4 * uh + 3].v;
      _MM_TRANSPOSE4_PS(a, b, cc, d);
      row[uh][4 * g] = a; row[uh][4 * g + 1] = b; row[uh][4 * g + 2] = cc; row[uh][4 * g + 3] = d;
    }
  for (int uh = 0; uh < 2; ++uh)
    aan8(row[uh][0], row[uh][1], row[uh][2], row[uh][3], row[uh][4], row[uh][5], row[uh][6], row[uh][7], c);
  const __m128 half = _mm_set1_ps(0.5f), sign = _mm_castsi128_ps(_mm_set1_epi32((int)0x80000000u));
  for (int v = 0; v < 8; ++v)
    for (int uh = 0; uh < 2; ++uh) {
      const __m128 qv = _mm_mul_ps(row[uh][v].v, _mm_loadu_ps(rq + 8 * v + 4 * uh));
      const __m128 h = _mm_or_ps(half, _mm_and_ps(qv, sign));   // +-0.5 with the sign of qv: round half away from zero
      _mm_storeu_si128(reinterpret_cast<__m128i*>(out + 8 * v + 4 * uh), _mm_cvttps_epi32(_mm_add_ps(qv, h)));
    }
}
#endif

inline void marker(std::vector<uint8_t>& o, uint8_t m, const std::vector<uint8_t>& payload) {
  o.push_back(0xFF); o.push_back(m);
  const size_t n = payload.size() + 2;
  o.push_back((uint8_t)(n >> 8)); o.push_back((uint8_t)n);
  o.insert(o.end(), payload.begin(), payload.end());
}

}  // namespace detail

// Quantisation tables of a quality setting: q[t] = the table as stored in the file (t = 0 luma, 1 chroma; natural
// order), rq[t] = reciprocal quantiser including the scaling of the AAN transform's outputs.
struct Tables {
  uint8_t q[2][64];
  float rq[2][64];
};
inline Tables make_tables(int quality) {
  using namespace detail;
  Tables t;
  quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int i = 0; i < 64; ++i) {
    int a = (kQLuma[i] * scale + 50) / 100, b = (kQChroma[i] * scale + 50) / 100;
    t.q[0][i] = (uint8_t)(a < 1 ? 1 : (a > 255 ? 255 : a));
    t.q[1][i] = (uint8_t)(b < 1 ? 1 : (b > 255 ? 255 : b));
  }
  for (int i = 0; i < 64; ++i) {
    const float sc = 1.0f / (kAan[i >> 3] * kAan[i & 7] * 8.0f);
    t.rq[0][i] = sc / t.q[0][i];
    t.rq[1][i] = sc / t.q[1][i];
  }
  return t;
}

namespace detail {

inline int bitsize(int v) { const unsigned a = (unsigned)(v < 0 ? -v : v); return a ? 32 - __builtin_clz(a) : 0; }

struct HuffSet {
  Huff dc[2], ac[2];
  // luma AC symbols with their value bits already appended, for the small values nearly every coefficient has:
  // fast[run][v + 128] = (bits << 8) | length for -127 <= v <= 127, v != 0 (length <= 16 + 7)
  uint32_t fast[16][256];
  HuffSet() {
    dc[0].build(kDcLumaBits, kDcVals); dc[1].build(kDcChromaBits, kDcVals);
    ac[0].build(kAcLumaBits, kAcLumaVals); ac[1].build(kAcChromaBits, kAcChromaVals);
    for (int run = 0; run < 16; ++run)
      for (int v = -127; v <= 127; ++v) {
        if (v == 0) { fast[run][128] = 0; continue; }
        const int sz = bitsize(v), rs = (run << 4) | sz;
        const uint32_t vb = (uint32_t)(v < 0 ? v - 1 : v) & ((1u << sz) - 1u);
        fast[run][v + 128] = ((((uint32_t)ac[0].code[rs] << sz) | vb) << 8) | (uint32_t)(ac[0].len[rs] + sz);
      }
    for (int run = 0; run < 16; ++run) fast[run][0] = 0;   // (-128 is not tabulated)
  }
};
inline const HuffSet& huffman() {   // built once; initialisation of a function-local static is thread-safe
  static const HuffSet hs;
  return hs;
}

// SOI, APP0, DQT, SOF0, DHT and SOS of a baseline file
inline void write_headers(std::vector<uint8_t>& o, int width, int height, int channels, const uint8_t (*q)[64]) {
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
}

// One block's symbols: the DC difference, then (run, size) symbols of the AC terms in zigzag order; code and value bits go
// out together.  zz = the 64 quantised coefficients in zigzag order, nz = bit k set iff zz[k] != 0 (bit 0 is ignored).
inline void put_block(BitWriter& bw, const Huff& hdc, const Huff& hac, const int16_t* zz, uint64_t nz, int& pred,
                      const uint32_t (*fast)[256] = nullptr) {
  const int diff = zz[0] - pred;
  pred = zz[0];
  {
    const int sz = bitsize(diff);
    const uint32_t vb = (uint32_t)(diff < 0 ? diff - 1 : diff) & ((1u << sz) - 1u);
    bw.put(((uint32_t)hdc.code[sz] << sz) | vb, hdc.len[sz] + sz);
  }
  nz &= ~(uint64_t)1;
  int prev = 0;
  while (nz) {   // the map of the non-zero terms lets the loop jump from one to the next
    const int i = __builtin_ctzll(nz);
    nz &= nz - 1;
    int run = i - prev - 1;
    prev = i;
    while (run > 15) { bw.put(hac.code[0xF0], hac.len[0xF0]); run -= 16; }
    const int v = zz[i];
    if (fast && (unsigned)(v + 127) <= 254u) {   // one look-up: code and value bits together
      const uint32_t e = fast[run][v + 128];
      bw.put(e >> 8, (int)(e & 0xFFu));
      continue;
    }
    const int sz = bitsize(v), rs = (run << 4) | sz;
    const uint32_t vb = (uint32_t)(v < 0 ? v - 1 : v) & ((1u << sz) - 1u);
    bw.put(((uint32_t)hac.code[rs] << sz) | vb, hac.len[rs] + sz);
  }
  if (prev != 63) bw.put(hac.code[0x00], hac.len[0x00]);  // EOB
}

constexpr size_t kWorstBlock = 512;   // 64 coefficients x 26 bits, every byte stuffed, rounded up

}  // namespace detail

// A grey image whose blocks were transformed and quantised elsewhere (nm03_outputs.export_blocks: the GPU does it next to
// the pipeline): zz = [blocks][64] quantised coefficients in zigzag order, blocks in raster order, nz = [blocks] maps of
// the non-zero terms.  width and height are multiples of 8.  Produces the bytes encode() produces for the same image.
inline std::vector<uint8_t> encode_blocks(const int16_t* zz, const uint64_t* nz, int width, int height, int quality = 90) {
  using namespace detail;
  if (width <= 0 || height <= 0 || (width & 7) || (height & 7)) throw std::runtime_error("jpeg: block input needs multiples of 8");
  const Tables tb = make_tables(quality);
  const HuffSet& hf = huffman();
  std::vector<uint8_t> o;
  o.reserve((size_t)width * height / 4 + 1024);
  write_headers(o, width, height, 1, tb.q);
  const int blocks_per_row = width / 8;
  size_t pos = o.size();
  int pred = 0;
  BitWriter bw(nullptr);
  size_t b = 0;
  for (int by = 0; by < height; by += 8) {
    if (o.size() - pos < (size_t)blocks_per_row * kWorstBlock + 64) o.resize(pos + (size_t)blocks_per_row * kWorstBlock * 2 + 64);
    bw.p = o.data() + pos;
    for (int bx = 0; bx < blocks_per_row; ++bx, ++b) put_block(bw, hf.dc[0], hf.ac[0], zz + 64 * b, nz[b], pred, hf.fast);
    pos = (size_t)(bw.p - o.data());
  }
  if (o.size() - pos < 32) o.resize(pos + 32);
  bw.p = o.data() + pos;
  bw.flush();
  pos = (size_t)(bw.p - o.data());
  o.resize(pos);
  o.push_back(0xFF); o.push_back(0xD9);  // EOI
  return o;
}

// level shift, transform, quantisation and zigzag of one 8x8 block of a grey image, as encode() does them: the CPU
// statement of what nm03_outputs.export_blocks holds (tests compare the two)
inline void block_coefficients(const uint8_t* px, size_t stride, const float* rq, int16_t* zz, uint64_t* nz) {
  using namespace detail;
  int q[64];
#if defined(__SSE2__)
  quantise_sse2(px, stride, rq, q);
#else
  float blk[64];
  for (int y = 0; y < 8; ++y) for (int x = 0; x < 8; ++x) blk[y * 8 + x] = (float)px[(size_t)y * stride + x] - 128.0f;
  quantise_scalar(blk, rq, q);
#endif
  uint64_t m = 0;
  for (int i = 0; i < 64; ++i) { zz[kUnzigzag[i]] = (int16_t)q[i]; }
#if defined(__SSE2__)
  for (int k = 0; k < 4; ++k) {
    const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(zz + 16 * k));
    const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(zz + 16 * k + 8));
    const int zero16 = _mm_movemask_epi8(_mm_cmpeq_epi8(_mm_packs_epi16(a, b), _mm_setzero_si128()));
    m |= (uint64_t)(uint16_t)~zero16 << (16 * k);
  }
#else
  for (int i = 0; i < 64; ++i) m |= (uint64_t)(zz[i] != 0) << i;
#endif
  *nz = m;
}

// pixels: height*width*channels bytes, channels = 1 (grey) or 3 (RGB interleaved); quality 1..100.
inline std::vector<uint8_t> encode(const uint8_t* pixels, int width, int height, int channels, int quality = 90) {
  using namespace detail;
  if (width <= 0 || height <= 0 || (channels != 1 && channels != 3)) throw std::runtime_error("jpeg: bad image");
  const Tables tb = make_tables(quality);
  const float (*rq)[64] = tb.rq;
  const HuffSet& hf = huffman();
  const Huff* hdc = hf.dc;
  const Huff* hac = hf.ac;
  std::vector<uint8_t> o;
  o.reserve((size_t)width * height / 4 + 1024);
  write_headers(o, width, height, channels, tb.q);
  // entropy-coded segment, written straight into the tail of `o` (grown by block row)
  const int blocks_per_row = ((width + 7) / 8) * channels;
  size_t pos = o.size();
  int pred[3] = {0, 0, 0};
  BitWriter bw(nullptr);
  for (int by = 0; by < height; by += 8) {
    if (o.size() - pos < (size_t)blocks_per_row * kWorstBlock + 64) o.resize(pos + (size_t)blocks_per_row * kWorstBlock * 2 + 64);
    bw.p = o.data() + pos;
    for (int bx = 0; bx < width; bx += 8)
      for (int c = 0; c < channels; ++c) {
        const int t = c ? 1 : 0;
        int q[64];                        // quantised coefficients, natural order
        bool dc_only = false;
        if (channels == 1 && by + 8 <= height && bx + 8 <= width) {
          const uint8_t* px = pixels + (size_t)by * width + bx;
          // constant block (most of a rendered mask): only the DC term survives
          uint64_t first, rowv;
          std::memcpy(&first, px, 8);
          bool flat = first == (first & 0xFFu) * 0x0101010101010101ull;
          for (int y = 1; y < 8 && flat; ++y) { std::memcpy(&rowv, px + (size_t)y * width, 8); flat = rowv == first; }
          if (flat) {
            const float qv = ((float)px[0] - 128.0f) * 64.0f * rq[t][0];   // the scaled transform's DC output is the plain sum
            q[0] = (int)(qv + (qv >= 0.0f ? 0.5f : -0.5f));
            dc_only = true;
          } else {
#if defined(__SSE2__)
            quantise_sse2(px, (size_t)width, rq[t], q);
#else
            float blk[64];
            for (int y = 0; y < 8; ++y) for (int x = 0; x < 8; ++x) blk[y * 8 + x] = (float)px[(size_t)y * width + x] - 128.0f;
            quantise_scalar(blk, rq[t], q);
#endif
          }
        } else {
          float blk[64];
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
          bool flat = true;
          for (int i = 1; i < 64 && flat; ++i) flat = blk[i] == blk[0];
          if (flat) {
            const float qv = blk[0] * 64.0f * rq[t][0];
            q[0] = (int)(qv + (qv >= 0.0f ? 0.5f : -0.5f));
            dc_only = true;
          } else {
            quantise_scalar(blk, rq[t], q);
          }
        }
        // DC difference, then (run, size) symbols of the AC terms in zigzag order; code and value bits go out together
        const int diff = q[0] - pred[c];
        pred[c] = q[0];
        {
          const int sz = bitsize(diff);
          const uint32_t vb = (uint32_t)(diff < 0 ? diff - 1 : diff) & ((1u << sz) - 1u);
          bw.put(((uint32_t)hdc[t].code[sz] << sz) | vb, hdc[t].len[sz] + sz);
        }
        if (dc_only) { bw.put(hac[t].code[0x00], hac[t].len[0x00]); continue; }
        // AC terms in zigzag order; a 64-bit map of the non-zero ones lets the loop jump from one to the next
        alignas(16) int16_t zz[64];
        for (int i = 0; i < 64; ++i) zz[kUnzigzag[i]] = (int16_t)q[i];
        uint64_t nz = 0;
#if defined(__SSE2__)
        for (int k = 0; k < 4; ++k) {
          const __m128i a = _mm_load_si128(reinterpret_cast<const __m128i*>(zz + 16 * k));
          const __m128i b = _mm_load_si128(reinterpret_cast<const __m128i*>(zz + 16 * k + 8));
          const int zero16 = _mm_movemask_epi8(_mm_cmpeq_epi8(_mm_packs_epi16(a, b), _mm_setzero_si128()));
          nz |= (uint64_t)(uint16_t)~zero16 << (16 * k);
        }
#else
        for (int i = 0; i < 64; ++i) nz |= (uint64_t)(zz[i] != 0) << i;
#endif
        nz &= ~(uint64_t)1;
        int prev = 0;
        while (nz) {
          const int i = __builtin_ctzll(nz);
          nz &= nz - 1;
          int run = i - prev - 1;
          prev = i;
          while (run > 15) { bw.put(hac[t].code[0xF0], hac[t].len[0xF0]); run -= 16; }
          const int v = zz[i];
          const int sz = bitsize(v), rs = (run << 4) | sz;
          const uint32_t vb = (uint32_t)(v < 0 ? v - 1 : v) & ((1u << sz) - 1u);
          bw.put(((uint32_t)hac[t].code[rs] << sz) | vb, hac[t].len[rs] + sz);
        }
        if (prev != 63) bw.put(hac[t].code[0x00], hac[t].len[0x00]);  // EOB
      }
    pos = (size_t)(bw.p - o.data());
  }
  if (o.size() - pos < 32) o.resize(pos + 32);
  bw.p = o.data() + pos;
  bw.flush();
  pos = (size_t)(bw.p - o.data());
  o.resize(pos);
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
