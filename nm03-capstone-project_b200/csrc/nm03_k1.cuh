// nm03_k1.cuh — K1, the fused preprocess kernel (templates).  Included by nm03_k1_w.cu, which is compiled once per
// median window (make builds the windows in parallel), and by nm03_kernels.cu for the few non-template helpers.
#pragma once
#include "nm03_device.h"

#include <cstdio>
#include <cstring>

#include <cooperative_groups.h>
#include <cuda.h>
#include <cuda_runtime.h>

namespace cg = cooperative_groups;

#define NM03_MIN(a, b) __vmins2((a), (b))
#define NM03_MAX(a, b) __vmaxs2((a), (b))
#define NM03_MIN3(a, b, c) __vimin3_s16x2((a), (b), (c))
#define NM03_MAX3(a, b, c) __vimax3_s16x2((a), (b), (c))
#include "median_net.inc"

namespace nm03 {

// =================================================================================================
// K1
// =================================================================================================
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// Raw sample as an order-preserving signed 16-bit key.  uint16 input is biased by 0x8000.  In
// zero-border mode samples outside the image become the sentinel -32768 (maps to 0.0f later) and
// real data is saturated at -32767 (the host has checked that this does not change f()).
template <bool ZERO>
__device__ __forceinline__ uint32_t fetch_key(const uint16_t* __restrict__ img, int w, int h, int x, int y,
                                              uint32_t bias) {
  if (ZERO) {
    if (x < 0 || y < 0 || x >= w || y >= h) return 0x8000u;
    uint32_t v = (img[(size_t)y * w + x] ^ bias) & 0xffffu;
    return v == 0x8000u ? 0x8001u : v;
  }
  return (img[(size_t)clampi(y, 0, h - 1) * w + clampi(x, 0, w - 1)] ^ bias) & 0xffffu;
}

// f(key) = clip(normalise(raw))  — oracle nm03o_normalize / nm03o_clip, same operation order.
__device__ __forceinline__ float key_to_float(int key, const PreParams& p) {
  if (p.border_zero && key == -32768) return 0.0f;
  const float v = p.input_unsigned ? (float)((key ^ 0x8000) & 0xffff) : (float)key;
  const float t = __fdiv_rn(__fsub_rn(v, p.norm_min), p.norm_range);
  const float o = __fadd_rn(__fmul_rn(t, p.norm_scale), p.norm_low);
  return fminf(fmaxf(o, p.clip_min), p.clip_max);
}

// lut[idx] through the read-only path; the address is one multiply-add (base + 4 * idx)
__device__ __forceinline__ float lut_at(const float* __restrict__ lut, uint32_t idx) {
  unsigned long long a;
  asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(a) : "r"(idx), "l"(lut));
  return __ldg(reinterpret_cast<const float*>(a));
}

// NW consecutive packed words of one tile row -> registers, with the widest shared-memory loads the alignment of a
// run allows (a run starts every RUN words: 16-byte loads when RUN % 4 == 0, 8-byte loads when RUN is even)
template <int NW, int RUN_, int K_>
__device__ __forceinline__ void load_window_row(const uint32_t* __restrict__ src, uint32_t (&v)[NW][K_], int r) {
  int c = 0;
  if constexpr (RUN_ % 4 == 0) {
#pragma unroll
    for (; c + 4 <= NW; c += 4) {
      const uint4 q = *reinterpret_cast<const uint4*>(src + c);
      v[c][r] = q.x; v[c + 1][r] = q.y; v[c + 2][r] = q.z; v[c + 3][r] = q.w;
    }
  }
  if constexpr (RUN_ % 2 == 0) {
#pragma unroll
    for (; c + 2 <= NW; c += 2) {
      const uint2 q = *reinterpret_cast<const uint2*>(src + c);
      v[c][r] = q.x; v[c + 1][r] = q.y;
    }
  }
#pragma unroll
  for (; c < NW; ++c) v[c][r] = src[c];
}

template <int K>
struct MedianRun;
template <int K>
struct MedianRunLen { static constexpr int value = 4; };   // windows without a generated network: runs of 4 outputs
#define NM03_DEFINE_MEDIAN_RUN(K_)                                                                              \
  template <>                                                                                                   \
  struct MedianRun<K_> {                                                                                        \
    static constexpr int RUN = NM03_MEDIAN_RUNLEN_##K_;                                                         \
    static __device__ __forceinline__ void run(uint32_t (&v)[RUN + K_ - 1][K_], uint32_t (&o)[RUN]) {           \
      NM03_MEDIAN_RUN_##K_(v, o);                                                                               \
    }                                                                                                           \
  };
NM03_DEFINE_MEDIAN_RUN(3)
NM03_DEFINE_MEDIAN_RUN(5)
NM03_DEFINE_MEDIAN_RUN(7)
NM03_DEFINE_MEDIAN_RUN(9)
#undef NM03_DEFINE_MEDIAN_RUN
// 1x1 "window": the median of one sample is the sample.  Second pass of the two-pass path for windows above 15x15
// (launch_median_big computes the window median of the raw samples first; median and f() commute).
template <>
struct MedianRun<1> {
  static constexpr int RUN = 8;
  static __device__ __forceinline__ void run(uint32_t (&v)[RUN][1], uint32_t (&o)[RUN]) {
#pragma unroll
    for (int i = 0; i < RUN; ++i) o[i] = v[i][0];
  }
};
template <> struct MedianRunLen<1> { static constexpr int value = MedianRun<1>::RUN; };
template <> struct MedianRunLen<3> { static constexpr int value = MedianRun<3>::RUN; };
template <> struct MedianRunLen<5> { static constexpr int value = MedianRun<5>::RUN; };
template <> struct MedianRunLen<7> { static constexpr int value = MedianRun<7>::RUN; };
template <> struct MedianRunLen<9> { static constexpr int value = MedianRun<9>::RUN; };

// Unsharp mask for 4 horizontally adjacent pixels, accumulating exactly in the oracle's order
// (tap x outer, tap y inner, from 0.0f; nm03o_sharpen).  M points at the median tile row of the
// pixels' centre minus HS rows, column 4*xq of the 12-column frame.
template <int HS, bool FMA>
__device__ __forceinline__ void blur4(const float* __restrict__ Mrow0, int mws, const PreParams& p, float (&acc)[4]) {
  constexpr int m = 2 * HS + 1;
#pragma unroll
  for (int j = 0; j < 4; ++j) acc[j] = 0.0f;
#pragma unroll
  for (int cgp = 0; cgp < 3; ++cgp) {
    float4 t[m];
#pragma unroll
    for (int yi = 0; yi < m; ++yi) t[yi] = *reinterpret_cast<const float4*>(Mrow0 + yi * mws + 4 * cgp);
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const int c = 4 * cgp + cc;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int xi = c - j - (K1_HS_MAX - HS);
        if (xi < 0 || xi >= m) continue;
#pragma unroll
        for (int yi = 0; yi < m; ++yi) {
          const float v = cc == 0 ? t[yi].x : cc == 1 ? t[yi].y : cc == 2 ? t[yi].z : t[yi].w;
          const float wgt = p.mask[xi + yi * m];
          if (FMA) acc[j] = __fmaf_rn(wgt, v, acc[j]);
          else acc[j] = __fadd_rn(acc[j], __fmul_rn(wgt, v));
        }
      }
    }
  }
}

// Exact blur of ONE pixel in the oracle's order (slow path of the fast mode; rare).
template <int HS>
__device__ __noinline__ float blur1_exact(const float* __restrict__ Mc, const PreParams& p) {
  constexpr int m = 2 * HS + 1;
  float acc = 0.0f;
#pragma unroll 1
  for (int xi = 0; xi < m; ++xi) {
#pragma unroll
    for (int yi = 0; yi < m; ++yi) {
      const float v = Mc[(yi - HS) * K1_MW + (xi - HS)];
      const float wgt = p.mask[xi + yi * m];
      if (p.use_fma) acc = __fmaf_rn(wgt, v, acc);
      else acc = __fadd_rn(acc, __fmul_rn(wgt, v));
    }
  }
  return acc;
}

// Separable blur of ONE pixel, times -gain, with exactly the operations (and order) of the straight-line fast path of
// phase C: per row a chain of FMAs over the taps from 0.0f, then a chain over the rows with the taps scaled by -gain.  Used off the hot path to redo the
// band test of a pixel with its own error bound.
template <int HE>
__device__ __noinline__ float blur1_fast(const float* __restrict__ Mc, const PreParams& p) {
  float blur = 0.0f;
#pragma unroll 1
  for (int t = -HE; t <= HE; ++t) {
    float a = 0.0f;
#pragma unroll
    for (int u = -HE; u <= HE; ++u) a = __fmaf_rn(p.gx[u + K1_HS_MAX], Mc[t * K1_MW + u], a);
    blur = __fmaf_rn(p.gyn[t + K1_HS_MAX], a, blur);   // taps scaled by -gain: the result is -gain * blur
  }
  return blur;
}

// ---- TMA helpers (cp.async.bulk.tensor + mbarrier, PTX ISA 8.x; SASS: UTMALDG / SYNCS) ------------
__device__ __forceinline__ uint32_t smem_u32(const void* ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

#ifdef NM03_TUNE_STAMPS   // tuning builds only: thread 0's clock per phase, summed over all work items of the process
static __device__ unsigned long long g_stamps[12];
#define NM03_STAMP(i) do { if (tid == 0) { const long long t_ = clock64(); atomicAdd(&g_stamps[i], (unsigned long long)(t_ - t_prev)); t_prev = t_; } } while (0)
#else
#define NM03_STAMP(i) do { } while (0)
#endif

// The small per-CTA state of K1.  It lives at the end of the dynamic shared memory block, so the kernel has NO static
// shared memory: the dynamic block then starts at the CTA's (1 KiB-aligned) shared window, the 128-byte alignment of the
// TMA destination needs no run-time padding, and every tile address is a link-time constant (the padding version spent
// ~560 k uniform-datapath instructions per launch on recomputing them).
template <int K>
struct K1Ctl {
  static constexpr bool NET = K <= 9;
  static constexpr int PR = K1_MHP + 2 * (K / 2), PWS = k1_pws(K), NRUN = K1_MW / MedianRunLen<K>::value;
  unsigned long long bar;                                        // mbarrier of the TMA copy
  int s_next, n_work;
  uint32_t s_diff;
  int pad_;
  int s_dec[2][4];
  uint32_t blkmax[NET ? (PR + 7) / 8 : 1][NET ? PWS / 8 : 1];
  uint16_t worklist[NET ? K1_MHP * NRUN : 1];
};

template <int K>
constexpr size_t k1_smem_bytes() {
  constexpr int R = K / 2;
  return (size_t)(K1_MHP + 2 * R) * k1_pws(K) * 4 + (size_t)K1_MH * K1_MW * 4 + (size_t)(K1_MH + 2 * R) * k1_pws(K) * 2 + sizeof(K1Ctl<K>);
}

template <int K, int HS, int MODE>
__global__ void __launch_bounds__(K1_THREADS, K1_MIN_CTAS) k1_preprocess(const __grid_constant__ PreParams p,
                                                                         const __grid_constant__ CUtensorMap tmap) {
  constexpr int R = K / 2;
  constexpr int PWS = k1_pws(K);           // packed raw tile row stride in words
  constexpr int XB = k1_xbox(K);           // the TMA box starts XB pixels left of the output tile
  constexpr bool NET = K <= 9;             // generated min/max network; larger windows use the generic rank search
  constexpr int PR = K1_MHP + 2 * R;       // packed raw rows
  constexpr int RW = K1_MW + 2 * R;        // raw columns actually needed
  constexpr int RUN = MedianRunLen<K>::value;   // outputs (packed registers) per work item
  constexpr int NRUN = K1_MW / RUN;
  // windows above 9x9 keep their keys in UNSIGNED order (signed key ^ 0x8000): the rank search compares with the
  // unsigned packed min
  constexpr uint32_t KFLIP = NET ? 0u : 0x8000u;
  static_assert(RW <= PWS, "packed tile stride too small");
  static_assert(K1_MW % RUN == 0 && RUN % 2 == 0, "median tile width must be a multiple of the (even) run length");

  extern __shared__ __align__(128) uint32_t smem_raw[];
  // Layout: P | M | T | control block.  P and M are multiples of 128 bytes, so T (the TMA destination) is 128-byte
  // aligned whenever the block is — which it is, since this kernel has no static shared memory (checked below; a
  // misplaced block must fail loudly, not corrupt tiles).  All pointers are derived from the __shared__ array by
  // element offsets, so the compiler keeps the shared address space (LDS/STS with 32-bit addresses).
  uint32_t* smem = smem_raw;
  if (smem_u32(smem_raw) & 127u) __trap();
  uint32_t* P = smem;                                         // [PR][PWS] packed keys (row q | row q+36 <<16)
  float* M = reinterpret_cast<float*>(smem + PR * PWS);    // [K1_MH][K1_MW] f(median)

  constexpr int RH = K1_MH + 2 * R;
  constexpr int OFF = XB - K1_HS_MAX - R;  // TMA box starts at x0-XB (16-byte rule): raw column col sits at box column col + OFF
  static_assert(OFF >= 0 && RW + OFF <= PWS, "halo does not fit the aligned TMA box");
  uint16_t* T = reinterpret_cast<uint16_t*>(M + K1_MH * K1_MW);   // [RH][PWS] raw samples of the NEXT work item (TMA destination)
  static_assert(((size_t)PR * PWS * 4) % 128 == 0 && ((size_t)K1_MH * K1_MW * 4) % 128 == 0 && ((size_t)RH * PWS * 2) % 8 == 0, "tile sizes must keep the alignment of what follows them");
  K1Ctl<K>& ctl = *reinterpret_cast<K1Ctl<K>*>(T + RH * PWS);
  uint64_t& bar = reinterpret_cast<uint64_t&>(ctl.bar);
  int& s_next = ctl.s_next;

  const int tid = threadIdx.x;
  const int w = p.w, h = p.h;
  const uint32_t bias = p.input_unsigned ? 0x8000u : 0u;
  const int total_blocks = p.n * p.tiles_x * p.tiles_y;
  // work item `blk` -> (slice, tile origin); central tiles come first (see PreParams::tile_order)
  auto decode = [&](int blk, int& slice_, int& x0_, int& y0_) {
    int tile_;
    if (p.tile_order_on) { slice_ = blk % p.n; tile_ = p.tile_order[blk / p.n]; }
    else { tile_ = blk % (p.tiles_x * p.tiles_y); slice_ = blk / (p.tiles_x * p.tiles_y); }
    x0_ = (tile_ % p.tiles_x) * K1_TW;
    y0_ = (tile_ / p.tiles_x) * K1_TH;
  };
  // Persistent CTA.  Work items are handed out in launch order by an atomic counter (p.sched[0]), so
  // the balance of the hardware block scheduler is kept while the raw tile of the NEXT item is fetched
  // by TMA during the computation of the current one.  Without a counter (sched == null) items are
  // taken with a fixed stride.
  auto fetch = [&](int prev) -> int {   // one thread
    return p.sched ? (int)atomicAdd(&p.sched[0], 1u) : prev + (int)gridDim.x;
  };
  // decode() divides by run-time values; thread 0 does it once per work item, one item ahead, and hands the result
  // to the CTA through shared memory (done by every thread it cost ~1900 cycles per item on the uniform datapath)
  auto& s_dec = ctl.s_dec;
  auto issue_tma = [&](int s_, int x_, int y_) {   // one thread: bulk tensor copy of the raw halo tile of a work item
    mbar_expect_tx(&bar, RH * PWS * 2);
    tma_load_3d(T, &tmap, &bar, x_ - XB, y_ - K1_HS_MAX - R, s_);
  };

  // let a dependent kernel (K23, launched with programmatic serialization) be scheduled while this
  // grid drains; it still waits in cudaGridDependencySynchronize() until this grid has completed
  cudaTriggerProgrammaticLaunchCompletion();

  if (tid == 0) {
    if (p.use_tma) mbar_init(&bar, 1);
    const int first = p.sched ? (int)atomicAdd(&p.sched[0], 1u) : (int)blockIdx.x;
    s_next = first;
    if (first < total_blocks) decode(first, s_dec[0][0], s_dec[0][1], s_dec[0][2]);
  }
  __syncthreads();
  int blk = s_next;
  __syncthreads();   // s_next is rewritten inside the loop: everybody must have read the first value
  if (p.use_tma && tid == 0 && blk < total_blocks) issue_tma(s_dec[0][0], s_dec[0][1], s_dec[0][2]);

  for (int iter = 0; blk < total_blocks; ++iter) {
#ifdef NM03_TUNE_STAMPS
  long long t_prev = clock64();
  if (tid == 0) atomicAdd(&g_stamps[11], 1ull);
#endif
  const int slice = s_dec[iter & 1][0], x0 = s_dec[iter & 1][1], y0 = s_dec[iter & 1][2];
  const uint16_t* __restrict__ img = reinterpret_cast<const uint16_t*>(p.in) + (size_t)slice * w * h;
  // the next work item is asked for now and used after phase A: the counter's round trip to L2 stays off the critical path
  int fetched = 0;
  if (tid == 0) fetched = fetch(blk);

  // ---- phase A: raw halo tile -> packed keys --------------------------------------------------
  const int gx0 = x0 - K1_HS_MAX - R, gy0 = y0 - K1_HS_MAX - R;
  if (p.use_tma) {
    // The (MH+2R) x 144 int16 halo tile arrives in shared memory by one bulk tensor copy; rows/columns
    // outside the image arrive as zeros.
    NM03_STAMP(9);
    mbar_wait(&bar, iter & 1);
    NM03_STAMP(0);
    const bool interior = gx0 >= 0 && gy0 >= 0 && gx0 + RW <= w && gy0 + RH <= h;
    if (!p.border_zero) {
      // 4 packed words per thread: pixels i..i+3 of row q (low halves) and of row q+36 (high halves);
      // the OFF-pixel shift is folded into the byte permutes
      const uint32_t bias2 = (bias ^ KFLIP) | ((bias ^ KFLIP) << 16);
      const uint32_t* Tw = reinterpret_cast<const uint32_t*>(T);
      // a thread keeps its column group and walks down the rows: no division in the loop, constant pointer steps
      constexpr int CPR = PWS / 4;                 // column groups per row
      constexpr int RPI = K1_THREADS / CPR;        // rows in flight
      const int qg = tid / CPR, c4 = tid - qg * CPR;
      if (qg < RPI) {
        const int w0 = min((4 * c4 + OFF) >> 1, PWS / 2 - 2);   // first 32-bit word holding pixel 4*c4 + OFF (columns past RW are never used)
        const bool last = w0 + 2 >= PWS / 2;                     // the word past the row end is never needed for real columns
        const uint32_t* ra = Tw + qg * (PWS / 2) + w0;
        uint32_t* dst = P + qg * PWS + 4 * c4;
#pragma unroll 3
        for (int q = qg; q < PR; q += RPI, ra += RPI * (PWS / 2), dst += RPI * PWS) {
          const uint32_t* rb = ra + K1_MHP * (PWS / 2);
          const uint32_t a0 = ra[0], a1 = ra[1], a2 = last ? 0u : ra[2];
          const uint32_t b0 = rb[0], b1 = rb[1], b2 = last ? 0u : rb[2];
          uint4 o;
          if ((OFF & 1) == 0) {
            o.x = __byte_perm(a0, b0, 0x5410); o.y = __byte_perm(a0, b0, 0x7632);
            o.z = __byte_perm(a1, b1, 0x5410); o.w = __byte_perm(a1, b1, 0x7632);
          } else {
            o.x = __byte_perm(a0, b0, 0x7632); o.y = __byte_perm(a1, b1, 0x5410);
            o.z = __byte_perm(a1, b1, 0x7632); o.w = __byte_perm(a2, b2, 0x5410);
          }
          if (bias2) { o.x ^= bias2; o.y ^= bias2; o.z ^= bias2; o.w ^= bias2; }   // signed input, network windows: nothing to flip
          *reinterpret_cast<uint4*>(dst) = o;
        }
      }
    }
    if (!interior || p.border_zero) {
      // border tile: patch the cells whose samples lie outside the image — replicate the image edge
      // (clamp-to-edge) or insert the zero-border sentinel; zero mode also saturates real samples
      if (!p.border_zero) __syncthreads();  // the vector pass above wrote other threads' cells
      auto key = [&](int col, int row) -> uint32_t {
        const int gx = gx0 + col, gy = gy0 + row;
        if (p.border_zero) {
          if (gx < 0 || gy < 0 || gx >= w || gy >= h) return 0x8000u ^ KFLIP;
          const uint32_t v = (T[row * PWS + min(col + OFF, PWS - 1)] ^ bias) & 0xffffu;
          return (v == 0x8000u ? 0x8001u : v) ^ KFLIP;
        }
        const int cx = clampi(clampi(gx, 0, w - 1) - gx0 + OFF, 0, PWS - 1), cy = clampi(clampi(gy, 0, h - 1) - gy0, 0, RH - 1);
        return ((T[cy * PWS + cx] ^ bias) & 0xffffu) ^ KFLIP;
      };
      // columns [0, cl) and [cr, PWS) of the tile are outside the image; a packed row q holds image
      // rows gy0+q and gy0+q+36
      const int cl = p.border_zero ? PWS : max(0, -gx0), cr = p.border_zero ? PWS : min(PWS, w - gx0);
      const int warp = tid >> 5, ln = tid & 31;
      auto row_outside = [&](int q) { const int gya = gy0 + q, gyb = gya + K1_MHP; return gya < 0 || gyb >= h || gya >= h || gyb < 0 || p.border_zero; };
      // rows with a half above or below the image: the whole row, one warp per row
      for (int q = warp; q < PR; q += K1_THREADS / 32)
        if (row_outside(q))
          for (int col = ln; col < PWS; col += 32) P[q * PWS + col] = key(col, q) | (key(col, q + K1_MHP) << 16);
      // the other rows: only the few columns left / right of the image, eight lanes per row (a side strip is 7-9 columns
      // wide for whole tiles, so a full warp per row would leave three quarters of its lanes idle)
      if (cl > 0 || cr < PWS) {
        const int grp = tid >> 3, gl = tid & 7;
        for (int q = grp; q < PR; q += K1_THREADS / 8) {
          if (row_outside(q)) continue;
          for (int col = gl; col < cl; col += 8) P[q * PWS + col] = key(col, q) | (key(col, q + K1_MHP) << 16);
          for (int col = cr + gl; col < PWS; col += 8) P[q * PWS + col] = key(col, q) | (key(col, q + K1_MHP) << 16);
        }
      }
    }
  } else {
    for (int idx = tid; idx < PR * PWS; idx += K1_THREADS) {
      const int q = idx / PWS, col = idx - q * PWS;
      uint32_t a, b;
      if (p.border_zero) {
        a = fetch_key<true>(img, w, h, gx0 + col, gy0 + q, bias);
        b = fetch_key<true>(img, w, h, gx0 + col, gy0 + q + K1_MHP, bias);
      } else {
        a = fetch_key<false>(img, w, h, gx0 + col, gy0 + q, bias);
        b = fetch_key<false>(img, w, h, gx0 + col, gy0 + q + K1_MHP, bias);
      }
      P[idx] = (a | (b << 16)) ^ (KFLIP | (KFLIP << 16));
    }
  }
  NM03_STAMP(1);
  if (tid == 0) s_next = fetched;   // the previous value of s_next was read by everybody before the last barrier
  __syncthreads();
  const int next_blk = s_next;
  if (tid == 0 && next_blk < total_blocks) {   // nobody waits for this: the others go on with the skip list
    int* d = s_dec[(iter + 1) & 1];            // read at the top of the next iteration, many barriers from here
    decode(next_blk, d[0], d[1], d[2]);
    if (p.use_tma) issue_tma(d[0], d[1], d[2]);   // T is free again
  }

  NM03_STAMP(2);
  // ---- phase B: median of the K x K window on packed keys, then f() ---------------------------
  // Work items (packed row, run of 8 columns).  An item whose whole input footprint lies at or below
  // the clip floor has a known result (skip_val) and is not sorted: 8x8-block maxima of the packed
  // tile decide that, and the remaining items are compacted into a list so that warps stay full.
  auto& blkmax = ctl.blkmax;
  auto& worklist = ctl.worklist;
  int& n_work = ctl.n_work;
  bool skip_list = NET && p.skip_on;   // items come from the compacted list (false: every item of the tile, in order)
  int n_uniform = skip_list ? -1 : K1_MHP * NRUN;   // >= 0: the number of items is known without the list (all of them, or none)
  if constexpr (NET) if (p.skip_on) {
    if (tid == 0) n_work = 0;
    constexpr int NBLK = ((PR + 7) / 8) * (PWS / 8);
    static_assert(2 * NBLK <= K1_THREADS, "two threads per block of maxima, all blocks in one trip (the vote below counts threads)");
    bool above = false;
    for (int base = 0; base < 2 * NBLK; base += K1_THREADS) {   // two threads per 8x8 block, four rows each (block-uniform trips)
      const int bidx = (base + tid) >> 1, half = tid & 1;
      const int g = bidx / (PWS / 8), c8 = bidx - g * (PWS / 8);
      uint32_t m = 0x80008000u;  // (-32768, -32768)
      if (bidx < NBLK) {
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const int q = 8 * g + 4 * half + qq;
          if (q < PR) {
            const uint4 a = *reinterpret_cast<const uint4*>(P + q * PWS + 8 * c8);
            const uint4 c = *reinterpret_cast<const uint4*>(P + q * PWS + 8 * c8 + 4);
            m = __vimax3_s16x2(m, __vmaxs2(a.x, a.y), __vmaxs2(a.z, a.w));
            m = __vimax3_s16x2(m, __vmaxs2(c.x, c.y), __vmaxs2(c.z, c.w));
          }
        }
      }
      m = __vmaxs2(m, __shfl_xor_sync(0xffffffffu, m, 1));
      if (bidx < NBLK && half == 0) { blkmax[g][c8] = m; above = __vmaxs2(m, p.skip_key2) != p.skip_key2; }
    }
    NM03_STAMP(10);
    // Whole-tile shortcuts: a tile in which EVERY 8x8 block holds a sample above the floor needs every window sorted (no
    // list: brain interior), a tile in which NO block does needs none (air: the median tile is the floor value); only
    // mixed tiles build the list.  One counting barrier publishes the block maxima and takes the vote (the thread that
    // wrote a block's maximum votes for it).
    const int n_above = __syncthreads_count(above);
    if (n_above == 0) {
      skip_list = false;
      n_uniform = 0;
      const float4 f4 = make_float4(p.skip_val, p.skip_val, p.skip_val, p.skip_val);
      for (int i = tid; i < K1_MH * K1_MW / 4; i += K1_THREADS) reinterpret_cast<float4*>(M)[i] = f4;
    } else if (n_above == NBLK) {
      skip_list = false;
      n_uniform = K1_MHP * NRUN;
    } else {
    const int lane = tid & 31;
    // every thread rates its (up to ROUNDS) items first, then each warp reserves room in the list once
    constexpr int ROUNDS = (K1_MHP * NRUN + K1_THREADS - 1) / K1_THREADS;
    uint32_t bal[ROUNDS];
    bool need[ROUNDS];
    int total = 0;
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
      const int it = r * K1_THREADS + tid;
      need[r] = false;
      if (it < K1_MHP * NRUN) {
        const int prow = it / NRUN, run = it - prow * NRUN;
        // footprint: packed rows prow .. prow+K-1, columns run*RUN .. run*RUN + RUN+K-2
        const int g0 = prow >> 3, g1 = (prow + K - 1) >> 3;
        const int c0 = (run * RUN) >> 3, c1 = (run * RUN + RUN + K - 2) >> 3;
        // the footprint spans at most two block rows and two block columns (checked by the static_assert below)
        const uint32_t m = __vmaxs2(__vmaxs2(blkmax[g0][c0], blkmax[g0][c1]), __vmaxs2(blkmax[g1][c0], blkmax[g1][c1]));
        static_assert(K - 1 <= 8 && (RUN + K - 2) % 8 + 0 >= 0 && RUN + K - 1 <= 16 && (8 % RUN == 0 || RUN % 8 == 0),
                      "work item footprint must fit in 2x2 blocks of 8x8 packed words");
        need[r] = __vmaxs2(m, p.skip_key2) != p.skip_key2;        // some sample above the floor, in either half
        if (!need[r]) {
          float* mlo = M + prow * K1_MW + run * RUN;
          const float4 f4 = make_float4(p.skip_val, p.skip_val, p.skip_val, p.skip_val);
#pragma unroll
          if constexpr (RUN % 4 == 0) {
#pragma unroll
            for (int i = 0; i < RUN; i += 4) {
              *reinterpret_cast<float4*>(mlo + i) = f4;
              *reinterpret_cast<float4*>(mlo + K1_MHP * K1_MW + i) = f4;
            }
          } else {
#pragma unroll
            for (int i = 0; i < RUN; i += 2) {
              *reinterpret_cast<float2*>(mlo + i) = make_float2(f4.x, f4.y);
              *reinterpret_cast<float2*>(mlo + K1_MHP * K1_MW + i) = make_float2(f4.x, f4.y);
            }
          }
        }
      }
      bal[r] = __ballot_sync(0xffffffffu, need[r]);
      total += __popc(bal[r]);
    }
    int slot = 0;
    if (lane == 0 && total) slot = atomicAdd(&n_work, total);
    slot = __shfl_sync(0xffffffffu, slot, 0);
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
      if (need[r]) worklist[slot + __popc(bal[r] & ((1u << lane) - 1u))] = (uint16_t)(r * K1_THREADS + tid);
      slot += __popc(bal[r]);
    }
    __syncthreads();   // the list and its length are complete (tiles that took a shortcut above need no barrier here)
    }
  }
  NM03_STAMP(3);
  const int n_items = n_uniform >= 0 ? n_uniform : n_work;
  if (tid == 0 && p.sched) atomicAdd(reinterpret_cast<unsigned long long*>(p.sched + 2), (unsigned long long)n_items);   // nm03_profile_work
  // windows above 9x9: every key of the tile shares the bits above the highest bit in which any two of them differ
  // (12-bit data in a 16-bit word: the top four), so the rank search below starts there
  uint32_t& s_diff = ctl.s_diff;
  int hb = -1;
  uint32_t prefix2 = 0u;
  if constexpr (!NET) {
    if (tid == 0) s_diff = 0u;
    __syncthreads();
    const uint32_t ref = P[0] & 0xffffu, ref2 = ref | (ref << 16);
    uint32_t d = 0u;
    for (int i = tid; i < PR * (PWS / 4); i += K1_THREADS) {
      const uint4 q4 = *reinterpret_cast<const uint4*>(P + 4 * i);
      d |= (q4.x ^ ref2) | (q4.y ^ ref2) | (q4.z ^ ref2) | (q4.w ^ ref2);
    }
    d = __reduce_or_sync(0xffffffffu, (d | (d >> 16)) & 0xffffu);
    if ((tid & 31) == 0 && d) atomicOr(&s_diff, d);
    __syncthreads();
    hb = 31 - __clz((int)s_diff);                       // -1: all keys equal
    const uint32_t low = hb < 0 ? 0u : ((2u << hb) - 1u);
    prefix2 = (ref & ~low) | ((ref & ~low) << 16);
  }
  for (int wi = tid; wi < n_items; wi += K1_THREADS) {
    int it = wi;
    if constexpr (NET) { if (skip_list) it = worklist[wi]; }
    const int prow = it / NRUN, run = it - prow * NRUN;
    uint32_t o[RUN];
    const uint32_t* src = P + prow * PWS + run * RUN;
    if constexpr (NET) {
      uint32_t v[RUN + K - 1][K];
#pragma unroll
      for (int r = 0; r < K; ++r) load_window_row<RUN + K - 1, RUN, K>(src + r * PWS, v, r);
#ifdef NM03_TUNE_SKIP_B   // tuning builds only (tools/README.md): the window centre instead of the median
#pragma unroll
      for (int i = 0; i < RUN; ++i) o[i] = v[i + R][R];
#else
      MedianRun<K>::run(v, o);
#endif
    } else {
      // Generic exact median for any odd window (11, 13, 15): most-significant-bit-first search for the value of rank
      // (K*K+1)/2.  For a trial value t the number of window samples below t is counted with packed operations —
      // min(x,t) ^ t is zero in a 16-bit half iff x >= t there, and min(.,1) turns that into 0/1 — for both pixels of a
      // packed word at once and for RUN adjacent outputs that share the loaded rows.  Slower than the networks
      // (about 16 * K*K * 1.75 instructions per pixel) but data-independent and exact, ties included.
      constexpr uint32_t RANK = (K * K + 1) / 2;
      constexpr int NC = RUN + K - 1, NC4 = (NC + 3) / 4;
      static_assert(RUN == 4 && 4 * (K1_MW / RUN - 1) + 4 * NC4 <= PWS, "row loads of the rank search must stay inside the packed row");
      uint32_t cand[RUN];
#pragma unroll
      for (int i = 0; i < RUN; ++i) cand[i] = prefix2;
      for (int b = hb; b >= 0; --b) {
        const uint32_t bit2 = (1u << b) * 0x00010001u;
        uint32_t t[RUN], cnt[RUN];
#pragma unroll
        for (int i = 0; i < RUN; ++i) { t[i] = cand[i] | bit2; cnt[i] = 0u; }
#pragma unroll 1
        for (int r = 0; r < K; ++r) {
          uint32_t x[4 * NC4];
#pragma unroll
          for (int c = 0; c < NC4; ++c) {
            const uint4 q4 = *reinterpret_cast<const uint4*>(src + r * PWS + 4 * c);
            x[4 * c] = q4.x; x[4 * c + 1] = q4.y; x[4 * c + 2] = q4.z; x[4 * c + 3] = q4.w;
          }
#pragma unroll
          for (int c = 0; c < NC; ++c)
#pragma unroll
            for (int i = 0; i < RUN; ++i)
              if (c >= i && c < i + K) cnt[i] += __vminu2(__vminu2(x[c], t[i]) ^ t[i], 0x00010001u);
        }
#pragma unroll
        for (int i = 0; i < RUN; ++i)   // fewer than RANK samples below t: the median is at least t
          cand[i] |= ((cnt[i] & 0xffffu) < RANK ? (1u << b) : 0u) | ((cnt[i] >> 16) < RANK ? (0x10000u << b) : 0u);
      }
#pragma unroll
      for (int i = 0; i < RUN; ++i) o[i] = cand[i] ^ 0x80008000u;   // back to the signed key order the table is indexed by
    }
    float lo[RUN], hi[RUN];
    if (p.lut) {
#pragma unroll
      for (int i = 0; i < RUN; ++i) {
        lo[i] = lut_at(p.lut, o[i] & 0xffffu);
        hi[i] = lut_at(p.lut, o[i] >> 16);
      }
    } else {
#pragma unroll
      for (int i = 0; i < RUN; ++i) {
        lo[i] = key_to_float((int)(short)(o[i] & 0xffffu), p);
        hi[i] = key_to_float((int)(short)(o[i] >> 16), p);
      }
    }
    float* mlo = M + prow * K1_MW + run * RUN;
    float* mhi = mlo + K1_MHP * K1_MW;
#pragma unroll
    if constexpr (RUN % 4 == 0) {
#pragma unroll
      for (int i = 0; i < RUN; i += 4) {
        *reinterpret_cast<float4*>(mlo + i) = make_float4(lo[i], lo[i + 1], lo[i + 2], lo[i + 3]);
        *reinterpret_cast<float4*>(mhi + i) = make_float4(hi[i], hi[i + 1], hi[i + 2], hi[i + 3]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < RUN; i += 2) {
        *reinterpret_cast<float2*>(mlo + i) = make_float2(lo[i], lo[i + 1]);
        *reinterpret_cast<float2*>(mhi + i) = make_float2(hi[i], hi[i + 1]);
      }
    }
  }
  NM03_STAMP(4);
  __syncthreads();
  NM03_STAMP(5);

  // ---- phase B': the sharpen stage samples the MEDIAN IMAGE with clamp-to-edge (or zero), so
  // median-tile entries whose coordinates fall outside the image are replaced (not recomputed from
  // replicated raw samples).
  const int mx0 = x0 - K1_HS_MAX, my0 = y0 - K1_HS_MAX;
  if (mx0 < 0 || my0 < 0 || mx0 + K1_MW > w || my0 + K1_MH > h) {
    // columns [0, cl) and [cr, MW) and rows [0, rt) and [rb, MH) of the tile lie outside the image
    const int cl = max(0, -mx0), cr = min(K1_MW, w - mx0), rt = max(0, -my0), rb = min(K1_MH, h - my0);
    const int warp = tid >> 5, ln = tid & 31;
    auto fill = [&](int my, int mx, int cy) {
      float val = 0.0f;
      if (!p.border_zero && rb > rt && cr > cl) val = M[cy * K1_MW + clampi(mx, cl, cr - 1)];  // reads in-image cells only
      M[my * K1_MW + mx] = val;                                                                 // writes outside cells only
    };
    // rows above / below the image: whole rows, one warp per row; the other rows: the few columns beside the image,
    // eight lanes per row (the strip is at most K1_HS_MAX columns wide for whole tiles)
    for (int my = warp; my < K1_MH; my += K1_THREADS / 32)
      if (my < rt || my >= rb)
        for (int mx = ln; mx < K1_MW; mx += 32) fill(my, mx, clampi(my, rt, rb - 1));
    if (cl > 0 || cr < K1_MW) {
      const int grp = tid >> 3, gl = tid & 7;
      for (int my = rt + grp; my < rb; my += K1_THREADS / 8) {
        for (int mx = gl; mx < cl; mx += 8) fill(my, mx, my);
        for (int mx = cr + gl; mx < K1_MW; mx += 8) fill(my, mx, my);
      }
    }
    __syncthreads();
  }

  NM03_STAMP(6);
  // ---- phase C: unsharp mask, range predicate, stores ------------------------------------------
  const int lane = tid & 31;
  if (MODE == K1_EXACT) {
  // 16-byte stores of the float planes need rows that start on 16 bytes: a width that is a multiple of 4 AND base
  // pointers aligned to 16 bytes (a device-entry caller may pass a view at any 4-byte offset)
  const bool vec4 = (w & 3) == 0 && (((uintptr_t)p.sharpened | (uintptr_t)p.median) & 15) == 0;
  // every pixel in the oracle's exact operation order (one warp = one 128-pixel output row)
  for (int it = tid; it < K1_TH * (K1_TW / 4); it += K1_THREADS) {
    const int oyl = it / (K1_TW / 4), xq = it - oyl * (K1_TW / 4);
    const int oy = y0 + oyl, ox = x0 + 4 * xq;
    const bool row_in = oy < h;   // rows below the image are computed like the others (the tile holds them) and not stored:
                                  // the lanes of a warp may sit on different rows, and all of them take part in the shuffles
    const float* mrow0 = M + (oyl + K1_HS_MAX - HS) * K1_MW + 4 * xq;
    float blur[4];
    if (p.use_fma) blur4<HS, true>(mrow0, K1_MW, p, blur);
    else blur4<HS, false>(mrow0, K1_MW, p, blur);
    const float4 c4 = *reinterpret_cast<const float4*>(M + (oyl + K1_HS_MAX) * K1_MW + 4 * xq + K1_HS_MAX);
    const float cv[4] = {c4.x, c4.y, c4.z, c4.w};
    float sv[4];
    uint32_t nib = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d = __fsub_rn(cv[j], blur[j]);
      sv[j] = __fadd_rn(cv[j], __fmul_rn(p.gain, d));
      if (ox + j < w && sv[j] >= p.rg_min && sv[j] <= p.rg_max) nib |= 1u << j;
    }
    const size_t rowoff = ((size_t)slice * h + oy) * w;
    if (p.sharpened && row_in) {
      if (vec4 && ox + 3 < w) *reinterpret_cast<float4*>(p.sharpened + rowoff + ox) = make_float4(sv[0], sv[1], sv[2], sv[3]);
      else
#pragma unroll
        for (int j = 0; j < 4; ++j) if (ox + j < w) p.sharpened[rowoff + ox + j] = sv[j];
    }
    if (p.median && row_in) {
      if (vec4 && ox + 3 < w) *reinterpret_cast<float4*>(p.median + rowoff + ox) = c4;
      else
#pragma unroll
        for (int j = 0; j < 4; ++j) if (ox + j < w) p.median[rowoff + ox + j] = cv[j];
    }
    // 8 lanes x 4 bits -> one 32-pixel word of the predicate bit-plane
    uint32_t word = nib << (4 * (lane & 7));
    word |= __shfl_xor_sync(0xffffffffu, word, 1);
    word |= __shfl_xor_sync(0xffffffffu, word, 2);
    word |= __shfl_xor_sync(0xffffffffu, word, 4);
    if (p.predicate && (lane & 7) == 0 && row_in) {
      const int wi = (x0 >> 5) + (xq >> 3);
      if (wi < p.wpr) p.predicate[((size_t)slice * h + oy) * p.wpr + wi] = word;
    }
  }
  } else {
    // Fast predicate: separable (2*HE+1)-tap blur shared down a column of 8 outputs; a pixel whose
    // fast value lies within the proven error bound of a threshold is redone exactly.
    constexpr int HE = MODE == K1_FAST_HE2 ? 2 : HS;
    constexpr int NT = 2 * HE + 1;
    constexpr int RB = 4;                       // output rows per thread (8 rows: fewer FFMAs but more code and registers; 918 vs 932 us at 480 slices)
#ifdef NM03_TUNE_SKIP_C   // tuning builds only: no sharpen/predicate work, an all-zero predicate plane
    for (int it = tid; it < K1_TH * 4; it += K1_THREADS) {
      const int oy = y0 + (it >> 2), wi = (x0 >> 5) + (it & 3);
      if (p.predicate && oy < h && wi < p.wpr) p.predicate[((size_t)slice * h + oy) * p.wpr + wi] = 0u;
    }
    if (false)
#endif
    for (int it = tid; it < (K1_TH / RB) * (K1_TW / 4); it += K1_THREADS) {
      const int band = it / (K1_TW / 4), xq = it - band * (K1_TW / 4);
      const int oyl0 = band * RB;
      const int ox = x0 + 4 * xq;
      const uint32_t xmask = ox + 3 < w ? 15u : (ox < w ? (1u << (w - ox)) - 1u : 0u);   // pixels inside the image
      float hsum[RB + 2 * HE][4], cv[RB][4];
#pragma unroll
      for (int ri = 0; ri < RB + 2 * HE; ++ri) {
        const float* row = M + (oyl0 + K1_HS_MAX - HE + ri) * K1_MW + 4 * xq;
        const float4 q0 = *reinterpret_cast<const float4*>(row);
        const float4 q1 = *reinterpret_cast<const float4*>(row + 4);
        const float4 q2 = *reinterpret_cast<const float4*>(row + 8);
        const float r12[12] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float a = 0.0f;
#pragma unroll
          for (int t = 0; t < NT; ++t) a = __fmaf_rn(p.gx[t + (K1_HS_MAX - HE)], r12[K1_HS_MAX + j + t - HE], a);
          hsum[ri][j] = a;
          if (ri >= HE && ri < HE + RB) cv[ri - HE][j] = r12[K1_HS_MAX + j];
        }
      }
      uint32_t nibs[RB];
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        uint32_t nib = 0;
        float dmin = 3.0e38f;
#pragma unroll
        for (int j = 3; j >= 0; --j) {   // pixel 3 first: each sign bit is shifted in below the previous ones
          // vertical pass with the taps scaled by -gain: nb = -gain * blur, and the sharpened value
          // v + gain*(v - blur) = (1 + gain)*v + nb is one fused multiply-add (its rounding is covered by the band)
          float nb = 0.0f;
#pragma unroll
          for (int t = 0; t < NT; ++t) nb = __fmaf_rn(p.gyn[t + (K1_HS_MAX - HE)], hsum[r + t][j], nb);
          const float sv = __fmaf_rn(p.gain1, cv[r][j], nb);
          // signed distance to the interval: negative inside; decided by its sign unless it lies in the band.  The
          // straight-line code only tracks the smallest |d| of the four pixels and compares it once with the widest
          // band any pixel can have (eps_max >= eps for every |blur|, |v|); a pixel's own band is worked out off the
          // straight-line path, for the rows that pass this test.
          const float d = __fsub_rn(fabsf(__fsub_rn(sv, p.rg_mid)), p.rg_half);
          nib = __funnelshift_l(__float_as_uint(d), nib, 1);
          dmin = fminf(dmin, fabsf(d));
        }
        if (dmin <= p.eps_max) {   // rare: pixels inside their band are redone in the oracle's exact order
          for (int j = 0; j < 4; ++j) {
            const float* mc = M + (oyl0 + r + K1_HS_MAX) * K1_MW + 4 * xq + j + K1_HS_MAX;
            const float v = *mc;
            const float nb = blur1_fast<HE>(mc, p);   // bit-identical to the value computed above
            const float svf = __fmaf_rn(p.gain1, v, nb);
            const float eps = __fmaf_rn(p.ean, fabsf(nb), __fmaf_rn(p.eb, fabsf(v), p.ec));
            const float d = __fsub_rn(fabsf(__fsub_rn(svf, p.rg_mid)), p.rg_half);
            if (fabsf(d) <= eps) {
              const float sv = __fadd_rn(v, __fmul_rn(p.gain, __fsub_rn(v, blur1_exact<HS>(mc, p))));
              nib = (sv >= p.rg_min && sv <= p.rg_max) ? (nib | (1u << j)) : (nib & ~(1u << j));
            }
          }
        }
        nibs[r] = nib & xmask;
        if (p.median && y0 + oyl0 + r < h) {
          const size_t rowoff = ((size_t)slice * h + y0 + oyl0 + r) * w;
#pragma unroll
          for (int j = 0; j < 4; ++j) if (ox + j < w) p.median[rowoff + ox + j] = cv[r][j];
        }
      }
      // Eight lanes hold the eight nibbles of one 32-pixel word, for each of the thread's four rows.  Two rows travel in
      // the two 16-bit halves of one register through the first two exchange steps; the third step hands the low lanes the
      // complete word of the first row of the pair and the high lanes that of the second: 3 shuffles per row pair, not 6.
      static_assert(RB == 4, "the word assembly below pairs rows (0,1) and (2,3)");
      const int sub = lane & 3;
      uint32_t x01 = (nibs[0] << (4 * sub)) | (nibs[1] << (4 * sub + 16));
      uint32_t x23 = (nibs[2] << (4 * sub)) | (nibs[3] << (4 * sub + 16));
      x01 |= __shfl_xor_sync(0xffffffffu, x01, 1); x23 |= __shfl_xor_sync(0xffffffffu, x23, 1);
      x01 |= __shfl_xor_sync(0xffffffffu, x01, 2); x23 |= __shfl_xor_sync(0xffffffffu, x23, 2);
      const uint32_t y01 = __shfl_xor_sync(0xffffffffu, x01, 4), y23 = __shfl_xor_sync(0xffffffffu, x23, 4);
      if (p.predicate && sub == 0) {
        const bool up = (lane & 4) != 0;   // high lanes of the group: second row of each pair
        const uint32_t wa = up ? ((y01 >> 16) | (x01 & 0xffff0000u)) : ((x01 & 0xffffu) | (y01 << 16));
        const uint32_t wb = up ? ((y23 >> 16) | (x23 & 0xffff0000u)) : ((x23 & 0xffffu) | (y23 << 16));
        const int wi = (x0 >> 5) + (xq >> 3), oy = y0 + oyl0 + (up ? 1 : 0);
        if (wi < p.wpr) {
          uint32_t* dst = p.predicate + ((size_t)slice * h + oy) * p.wpr + wi;
          if (oy < h) dst[0] = wa;
          if (oy + 2 < h) dst[2 * (size_t)p.wpr] = wb;
        }
      }
    }
  }
  NM03_STAMP(7);
  __syncthreads();   // everybody is done with this item (and has read s_next) before the next one starts
  NM03_STAMP(8);
  blk = next_blk;
  }  // persistent loop over work items
  // the last CTA to leave resets the work counter for the next launch that uses it
  if (p.sched && tid == 0) {
    __threadfence();
    if (atomicAdd(&p.sched[1], 1u) == gridDim.x - 1) {
      p.sched[0] = 0u; p.sched[1] = 0u; __threadfence();
#ifdef NM03_TUNE_STAMPS
      if (p.n >= 400) {
        const double n = (double)g_stamps[11];
        printf("K1 cycles per work item (thread 0): decode %.0f tma-wait %.0f pack %.0f next+sync %.0f blkmax %.0f worklist %.0f B %.0f B-barrier %.0f B' %.0f C %.0f end-barrier %.0f (items %.0f)\n",
               g_stamps[9] / n, g_stamps[0] / n, g_stamps[1] / n, g_stamps[2] / n, g_stamps[10] / n, g_stamps[3] / n, g_stamps[4] / n, g_stamps[5] / n, g_stamps[6] / n, g_stamps[7] / n, g_stamps[8] / n, n);
      }
#endif
    }
  }
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) f = nullptr;
    return (EncodeTiledFn)f;
  }();
  return fn;
}

// tensor map over the [n][h][w] 16-bit batch with a (144, MH+2R, 1) box; false if TMA cannot be used
static bool make_tmap(const PreParams& p, int R, CUtensorMap* map) {
  if ((p.w & 7) || ((uintptr_t)p.in & 15) || !encode_tiled_fn()) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)p.w, (cuuint64_t)p.h, (cuuint64_t)p.n};
  const cuuint64_t strides[2] = {(cuuint64_t)p.w * 2, (cuuint64_t)p.w * p.h * 2};
  const cuuint32_t box[3] = {(cuuint32_t)k1_pws(2 * R + 1), (cuuint32_t)(K1_MH + 2 * R), 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return encode_tiled_fn()(map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, const_cast<void*>(p.in), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int K, int HS, int MODE>
static cudaError_t launch_k1m(const PreParams& p0, cudaStream_t s) {
  constexpr int R = K / 2;
  const size_t smem = k1_smem_bytes<K>();
  PreParams p = p0;
  CUtensorMap map;
  memset(&map, 0, sizeof(map));
  p.use_tma = p0.use_tma && make_tmap(p, R, &map);
  // persistent grid: at most one CTA per resident slot (K1_MIN_CTAS per SM), each loops over work items
  static int sm_count[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && sm_count[dev] == 0) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
  const int slots = K1_MIN_CTAS * ((dev >= 0 && dev < 64 && sm_count[dev] > 0) ? sm_count[dev] : 148);
  const int total = p.n * p.tiles_x * p.tiles_y;
  const int grid = total < slots ? total : slots;
  k1_preprocess<K, HS, MODE><<<grid, K1_THREADS, smem, s>>>(p, map);
  return cudaGetLastError();
}

template <int K, int HS>
static cudaError_t launch_k1(const PreParams& p, int mode, cudaStream_t s) {
  if (mode == K1_FAST_HE2 && HS > 2) return launch_k1m<K, HS, (HS > 2 ? K1_FAST_HE2 : K1_FAST_FULL)>(p, s);
  if (mode != K1_EXACT) return launch_k1m<K, HS, K1_FAST_FULL>(p, s);
  return launch_k1m<K, HS, K1_EXACT>(p, s);
}

template <int K, int HS>
static cudaError_t config_k1() {
  constexpr int R = K / 2;
  const int smem = (int)k1_smem_bytes<K>();
  cudaError_t e = cudaFuncSetAttribute(k1_preprocess<K, HS, K1_EXACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_preprocess<K, HS, K1_FAST_FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess && HS > 2)
    e = cudaFuncSetAttribute(k1_preprocess<K, HS, (HS > 2 ? K1_FAST_HE2 : K1_FAST_FULL)>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  return e;
}

}  // namespace nm03
