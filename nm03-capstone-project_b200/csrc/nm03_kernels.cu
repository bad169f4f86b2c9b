// nm03_kernels.cu — sm_100a kernels of the per-slice brain-MRI pipeline.
//
//   K1  k1_preprocess   IntensityNormalization + IntensityClipping + VectorMedianFilter +
//                       ImageSharpening (+ the region-growing range test) in one pass over the raw
//                       int16 slice  (reference main_sequential.cpp:195-210, :233).
//   K23 k23_segment     SeededRegionGrowing + ImageCaster + Dilation/Erosion on bit-packed planes
//                       held in (distributed) shared memory, one thread-block cluster per slice
//                       (reference main_sequential.cpp:232-252, test_pipeline.cpp:119-125).
//
// Nothing here is a dense contraction, so no tensor cores: K1 is bound by the integer min/max
// pipe (packed VIMNMX.S16x2 networks), K23 by shared-memory latency.  Built with -fmad=false; the
// places where rounding must follow the oracle use explicit _rn intrinsics.
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

// f() for all 65536 keys, computed once per parameter set with exactly the operations above; K1 then
// replaces ~30 instructions per pixel (IEEE division with its special-case scaffolding) by one
// cached load.  lut index = the key's 16 bits as an unsigned number.
__global__ void k_build_lut(const __grid_constant__ PreParams p, float* __restrict__ lut) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 65536) lut[i] = key_to_float((int)(short)i, p);
}

cudaError_t launch_build_lut(const PreParams& p, float* lut, cudaStream_t s) {
  k_build_lut<<<256, 256, 0, s>>>(p, lut);
  return cudaGetLastError();
}

// NM03_K1_WS=1 selects the warp-specialised K1 (k1_ws) where it applies
static const bool g_k1_ws = [] { const char* e = getenv("NM03_K1_WS"); return e && e[0] == '1'; }();

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

// Separable blur of ONE pixel with exactly the operations (and order) of the straight-line fast path of phase C:
// per row a chain of FMAs over the taps from 0.0f, then a chain over the rows.  Used off the hot path to redo the
// band test of a pixel with its own error bound.
template <int HE>
__device__ __noinline__ float blur1_fast(const float* __restrict__ Mc, const PreParams& p) {
  float blur = 0.0f;
#pragma unroll 1
  for (int t = -HE; t <= HE; ++t) {
    float a = 0.0f;
#pragma unroll
    for (int u = -HE; u <= HE; ++u) a = __fmaf_rn(p.gx[u + K1_HS_MAX], Mc[t * K1_MW + u], a);
    blur = __fmaf_rn(p.gx[t + K1_HS_MAX], a, blur);
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
__device__ unsigned long long g_stamps[12];
#define NM03_STAMP(i) do { if (tid == 0) { const long long t_ = clock64(); atomicAdd(&g_stamps[i], (unsigned long long)(t_ - t_prev)); t_prev = t_; } } while (0)
#else
#define NM03_STAMP(i) do { } while (0)
#endif

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

  extern __shared__ __align__(16) uint32_t smem_raw[];
  // TMA destination: 128-byte aligned.  The padding is added as an element offset to the __shared__ array, so the
  // compiler keeps the shared address space (LDS/STS with 32-bit addresses); rounding the pointer through uintptr_t
  // turned every tile access into a generic LD.E/ST.E with 64-bit address arithmetic.
  uint32_t* smem = smem_raw + (((128u - (smem_u32(smem_raw) & 127u)) & 127u) >> 2);
  uint32_t* P = smem;                                         // [PR][PWS] packed keys (row q | row q+36 <<16)
  float* M = reinterpret_cast<float*>(smem + PR * PWS);    // [K1_MH][K1_MW] f(median)

  constexpr int RH = K1_MH + 2 * R;
  constexpr int OFF = XB - K1_HS_MAX - R;  // TMA box starts at x0-XB (16-byte rule): raw column col sits at box column col + OFF
  static_assert(OFF >= 0 && RW + OFF <= PWS, "halo does not fit the aligned TMA box");
  uint16_t* T = reinterpret_cast<uint16_t*>(M + K1_MH * K1_MW);   // [RH][PWS] raw samples of the NEXT work item (TMA destination)
  __shared__ __align__(8) uint64_t bar;
  __shared__ int s_next;

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
  __shared__ int s_dec[2][4];
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
#pragma unroll 3
      for (int idx = tid; idx < PR * (PWS / 4); idx += K1_THREADS) {
        const int q = idx / (PWS / 4), c4 = idx - q * (PWS / 4);
        const int w0 = min((4 * c4 + OFF) >> 1, PWS / 2 - 2);   // first 32-bit word holding pixel 4*c4 + OFF (columns past RW are never used)
        const uint32_t* ra = Tw + q * (PWS / 2) + w0;
        const uint32_t* rb = Tw + (q + K1_MHP) * (PWS / 2) + w0;
        const bool last = w0 + 2 >= PWS / 2;             // the word past the row end is never needed for real columns
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
        o.x ^= bias2; o.y ^= bias2; o.z ^= bias2; o.w ^= bias2;
        *reinterpret_cast<uint4*>(P + q * PWS + 4 * c4) = o;
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
  __shared__ uint32_t blkmax[NET ? (PR + 7) / 8 : 1][NET ? PWS / 8 : 1];
  __shared__ uint16_t worklist[NET ? K1_MHP * NRUN : 1];
  __shared__ int n_work;
  if constexpr (NET) if (p.skip_on) {
    if (tid == 0) n_work = 0;
    constexpr int NBLK = ((PR + 7) / 8) * (PWS / 8);
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
      if (bidx < NBLK && half == 0) blkmax[g][c8] = m;
    }
    __syncthreads();
    NM03_STAMP(10);
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
    __syncthreads();
  }
  NM03_STAMP(3);
  const int n_items = (NET && p.skip_on) ? n_work : K1_MHP * NRUN;
  if (tid == 0 && p.sched) atomicAdd(reinterpret_cast<unsigned long long*>(p.sched + 2), (unsigned long long)n_items);   // nm03_profile_work
  // windows above 9x9: every key of the tile shares the bits above the highest bit in which any two of them differ
  // (12-bit data in a 16-bit word: the top four), so the rank search below starts there
  __shared__ uint32_t s_diff;
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
    if constexpr (NET) { if (p.skip_on) it = worklist[wi]; }
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
        lo[i] = __ldg(p.lut + (o[i] & 0xffffu));
        hi[i] = __ldg(p.lut + (o[i] >> 16));
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
    if (oy >= h) continue;  // warp-uniform: a warp owns one row
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
    if (p.sharpened) {
      if (vec4 && ox + 3 < w) *reinterpret_cast<float4*>(p.sharpened + rowoff + ox) = make_float4(sv[0], sv[1], sv[2], sv[3]);
      else
#pragma unroll
        for (int j = 0; j < 4; ++j) if (ox + j < w) p.sharpened[rowoff + ox + j] = sv[j];
    }
    if (p.median) {
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
    if (p.predicate && (lane & 7) == 0) {
      const int wi = (x0 >> 5) + (lane >> 3);
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
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const int oy = y0 + oyl0 + r;
        uint32_t nib = 0, unc = 0;
#pragma unroll
        for (int j = 3; j >= 0; --j) {   // pixel 3 first: each sign bit is shifted in below the previous ones
          float blur = 0.0f;
#pragma unroll
          for (int t = 0; t < NT; ++t) blur = __fmaf_rn(p.gx[t + (K1_HS_MAX - HE)], hsum[r + t][j], blur);
          const float v = cv[r][j];
          const float sv = __fadd_rn(v, __fmul_rn(p.gain, __fsub_rn(v, blur)));
          // signed distance to the interval: negative inside; decided by its sign unless it lies in the band.  The
          // straight-line code tests against the widest band any pixel can have (eps_max >= eps for every |blur|, |v|);
          // the pixel's own band is only worked out for the few that pass this test.
          const float d = __fsub_rn(fabsf(__fsub_rn(sv, p.rg_mid)), p.rg_half);
          nib = __funnelshift_l(__float_as_uint(d), nib, 1);
          unc = (unc << 1) | (fabsf(d) <= p.eps_max ? 1u : 0u);
        }
        if (unc) {   // rare: pixels inside their band are redone in the oracle's exact order, off the straight-line path
          for (int j = 0; j < 4; ++j)
            if (unc >> j & 1u) {
              const float* mc = M + (oyl0 + r + K1_HS_MAX) * K1_MW + 4 * xq + j + K1_HS_MAX;
              const float v = *mc;
              const float blur = blur1_fast<HE>(mc, p);   // bit-identical to the value computed above
              const float svf = __fadd_rn(v, __fmul_rn(p.gain, __fsub_rn(v, blur)));
              const float eps = __fmaf_rn(p.ea, fabsf(blur), __fmaf_rn(p.eb, fabsf(v), p.ec));
              const float d = __fsub_rn(fabsf(__fsub_rn(svf, p.rg_mid)), p.rg_half);
              if (fabsf(d) <= eps) {
                const float sv = __fadd_rn(v, __fmul_rn(p.gain, __fsub_rn(v, blur1_exact<HS>(mc, p))));
                nib = (sv >= p.rg_min && sv <= p.rg_max) ? (nib | (1u << j)) : (nib & ~(1u << j));
              }
            }
        }
        nib &= xmask;
        uint32_t word = nib << (4 * (lane & 7));
        word |= __shfl_xor_sync(0xffffffffu, word, 1);
        word |= __shfl_xor_sync(0xffffffffu, word, 2);
        word |= __shfl_xor_sync(0xffffffffu, word, 4);
        if (p.predicate && (lane & 7) == 0 && oy < h) {
          const int wi = (x0 >> 5) + (lane >> 3);
          if (wi < p.wpr) p.predicate[((size_t)slice * h + oy) * p.wpr + wi] = word;
        }
        if (p.median && oy < h) {
          const size_t rowoff = ((size_t)slice * h + oy) * w;
#pragma unroll
          for (int j = 0; j < 4; ++j) if (ox + j < w) p.median[rowoff + ox + j] = cv[r][j];
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

// =================================================================================================
// K1, warp-specialised variant (fast predicate modes, TMA staging, clamp border)
// =================================================================================================
// One CTA per SM, three groups of warps that work on different tiles at the same time:
//   loader  (4 warps): raw TMA tile -> packed keys, block maxima, list of windows that need sorting
//   median  (8 warps): the min/max network and f() -> median tile (the only group that needs many registers)
//   sharpen (8 warps): border composition, separable unsharp mask, predicate bits -> global memory
// Packed-key, median and raw tiles are double-buffered; full/empty mbarriers pass them between the groups.
// In k1_preprocess the same three phases run one after the other in every CTA: the median phase keeps the
// min/max pipe busy with 8 warps, the other two are latency-bound and leave that pipe idle.
constexpr int WS_B_WARPS = 8, WS_P_WARPS = 4, WS_C_WARPS = 8;
constexpr int WS_B_THREADS = 32 * WS_B_WARPS, WS_P_THREADS = 32 * WS_P_WARPS, WS_C_THREADS = 32 * WS_C_WARPS;
constexpr int WS_THREADS = WS_B_THREADS + WS_P_THREADS + WS_C_THREADS;   // 640
constexpr int WS_RING = 8;         // work-item descriptors in flight (the loader runs at most 4 items ahead of the sharpen group)
constexpr int WS_REGS_B = 136, WS_REGS_P = 56, WS_REGS_C = 72;   // 256*136 + 128*56 + 256*72 <= 640*96

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void group_bar(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

#ifdef NM03_TUNE_STAMPS
__device__ unsigned long long g_ws_stamps[8];
#define WS_STAMP(cond, i) do { if (cond) { const long long t_ = clock64(); atomicAdd(&g_ws_stamps[i], (unsigned long long)(t_ - ws_prev)); ws_prev = t_; } } while (0)
#else
#define WS_STAMP(cond, i) do { } while (0)
#endif

template <int K>
constexpr size_t ws_smem_bytes() {
  constexpr int R = K / 2;
  return 2 * ((((size_t)(K1_MH + 2 * R) * K1_PWS * 2 + 127) / 128) * 128 + (size_t)(K1_MHP + 2 * R) * K1_PWS * 4 + (size_t)K1_MH * K1_MW * 4 +
              (size_t)((K1_MHP * (K1_MW / MedianRun<K>::RUN) * 2 + 15) / 16) * 16) + 128;
}

template <int K, int HS, int MODE>
__global__ void __launch_bounds__(WS_THREADS, 1) k1_ws(const __grid_constant__ PreParams p, const __grid_constant__ CUtensorMap tmap) {
  constexpr int R = K / 2;
  constexpr int PR = K1_MHP + 2 * R;
  constexpr int RW = K1_MW + 2 * R;
  constexpr int RUN = MedianRun<K>::RUN;
  constexpr int NRUN = K1_MW / RUN;
  constexpr int NITEMS = K1_MHP * NRUN;
  constexpr int RH = K1_MH + 2 * R;
  constexpr int OFF = 8 - K1_HS_MAX - R;
  static_assert(RW + OFF <= K1_PWS && OFF >= 0, "halo does not fit the aligned TMA box");
  static_assert(MODE != K1_EXACT, "the warp-specialised kernel implements the fast predicate modes only");

  extern __shared__ __align__(16) uint32_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>(smem_raw + (((128u - (smem_u32(smem_raw) & 127u)) & 127u) >> 2));   // keeps the shared address space
  constexpr size_t T_BYTES = (((size_t)RH * K1_PWS * 2 + 127) / 128) * 128, P_BYTES = (size_t)PR * K1_PWS * 4, M_BYTES = (size_t)K1_MH * K1_MW * 4;
  constexpr size_t W_BYTES = (size_t)((NITEMS * 2 + 15) / 16) * 16;
  static_assert(T_BYTES % 128 == 0 && P_BYTES % 16 == 0 && M_BYTES % 16 == 0, "tile buffers must keep their alignment");
  // buffer b of each kind (computed, not looked up: an indexed pointer array would live in local memory)
  auto Tbuf = [&](int b) { return reinterpret_cast<uint16_t*>(base + b * T_BYTES); };
  auto Pbuf = [&](int b) { return reinterpret_cast<uint32_t*>(base + 2 * T_BYTES + b * P_BYTES); };
  auto Mbuf = [&](int b) { return reinterpret_cast<float*>(base + 2 * T_BYTES + 2 * P_BYTES + b * M_BYTES); };
  auto Wbuf = [&](int b) { return reinterpret_cast<uint16_t*>(base + 2 * (T_BYTES + P_BYTES + M_BYTES) + b * W_BYTES); };
  __shared__ __align__(8) uint64_t bar_tma[2], bar_pfull[2], bar_pempty[2], bar_mfull[2], bar_mempty[2];
  __shared__ int s_dec[WS_RING][4];   // slice, x0, y0, valid
  __shared__ int s_nwork[2], s_nskip[2];
  __shared__ uint32_t blkmax[(PR + 7) / 8][K1_PWS / 8];

  const int tid = threadIdx.x;
  const int w = p.w, h = p.h;
  const uint32_t bias = p.input_unsigned ? 0x8000u : 0u;
  const int total_blocks = p.n * p.tiles_x * p.tiles_y;
#ifdef NM03_TUNE_STAMPS
  long long ws_prev = clock64();
#endif

  cudaTriggerProgrammaticLaunchCompletion();   // as in k1_preprocess: lets K23's CTAs become resident while this grid drains
  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bar_tma[b], 1);
      mbar_init(&bar_pfull[b], WS_P_THREADS);
      mbar_init(&bar_pempty[b], WS_B_THREADS);
      mbar_init(&bar_mfull[b], WS_B_THREADS);
      mbar_init(&bar_mempty[b], WS_C_THREADS);
    }
  }
  __syncthreads();

  if (tid >= WS_B_THREADS && tid < WS_B_THREADS + WS_P_THREADS) {
    // ================= loader group =================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(WS_REGS_P));
    const int t = tid - WS_B_THREADS, warp = t >> 5, ln = t & 31;
    auto decode = [&](int blk, int* d) {   // work item -> slice, tile origin (central tiles first, see PreParams::tile_order)
      int tile_;
      if (p.tile_order_on) { d[0] = blk % p.n; tile_ = p.tile_order[blk / p.n]; }
      else { tile_ = blk % (p.tiles_x * p.tiles_y); d[0] = blk / (p.tiles_x * p.tiles_y); }
      d[1] = (tile_ % p.tiles_x) * K1_TW;
      d[2] = (tile_ / p.tiles_x) * K1_TH;
    };
    auto announce = [&](int i, int blk) {   // leader: descriptor of item i, and its raw tile on the way
      int* d = s_dec[i % WS_RING];
      d[3] = blk < total_blocks;
      if (d[3]) {
        decode(blk, d);
        mbar_expect_tx(&bar_tma[i & 1], RH * K1_PWS * 2);
        tma_load_3d(Tbuf(i & 1), &tmap, &bar_tma[i & 1], d[1] - 8, d[2] - K1_HS_MAX - R, d[0]);
      }
    };
    int blk = 0;
    if (t == 0) {
      blk = p.sched ? (int)atomicAdd(&p.sched[0], 1u) : (int)blockIdx.x;
      announce(0, blk);
    }
    group_bar(1, WS_P_THREADS);
    for (int i = 0;; ++i) {
      const int b = i & 1, n = i >> 1;
      const int* d = s_dec[i % WS_RING];
      // (a full/empty barrier must never run two phases ahead of its waiters, so even the end marker waits its turn)
      if (n >= 1) mbar_wait(&bar_pempty[b], (n - 1) & 1);   // the median group is done with this packed buffer
      if (!d[3]) { mbar_arrive(&bar_pfull[b]); break; }     // no more work: the end marker travels down the pipeline
      const int x0 = d[1], y0 = d[2];
      if (t == 0) {   // next item: its raw buffer was last read while item i-1 was packed (a group barrier ago)
        blk = p.sched ? (int)atomicAdd(&p.sched[0], 1u) : blk + (int)gridDim.x;
        announce(i + 1, blk);
      }
      WS_STAMP(t == 0, 0);    // descriptor, next-item bookkeeping, wait for the median group's release
      mbar_wait(&bar_tma[b], n & 1);
      WS_STAMP(t == 0, 1);    // wait for the raw tile
      const uint16_t* T = Tbuf(b);
      uint32_t* P = Pbuf(b);
      const int gx0 = x0 - K1_HS_MAX - R, gy0 = y0 - K1_HS_MAX - R;
      {
        // 4 packed words per step: pixels i..i+3 of row q (low halves) and of row q+36 (high halves);
        // the OFF-pixel shift is folded into the byte permutes (same code as k1_preprocess)
        const uint32_t bias2 = bias | (bias << 16);
        const uint32_t* Tw = reinterpret_cast<const uint32_t*>(T);
#pragma unroll 3
        for (int idx = t; idx < PR * (K1_PWS / 4); idx += WS_P_THREADS) {
          const int q = idx / (K1_PWS / 4), c4 = idx - q * (K1_PWS / 4);
          const int w0 = (4 * c4 + OFF) >> 1;
          const uint32_t* ra = Tw + q * (K1_PWS / 2) + w0;
          const uint32_t* rb = Tw + (q + K1_MHP) * (K1_PWS / 2) + w0;
          const bool last = w0 + 2 >= K1_PWS / 2;
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
          o.x ^= bias2; o.y ^= bias2; o.z ^= bias2; o.w ^= bias2;
          *reinterpret_cast<uint4*>(P + q * K1_PWS + 4 * c4) = o;
        }
      }
      if (t == 0) { s_nwork[b] = 0; s_nskip[b] = 0; }
      group_bar(1, WS_P_THREADS);
      const bool interior = gx0 >= 0 && gy0 >= 0 && gx0 + RW <= w && gy0 + RH <= h;
      if (!interior) {
        // border tile: cells whose samples lie outside the image replicate the image edge (clamp-to-edge)
        auto key = [&](int col, int row) -> uint32_t {
          const int gx = gx0 + col, gy = gy0 + row;
          const int cx = clampi(clampi(gx, 0, w - 1) - gx0 + OFF, 0, K1_PWS - 1), cy = clampi(clampi(gy, 0, h - 1) - gy0, 0, RH - 1);
          return (T[cy * K1_PWS + cx] ^ bias) & 0xffffu;
        };
        const int cl = max(0, -gx0), cr = min(K1_PWS, w - gx0);
        for (int q = warp; q < PR; q += WS_P_WARPS) {
          const int gya = gy0 + q, gyb = gya + K1_MHP;
          if (gya < 0 || gyb >= h || gya >= h || gyb < 0) {
            for (int col = ln; col < K1_PWS; col += 32) P[q * K1_PWS + col] = key(col, q) | (key(col, q + K1_MHP) << 16);
          } else {
            for (int col = ln; col < cl; col += 32) P[q * K1_PWS + col] = key(col, q) | (key(col, q + K1_MHP) << 16);
            for (int col = cr + ln; col < K1_PWS; col += 32) P[q * K1_PWS + col] = key(col, q) | (key(col, q + K1_MHP) << 16);
          }
        }
        group_bar(1, WS_P_THREADS);
      }
      // windows to sort: the front of the list; windows that lie at or below the clip floor: its back
      uint16_t* wl = Wbuf(b);
      if (p.skip_on) {
        constexpr int NBLK = ((PR + 7) / 8) * (K1_PWS / 8);
        for (int base2 = 0; base2 < 2 * NBLK; base2 += WS_P_THREADS) {   // two threads per 8x8 block, four rows each
          const int bidx = (base2 + t) >> 1, half = t & 1;
          const int g = bidx / (K1_PWS / 8), c8 = bidx - g * (K1_PWS / 8);
          uint32_t m = 0x80008000u;
          if (bidx < NBLK) {
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              const int q = 8 * g + 4 * half + qq;
              if (q < PR) {
                const uint4 a = *reinterpret_cast<const uint4*>(P + q * K1_PWS + 8 * c8);
                const uint4 c = *reinterpret_cast<const uint4*>(P + q * K1_PWS + 8 * c8 + 4);
                m = __vimax3_s16x2(m, __vmaxs2(a.x, a.y), __vmaxs2(a.z, a.w));
                m = __vimax3_s16x2(m, __vmaxs2(c.x, c.y), __vmaxs2(c.z, c.w));
              }
            }
          }
          m = __vmaxs2(m, __shfl_xor_sync(0xffffffffu, m, 1));
          if (bidx < NBLK && half == 0) blkmax[g][c8] = m;
        }
        group_bar(1, WS_P_THREADS);
        for (int base2 = 0; base2 < NITEMS; base2 += WS_P_THREADS) {   // trip count is group-uniform
          const int it = base2 + t;
          bool need = false;
          if (it < NITEMS) {
            const int prow = it / NRUN, run = it - prow * NRUN;
            const int g0 = prow >> 3, g1 = (prow + K - 1) >> 3;
            const int c0 = (run * RUN) >> 3, c1 = (run * RUN + RUN + K - 2) >> 3;
            const uint32_t m = __vmaxs2(__vmaxs2(blkmax[g0][c0], blkmax[g0][c1]), __vmaxs2(blkmax[g1][c0], blkmax[g1][c1]));
            need = __vmaxs2(m, p.skip_key2) != p.skip_key2;
          }
          const bool skip = it < NITEMS && !need;
          const uint32_t bal = __ballot_sync(0xffffffffu, need), sbal = __ballot_sync(0xffffffffu, skip);
          int slot = 0, sslot = 0;
          if (ln == 0 && bal) slot = atomicAdd(&s_nwork[b], __popc(bal));
          if (ln == 0 && sbal) sslot = atomicAdd(&s_nskip[b], __popc(sbal));
          slot = __shfl_sync(0xffffffffu, slot, 0);
          sslot = __shfl_sync(0xffffffffu, sslot, 0);
          const uint32_t below = (1u << ln) - 1u;
          if (need) wl[slot + __popc(bal & below)] = (uint16_t)it;
          if (skip) wl[NITEMS - 1 - (sslot + __popc(sbal & below))] = (uint16_t)it;
        }
      } else {
        for (int it = t; it < NITEMS; it += WS_P_THREADS) wl[it] = (uint16_t)it;
        if (t == 0) s_nwork[b] = NITEMS;
      }
      mbar_arrive(&bar_pfull[b]);
      WS_STAMP(t == 0, 2);    // pack, border, block maxima, lists
    }
  }
  else if (tid < WS_B_THREADS) {
    // ================= median group =================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(WS_REGS_B));
    for (int i = 0;; ++i) {
      const int b = i & 1, n = i >> 1;
      mbar_wait(&bar_pfull[b], n & 1);
      if (n >= 1) mbar_wait(&bar_mempty[b], (n - 1) & 1);   // the sharpen group is done with this median buffer
      if (!s_dec[i % WS_RING][3]) { mbar_arrive(&bar_mfull[b]); break; }
      WS_STAMP(tid == 0, 3);  // wait for packed keys and a free median buffer
      const uint32_t* P = Pbuf(b);
      float* M = Mbuf(b);
      const uint16_t* wl = Wbuf(b);
      const int nw = s_nwork[b];
      for (int wi = tid; wi < nw; wi += WS_B_THREADS) {
        const int it = wl[wi];
        const int prow = it / NRUN, run = it - prow * NRUN;
        uint32_t v[RUN + K - 1][K], o[RUN];
        const uint32_t* src = P + prow * K1_PWS + run * RUN;
#pragma unroll
        for (int r = 0; r < K; ++r) load_window_row<RUN + K - 1, RUN, K>(src + r * K1_PWS, v, r);
        MedianRun<K>::run(v, o);
        float lo[RUN], hi[RUN];
        if (p.lut) {
#pragma unroll
          for (int j = 0; j < RUN; ++j) {
            lo[j] = __ldg(p.lut + (o[j] & 0xffffu));
            hi[j] = __ldg(p.lut + (o[j] >> 16));
          }
        } else {
#pragma unroll
          for (int j = 0; j < RUN; ++j) {
            lo[j] = key_to_float((int)(short)(o[j] & 0xffffu), p);
            hi[j] = key_to_float((int)(short)(o[j] >> 16), p);
          }
        }
        float* mlo = M + prow * K1_MW + run * RUN;
        float* mhi = mlo + K1_MHP * K1_MW;
        if constexpr (RUN % 4 == 0) {
#pragma unroll
          for (int j = 0; j < RUN; j += 4) {
            *reinterpret_cast<float4*>(mlo + j) = make_float4(lo[j], lo[j + 1], lo[j + 2], lo[j + 3]);
            *reinterpret_cast<float4*>(mhi + j) = make_float4(hi[j], hi[j + 1], hi[j + 2], hi[j + 3]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < RUN; j += 2) {
            *reinterpret_cast<float2*>(mlo + j) = make_float2(lo[j], lo[j + 1]);
            *reinterpret_cast<float2*>(mhi + j) = make_float2(hi[j], hi[j + 1]);
          }
        }
      }
      for (int wi = tid; wi < NITEMS - nw; wi += WS_B_THREADS) {   // windows at or below the clip floor: known result
        const int it = wl[NITEMS - 1 - wi];
        const int prow = it / NRUN, run = it - prow * NRUN;
        float* mlo = M + prow * K1_MW + run * RUN;
        float* mhi = mlo + K1_MHP * K1_MW;
#pragma unroll
        for (int j = 0; j < RUN; j += 2) {
          *reinterpret_cast<float2*>(mlo + j) = make_float2(p.skip_val, p.skip_val);
          *reinterpret_cast<float2*>(mhi + j) = make_float2(p.skip_val, p.skip_val);
        }
      }
      mbar_arrive(&bar_pempty[b]);
      mbar_arrive(&bar_mfull[b]);
      WS_STAMP(tid == 0, 4);  // sorting networks, f(), stores
    }
  }
  else {
    // ================= sharpen group =================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(WS_REGS_C));
    const int t = tid - WS_B_THREADS - WS_P_THREADS, warp = t >> 5, lane = t & 31;
    constexpr int HE = MODE == K1_FAST_HE2 ? 2 : HS;
    constexpr int NT = 2 * HE + 1;
    constexpr int RB = 2;   // output rows per thread: this group has few registers and is off the critical path
    for (int i = 0;; ++i) {
      const int b = i & 1, n = i >> 1;
      mbar_wait(&bar_mfull[b], n & 1);
      const int* d = s_dec[i % WS_RING];
      if (!d[3]) break;
      WS_STAMP(t == 0, 5);    // wait for a median tile
      const int slice = d[0], x0 = d[1], y0 = d[2];
      float* M = Mbuf(b);
      // the sharpen stage samples the MEDIAN IMAGE with clamp-to-edge: entries outside the image are replaced
      const int mx0 = x0 - K1_HS_MAX, my0 = y0 - K1_HS_MAX;
      if (mx0 < 0 || my0 < 0 || mx0 + K1_MW > w || my0 + K1_MH > h) {
        const int cl = max(0, -mx0), cr = min(K1_MW, w - mx0), rt = max(0, -my0), rb = min(K1_MH, h - my0);
        auto fill = [&](int my, int mx, int cy) {
          float val = 0.0f;
          if (rb > rt && cr > cl) val = M[cy * K1_MW + clampi(mx, cl, cr - 1)];   // reads in-image cells only
          M[my * K1_MW + mx] = val;                                               // writes outside cells only
        };
        for (int my = warp; my < K1_MH; my += WS_C_WARPS) {
          const int cy = clampi(my, rt, rb - 1);
          if (my < rt || my >= rb) {
            for (int mx = lane; mx < K1_MW; mx += 32) fill(my, mx, cy);
          } else {
            for (int mx = lane; mx < cl; mx += 32) fill(my, mx, cy);
            for (int mx = cr + lane; mx < K1_MW; mx += 32) fill(my, mx, cy);
          }
        }
        group_bar(2, WS_C_THREADS);
      }
      for (int it = t; it < (K1_TH / RB) * (K1_TW / 4); it += WS_C_THREADS) {
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
            for (int tt = 0; tt < NT; ++tt) a = __fmaf_rn(p.gx[tt + (K1_HS_MAX - HE)], r12[K1_HS_MAX + j + tt - HE], a);
            hsum[ri][j] = a;
            if (ri >= HE && ri < HE + RB) cv[ri - HE][j] = r12[K1_HS_MAX + j];
          }
        }
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          const int oy = y0 + oyl0 + r;
          uint32_t nib = 0, unc = 0;
#pragma unroll
          for (int j = 3; j >= 0; --j) {   // pixel 3 first: each sign bit is shifted in below the previous ones
            float blur = 0.0f;
#pragma unroll
            for (int tt = 0; tt < NT; ++tt) blur = __fmaf_rn(p.gx[tt + (K1_HS_MAX - HE)], hsum[r + tt][j], blur);
            const float v = cv[r][j];
            const float sv = __fadd_rn(v, __fmul_rn(p.gain, __fsub_rn(v, blur)));
            const float eps = __fmaf_rn(p.ea, fabsf(blur), __fmaf_rn(p.eb, fabsf(v), p.ec));
            const float dd = __fsub_rn(fabsf(__fsub_rn(sv, p.rg_mid)), p.rg_half);   // signed distance to the interval
            nib = __funnelshift_l(__float_as_uint(dd), nib, 1);
            unc = (unc << 1) | (fabsf(dd) <= eps ? 1u : 0u);
          }
          if (unc) {   // rare: pixels inside the band are redone in the oracle's exact order
            for (int j = 0; j < 4; ++j)
              if (unc >> j & 1u) {
                const float* mc = M + (oyl0 + r + K1_HS_MAX) * K1_MW + 4 * xq + j + K1_HS_MAX;
                const float v = *mc;
                const float sv = __fadd_rn(v, __fmul_rn(p.gain, __fsub_rn(v, blur1_exact<HS>(mc, p))));
                nib = (sv >= p.rg_min && sv <= p.rg_max) ? (nib | (1u << j)) : (nib & ~(1u << j));
              }
          }
          nib &= xmask;
          uint32_t word = nib << (4 * (lane & 7));
          word |= __shfl_xor_sync(0xffffffffu, word, 1);
          word |= __shfl_xor_sync(0xffffffffu, word, 2);
          word |= __shfl_xor_sync(0xffffffffu, word, 4);
          if (p.predicate && (lane & 7) == 0 && oy < h) {
            const int wi = (x0 >> 5) + (lane >> 3);
            if (wi < p.wpr) p.predicate[((size_t)slice * h + oy) * p.wpr + wi] = word;
          }
          if (p.median && oy < h) {
            const size_t rowoff = ((size_t)slice * h + oy) * w;
#pragma unroll
            for (int j = 0; j < 4; ++j) if (ox + j < w) p.median[rowoff + ox + j] = cv[r][j];
          }
        }
      }
      mbar_arrive(&bar_mempty[b]);
      WS_STAMP(t == 0, 6);    // border composition, sharpen, predicate
    }
  }
  __syncthreads();
  // the last CTA to leave resets the work counter for the next launch that uses it
  if (p.sched && tid == 0) {
    __threadfence();
    if (atomicAdd(&p.sched[1], 1u) == gridDim.x - 1) {
      p.sched[0] = 0u; p.sched[1] = 0u; __threadfence();
#ifdef NM03_TUNE_STAMPS
      if (p.n >= 400)
        printf("k1_ws cycles summed over CTAs: loader wait-release %llu wait-tma %llu work %llu | median wait %llu work %llu | sharpen wait %llu work %llu\n",
               g_ws_stamps[0], g_ws_stamps[1], g_ws_stamps[2], g_ws_stamps[3], g_ws_stamps[4], g_ws_stamps[5], g_ws_stamps[6]);
#endif
    }
  }
}

template <int K, int HS, int MODE>
static cudaError_t launch_k1_ws(const PreParams& p, const CUtensorMap& map, cudaStream_t s) {
  static int sm_count[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && sm_count[dev] == 0) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
  const int slots = (dev >= 0 && dev < 64 && sm_count[dev] > 0) ? sm_count[dev] : 148;
  const int total = p.n * p.tiles_x * p.tiles_y;
  k1_ws<K, HS, MODE><<<total < slots ? total : slots, WS_THREADS, ws_smem_bytes<K>(), s>>>(p, map);
  return cudaGetLastError();
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
  const size_t smem = (size_t)(K1_MHP + 2 * R) * k1_pws(K) * 4 + (size_t)K1_MH * K1_MW * 4 + (size_t)(K1_MH + 2 * R) * k1_pws(K) * 2 + 128;
  PreParams p = p0;
  CUtensorMap map;
  memset(&map, 0, sizeof(map));
  p.use_tma = p0.use_tma && make_tmap(p, R, &map);
  if constexpr (MODE != K1_EXACT && K <= 9) {
    if (g_k1_ws && p.use_tma && !p.border_zero) return launch_k1_ws<K, HS, MODE>(p, map, s);
  }
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
  if constexpr (K <= 9) {
    cudaError_t e2 = cudaFuncSetAttribute(k1_ws<K, HS, K1_FAST_FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws_smem_bytes<K>());
    if (e2 == cudaSuccess && HS > 2)
      e2 = cudaFuncSetAttribute(k1_ws<K, HS, (HS > 2 ? K1_FAST_HE2 : K1_FAST_FULL)>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws_smem_bytes<K>());
    if (e2 != cudaSuccess) return e2;
  }
  const int smem = (int)((size_t)(K1_MHP + 2 * R) * k1_pws(K) * 4 + (size_t)K1_MH * K1_MW * 4 + (size_t)(K1_MH + 2 * R) * k1_pws(K) * 2 + 128);
  cudaError_t e = cudaFuncSetAttribute(k1_preprocess<K, HS, K1_EXACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_preprocess<K, HS, K1_FAST_FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess && HS > 2)
    e = cudaFuncSetAttribute(k1_preprocess<K, HS, (HS > 2 ? K1_FAST_HE2 : K1_FAST_FULL)>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  return e;
}

#define NM03_K1_DISPATCH(FN, ...)                                                                   \
  switch (median_window * 10 + sharpen_half) {                                                      \
    case 31: return FN<3, 1>(__VA_ARGS__); case 32: return FN<3, 2>(__VA_ARGS__);                   \
    case 33: return FN<3, 3>(__VA_ARGS__); case 34: return FN<3, 4>(__VA_ARGS__);                   \
    case 51: return FN<5, 1>(__VA_ARGS__); case 52: return FN<5, 2>(__VA_ARGS__);                   \
    case 53: return FN<5, 3>(__VA_ARGS__); case 54: return FN<5, 4>(__VA_ARGS__);                   \
    case 71: return FN<7, 1>(__VA_ARGS__); case 72: return FN<7, 2>(__VA_ARGS__);                   \
    case 73: return FN<7, 3>(__VA_ARGS__); case 74: return FN<7, 4>(__VA_ARGS__);                   \
    case 91: return FN<9, 1>(__VA_ARGS__); case 92: return FN<9, 2>(__VA_ARGS__);                   \
    case 93: return FN<9, 3>(__VA_ARGS__); case 94: return FN<9, 4>(__VA_ARGS__);                   \
    case 111: return FN<11, 1>(__VA_ARGS__); case 112: return FN<11, 2>(__VA_ARGS__);               \
    case 113: return FN<11, 3>(__VA_ARGS__); case 114: return FN<11, 4>(__VA_ARGS__);               \
    case 131: return FN<13, 1>(__VA_ARGS__); case 132: return FN<13, 2>(__VA_ARGS__);               \
    case 133: return FN<13, 3>(__VA_ARGS__); case 134: return FN<13, 4>(__VA_ARGS__);               \
    case 151: return FN<15, 1>(__VA_ARGS__); case 152: return FN<15, 2>(__VA_ARGS__);               \
    case 153: return FN<15, 3>(__VA_ARGS__); case 154: return FN<15, 4>(__VA_ARGS__);               \
    default: return cudaErrorInvalidValue;                                                          \
  }

int median_network_ops(int median_window) {
  switch (median_window) {
    case 3: return NM03_MEDIAN_OPS_3;
    case 5: return NM03_MEDIAN_OPS_5;
    case 7: return NM03_MEDIAN_OPS_7;
    case 9: return NM03_MEDIAN_OPS_9;
    default: return 0;
  }
}

cudaError_t launch_preprocess(const PreParams& p, int median_window, int sharpen_half, int mode, cudaStream_t s) {
  NM03_K1_DISPATCH(launch_k1, p, mode, s)
}

// =================================================================================================
// K23
// =================================================================================================
// Horizontal flood inside one 32-pixel word: spread the bits of c (c subset of p) along the runs of
// ones of p, in both directions.  Adding c to p makes the carry ripple to the end of each run.
__device__ __forceinline__ uint32_t hfill(uint32_t c, uint32_t p) {
  const uint32_t up = ((p + c) ^ p) & p;
  const uint32_t rp = __brev(p), rc = __brev(c);
  const uint32_t dn = __brev(((rp + rc) ^ rp) & rp);
  return up | dn | c;
}

__device__ __forceinline__ uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

// write 32 mask pixels (one word) as uint8 {0,1}
__device__ __forceinline__ void store_word_u8(uint8_t* __restrict__ row, int x, int w, uint32_t bits, bool vec) {
  if (vec && x + 32 <= w) {
    uint4 a, b;
    a.x = spread4(bits & 15u); a.y = spread4((bits >> 4) & 15u); a.z = spread4((bits >> 8) & 15u); a.w = spread4((bits >> 12) & 15u);
    b.x = spread4((bits >> 16) & 15u); b.y = spread4((bits >> 20) & 15u); b.z = spread4((bits >> 24) & 15u); b.w = spread4(bits >> 28);
    *reinterpret_cast<uint4*>(row + x) = a;
    *reinterpret_cast<uint4*>(row + x + 16) = b;
  } else {
    for (int i = 0; i < 32 && x + i < w; ++i) row[x + i] = (bits >> i) & 1u;
  }
}

// One sweep pass of the flood fill over the rows this CTA owns.  A thread owns one 32-pixel word
// column over a segment of rows and walks it down, then up; the freshly computed word of the previous
// row travels in a register, the other vertical neighbour and the two horizontal neighbour words are
// read from shared memory as they are (chaotic relaxation of a monotone operator: any interleaving
// converges to the same fixed point, and a pass without any change proves it has been reached).
// MASKED = image-border pixels may be labelled but never act as sources (rg_border_expand = 0).
// Rows of both bit planes are stored with an ODD stride in shared memory, so a warp whose lanes walk
// along the words of a row and a warp whose lanes walk down the rows of one word column are both
// free of bank conflicts.
__host__ __device__ inline int seg_row_stride(int wpr) { return wpr | 1; }

// Activity flags: an item (row segment x word column) is processed in a pass only if it, or one of
// the items it reads from, changed since it was last looked at; a changing item wakes itself and its
// neighbours for the next pass.  Work is therefore proportional to the grown region, not to the image.
template <bool CONN8, bool MASKED>
__device__ __forceinline__ int fill_pass(uint32_t* __restrict__ Mm, const uint32_t* __restrict__ Pp,
                                         const SegParams& p, int nrows, int y_begin, int tid,
                                         uint8_t* __restrict__ act_now, uint8_t* __restrict__ act_next) {
  const int wpr = p.wpr, ws = seg_row_stride(p.wpr);
#define MROW(r) (Mm + (size_t)((r) + SEG_HALO) * ws)
  const int lastk = (p.w - 1) >> 5;
  const uint32_t lastbit = 1u << ((p.w - 1) & 31);
  int changed = 0;
  const int nseg = (nrows + p.seg_rows - 1) / p.seg_rows;
  for (int it = tid; it < nseg * wpr; it += SEG_THREADS) {
    if (!act_now[it]) continue;
    act_now[it] = 0;
    int item_changed = 0;
    const int sgi = it / wpr, k = it - sgi * wpr;
    const int ra = sgi * p.seg_rows, rb = min(nrows, ra + p.seg_rows);
    const bool has_l = k > 0, has_r = k < wpr - 1;
    uint32_t colmask = 0xffffffffu, lmask = 0xffffffffu, rmask = 0xffffffffu;  // source masks: own, left, right word
    if (MASKED) {
      if (k == 0) colmask &= ~1u;
      if (k == lastk) colmask &= ~lastbit;
      if (k - 1 == 0) lmask &= ~1u;
      if (k - 1 == lastk) lmask &= ~lastbit;
      if (k + 1 == lastk) rmask &= ~lastbit;
    }
    auto row_is_source = [&](int rr) { return !MASKED || (y_begin + rr > 0 && y_begin + rr < p.h - 1); };
    // words k-1, k, k+1 of one mask row: c = raw centre word, s/l/r = the three words as SOURCES
    struct Row3 { uint32_t c, s, l, r; };
    auto load_row = [&](int rr) {
      const uint32_t* row = MROW(rr);
      Row3 t;
      t.c = row[k];
      t.l = has_l ? row[k - 1] & lmask : 0u;
      t.r = has_r ? row[k + 1] & rmask : 0u;
      t.s = t.c & colmask;
      if (MASKED && !row_is_source(rr)) { t.s = 0u; t.l = 0u; t.r = 0u; }
      return t;
    };
#pragma unroll 1
    for (int dir = 0; dir < 2; ++dir) {
      const int step = dir == 0 ? 1 : -1;
      int r = dir == 0 ? ra : rb - 1;
      // software pipeline: rows r-step (behind), r (here), r+step (ahead) live in registers and the
      // row after that is fetched while this one is computed, so a step's critical path is ALU only
      Row3 behind = load_row(r - step), here = load_row(r), ahead = load_row(r + step);
      uint32_t pw = Pp[r * ws + k];
      for (int cnt = ra; cnt < rb; ++cnt, r += step) {
        const Row3 next = load_row(r + 2 * step);   // rows up to +-2 outside the segment exist (halo >= 2)
        const uint32_t pw_next = (cnt + 1 < rb) ? Pp[(r + step) * ws + k] : 0u;
        const uint32_t m = here.c;
        uint32_t vs = behind.s | ahead.s;
        if (CONN8) vs |= (vs << 1) | (vs >> 1) | ((behind.l | ahead.l) >> 31) | ((behind.r | ahead.r) << 31);
        const uint32_t hz = (here.l >> 31) | (here.r << 31);
        const uint32_t cand = m | ((vs | hz) & pw);
        uint32_t nw;
        const bool src_here = row_is_source(r);
        if (MASKED) {
          const uint32_t rowmask = src_here ? colmask : 0u;
          const uint32_t f = hfill(cand & rowmask, pw & rowmask);
          nw = m | f | (pw & ~rowmask & (cand | (f << 1) | (f >> 1)));
        } else {
          nw = hfill(cand, pw);
        }
        if (nw != m) { MROW(r)[k] = nw; item_changed = 1; }
        behind = here;
        behind.c = nw;
        behind.s = (MASKED && !src_here) ? 0u : nw & colmask;
        here = ahead;
        ahead = next;
        pw = pw_next;
      }
    }
    if (item_changed) {
      changed = 1;
      act_next[it] = 1;
      if (has_l) act_next[it - 1] = 1;
      if (has_r) act_next[it + 1] = 1;
      if (sgi > 0) act_next[it - wpr] = 1;
      if (sgi < nseg - 1) act_next[it + wpr] = 1;
      if (CONN8) {
        if (sgi > 0) { if (has_l) act_next[it - wpr - 1] = 1; if (has_r) act_next[it - wpr + 1] = 1; }
        if (sgi < nseg - 1) { if (has_l) act_next[it + wpr - 1] = 1; if (has_r) act_next[it + wpr + 1] = 1; }
      }
    }
  }
#undef MROW
  return changed;
}

__global__ void __launch_bounds__(SEG_THREADS, 1) k23_segment(const __grid_constant__ SegParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) uint32_t smem[];
  const int wpr = p.wpr, ws = seg_row_stride(p.wpr), rows = p.rows_per, C = p.cluster;
  const int rank = C > 1 ? (int)cluster.block_rank() : 0;
  const int slice = blockIdx.x / C;
  const int tid = threadIdx.x;
  // planes: predicate rows owned by this CTA; mask rows with SEG_HALO halo rows above and below
  uint32_t* Pp = smem;                                   // [rows][wpr]
  uint32_t* Mm = smem + (size_t)rows * ws;               // [rows + 2*HALO][ws]  (ws = odd row stride)
  const int nseg_max = (rows + p.seg_rows - 1) / p.seg_rows;
  uint8_t* act0 = reinterpret_cast<uint8_t*>(Mm + (size_t)(rows + 2 * SEG_HALO) * ws);   // [nseg][wpr] x 2
  uint8_t* act1 = act0 + (size_t)nseg_max * wpr;
  __shared__ int s_flag[3];                              // rotating "somebody changed" flags (see below)
#define MROW(r) (Mm + (size_t)((r) + SEG_HALO) * ws)
  const int y_begin = rank * rows;                       // first image row owned
  const int nrows = max(0, min(rows, p.h - y_begin));    // rows actually inside the image
  const uint32_t* gp = p.pred + (size_t)slice * p.h * wpr;
  const uint32_t lastmask = (p.w & 31) ? ((1u << (p.w & 31)) - 1u) : 0xffffffffu;

  for (int i = tid; i < (rows + 2 * SEG_HALO) * ws; i += SEG_THREADS) Mm[i] = 0u;
  for (int i = tid; i < 2 * nseg_max * wpr; i += SEG_THREADS) act0[i] = 0;
  // Launched with programmatic stream serialization: everything above overlaps the tail of the kernel
  // that produces the predicate plane; its memory is only touched after this point.
  cudaGridDependencySynchronize();
  for (int i = tid; i < rows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    uint32_t v = r < nrows ? __ldcg(gp + (size_t)(y_begin + r) * wpr + k) : 0u;
    if (k == wpr - 1) v &= lastmask;
    Pp[r * ws + k] = v;
  }
  if (tid < 3) s_flag[tid] = 0;
  __syncthreads();

  int passes = 0;
  if (p.do_fill) {
    // ---- seeds ------------------------------------------------------------------------------
    // a seed is a change nobody has looked at yet: wake its item AND the items around it (the growth
    // may start in a neighbouring segment or word column although the seed's own item stays as it is)
    auto wake = [&](int r, int k) {
      const int nsg = (nrows + p.seg_rows - 1) / p.seg_rows, sg = r / p.seg_rows;
      for (int a = max(0, sg - 1); a <= min(nsg - 1, sg + 1); ++a)
        for (int b = max(0, k - 1); b <= min(wpr - 1, k + 1); ++b) act0[a * wpr + b] = 1;
    };
    for (int i = tid; i < p.n_seeds; i += SEG_THREADS) {
      const int sx = p.seeds[i].x, sy = p.seeds[i].y;
      if (!p.host_mode) {
        const int r = sy - y_begin;
        if (r >= 0 && r < nrows) {
          atomicOr(&MROW(r)[sx >> 5], (1u << (sx & 31)) & Pp[r * ws + (sx >> 5)]);
          wake(r, sx >> 5);
        }
      } else {
        // FAST executeOnHost: the seed's neighbours are tested, not the seed itself
        const int ddx[8] = {0, 1, -1, 0, 1, -1, 1, -1}, ddy[8] = {1, 0, 0, -1, 1, 1, -1, -1};
        for (int d = 0; d < (p.conn8 ? 8 : 4); ++d) {
          const int nx = sx + ddx[d], ny = sy + ddy[d];
          const int r = ny - y_begin;
          if (nx < 0 || nx >= p.w || r < 0 || r >= nrows) continue;
          atomicOr(&MROW(r)[nx >> 5], (1u << (nx & 31)) & Pp[r * ws + (nx >> 5)]);
          wake(r, nx >> 5);
        }
      }
    }
    if (C > 1) cluster.sync();  // every CTA's planes and flags are initialised before anyone touches them
    else __syncthreads();

    // ---- fill: sweep passes until nothing changes anywhere in the cluster --------------------
    // Cluster protocol, one cluster.sync per pass: a CTA that changed something in pass q stores 1
    // into flag[q % 3] of EVERY CTA (distributed shared memory); after the sync all CTAs read the same
    // OR.  flag[(q+2) % 3] is cleared for later use — it is next written in pass q+2, which no CTA can
    // reach before this CTA has passed the sync of pass q+1, i.e. after the clear.
    const bool masked = !p.border_expand && !p.host_mode;
    int hchg = 0;
    for (int q = 0;; ++q) {
      ++passes;
      int changed;
      uint8_t* an = (q & 1) ? act1 : act0;
      uint8_t* ax = (q & 1) ? act0 : act1;
      if (p.conn8) changed = masked ? fill_pass<true, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<true, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
      else changed = masked ? fill_pass<false, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<false, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
      // the seeds themselves are progress the neighbouring CTAs have not seen: the first round always
      // continues, so that every halo row is refreshed (and its readers woken) at least once
      changed = __syncthreads_or(changed | hchg) | (q == 0 && C > 1);
      if (C == 1) {
        if (!changed) break;
        continue;
      }
      if (changed && tid < C) *cluster.map_shared_rank(&s_flag[q % 3], tid) = 1;
      cluster.sync();
      const int any = s_flag[q % 3];
      if (tid == 0) s_flag[(q + 2) % 3] = 0;
      if (!any) break;
      // refresh the one halo row each side from the neighbouring CTAs; a change there is progress
      hchg = 0;
      for (int k = tid; k < 2 * wpr; k += SEG_THREADS) {
        const int side = k / wpr, kk = k - side * wpr;
        const int nb = side == 0 ? rank - 1 : rank + 1;
        if (nb < 0 || nb >= C) continue;
        const uint32_t* remote = cluster.map_shared_rank(Mm, nb);
        const uint32_t v = remote[(size_t)((side == 0 ? rows - 1 : 0) + SEG_HALO) * ws + kk];
        uint32_t* mine = MROW(side == 0 ? -1 : rows) + kk;
        if (*mine != v) {
          *mine = v;
          hchg = 1;
          // the boundary segment next to that halo row has new input: wake it (and its word neighbours)
          const int nsg = (nrows + p.seg_rows - 1) / p.seg_rows;
          if (nsg > 0) {
            uint8_t* row = ax + (side == 0 ? 0 : nsg - 1) * wpr;
            row[kk] = 1;
            if (kk > 0) row[kk - 1] = 1;
            if (kk < wpr - 1) row[kk + 1] = 1;
          }
        }
      }
      __syncthreads();
    }
  } else {
    for (int i = tid; i < rows * wpr; i += SEG_THREADS) { const int r = i / wpr, k = i - r * wpr; MROW(r)[k] = Pp[r * ws + k]; }
    __syncthreads();
  }
  if (p.iterations && rank == 0 && tid == 0) p.iterations[slice] = passes;

  // ---- halo rows of the final mask for the structuring element --------------------------------
  const int nmax = max((p.dilated || p.dilated_bits) ? p.dil_n : 0, p.eroded ? p.ero_n : 0);
  if (C > 1) {
    cluster.sync();  // all masks final (also orders the do_fill = 0 copy)
    for (int i = tid; i < 2 * nmax * wpr; i += SEG_THREADS) {
      const int side = i / (nmax * wpr), j = i - side * nmax * wpr;
      const int hr = j / wpr, kk = j - hr * wpr;
      const int nb = side == 0 ? rank - 1 : rank + 1;
      if (nb < 0 || nb >= C) continue;
      const uint32_t* remote = cluster.map_shared_rank(Mm, nb);
      if (side == 0) MROW(-1 - hr)[kk] = remote[(size_t)(rows - 1 - hr + SEG_HALO) * ws + kk];
      else MROW(rows + hr)[kk] = remote[(size_t)(hr + SEG_HALO) * ws + kk];
    }
    cluster.sync();  // nobody exits (or overwrites) while a neighbour still reads its rows
  } else {
    __syncthreads();
  }

  // ---- ImageCaster is the identity; dilation / erosion with the row-run form of the SE ---------
  const bool vec = (p.w & 15) == 0 && ((((size_t)p.dilated) | ((size_t)p.eroded) | ((size_t)p.rg_mask)) & 15) == 0;
  for (int i = tid; i < nrows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    const int y = y_begin + r;
    const size_t rowoff = ((size_t)slice * p.h + y) * p.w;
    const uint32_t m = MROW(r)[k];
    const bool has_l = k > 0, has_r = k < wpr - 1;
    if (p.rg_mask) store_word_u8(p.rg_mask + rowoff, k * 32, p.w, m, vec);
    if (p.dilated || p.dilated_bits) {
      uint32_t acc = m;
      if (p.dil_n == 1 && p.dil_hw[0] == 0 && p.dil_hw[1] == 1) {   // Dilation(3) with the disc SE: the 5-pixel plus
        const uint32_t* row = MROW(r);
        acc = m | __funnelshift_r(m, has_r ? row[k + 1] : 0u, 1) | __funnelshift_l(has_l ? row[k - 1] : 0u, m, 1);
        if (y > 0) acc |= MROW(r - 1)[k];
        if (y < p.h - 1) acc |= MROW(r + 1)[k];
      } else if (p.dil_n > 0) {
        acc = 0u;
        for (int b = -p.dil_n; b <= p.dil_n; ++b) {
          const int yy = y + b;
          if (yy < 0 || yy >= p.h) continue;
          const uint32_t* row = MROW(r + b);
          const uint32_t c = row[k], l = has_l ? row[k - 1] : 0u, rr = has_r ? row[k + 1] : 0u;
          const int hw = p.dil_hw[b + p.dil_n];
          acc |= c;
          for (int s2 = 1; s2 <= hw; ++s2) acc |= __funnelshift_r(c, rr, s2) | __funnelshift_l(l, c, s2);
        }
      }
      if (p.dilated) store_word_u8(p.dilated + rowoff, k * 32, p.w, acc, vec);
      if (p.dilated_bits)   // dilation may have set bits beyond the last pixel of the row: they are not part of the image
        p.dilated_bits[((size_t)slice * p.h + y) * wpr + k] = (k == wpr - 1 && (p.w & 31)) ? acc & ((1u << (p.w & 31)) - 1u) : acc;
    }
    if (p.eroded) {
      uint32_t acc = m;
      if (p.ero_n > 0) {
        for (int b = -p.ero_n; b <= p.ero_n && acc; ++b) {
          const int yy = y + b;
          if (yy < 0 || yy >= p.h) { acc = 0u; break; }
          const uint32_t* row = MROW(r + b);
          const uint32_t c = row[k], l = has_l ? row[k - 1] : 0u, rr = has_r ? row[k + 1] : 0u;
          const int hw = p.ero_hw[b + p.ero_n];
          acc &= c;
          for (int s2 = 1; s2 <= hw; ++s2) acc &= __funnelshift_r(c, rr, s2) & __funnelshift_l(l, c, s2);
        }
      }
      store_word_u8(p.eroded + rowoff, k * 32, p.w, acc, vec);
    }
  }
#undef MROW
}

// ---- global-memory fallback for slices that exceed one cluster's shared memory ------------------------
// The mask lives bit-packed in global memory.  One CTA takes a band of rows, loads it (plus one halo
// row each side) into shared memory, grows it to its local fixed point with the same sweep passes as
// K23, and writes it back; information crosses a band boundary once per launch.  The host relaunches
// until a launch reports no change through the global flag.
__global__ void __launch_bounds__(SEG_THREADS, 1) k2_band_fill(const __grid_constant__ SegParams p, uint32_t* __restrict__ gmask,
                                                               int* __restrict__ gflag, int first_launch) {
  extern __shared__ __align__(16) uint32_t smem[];
  const int wpr = p.wpr, ws = seg_row_stride(p.wpr), rows = p.rows_per;
  const int nbands = (p.h + rows - 1) / rows;
  const int slice = blockIdx.x / nbands, band = blockIdx.x - slice * nbands;
  const int tid = threadIdx.x;
  uint32_t* Pp = smem;
  uint32_t* Mm = smem + (size_t)rows * ws;
  const int nseg_max = (rows + p.seg_rows - 1) / p.seg_rows;
  uint8_t* act0 = reinterpret_cast<uint8_t*>(Mm + (size_t)(rows + 2 * SEG_HALO) * ws);
  uint8_t* act1 = act0 + (size_t)nseg_max * wpr;
#define MROW(r) (Mm + (size_t)((r) + SEG_HALO) * ws)
  const int y_begin = band * rows;
  const int nrows = min(rows, p.h - y_begin);
  const uint32_t* gp = p.pred + (size_t)slice * p.h * wpr;
  uint32_t* gm = gmask + (size_t)slice * p.h * wpr;
  const uint32_t lastmask = (p.w & 31) ? ((1u << (p.w & 31)) - 1u) : 0xffffffffu;
  for (int i = tid; i < (rows + 2 * SEG_HALO) * ws; i += SEG_THREADS) Mm[i] = 0u;
  for (int i = tid; i < 2 * nseg_max * wpr; i += SEG_THREADS) act0[i] = 1;   // everything is looked at once per launch
  __syncthreads();
  for (int i = tid; i < rows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    uint32_t v = r < nrows ? gp[(size_t)(y_begin + r) * wpr + k] : 0u;
    if (k == wpr - 1) v &= lastmask;
    Pp[r * ws + k] = v;
    if (!first_launch && r < nrows) MROW(r)[k] = gm[(size_t)(y_begin + r) * wpr + k];
  }
  if (!first_launch)
    for (int k = tid; k < 2 * wpr; k += SEG_THREADS) {   // halo rows written by the neighbouring bands in earlier launches
      const int side = k / wpr, kk = k - side * wpr;
      const int y = side == 0 ? y_begin - 1 : y_begin + nrows;
      if (y >= 0 && y < p.h) MROW(side == 0 ? -1 : nrows)[kk] = gm[(size_t)y * wpr + kk];
    }
  __syncthreads();
  if (first_launch) {
    for (int i = tid; i < p.n_seeds; i += SEG_THREADS) {
      const int sx = p.seeds[i].x, sy = p.seeds[i].y;
      if (!p.host_mode) {
        const int r = sy - y_begin;
        if (r >= 0 && r < nrows) atomicOr(&MROW(r)[sx >> 5], (1u << (sx & 31)) & Pp[r * ws + (sx >> 5)]);
      } else {
        const int ddx[8] = {0, 1, -1, 0, 1, -1, 1, -1}, ddy[8] = {1, 0, 0, -1, 1, 1, -1, -1};
        for (int d = 0; d < (p.conn8 ? 8 : 4); ++d) {
          const int nx = sx + ddx[d], ny = sy + ddy[d];
          const int r = ny - y_begin;
          if (nx < 0 || nx >= p.w || r < 0 || r >= nrows) continue;
          atomicOr(&MROW(r)[nx >> 5], (1u << (nx & 31)) & Pp[r * ws + (nx >> 5)]);
        }
      }
    }
    __syncthreads();
  }
  const bool masked = !p.border_expand && !p.host_mode;
  int ever = first_launch;   // the seeds themselves are a change
  for (int q = 0;; ++q) {
    uint8_t* an = (q & 1) ? act1 : act0;
    uint8_t* ax = (q & 1) ? act0 : act1;
    int changed;
    if (p.conn8) changed = masked ? fill_pass<true, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<true, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
    else changed = masked ? fill_pass<false, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<false, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
    changed = __syncthreads_or(changed);
    if (!changed) break;
    ever = 1;
  }
  if (ever && !first_launch && tid == 0) *gflag = 1;
  if (first_launch && tid == 0) *gflag = 1;           // a second launch always runs so that the halos are seen once
  for (int i = tid; i < nrows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    gm[(size_t)(y_begin + r) * wpr + k] = MROW(r)[k];
  }
#undef MROW
}

// cast + dilation / erosion from the global bit-packed mask (same row-run form as in K23)
__global__ void k3_global_morph(const __grid_constant__ SegParams p, const uint32_t* __restrict__ gmask, int passes) {
  const int wpr = p.wpr;
  const long long total = (long long)p.n * p.h * wpr;
  const bool vec = (p.w & 15) == 0 && ((((size_t)p.dilated) | ((size_t)p.eroded) | ((size_t)p.rg_mask)) & 15) == 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % wpr);
    const long long row = i / wpr;
    const int y = (int)(row % p.h), slice = (int)(row / p.h);
    if (p.iterations && k == 0 && y == 0) p.iterations[slice] = passes;
    const uint32_t* gm = gmask + (size_t)slice * p.h * wpr;
    auto word = [&](int yy, int kk) -> uint32_t { return (yy < 0 || yy >= p.h || kk < 0 || kk >= wpr) ? 0u : gm[(size_t)yy * wpr + kk]; };
    const size_t rowoff = ((size_t)slice * p.h + y) * p.w;
    const uint32_t m = word(y, k);
    if (p.rg_mask) store_word_u8(p.rg_mask + rowoff, k * 32, p.w, m, vec);
    if (p.dilated || p.dilated_bits) {
      uint32_t acc = p.dil_n > 0 ? 0u : m;
      for (int b = -p.dil_n; b <= p.dil_n && p.dil_n > 0; ++b) {
        if (y + b < 0 || y + b >= p.h) continue;
        const uint32_t c = word(y + b, k), l = word(y + b, k - 1), rr = word(y + b, k + 1);
        const int hw = p.dil_hw[b + p.dil_n];
        acc |= c;
        for (int s2 = 1; s2 <= hw; ++s2) acc |= __funnelshift_r(c, rr, s2) | __funnelshift_l(l, c, s2);
      }
      if (p.dilated) store_word_u8(p.dilated + rowoff, k * 32, p.w, acc, vec);
      if (p.dilated_bits)
        p.dilated_bits[((size_t)slice * p.h + y) * wpr + k] = (k == wpr - 1 && (p.w & 31)) ? acc & ((1u << (p.w & 31)) - 1u) : acc;
    }
    if (p.eroded) {
      uint32_t acc = m;
      for (int b = -p.ero_n; b <= p.ero_n && p.ero_n > 0 && acc; ++b) {
        if (y + b < 0 || y + b >= p.h) { acc = 0u; break; }
        const uint32_t c = word(y + b, k), l = word(y + b, k - 1), rr = word(y + b, k + 1);
        const int hw = p.ero_hw[b + p.ero_n];
        acc &= c;
        for (int s2 = 1; s2 <= hw; ++s2) acc &= __funnelshift_r(c, rr, s2) & __funnelshift_l(l, c, s2);
      }
      store_word_u8(p.eroded + rowoff, k * 32, p.w, acc, vec);
    }
  }
}

cudaError_t launch_band_fill(const SegParams& p, uint32_t* gmask, int* gflag, int first_launch, cudaStream_t s) {
  const int nbands = (p.h + p.rows_per - 1) / p.rows_per;
  k2_band_fill<<<p.n * nbands, SEG_THREADS, seg_smem_bytes(p.rows_per, p.wpr), s>>>(p, gmask, gflag, first_launch);
  return cudaGetLastError();
}

cudaError_t launch_global_morph(const SegParams& p, const uint32_t* gmask, int passes, cudaStream_t s) {
  const long long total = (long long)p.n * p.h * p.wpr;
  long long g = (total + 255) / 256;
  k3_global_morph<<<(int)(g > 148 * 8 ? 148 * 8 : (g < 1 ? 1 : g)), 256, 0, s>>>(p, gmask, passes);
  return cudaGetLastError();
}

size_t seg_smem_bytes(int rows_per, int wpr) {
  // predicate + mask planes (odd row stride), plus two activity-flag arrays (at most one byte per row and word column)
  return ((size_t)rows_per + (size_t)rows_per + 2 * SEG_HALO) * seg_row_stride(wpr) * 4 + 2 * (size_t)rows_per * wpr;
}

cudaError_t launch_segment(const SegParams& p, cudaStream_t s) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(p.n * p.cluster);
  cfg.blockDim = dim3(SEG_THREADS);
  cfg.dynamicSmemBytes = seg_smem_bytes(p.rows_per, p.wpr);
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, k23_segment, p);
}

// =================================================================================================
// helpers for the per-operator entry points
// =================================================================================================
__global__ void k_threshold_pack(const float* __restrict__ img, int n, int w, int h, float lo, float hi,
                                 uint32_t* __restrict__ bits) {
  const int wpr = (w + 31) >> 5;
  const long long total = (long long)n * h * wpr * 32;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long word = i >> 5;
    const int x = (int)(word % wpr) * 32 + (int)(i & 31);
    const long long row = word / wpr;
    bool in = false;
    if (x < w) { const float v = img[row * w + x]; in = v >= lo && v <= hi; }
    const uint32_t b = __ballot_sync(0xffffffffu, in);
    if ((i & 31) == 0) bits[word] = b;
  }
}

__global__ void k_pack_u8(const uint8_t* __restrict__ mask, int n, int w, int h, uint32_t* __restrict__ bits) {
  const int wpr = (w + 31) >> 5;
  const long long total = (long long)n * h * wpr * 32;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long word = i >> 5;
    const int x = (int)(word % wpr) * 32 + (int)(i & 31);
    const long long row = word / wpr;
    const bool in = x < w && mask[row * w + x] == 1;
    const uint32_t b = __ballot_sync(0xffffffffu, in);
    if ((i & 31) == 0) bits[word] = b;
  }
}

static int grid_for(long long total) {
  long long g = (total + 255) / 256;
  return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g));
}

cudaError_t launch_threshold_pack(const float* img, int n, int w, int h, float lo, float hi, uint32_t* bits, cudaStream_t s) {
  const long long total = (long long)n * h * ((w + 31) >> 5) * 32;
  k_threshold_pack<<<grid_for(total), 256, 0, s>>>(img, n, w, h, lo, hi, bits);
  return cudaGetLastError();
}

cudaError_t launch_pack_u8(const uint8_t* mask, int n, int w, int h, uint32_t* bits, cudaStream_t s) {
  const long long total = (long long)n * h * ((w + 31) >> 5) * 32;
  k_pack_u8<<<grid_for(total), 256, 0, s>>>(mask, n, w, h, bits);
  return cudaGetLastError();
}

cudaError_t configure_kernels() {
  cudaError_t e;
  int median_window, sharpen_half;
  for (median_window = 3; median_window <= 15; median_window += 2)
    for (sharpen_half = 1; sharpen_half <= 4; ++sharpen_half) {
      e = [&]() -> cudaError_t { NM03_K1_DISPATCH(config_k1) }();
      if (e != cudaSuccess) return e;
    }
  e = cudaFuncSetAttribute(k23_segment, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 64);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k2_band_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 64);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k23_segment, cudaFuncAttributeNonPortableClusterSizeAllowed, 0);
}

}  // namespace nm03
