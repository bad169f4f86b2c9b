// nm03_device.h — parameter blocks shared by the kernels (nm03_kernels.cu) and the C-ABI host
// code (nm03_api.cu).  Everything is passed to the kernels by value (kernel parameter space).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nm03 {

// ---- K1: fused normalise + clip + median + sharpen + range predicate --------------------------
// Output tile of one CTA.  The median is needed on the tile plus the sharpen halo (HS_MAX each
// side); raw samples on that plus the median radius.
#ifndef NM03_K1_TH
#define NM03_K1_TH 64
#endif
#ifndef NM03_K1_THREADS
#define NM03_K1_THREADS 256
#endif
#ifndef NM03_K1_MIN_CTAS
#define NM03_K1_MIN_CTAS 2
#endif
#ifndef NM03_K1_TW
#define NM03_K1_TW 128
#endif
constexpr int K1_TW = NM03_K1_TW;                  // multiple of 32 (one predicate word = 32 pixels)
constexpr int K1_TH = NM03_K1_TH;
constexpr int K1_HS_MAX = 4;                      // sharpen mask up to 9x9
constexpr int K1_MW = K1_TW + 2 * K1_HS_MAX;      // 136 median columns
constexpr int K1_MH = K1_TH + 2 * K1_HS_MAX;      // 72 median rows
constexpr int K1_MHP = K1_MH / 2;                 // 36 packed median rows (row p and row p+36 share a register)
constexpr int K1_PWS = K1_MW + 8;                 // packed raw tile row stride in words (>= MW + 2R + OFF for R <= 4, multiple of 8)
static_assert(K1_TW % 32 == 0 && K1_TH % 8 == 0, "tile shape");
// windows above 9x9 need a wider halo: their packed tile has a longer row and their TMA box starts further left
__host__ __device__ constexpr int k1_pws(int K) { return K <= 9 ? K1_PWS : K1_MW + 24; }
__host__ __device__ constexpr int k1_xbox(int K) { return K <= 9 ? 8 : 16; }
constexpr int K1_THREADS = NM03_K1_THREADS;
constexpr int K1_MIN_CTAS = NM03_K1_MIN_CTAS;                   // resident CTAs per SM the register budget is set for

struct PreParams {
  const void* in;        // [n][h][w] int16 / uint16
  float* median;         // optional [n][h][w]
  float* sharpened;      // optional [n][h][w]
  uint32_t* predicate;   // optional [n][h][wpr] bit-plane
  int w, h, n, wpr;
  int tiles_x, tiles_y;
  float norm_min, norm_range, norm_scale, norm_low;  // t=(v-min)/range ; t*scale+low
  float clip_min, clip_max;
  float gain, rg_min, rg_max;
  int border_zero, use_fma, input_unsigned;
  unsigned int* sched;   // optional [4]: work counter + finished-CTA counter of the persistent grid (zero between launches),
                         // then a 64-bit running count of the median work items that were sorted (nm03_profile_work)
  const float* lut;      // optional [65536]: f(key) for every 16-bit key (bit-identical to the arithmetic), else null
  // windows whose samples all have key <= skip_key map to f() == skip_val, so their median is known
  // without sorting (clip floor; 0 = off).  skip_key2 = the key replicated in both 16-bit halves.
  int skip_on;
  uint32_t skip_key2;
  float skip_val;
  int use_tma;           // stage the raw halo tile with a TMA bulk tensor copy (needs w % 8 == 0, 16 B aligned input)
  // launch order: block b works on tile tile_order[b / n] of slice b % n, central (busiest) tiles first,
  // so the last, partially filled wave holds the cheap corner tiles (tile_order_on = 0: identity)
  int tile_order_on;
  uint16_t tile_order[256];
  float mask[81];        // normalised Gaussian, index x+hs + (y+hs)*m, m = 2*hs+1
  // fast predicate path (no sharpened plane requested): separable blur with taps gx[-he..he] and a
  // proven bound eps = ea*|blur| + eb*|v| + ec on |sharpened_fast - sharpened_exact|; pixels closer
  // than eps to a threshold are recomputed in the oracle's exact order.
  float gx[9];
  float ea, eb, ec;
  float eps_max;                 // upper bound of eps over every |blur|, |v| these parameters can produce
  float gyn[9];                  // vertical taps of the fast path scaled by -gain: fl(-gain * gx[t])
  float gain1;                   // fl(1 + gain)
  float ean;                     // ea / |gain|: the band in terms of |gain * blur| (what the fast path holds)
  float rg_mid, rg_half;         // fast predicate: inside [rg_min, rg_max]  <=>  |s - rg_mid| - rg_half <= 0 (up to the band)
};

enum { K1_EXACT = 0, K1_FAST_FULL = 1, K1_FAST_HE2 = 2 };  // sharpen evaluation mode

// ---- K2+K3: flood fill + morphology on bit-packed planes, one thread-block cluster per slice ----
#ifndef NM03_SEG_THREADS
#define NM03_SEG_THREADS 256
#endif
constexpr int SEG_THREADS = NM03_SEG_THREADS;
constexpr int SEG_HALO = 7;            // rows of the neighbouring CTA's mask kept for the SE (size <= 15)
constexpr int SEG_MAX_SEEDS = 512;
constexpr int SEG_MAX_CLUSTER = 8;

struct SegParams {
  const uint32_t* pred;  // [n][h][wpr]: range predicate (do_fill=1) or the mask itself (do_fill=0)
  uint8_t* dilated;      // optional [n][h][w]
  uint32_t* dilated_bits;  // optional [n][h][wpr]: the dilated mask bit-packed (bit x%32 of word x/32)
  uint8_t* eroded;       // optional
  uint8_t* rg_mask;      // optional
  int32_t* iterations;   // optional [n]
  int w, h, n, wpr;
  int cluster;           // CTAs per slice; 0 = the slice exceeds cluster capacity: global-memory fallback, rows_per = band height
  int rows_per;          // rows owned by one CTA
  int seg_rows;          // rows swept by one thread per pass
  int do_fill, conn8, border_expand, host_mode;
  int dil_n, ero_n;      // SE radius N = size/2 (0 = stage off)
  int8_t dil_hw[16];     // horizontal half-width of the SE at row offset b (index b + N)
  int8_t ero_hw[16];
  int n_seeds;
  short2 seeds[SEG_MAX_SEEDS];
};

size_t seg_smem_bytes(int rows_per, int wpr);

// launchers (return cudaError_t of the launch)
cudaError_t launch_preprocess(const PreParams& p, int median_window, int sharpen_half, int mode, cudaStream_t s);
// windows above K1_MAX_WINDOW: window median of the raw 16-bit samples (clamp-to-edge) into `out` [n][h][w], same sample
// type as the input; K1 then runs on that plane with the 1x1 window (median_window = 1 above)
constexpr int K1_MAX_WINDOW = 15;      // largest window K1 evaluates itself (tile halo)
constexpr int BIG_MAX_WINDOW = 255;    // largest window of the two-pass path
cudaError_t launch_median_big(const void* in, int n, int w, int h, int window, int input_unsigned, void* out, cudaStream_t s);
cudaError_t launch_segment(const SegParams& p, cudaStream_t s);
// global-memory fallback (SegParams::cluster == 0): band-wise fill pass over the bit-packed mask, then morphology
cudaError_t launch_band_fill(const SegParams& p, uint32_t* gmask, int* gflag, int first_launch, cudaStream_t s);
cudaError_t launch_global_morph(const SegParams& p, const uint32_t* gmask, int passes, cudaStream_t s);
cudaError_t launch_threshold_pack(const float* img, int n, int w, int h, float lo, float hi, uint32_t* bits,
                                  cudaStream_t s);
cudaError_t launch_pack_u8(const uint8_t* mask, int n, int w, int h, uint32_t* bits, cudaStream_t s);
cudaError_t launch_build_lut(const PreParams& p, float* lut, cudaStream_t s);  // fills lut[65536] with f(key)
int median_network_ops(int median_window);   // min/max instructions of one generated median work item (0 = no network)
cudaError_t configure_kernels();  // opt in to large dynamic shared memory; once per device

}  // namespace nm03
