/*
 * nm03_oracle.h — CPU ORACLE for the per-slice brain-MRI pipeline.   ** TEST INFRASTRUCTURE ONLY **
 *
 * This is a plain C++17 restatement of the arithmetic that the FAST framework performs for the
 * reference's hot path (reference: src/sequential/main_sequential.cpp:175-252,
 * src/parallel/main_parallel.cpp:78-161, src/test/test_pipeline.cpp:33-125).  It exists so that the
 * CUDA path can be checked bit-for-bit; nothing in the product links, imports or calls it.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * PARITY UNPINNED.  The arithmetic lives in the third-party FAST library (github.com/smistad/FAST,
 * pulled in by `find_package(FAST REQUIRED)`, reference CMakeLists.txt:9; no version is pinned, the
 * T::create() factory style bounds it to FAST >= 4.x).  FAST's source is not under /root/reference
 * and is not installable offline, and the reference ships no golden vectors, known-answer tests or
 * fixtures for this path (its only "test" is a GUI prototype judged by eye).  The oracle therefore
 * follows the reference's call sites and literals plus FAST's published per-filter semantics as
 * recorded in SURVEY.md §8c; every detail that cannot be verified offline is a switch below, and
 * the CUDA path is tested against the oracle under every switch combination.  Secondary pins that
 * do exist: scipy.ndimage (median/label/binary morphology) cross-checks and hand-derived
 * known-answer vectors in tests/.
 *
 * All floating point is fp32, round-to-nearest-even, built with -ffp-contract=off.
 */
#ifndef NM03_ORACLE_H
#define NM03_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* switches for the details SURVEY.md §8c marks MED/LOW confidence (defaults marked *) */
enum { NM03O_BORDER_CLAMP = 0 /* * */, NM03O_BORDER_ZERO = 1 };
enum { NM03O_RG_OPENCL = 0 /* * */, NM03O_RG_HOST = 1 };
enum { NM03O_SE_DISC = 0 /* * */, NM03O_SE_SQUARE = 1 };

typedef struct nm03o_params {
  /* IntensityNormalization::create(0.5f, 2.5f, 0.0f, 10000.0f)   main_sequential.cpp:195-196 */
  float norm_low, norm_high, norm_min, norm_max;
  /* IntensityClipping::create(0.68f, 4000.0f)                     main_sequential.cpp:200 */
  float clip_min, clip_max;
  /* VectorMedianFilter::create(7)                                 main_sequential.cpp:204 */
  int median_window;
  /* ImageSharpening::create(2.0f, 0.5f, 9)                        main_sequential.cpp:208 */
  float sharpen_gain, sharpen_sigma;
  int sharpen_mask;
  /* SeededRegionGrowing::create(0.74f, 0.91f, seeds)              main_sequential.cpp:232-233 */
  float rg_min, rg_max;
  /* Dilation::create(3) main_sequential.cpp:250 ; Erosion::create(3) test_pipeline.cpp:119 (0 = off) */
  int dilate_size, erode_size;
  /* oracle switches */
  int border_mode;      /* NM03O_BORDER_*  sampler addressing of median and sharpen */
  int sharpen_fma;      /* 0* = mul then add, 1 = fmaf(mask, v, sum) */
  int rg_mode;          /* NM03O_RG_* */
  int rg_connectivity;  /* 4* or 8 */
  int rg_border_expand; /* 1* = image-border pixels expand like any other, 0 = they do not */
  int se_shape;         /* NM03O_SE_* */
  int input_unsigned;   /* 0* = int16, 1 = uint16 */
} nm03o_params;

void nm03o_default_params(nm03o_params* p);

/* a7: adaptive seeds (main_sequential.cpp:214-241).  Writes x,y pairs; returns the count, or the
 * needed count if cap is too small, or -1 if w or h < 10 (the reference loop would not terminate). */
int nm03o_make_seeds(int w, int h, int* xy, int cap);

/* per-operator functions; all return 0 on success, <0 on the error FAST would throw */
int nm03o_normalize(const void* in, int is_unsigned, long n, float low, float high, float mn, float mx, float* out);
int nm03o_clip(const float* in, long n, float lo, float hi, float* out);
int nm03o_median(const float* in, int w, int h, int k, int border_mode, float* out);
int nm03o_gauss_mask(float sigma, int m, float* mask /* m*m */);
int nm03o_sharpen(const float* in, int w, int h, float gain, float sigma, int m, int border_mode, int use_fma, float* out);
int nm03o_region_grow(const float* in, int w, int h, float lo, float hi, const int* seeds_xy, int nseeds,
                      int mode, int conn, int border_expand, uint8_t* out, int* iterations);
int nm03o_dilate(const uint8_t* in, int w, int h, int size, int se_shape, uint8_t* out);
int nm03o_erode(const uint8_t* in, int w, int h, int size, int se_shape, uint8_t* out);

/* optional outputs may be NULL */
typedef struct nm03o_outputs {
  float* median;     /* after normalise+clip+median, w*h */
  float* sharpened;  /* w*h */
  uint8_t* rg_mask;  /* region growing result {0,1} */
  uint8_t* dilated;  /* final result (main_sequential.cpp:250-252) */
  uint8_t* eroded;   /* test path (test_pipeline.cpp:119-121), needs erode_size>0 */
  int* rg_iterations;
} nm03o_outputs;

/* whole pipeline for one slice, stage by stage exactly as processSingleImage() does.
 * Returns 0, or <0: -2 = dimensions too small (main_sequential.cpp:189-192). */
int nm03o_pipeline(const void* in, int w, int h, const nm03o_params* p, const nm03o_outputs* out);

/* CPU baseline structure: restatement of main_parallel.cpp:330-347 — batches of `batch` slices,
 * `#pragma omp parallel for schedule(auto) reduction(+:ok)` over the batch with `threads` threads.
 * dilated: n*w*h bytes.  Returns the success count. */
int nm03o_run_batch_omp(const void* slices, int n, int w, int h, const nm03o_params* p, int batch, int threads,
                        uint8_t* dilated);
int nm03o_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
