/*
 * nm03.h — C ABI of the B200-native per-slice brain-MRI pipeline (libnm03.so).
 *
 * This is the drop-in boundary for the ONE hot path of calebhabesh/NM03-Capstone-Project:
 * processSingleImage() — import → IntensityNormalization → IntensityClipping → VectorMedianFilter →
 * ImageSharpening → SeededRegionGrowing (adaptive seeds) → ImageCaster → Dilation (+ Erosion in the
 * test path).  Reference call sites: src/sequential/main_sequential.cpp:170-272,
 * src/parallel/main_parallel.cpp:66-170, src/test/test_pipeline.cpp:33-125.  In the reference each
 * stage is a FAST ProcessObject (`T::create(..)->connect(..)->update()`); the entry points below are
 * what a binding of that operator API for this path calls instead.  The C++17 shim
 * (nm03-capstone-project_b200/host/nm03.hpp) re-creates the FAST-shaped classes on top of this ABI.
 *
 * Conventions: plain pointers and sizes; every function returns 0 (NM03_OK) or a negative
 * nm03_status; nothing throws across the ABI; nm03_last_error() gives a thread-local message.
 * All buffers are caller-owned.  Images are row-major, one channel.  Entry points taking host
 * buffers copy to/from the GPU themselves; `_device` entry points take device pointers that live
 * on the context's GPU.  A context may be used from many host threads at once (the reference runs
 * 16 OpenMP threads, main_parallel.cpp:336-343,401): calls are serialised per stream slot inside.
 * There is no CPU fallback: without a CUDA device nm03_init fails.
 */
#ifndef NM03_H
#define NM03_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum nm03_status {
  NM03_OK = 0,
  NM03_ERR_INVALID = -1,       /* bad argument (null pointer, negative size ...) */
  NM03_ERR_TOO_SMALL = -2,     /* width or height < 100: main_sequential.cpp:189-192 */
  NM03_ERR_PARAM = -3,         /* FAST would throw: even / <=2 morphology size, even median window ... */
  NM03_ERR_NO_SEEDS = -4,      /* SeededRegionGrowing without seed points */
  NM03_ERR_SEED_OOB = -5,      /* seed point outside the image */
  NM03_ERR_UNSUPPORTED = -6,   /* valid in FAST, not implemented here (stated in the message) */
  NM03_ERR_TOO_LARGE = -7,     /* batch or slice exceeds a fixed capacity (built-in scratch of the _noalloc entry point; rows too wide for the band fallback) */
  NM03_ERR_CUDA = -8,          /* CUDA runtime error (message has the detail) */
  NM03_ERR_NO_DEVICE = -9,
  NM03_ERR_NOMEM = -10,
  NM03_ERR_BUSY = -11          /* nm03_process_batch_async: every stream slot of the GPU holds an unfinished ticket; wait for one first */
} nm03_status;

/* switch values for the details of FAST that cannot be verified offline */
enum { NM03_BORDER_CLAMP = 0, NM03_BORDER_ZERO = 1 };
enum { NM03_RG_OPENCL = 0, NM03_RG_HOST = 1 };
enum { NM03_SE_DISC = 0, NM03_SE_SQUARE = 1 };
enum { NM03_MORPH_DILATE = 0, NM03_MORPH_ERODE = 1 };

/* Every literal of the reference pipeline in one POD (the reference hard-codes them). */
typedef struct nm03_params {
  float norm_low, norm_high, norm_min, norm_max; /* IntensityNormalization::create(0.5f,2.5f,0.0f,10000.0f) main_sequential.cpp:195-196 */
  float clip_min, clip_max;                      /* IntensityClipping::create(0.68f,4000.0f)     main_sequential.cpp:200 */
  int median_window;                             /* VectorMedianFilter::create(7)                main_sequential.cpp:204; any odd size 3..255 (above 15: two passes, clamp-to-edge border only) */
  float sharpen_gain, sharpen_sigma;             /* ImageSharpening::create(2.0f,0.5f,9)         main_sequential.cpp:208 */
  int sharpen_mask;
  float rg_min, rg_max;                          /* SeededRegionGrowing::create(0.74f,0.91f,..)  main_sequential.cpp:232-233 */
  int dilate_size;                               /* Dilation::create(3)                          main_sequential.cpp:250 */
  int erode_size;                                /* Erosion::create(3) test_pipeline.cpp:119; 0 = off */
  int border_mode;                               /* NM03_BORDER_*: sampler addressing of median + sharpen */
  int sharpen_fma;                               /* 0 = multiply then add, 1 = fused multiply-add */
  int rg_mode;                                   /* NM03_RG_* */
  int rg_connectivity;                           /* 4 or 8 */
  int rg_border_expand;                          /* 1 = image-border pixels expand, 0 = they do not */
  int se_shape;                                  /* NM03_SE_* */
  int input_unsigned;                            /* 0 = int16 pixels, 1 = uint16 pixels */
} nm03_params;

/* Greys and border radius of the result's rendering for nm03_outputs.export_seg_blocks: SegmentationRenderer::create(
 * labelColors, 0.6f, 1.0f, 2) over black (main_sequential.cpp:261) gives background 0, interior 153, border 255, radius 2. */
typedef struct nm03_export_style {
  uint8_t background, interior, border;
  int border_radius;                             /* 0..7 */
} nm03_export_style;

/* Optional per-slice outputs of nm03_process_batch*; any pointer may be NULL.  Each is a dense
 * [n][h][w] array (host pointers for nm03_process_batch, device pointers for the _device variant). */
typedef struct nm03_outputs {
  uint8_t* dilated;    /* final result: dilation->getOutputData<Image>(0)  main_parallel.cpp:161 */
  uint8_t* eroded;     /* erosion branch off the caster, test_pipeline.cpp:119-121 (needs erode_size>0) */
  uint8_t* rg_mask;    /* regionGrowing output, {0,1} */
  float* sharpened;    /* sharpen output (test_pipeline.cpp:71-73) */
  float* median;       /* medianfilter output */
  int32_t* rg_iterations; /* [n] passes the flood fill needed to converge */
  /* The final result once more, bit-packed: [n][h][ceil(w/32)] 32-bit words, bit x%32 of word x/32 of a row = pixel x
   * (unused high bits of the last word are 0).  No counterpart in the reference (FAST images are one byte per label);
   * it is the form the segmentation kernel holds the mask in, and it crosses the host link at 1/8 of the bytes of
   * `dilated`.  May be requested with or without `dilated`. */
  uint32_t* dilated_bits;
  /* Host export helper (no counterpart in the reference, whose renderers and JPEG writer live in FAST/Qt):
   * the grey rendering of every INPUT slice — ImageRenderer's window/level on a canvas of the slice's own size,
   * main_sequential.cpp:258-259 — level-shifted, transformed and quantised per 8x8 block for a baseline JPEG, so that the
   * host is left with the entropy coding (host/jpeg.hpp encode_blocks gives the bytes of its own encode()).
   * export_blocks: [n][(h/8)*(w/8)][64] quantised coefficients in zigzag order, blocks in raster order; export_nz:
   * [n][(h/8)*(w/8)] 64-bit maps, bit k set iff coefficient k is non-zero; export_rq: 64 floats on the HOST (also for the
   * _device variant), the reciprocal quantiser in natural order including the transform's scaling (jpeg::make_tables).
   * All three or none; w and h must be multiples of 8 (NM03_ERR_UNSUPPORTED otherwise); not offered by
   * nm03_process_batch_device_noalloc. */
  int16_t* export_blocks;
  uint64_t* export_nz;
  const float* export_rq;
  /* The same for the rendering of the RESULT (SegmentationRenderer of the dilated mask on a canvas of the slice's own
   * size, main_sequential.cpp:261-262): blocks and maps laid out as above, quantised with export_rq; export_seg_style (on
   * the HOST) holds the three greys and the border radius.  Needs dilated_bits, export_rq, export_seg_nz and the style. */
  int16_t* export_seg_blocks;
  uint64_t* export_seg_nz;
  const nm03_export_style* export_seg_style;
} nm03_outputs;

typedef struct nm03_ctx nm03_ctx;

/* ---- lifetime ------------------------------------------------------------------------------- */
/* Replaces FAST's implicit device manager (first ProcessObject::update()).  gpu_ids: CUDA device
 * ordinals, one worker (streams + pinned staging + scratch) per entry. */
int nm03_init(const int* gpu_ids, int n_gpus, nm03_ctx** out);
void nm03_shutdown(nm03_ctx* ctx);
const char* nm03_last_error(void);
int nm03_version(void);   /* 102: nm03_outputs.export_* */
int nm03_gpu_count(const nm03_ctx* ctx);
/* Number of CUDA devices visible to the process (0 if none); lets a driver build the gpu_ids list for
 * nm03_init.  No counterpart in the reference, which uses whatever single device FAST picks. */
int nm03_device_count(void);

/* Literals of main_sequential.cpp:196,200,204,208,233,250 (erosion off, default switches). */
void nm03_default_params(nm03_params* p);

/* Adaptive dimension-based seeds: main_sequential.cpp:214-241 / main_parallel.cpp:119-148 /
 * test_pipeline.cpp:79-106.  Writes (x,y) pairs, returns the count (the needed count if cap is
 * too small), NM03_ERR_TOO_SMALL if w or h < 10 (the reference's loop would not terminate). */
int nm03_make_seeds(int w, int h, int* xy, int cap);

/* ---- whole pipeline --------------------------------------------------------------------------- */
/* processSingleImage() for n slices of one size, host buffers in and out (main_parallel.cpp:66-170
 * applied to a batch, main_parallel.cpp:336-343).  slices: [n][h][w] int16 (or uint16 if
 * params->input_unsigned).  seeds_xy may be NULL = the adaptive seeds of nm03_make_seeds(w,h).
 * The call returns when every requested output is in the caller's buffers.  A lone caller gets its batch cut into a
 * few chunks that are pipelined over the stream slots (copy in / kernels / copy out overlap); when other host-buffer
 * calls are in flight on the same GPU the batch goes through whole, because the calls already overlap each other.
 * status (optional, [n]): every cause of failure here is a property of the whole call (size below 100, a parameter
 * FAST would reject, a seed outside the image, a CUDA error), so all n entries receive the call's return value.  The
 * reference's catch-and-continue (main_parallel.cpp:163-167) therefore happens at the granularity the caller batches
 * at: a driver that wants a bad slice not to take others with it groups slices by size, as nm03_driver does. */
int nm03_process_batch(nm03_ctx* ctx, int gpu, const void* slices, int n, int w, int h, const nm03_params* params,
                       const int* seeds_xy, int n_seeds, const nm03_outputs* out, int* status);

/* The same work without waiting for it (SURVEY §8b "async variant with stream ticket"): the copies and kernels of the
 * batch are queued on one of the GPU's 8 stream slots and the call returns a ticket; nm03_wait blocks until the
 * outputs are in the caller's buffers, reports the result and frees the ticket (also after an error).  Input and
 * output buffers must stay valid and untouched until then, and should be pinned (nm03_host_alloc) — with pageable
 * memory the copies fall back to staged, partly synchronous transfers.  One host thread can keep several batches in
 * flight this way (three saturate the host link), where the reference needs its 16 OpenMP threads
 * (main_parallel.cpp:336-343,401) to overlap anything.  At most 8 tickets per GPU can be outstanding: with all
 * slots taken the call returns NM03_ERR_BUSY instead of blocking (the caller may hold every ticket itself).  The blocking
 * entry points, by contrast, wait for a free slot: a thread that holds all eight tickets of a GPU must wait for one before
 * it calls them on that GPU.
 * nm03_wait may be called from any thread, once per ticket; every ticket must be waited for before nm03_shutdown (a ticket
 * holds one of the context's stream slots, and an unwaited one is leaked). */
typedef struct nm03_ticket nm03_ticket;
int nm03_process_batch_async(nm03_ctx* ctx, int gpu, const void* slices, int n, int w, int h, const nm03_params* params,
                             const int* seeds_xy, int n_seeds, const nm03_outputs* out, nm03_ticket** ticket);
int nm03_wait(nm03_ticket* ticket);

/* Same, but every buffer is already in the memory of GPU `gpu`; work is queued on `stream`
 * (a cudaStream_t passed as void*, NULL = the worker's own stream) and the call does not wait.
 * The scratch between the two kernels belongs to one of 8 slots per GPU; a calling stream keeps the
 * slot it used last, so a caller that rotates its batches over up to 8 streams gets them overlapped on
 * the GPU (the fill kernel of one batch under the median kernel of the next); with more streams
 * than slots a slot changes hands in stream order (event wait), never concurrently. */
int nm03_process_batch_device(nm03_ctx* ctx, int gpu, const void* d_slices, int n, int w, int h,
                              const nm03_params* params, const int* seeds_xy, int n_seeds,
                              const nm03_outputs* d_out, void* stream);

/* Whole pipeline on device buffers for callers that must not allocate.  This entry point has no
 * counterpart in the reference (FAST allocates freely inside every update()).
 *
 * Who needs it: a host thread that feeds one GPU stream while other streams of the same process are
 * busy.  cudaMalloc and cudaFree are not ordinary calls there: they take the driver's allocator lock
 * and wait for the whole device, so one allocation in a feeder thread stalls every other feeder's
 * copies and kernels.  The same holds for a process whose device memory has been handed out in full
 * to a pool owned by another component (an inference server, a framework's caching allocator): a
 * late cudaMalloc there simply fails.  Such callers cannot use nm03_process_batch_device, whose
 * predicate bit-plane (the scratch that carries the range test from the preprocess kernel to the
 * region-growing kernel) is grown with cudaMalloc on demand.
 *
 * Here that scratch is a fixed array inside the library image: NM03_NOALLOC_MAX_PRED_WORDS 32-bit words
 * (1 MiB) of device memory on every GPU that loads libnm03.so, held until the process exits.  The
 * call itself performs no cudaMalloc, cudaFree, stream/event creation or any other allocating
 * routine; inputs and outputs are the caller's device buffers, exactly as for
 * nm03_process_batch_device, and the work is queued on `stream` (NULL = the default stream) without
 * waiting for it.
 *
 * Limit: n * h * ceil(w/32) must not exceed NM03_NOALLOC_MAX_PRED_WORDS (32 slices of 512x512, 8 of
 * 1024x1024); larger batches are refused with NM03_ERR_TOO_LARGE and must be split by the caller.
 *
 * First call: the first call on each GPU looks up the device address of the array (with that GPU
 * current) and caches it.  Make that first call during start-up, before any stream capture begins.
 *
 * Sharing: every call on a GPU uses the same array, and it keeps whatever the previous call left.
 * Nothing in it has to start at zero — the preprocess kernel overwrites every word the
 * region-growing kernel will read — so no clearing pass is issued.  Two calls whose kernels can
 * overlap on the GPU (different streams without an ordering between them, or different host
 * threads) race on the array and produce wrong masks; order such calls on one stream or with
 * events. */
#define NM03_NOALLOC_MAX_PRED_WORDS (256 * 1024)
int nm03_process_batch_device_noalloc(nm03_ctx* ctx, int gpu, const void* d_slices, int n, int w, int h,
                                      const nm03_params* params, const int* seeds_xy, int n_seeds,
                                      const nm03_outputs* d_out, void* stream);

/* Pinned host memory for zero-copy staging of nm03_process_batch inputs/outputs. */
int nm03_host_alloc(void** p, size_t bytes);
int nm03_host_free(void* p);

/* ---- per-operator entry points (host buffers), one per FAST operator on the path ---------------- */
/* IntensityNormalization + IntensityClipping + VectorMedianFilter + ImageSharpening
 * (main_sequential.cpp:195-210) as ONE fused kernel.  Outputs optional: median (after median
 * filter), sharpened, predicate (bit-plane, ceil(w/32) words per row, bit x%32 of word x/32 = 1 iff
 * rg_min <= sharpened <= rg_max). */
int nm03_preprocess(nm03_ctx* ctx, int gpu, const void* slice, int w, int h, const nm03_params* params,
                    float* median, float* sharpened, uint32_t* predicate);
/* SeededRegionGrowing::create(lo,hi,seeds)->update()  (main_sequential.cpp:232-243). */
int nm03_region_grow(nm03_ctx* ctx, int gpu, const float* image, int w, int h, const nm03_params* params,
                     const int* seeds_xy, int n_seeds, uint8_t* mask, int* iterations);
/* Dilation::create(size) / Erosion::create(size) on a uint8 {0,1} image (main_sequential.cpp:246-252,
 * test_pipeline.cpp:119-125); ImageCaster(TYPE_UINT8) in front of them is the identity. */
int nm03_morphology(nm03_ctx* ctx, int gpu, const uint8_t* mask, int w, int h, int size, int se_shape, int op,
                    uint8_t* out);

/* ---- introspection for benchmarks ------------------------------------------------------------- */
/* Number of kernel launches this library has issued on `gpu` since nm03_init. */
long long nm03_kernel_launches(const nm03_ctx* ctx, int gpu);
/* Per-kernel device timing of the whole-pipeline entry points (CUDA events on the launching stream,
 * recorded around each kernel).  No counterpart in the reference, which ships no profiling hooks
 * (README.md:92-96 names external tools only).  nm03_profile_read waits for the recorded work,
 * returns the summed milliseconds of the preprocess and segment kernels and the number of
 * pipeline calls since the last read, and resets the sums. */
int nm03_profile_enable(nm03_ctx* ctx, int gpu, int on);
int nm03_profile_read(nm03_ctx* ctx, int gpu, double* preprocess_ms, double* segment_ms, long long* calls);
/* Work counter of the preprocess kernel, for the ALU roofline of benchmarks (no counterpart in the reference):
 * the number of median work items (one thread computing a run of adjacent outputs with the generated min/max
 * network) the kernel actually sorted on `gpu` since the last call — windows that take the clip-floor shortcut are
 * not counted.  Waits for the GPU, returns the count and resets it.  nm03_median_network_ops(window) = min/max
 * instructions one such work item executes (0 for a window without a generated network). */
long long nm03_profile_work(nm03_ctx* ctx, int gpu);
int nm03_median_network_ops(int median_window);

#ifdef __cplusplus
}
#endif
#endif /* NM03_H */
