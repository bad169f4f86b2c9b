// nm03_api.cu — the extern "C" boundary declared in include/nm03.h.
//
// One `Worker` per GPU holds a small pool of stream slots; a slot owns a stream, an event and the
// scratch bit-plane that carries the range predicate from K1 to K23.  Host threads (the reference
// runs 16, main_parallel.cpp:401) take a slot, queue copies + two kernels on it and release it;
// slots make the context re-entrant without any global lock on the data path.
#include "../../include/nm03.h"
#include "nm03_device.h"
#include "nm03_export.h"

#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <vector>

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) return fail(NM03_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

constexpr int kSlots = 8;          // stream slots per GPU
const bool g_chunks_forced = getenv("NM03_CHUNKS") != nullptr;
const int kChunksPerCall = [] { const char* e = getenv("NM03_CHUNKS"); const int v = e ? atoi(e) : 0; return v >= 1 && v <= 8 ? v : 4; }();  // a host-buffer batch is pipelined over this many of them (NM03_CHUNKS: tuning)

const bool g_no_tma = [] { const char* e = getenv("NM03_NO_TMA"); return e && e[0] == '1'; }();  // A/B measurements
const bool g_no_tile_order = [] { const char* e = getenv("NM03_NO_TILE_ORDER"); return e && e[0] == '1'; }();
const bool g_no_dynamic = [] { const char* e = getenv("NM03_NO_DYNAMIC"); return e && e[0] == '1'; }();
const bool g_no_skip = [] { const char* e = getenv("NM03_NO_SKIP"); return e && e[0] == '1'; }();  // A/B measurements
const bool g_force_global = [] { const char* e = getenv("NM03_FORCE_GLOBAL_FILL"); return e && e[0] == '1'; }();  // tests
const int g_seg_rows = [] { const char* e = getenv("NM03_SEG_ROWS"); return e ? atoi(e) : 0; }();
const int g_force_cluster = [] { const char* e = getenv("NM03_CLUSTER"); return e ? atoi(e) : 0; }();

// NM03_FORCE_EXACT_SHARPEN=1 makes K1 evaluate the unsharp mask exactly everywhere even when only the
// predicate is needed (A/B measurements; results are identical by construction)
// NM03_SPIN_WAIT=1: host waits spin on the GPU's completion flag instead of sleeping until the driver wakes the
// thread (lower wake-up latency for a lone caller, one busy core per waiting thread)
const bool g_spin_wait = [] { const char* e = getenv("NM03_SPIN_WAIT"); return e && e[0] == '1'; }();

// NM03_SPLIT=1: a device-resident batch of a lone single-stream caller is spread over library streams (fork / join on
// the caller's stream).  Off by default: measured slower than leaving the batch whole (see nm03_process_batch_device).
const bool g_no_split = [] { const char* e = getenv("NM03_SPLIT"); return !(e && e[0] == '1'); }();
const int g_split_parts = [] { const char* e = getenv("NM03_SPLIT_PARTS"); const int v = e ? atoi(e) : 0; return v >= 2 && v <= 7 ? v : 4; }();  // tuning

const bool g_force_exact = [] { const char* e = getenv("NM03_FORCE_EXACT_SHARPEN"); return e && e[0] == '1'; }();

struct Slot {
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;     // recorded after the last use of this slot's scratch
  cudaEvent_t fork = nullptr;     // recorded on a caller's stream before parts of its batch move to library streams
  std::atomic<bool> used{false};
  std::atomic<cudaStream_t> last_stream{nullptr};  // the stream `done` was last recorded on (both read by other callers' slots)
  std::atomic<cudaStream_t> forked_from{nullptr};  // caller stream whose batch this slot last carried a part of (null: none)
  uint32_t* pred = nullptr;       // scratch: predicate bit-plane [n][h][wpr]
  size_t pred_bytes = 0;
  void* stage_in = nullptr;       // scratch: device copy of host inputs
  size_t stage_in_bytes = 0;
  uint32_t* gmask = nullptr;      // scratch: bit-packed mask of the global-memory fill fallback
  size_t gmask_bytes = 0;
  int* dflag = nullptr;           // scratch: device "changed" flag of that fallback
  size_t dflag_bytes = 0;
  int* hflag = nullptr;           // pinned host copy of the flag
  void* stage_out = nullptr;      // scratch: device copy of outputs before D2H
  size_t stage_out_bytes = 0;
  unsigned int* sched = nullptr;  // scratch: K1's work counters (zero between launches; the kernel resets them)
  size_t sched_bytes = 0;
  float* lut = nullptr;           // scratch: f(key) table of K1 for the parameters in lut_key
  size_t lut_bytes = 0;
  void* bigmed = nullptr;         // scratch: raw-sample median plane [n][h][w] of windows above 15x15 (two-pass path)
  size_t bigmed_bytes = 0;
  void* win = nullptr;            // scratch: [n] (min, max) of each slice, the rendering window of nm03_outputs.export_blocks
  size_t win_bytes = 0;
  float lut_key[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  bool lut_valid = false;
  bool busy = false;
};

struct ProfRec {
  cudaEvent_t e[3] = {nullptr, nullptr, nullptr};  // before K1, between, after K23
};

// Everything of a pipeline call that depends only on (parameters, slice size, seed list, which K1 mode): the two kernel
// parameter blocks with all host-side preparation done (Gaussian masks, error bounds of the fast predicate, clip-floor
// key, tile order, validated seeds, structuring elements).  A caller repeats the same few combinations for every batch,
// so they are prepared once and looked up afterwards.
struct Plan {
  nm03_params params;
  int w = 0, h = 0, n_seeds = -1;          // n_seeds = -1: the adaptive seeds
  std::vector<int> seeds;
  bool exact = false, erode_out = false;
  int hs = 0, k1_mode = 0;
  nm03::PreParams pre;
  nm03::SegParams seg;
};

struct Worker {
  int device = 0;
  std::mutex plan_mu;
  std::vector<Plan> plans;                 // small most-recently-used list
  Slot slots[kSlots];
  std::atomic<int> host_calls{0};   // nm03_process_batch calls in flight on this GPU
  std::mutex mu;
  std::condition_variable cv;
  std::atomic<long long> launches{0};
  std::atomic<int> profiling{0};
  std::mutex prof_mu;
  std::vector<ProfRec> prof_pending, prof_free;
};

}  // namespace

struct nm03_ctx {
  std::vector<Worker*> workers;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

// `prefer` = the caller's own stream (device entry point): a free slot last used on that stream needs no
// cross-stream wait, a never-used one neither, so callers that alternate between a few streams each settle
// on a slot of their own and their batches overlap on the GPU.
Slot* acquire(Worker* wk, bool block = true, cudaStream_t prefer = nullptr) {
  std::unique_lock<std::mutex> lk(wk->mu);
  for (;;) {
    Slot* pick = nullptr;
    for (auto& s : wk->slots) {
      if (s.busy) continue;
      if (!prefer) { pick = &s; break; }
      if (s.used && s.last_stream == prefer) { pick = &s; break; }
      if (!pick || (!s.used && pick->used)) pick = &s;
    }
    if (pick) { pick->busy = true; return pick; }
    if (!block) return nullptr;
    wk->cv.wait(lk);
  }
}

void release(Worker* wk, Slot* s) {
  {
    std::lock_guard<std::mutex> lk(wk->mu);
    s->busy = false;
  }
  wk->cv.notify_one();
}

struct SlotLease {
  Worker* wk;
  Slot* s;
  explicit SlotLease(Worker* w, cudaStream_t prefer = nullptr) : wk(w), s(acquire(w, true, prefer)) {}
  ~SlotLease() { release(wk, s); }
};

// Work queued earlier on another stream may still be using this slot's scratch.
int order_after_last_use(Slot* s, cudaStream_t st) {
  if (s->used && s->last_stream != st) CK(cudaStreamWaitEvent(st, s->done, 0));
  return NM03_OK;
}

// Is a batch that was queued through another slot, on another stream, still running on this GPU?
bool others_in_flight(Worker* wk, const Slot* mine, cudaStream_t st) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;   // event queries are not allowed while a stream is being captured
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) { (void)cudaGetLastError(); return false; }
  for (auto& o : wk->slots) {
    if (&o == mine || !o.used.load(std::memory_order_relaxed) || o.last_stream.load(std::memory_order_relaxed) == st) continue;
    if (o.forked_from.load(std::memory_order_relaxed) == st) continue;   // a part of this caller's own previous batch: st waits for it anyway
    if (cudaEventQuery(o.done) == cudaErrorNotReady) {
      (void)cudaGetLastError();   // "not ready" is an answer, not a failure: keep it out of the next launch check
      return true;
    }
  }
  return false;
}

int mark_used(Slot* s, cudaStream_t st) {
  CK(cudaEventRecord(s->done, st));
  s->used = true;
  s->last_stream = st;
  s->forked_from = nullptr;
  return NM03_OK;
}

// grow-only scratch; growing waits for the slot's previous work first
int ensure(Slot* s, void** buf, size_t* have, size_t need) {
  if (*have >= need) return NM03_OK;
  if (s->used) CK(cudaEventSynchronize(s->done));
  if (*buf) CK(cudaFree(*buf));
  *buf = nullptr;
  *have = 0;
  if (buf == (void**)&s->lut) s->lut_valid = false;
  const size_t want = need + need / 4;
  cudaError_t e = cudaMalloc(buf, want);
  if (e != cudaSuccess) return fail(NM03_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
  *have = want;
  return NM03_OK;
}

int check_params(const nm03_params* p, int w, int h, bool whole_pipeline) {
  if (!p) return fail(NM03_ERR_INVALID, "params is null");
  if (w <= 0 || h <= 0) return fail(NM03_ERR_INVALID, "bad image size %dx%d", w, h);
  if (whole_pipeline && (w < 100 || h < 100))
    return fail(NM03_ERR_TOO_SMALL, "Image dimensions too small: %dx%d", w, h);  // main_sequential.cpp:189-192
  return NM03_OK;
}

int check_preprocess(const nm03_params* p, int* sharpen_half) {
  if (p->median_window < 3 || (p->median_window & 1) == 0)
    return fail(NM03_ERR_PARAM, "median window must be odd and >= 3, got %d", p->median_window);
  if (p->median_window > nm03::BIG_MAX_WINDOW)
    return fail(NM03_ERR_UNSUPPORTED, "median window %d not implemented (odd sizes 3..%d are)", p->median_window, nm03::BIG_MAX_WINDOW);
  if (p->median_window > nm03::K1_MAX_WINDOW && p->border_mode == NM03_BORDER_ZERO)
    return fail(NM03_ERR_UNSUPPORTED, "zero border mode is implemented for median windows up to %d, got %d", nm03::K1_MAX_WINDOW,
                p->median_window);
  int m = p->sharpen_mask;
  if (m == 0) m = (int)std::ceil(2.0f * p->sharpen_sigma) * 2 + 1;  // FAST's default mask size
  if (m < 1 || (m & 1) == 0) return fail(NM03_ERR_PARAM, "sharpen mask size must be odd, got %d", m);
  if (m < 3 || m > 9) return fail(NM03_ERR_UNSUPPORTED, "sharpen mask size %d not implemented (3..9 are)", m);
  if (!(p->norm_max != p->norm_min)) return fail(NM03_ERR_PARAM, "normalisation range is empty");
  *sharpen_half = (m - 1) / 2;
  return NM03_OK;
}

// host-side twin of the kernel's f(): used only to validate the zero-border sentinel trick
float f_host(float v, const nm03_params* p) {
  const float range = p->norm_max - p->norm_min, scale = p->norm_high - p->norm_low;
  const float t = (v - p->norm_min) / range;
  const float o = t * scale + p->norm_low;
  return fminf(fmaxf(o, p->clip_min), p->clip_max);
}

int fill_pre(nm03::PreParams& k, const nm03_params* p, int n, int w, int h, int sharpen_half) {
  k.w = w; k.h = h; k.n = n; k.wpr = (w + 31) / 32;
  k.tiles_x = (w + nm03::K1_TW - 1) / nm03::K1_TW;
  k.tiles_y = (h + nm03::K1_TH - 1) / nm03::K1_TH;
  k.norm_min = p->norm_min;
  k.norm_range = p->norm_max - p->norm_min;
  k.norm_scale = p->norm_high - p->norm_low;
  k.norm_low = p->norm_low;
  k.clip_min = p->clip_min; k.clip_max = p->clip_max;
  k.gain = p->sharpen_gain; k.rg_min = p->rg_min; k.rg_max = p->rg_max;
  k.border_zero = p->border_mode == NM03_BORDER_ZERO;
  k.use_fma = p->sharpen_fma != 0;
  k.input_unsigned = p->input_unsigned != 0;
  // central tiles first (heaviest on brain slices), see PreParams::tile_order
  k.tile_order_on = 0;
  {
    const int nt = k.tiles_x * k.tiles_y;
    if (nt <= 256 && nt > 1 && !g_no_tile_order) {
      std::vector<std::pair<double, int>> ord;
      for (int t = 0; t < nt; ++t) {
        const double cx = ((t % k.tiles_x) + 0.5) * nm03::K1_TW - 0.5 * w, cy = ((t / k.tiles_x) + 0.5) * nm03::K1_TH - 0.5 * h;
        ord.push_back({(cx / w) * (cx / w) + (cy / h) * (cy / h), t});
      }
      std::stable_sort(ord.begin(), ord.end());
      for (int t = 0; t < nt; ++t) k.tile_order[t] = (uint16_t)ord[t].second;
      k.tile_order_on = 1;
    }
  }
  k.lut = nullptr;
  k.sched = nullptr;
  k.use_tma = g_no_tma ? 0 : 1;
  // clip floor shortcut: the largest key K with f(K) == f(lowest key); f must be non-decreasing in the
  // key (both normalisation factors of one sign) and the zero-border sentinel must not be in play
  k.skip_on = 0; k.skip_key2 = 0; k.skip_val = 0.0f;
  if (!g_no_skip && !k.border_zero && (p->median_window <= 9 || p->median_window > nm03::K1_MAX_WINDOW) &&
      (double)k.norm_scale * (double)k.norm_range > 0.0) {   // network windows, and the 1x1 pass of windows above 15
    const int kmin = -32768, off = k.input_unsigned ? 32768 : 0;
    const float fmin_ = f_host((float)(kmin + off), p);
    int lo = kmin, hi = 32767;                        // invariant: f(lo) == fmin_
    if (f_host((float)(hi + off), p) == fmin_) lo = hi;
    while (hi - lo > 1) { const int mid = lo + (hi - lo) / 2; if (f_host((float)(mid + off), p) == fmin_) lo = mid; else hi = mid; }
    if (lo > kmin) {
      k.skip_on = 1;
      k.skip_key2 = ((uint32_t)lo & 0xffffu) | ((uint32_t)lo << 16);
      k.skip_val = fmin_;
    }
  }  // the launcher falls back to plain loads when the shape/alignment rules fail
  if (k.border_zero) {
    // outside-image samples are carried as the key -32768; real data is saturated to -32767, which is
    // only exact if f() does not distinguish the two lowest keys
    if (k.input_unsigned)
      return fail(NM03_ERR_UNSUPPORTED, "zero border mode with uint16 input is not implemented");
    if (f_host(-32768.0f, p) != f_host(-32767.0f, p))
      return fail(NM03_ERR_UNSUPPORTED, "zero border mode needs f(-32768) == f(-32767) for these parameters");
    // the sentinel sorts below every real key, which is only the position of 0.0 in the window if f() is
    // non-decreasing in the key and never negative
    if (!((double)k.norm_scale * (double)k.norm_range > 0.0) || f_host(-32767.0f, p) < 0.0f)
      return fail(NM03_ERR_UNSUPPORTED, "zero border mode needs a non-decreasing normalisation whose clipped values are >= 0 "
                  "(clip_min %g, scale %g, range %g)", (double)p->clip_min, (double)k.norm_scale, (double)k.norm_range);
  }
  // normalised Gaussian exactly as FAST's host code builds it: x outer, y inner
  const int m = 2 * sharpen_half + 1;
  float sum = 0.0f;
  for (int x = -sharpen_half; x <= sharpen_half; ++x)
    for (int y = -sharpen_half; y <= sharpen_half; ++y) {
      const float v = expf(-(float)(x * x + y * y) / (2.0f * p->sharpen_sigma * p->sharpen_sigma));
      k.mask[x + sharpen_half + (y + sharpen_half) * m] = v;
      sum += v;
    }
  for (int i = 0; i < m * m; ++i) k.mask[i] /= sum;
  for (int i = m * m; i < 81; ++i) k.mask[i] = 0.0f;
  for (int i = 0; i < 9; ++i) k.gx[i] = 0.0f;
  k.ea = k.eb = k.ec = k.eps_max = k.ean = k.gain1 = 0.0f;
  for (int i = 0; i < 9; ++i) k.gyn[i] = 0.0f;
  return NM03_OK;
}

// Fast predicate mode of K1 (used when the sharpened plane itself is not requested).
//
// The kernel evaluates blur_a = sum_y gx[y] * (sum_x gx[x] * v) over |x|,|y| <= he with FMAs, and
// sharpened_a = v + gain*(v - blur_a).  The oracle's value uses the full m x m mask accumulated tap by
// tap.  With u = 2^-24, gamma_n = n*u/(1-n*u), w = the fp32 mask, S_in = sum_inside w|v|,
//   |blur_a - blur_exact| <= [gamma_{m*m} + gamma_{2(2he+1)}(1+rho) + rho] * S_in
//                            + [(1+gamma_{m*m}) * D_out + D_abs] * Vmax
// where rho bounds |gx[x]gx[y] - w|/w over the significant inside taps, D_abs collects the others,
// D_out is the mask mass outside the (2he+1)^2 window and Vmax bounds |f|.  If f >= 0 everywhere then
// S_in <= blur_a * (1 + 2 rho + 4 gamma), otherwise S_in <= Vmax.  The three rounded operations after
// the blur add u*(6|gain|(|v|+|blur|) + 2|v|).  Everything is evaluated in double and inflated by 5 %.
// A pixel whose fast value is farther than eps = ea*|blur_a| + eb*|v| + ec from both thresholds has
// the same predicate bit as the oracle; the others are recomputed exactly in the kernel.
// Returns the chosen K1 mode.
int fill_fast(nm03::PreParams& k, const nm03_params* p, int sharpen_half) {
  const int hs = sharpen_half, m = 2 * hs + 1;
  // separable factors in fp32
  float g[9], gs = 0.0f;
  for (int x = -hs; x <= hs; ++x) { g[x + hs] = expf(-(float)(x * x) / (2.0f * p->sharpen_sigma * p->sharpen_sigma)); gs += g[x + hs]; }
  for (int x = -hs; x <= hs; ++x) k.gx[x + nm03::K1_HS_MAX] = g[x + hs] / gs;
  const double lo = f_host(p->input_unsigned ? 0.0f : -32768.0f, p), hi = f_host(p->input_unsigned ? 65535.0f : 32767.0f, p);
  const double vmax = std::fmax(std::fabs(lo), std::fabs(hi));
  const bool nonneg = lo >= 0.0 && hi >= 0.0;
  const double u = 1.0 / 16777216.0;
  auto gamma = [&](double n) { return n * u / (1.0 - n * u); };
  const double gain = std::fabs((double)p->sharpen_gain);
  int best_mode = nm03::K1_FAST_FULL;
  for (int attempt = 0; attempt < 2; ++attempt) {
    const int he = (attempt == 0 && hs > 2) ? 2 : hs;
    if (attempt == 1 && hs <= 2 && best_mode == nm03::K1_FAST_FULL && k.ea != 0.0f) break;
    double rho = 0.0, d_abs = 0.0, d_out = 0.0;
    for (int x = -hs; x <= hs; ++x)
      for (int y = -hs; y <= hs; ++y) {
        const double wxy = k.mask[x + hs + (y + hs) * m];
        if (std::abs(x) <= he && std::abs(y) <= he) {
          const double sep = (double)k.gx[x + nm03::K1_HS_MAX] * (double)k.gx[y + nm03::K1_HS_MAX];
          if (wxy >= 1e-12) rho = std::fmax(rho, std::fabs(sep - wxy) / wxy);
          else d_abs += std::fabs(sep - wxy);
        } else {
          d_out += wxy;
        }
      }
    const double gM = gamma(m * m), g2 = gamma(2.0 * (2 * he + 1));
    // (+ 2u: the kernel's vertical taps are fl(-gain * gx), one more rounding per tap than the analysis above assumes)
    double c1 = gM + g2 * (1.0 + rho) + rho + 2.0 * u;
    const double c2 = ((1.0 + gM) * d_out + d_abs) * vmax * 1.05;
    double ea, ec = gain * c2;
    if (nonneg) { c1 *= (1.0 + 2.0 * rho + 4.0 * g2) * 1.05; ea = gain * c1 + 6.3 * u * gain; }
    else { ec += gain * c1 * vmax * 1.05; ea = 6.3 * u * gain; }
    const double eb = u * (6.3 * gain + 2.1);
    // the truncated window is only worth it when it does not widen the band noticeably
    if (attempt == 0 && hs > 2 && ec > 4e-6) continue;
    // The kernel classifies with d = fl(fl(|s - mid|) - half), mid/half the rounded centre and half-width of
    // [rg_min, rg_max]; in exact arithmetic d = max(rg_min - s, s - rg_max).  Its rounding (two operations, two
    // rounded constants) is below 4u(|s| + |rg_min| + |rg_max|) with |s| <= (1+|gain|)|v| + |gain||blur|, and
    // is folded into the band, so outside the band the sign of d is the oracle's predicate.
    const double rlo = p->rg_min, rhi = p->rg_max;
    k.rg_mid = (float)(0.5 * rlo + 0.5 * rhi);
    k.rg_half = (float)(0.5 * rhi - 0.5 * rlo);
    if (!std::isfinite(k.rg_mid) || !std::isfinite(k.rg_half)) return nm03::K1_EXACT;
    const double ea2 = ea + 4.2 * u * gain, eb2 = eb + 4.2 * u * (1.0 + gain);
    const double ec2 = ec + 4.2 * u * (std::fabs(rlo) + std::fabs(rhi));
    k.ea = (float)(ea2 * 1.02) + 1e-12f;
    k.eb = (float)(eb2 * 1.02) + 1e-12f;
    k.ec = (float)(ec2 * 1.02) + 1e-12f;
    // The kernel forms the sharpened value as fma(fl(1+gain), v, nb) with nb = -gain*blur_a from the scaled taps: two
    // roundings (the constant, the result) instead of the three of sub/mul/add; both are below the 6.3u|gain|(|v|+|blur|)
    // + 2.1u|v| already in ea/eb (|1+gain||v| + |s| <= 2(1+|gain|)|v| + |gain||blur|).
    for (int x = 0; x < 9; ++x) k.gyn[x] = -p->sharpen_gain * k.gx[x];
    k.gain1 = 1.0f + p->sharpen_gain;
    k.ean = gain > 0.0 ? (float)((double)k.ea / gain * (1.0 + 4.0 * u)) + 1e-12f : 0.0f;
    if (!std::isfinite(k.gain1) || !std::isfinite(k.ean)) return nm03::K1_EXACT;
    // widest band of any pixel: |v| <= vmax and |blur_a| <= (sum |gx|)^2 * vmax (inflated for the rounding of the FMAs)
    double gsum = 0.0;
    for (int x = -he; x <= he; ++x) gsum += std::fabs((double)k.gx[x + nm03::K1_HS_MAX]);
    k.eps_max = (float)(((double)k.ea * gsum * gsum * 1.0001 + (double)k.eb) * vmax * 1.0001 + (double)k.ec) + 1e-12f;
    best_mode = (he < hs) ? nm03::K1_FAST_HE2 : nm03::K1_FAST_FULL;
    break;
  }
  return best_mode;
}

int se_halfwidths(int size, int se_shape, int* n_out, int8_t* hw) {
  if (size <= 2 || (size & 1) == 0)
    return fail(NM03_ERR_PARAM, "morphology size must be odd and > 2, got %d", size);  // FAST throws
  const int N = size / 2;
  if (N > nm03::SEG_HALO) return fail(NM03_ERR_UNSUPPORTED, "morphology size %d not implemented (<= 15 is)", size);
  for (int b = -N; b <= N; ++b) {
    int a = N;
    if (se_shape != NM03_SE_SQUARE)
      while (a * a + b * b > N * N) --a;  // disc: offsets with a*a + b*b <= N*N
    hw[b + N] = (int8_t)a;
  }
  *n_out = N;
  return NM03_OK;
}

// cluster size: the smallest power of two whose per-CTA planes fit in shared memory, widened while
// the batch alone cannot fill the 148 SMs
int pick_cluster(int n, int h, int wpr, int* rows_per, bool throughput) {
  const size_t cap = 200 * 1024;
  int c = 1;
  for (;; c *= 2) {
    const int rows = (h + c - 1) / c;
    if (!g_force_global && nm03::seg_smem_bytes(rows, wpr) <= cap) break;
    if (c == nm03::SEG_MAX_CLUSTER || g_force_global) {
      // beyond what 8 CTAs can hold: the global-memory fallback works on bands of rows
      int band = g_force_global ? 64 : 512;
      while (band >= 16 && nm03::seg_smem_bytes(band, wpr) > cap) band /= 2;
      if (band < 16) return fail(NM03_ERR_TOO_LARGE, "image rows of %d words are too wide for the band fallback", wpr);
      *rows_per = band < h ? band : h;
      return 0;
    }
  }
  // A wider cluster shortens the fill of a batch that has the GPU to itself (24 slices 512x512: 27 us with 4 CTAs per
  // slice, 35 us with one).  When other batches are in flight on other streams the latency is hidden anyway and what
  // counts is how many SM slots the clusters take away from their K1: then the smallest cluster that fits is kept
  // (4 streams x 24 slices: 438 000 slices/s with 1 CTA per slice, 411 000 with 4).
  while (!throughput && c < nm03::SEG_MAX_CLUSTER && n * c * 2 <= 148 && (h + 2 * c - 1) / (2 * c) >= 32) c *= 2;
  if (g_force_cluster > 0) {  // NM03_CLUSTER: tuning override (must still fit)
    int cc = g_force_cluster;
    if ((cc == 1 || cc == 2 || cc == 4 || cc == 8) && nm03::seg_smem_bytes((h + cc - 1) / cc, wpr) <= cap && (h + cc - 1) / cc >= nm03::SEG_HALO) c = cc;
  }
  *rows_per = (h + c - 1) / c;
  return c;
}

// the batch-dependent part of SegParams: cluster size (or the band height of the fallback) and rows per sweep
int seg_set_batch(nm03::SegParams& k, int n, bool throughput) {
  k.n = n;
  int rows = 0;
  const int c = pick_cluster(n, k.h, k.wpr, &rows, throughput);
  if (c < 0) return c;
  k.cluster = c;
  k.rows_per = rows;
  int seg = (rows * k.wpr + nm03::SEG_THREADS - 1) / nm03::SEG_THREADS;
  k.seg_rows = seg < 4 ? 4 : (seg > 16 ? 16 : seg);   // 8-16 rows per sweep measured best (tools/k_time.py)
  if (g_seg_rows > 0) k.seg_rows = g_seg_rows;   // NM03_SEG_ROWS: tuning override
  return NM03_OK;
}

// everything of SegParams that does not depend on the batch size
int fill_seg_static(nm03::SegParams& k, const nm03_params* p, int w, int h, const int* seeds_xy, int n_seeds, bool do_fill) {
  memset(&k, 0, sizeof(k));
  k.w = w; k.h = h; k.wpr = (w + 31) / 32;
  k.do_fill = do_fill;
  if (do_fill) {
    if (p->rg_connectivity != 4 && p->rg_connectivity != 8)
      return fail(NM03_ERR_PARAM, "region growing connectivity must be 4 or 8, got %d", p->rg_connectivity);
    k.conn8 = p->rg_connectivity == 8;
    k.border_expand = p->rg_border_expand != 0;
    k.host_mode = p->rg_mode == NM03_RG_HOST;
    int auto_seeds[2 * nm03::SEG_MAX_SEEDS];
    if (!seeds_xy) {
      n_seeds = nm03_make_seeds(w, h, auto_seeds, nm03::SEG_MAX_SEEDS);
      if (n_seeds < 0) return n_seeds;
      seeds_xy = auto_seeds;
    }
    if (n_seeds <= 0) return fail(NM03_ERR_NO_SEEDS, "no seed points");              // FAST throws
    if (n_seeds > nm03::SEG_MAX_SEEDS)
      return fail(NM03_ERR_UNSUPPORTED, "%d seed points (at most %d are supported)", n_seeds, nm03::SEG_MAX_SEEDS);
    for (int i = 0; i < n_seeds; ++i) {
      const int x = seeds_xy[2 * i], y = seeds_xy[2 * i + 1];
      if (x < 0 || y < 0 || x >= w || y >= h)
        return fail(NM03_ERR_SEED_OOB, "seed point (%d,%d) outside %dx%d image", x, y, w, h);  // FAST throws
      k.seeds[i].x = (short)x;
      k.seeds[i].y = (short)y;
    }
    k.n_seeds = n_seeds;
  }
  return NM03_OK;
}

int fill_seg(nm03::SegParams& k, const nm03_params* p, int n, int w, int h, const int* seeds_xy, int n_seeds,
             bool do_fill, bool throughput = false) {
  int rc;
  if ((rc = fill_seg_static(k, p, w, h, seeds_xy, n_seeds, do_fill))) return rc;
  return seg_set_batch(k, n, throughput);
}

int get_worker(nm03_ctx* ctx, int gpu, Worker** wk) {
  if (!ctx) return fail(NM03_ERR_INVALID, "context is null");
  if (gpu < 0 || gpu >= (int)ctx->workers.size())
    return fail(NM03_ERR_INVALID, "gpu index %d out of range (context has %zu)", gpu, ctx->workers.size());
  *wk = ctx->workers[gpu];
  return NM03_OK;
}

// K23 for `seg` on stream st: one cluster launch, or — for slices beyond cluster capacity — band-wise
// fill launches repeated until the device flag stays clear (this path waits for the stream)
int run_segment(Worker* wk, Slot* s, cudaStream_t st, const nm03::SegParams& seg) {
  if (seg.cluster > 0) {
    CK(nm03::launch_segment(seg, st));
    wk->launches += 1;
    return NM03_OK;
  }
  int rc;
  const size_t bytes = (size_t)seg.n * seg.h * seg.wpr * 4;
  if ((rc = ensure(s, (void**)&s->gmask, &s->gmask_bytes, bytes))) return rc;
  if ((rc = ensure(s, (void**)&s->dflag, &s->dflag_bytes, sizeof(int)))) return rc;
  if (!s->hflag) {
    cudaError_t e = cudaHostAlloc((void**)&s->hflag, sizeof(int), cudaHostAllocPortable);
    if (e != cudaSuccess) return fail(NM03_ERR_NOMEM, "cudaHostAlloc failed: %s", cudaGetErrorString(e));
  }
  int passes = 0;
  if (seg.do_fill) {
    const int nbands = (seg.h + seg.rows_per - 1) / seg.rows_per;
    for (;; ++passes) {
      CK(cudaMemsetAsync(s->dflag, 0, sizeof(int), st));
      CK(nm03::launch_band_fill(seg, s->gmask, s->dflag, passes == 0, st));
      wk->launches += 1;
      CK(cudaMemcpyAsync(s->hflag, s->dflag, sizeof(int), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      if (passes > 0 && *s->hflag == 0) { ++passes; break; }
      if (passes > 4 * (nbands + seg.wpr) * 64) return fail(NM03_ERR_CUDA, "region growing did not converge");
    }
  } else {
    CK(cudaMemcpyAsync(s->gmask, seg.pred, bytes, cudaMemcpyDeviceToDevice, st));
  }
  CK(nm03::launch_global_morph(seg, s->gmask, passes, st));
  wk->launches += 1;
  return NM03_OK;
}

// make sure the slot's f(key) table matches these parameters (rebuilt by a tiny kernel when they change)
// nm03_outputs.export_*: all three pointers or none, whole 8x8 blocks only
int check_export(const nm03_outputs* o, int w, int h) {
  if (!o->export_blocks && !o->export_nz && !o->export_rq) {
    if (o->export_seg_blocks || o->export_seg_nz || o->export_seg_style)
      return fail(NM03_ERR_INVALID, "export_seg_blocks is only offered together with export_blocks");
    return NM03_OK;
  }
  if (!o->export_blocks || !o->export_nz || !o->export_rq)
    return fail(NM03_ERR_INVALID, "export_blocks, export_nz and export_rq go together");
  if ((w & 7) || (h & 7))
    return fail(NM03_ERR_UNSUPPORTED, "export blocks need a width and height that are multiples of 8, got %dx%d", w, h);
  if (o->export_seg_blocks || o->export_seg_nz || o->export_seg_style) {
    if (!o->export_seg_blocks || !o->export_seg_nz || !o->export_seg_style || !o->dilated_bits)
      return fail(NM03_ERR_INVALID, "export_seg_blocks needs export_seg_nz, export_seg_style and dilated_bits");
    if (o->export_seg_style->border_radius < 0 || o->export_seg_style->border_radius > 7)
      return fail(NM03_ERR_UNSUPPORTED, "export border radius %d not implemented (0..7 are)", o->export_seg_style->border_radius);
  }
  return NM03_OK;
}

// the export half of a batch (see nm03_export.h): queued behind the pipeline on the same stream, reads the same input
int enqueue_export(Slot* s, cudaStream_t st, const void* d_slices, int n, int w, int h, int input_unsigned, const nm03_outputs* d_out) {
  if (!d_out->export_blocks) return NM03_OK;
  int rc;
  // the block kernels load and store 16 bytes at a time (always true of the library's own staging; a device-pointer caller
  // gets a clear refusal instead of a misaligned access)
  if (((uintptr_t)d_slices & 15) || ((uintptr_t)d_out->export_blocks & 15) || ((uintptr_t)d_out->export_nz & 7) ||
      (d_out->export_seg_blocks && (((uintptr_t)d_out->export_seg_blocks & 15) || ((uintptr_t)d_out->export_seg_nz & 7))))
    return fail(NM03_ERR_UNSUPPORTED, "export blocks need 16-byte aligned slices and block arrays (8-byte aligned maps)");
  if ((rc = ensure(s, &s->win, &s->win_bytes, (size_t)n * sizeof(int2)))) return rc;
  nm03::ExportParams e;
  e.in = d_slices; e.n = n; e.w = w; e.h = h; e.input_unsigned = input_unsigned;
  e.mm = (int2*)s->win;
  e.blocks = d_out->export_blocks;
  e.nz = (unsigned long long*)d_out->export_nz;
  memcpy(e.rq, d_out->export_rq, sizeof(e.rq));
  CK(nm03::launch_export_blocks(e, st));
  if (d_out->export_seg_blocks) {
    nm03::ExportSegParams g;
    g.bits = d_out->dilated_bits; g.n = n; g.w = w; g.h = h; g.wpr = (w + 31) / 32;
    const nm03_export_style& sty = *d_out->export_seg_style;
    g.radius = sty.border_radius;
    memset(g.hw, 0, sizeof(g.hw));
    for (int b = -g.radius; b <= g.radius; ++b) {   // disc: offsets with a*a + b*b <= R*R (SegmentationRenderer, host/nm03.hpp)
      int a = 0;
      while ((a + 1) * (a + 1) + b * b <= g.radius * g.radius) ++a;
      g.hw[b + g.radius] = (int8_t)a;
    }
    g.background = sty.background; g.interior = sty.interior; g.border = sty.border;
    g.blocks = d_out->export_seg_blocks;
    g.nz = (unsigned long long*)d_out->export_seg_nz;
    memcpy(g.rq, d_out->export_rq, sizeof(g.rq));
    CK(nm03::launch_export_seg_blocks(g, st));
  }
  return NM03_OK;
}

// K1 for any supported window.  Up to 15x15 the fused kernel does everything; above, the window median of the raw
// samples is taken first (generic kernel, slot scratch) and K1 runs on that plane with the 1x1 window.  The caller has
// ordered the stream after the slot's previous use.
int launch_preprocess_any(Slot* s, nm03::PreParams& pre, int median_window, int hs, int k1_mode, cudaStream_t st) {
  if (median_window <= nm03::K1_MAX_WINDOW) {
    CK(nm03::launch_preprocess(pre, median_window, hs, k1_mode, st));
    return NM03_OK;
  }
  int rc;
  if ((rc = ensure(s, &s->bigmed, &s->bigmed_bytes, (size_t)pre.n * pre.w * pre.h * 2))) return rc;
  CK(nm03::launch_median_big(pre.in, pre.n, pre.w, pre.h, median_window, pre.input_unsigned, s->bigmed, st));
  const void* in0 = pre.in;
  pre.in = s->bigmed;
  const cudaError_t e = nm03::launch_preprocess(pre, 1, hs, k1_mode, st);
  pre.in = in0;
  CK(e);
  return NM03_OK;
}

int ensure_lut(Worker* wk, Slot* s, cudaStream_t st, nm03::PreParams& pre) {
  int rc;
  if ((rc = ensure(s, (void**)&s->lut, &s->lut_bytes, 65536 * sizeof(float)))) return rc;
  const float key[8] = {pre.norm_min, pre.norm_range, pre.norm_scale, pre.norm_low, pre.clip_min, pre.clip_max,
                        (float)pre.border_zero, (float)pre.input_unsigned};
  if (!s->lut_valid || memcmp(key, s->lut_key, sizeof(key)) != 0) {
    pre.lut = nullptr;
    CK(nm03::launch_build_lut(pre, s->lut, st));
    wk->launches += 1;
    memcpy(s->lut_key, key, sizeof(key));
    s->lut_valid = true;
  }
  pre.lut = s->lut;
  return NM03_OK;
}

// queue K1 + K23 for n device-resident slices on `st`
// how the segment kernel's cluster size is chosen: for the shortest latency of this batch (wide clusters), for the
// least interference with other batches (smallest cluster that fits), or by looking at what else is in flight
enum SegPolicy { SEG_LATENCY = 0, SEG_LOOK = 1, SEG_THROUGHPUT = 2 };

int enqueue_pipeline(Worker* wk, Slot* s, cudaStream_t st, const void* d_slices, int n, int w, int h,
                     const nm03_params* p, const int* seeds_xy, int n_seeds, const nm03_outputs* d_out, SegPolicy policy) {
  int rc, hs = 0;
  if ((rc = check_params(p, w, h, true))) return rc;
  if ((rc = check_export(d_out, w, h))) return rc;
  const bool exact = d_out->sharpened || g_force_exact, erode_out = p->erode_size > 0 && d_out->eroded;
  nm03::PreParams pre;
  nm03::SegParams seg;
  int k1_mode = nm03::K1_EXACT;
  bool hit = false;
  {
    std::lock_guard<std::mutex> lk(wk->plan_mu);
    for (size_t i = 0; i < wk->plans.size(); ++i) {
      Plan& c = wk->plans[i];
      if (c.w != w || c.h != h || c.exact != exact || c.erode_out != erode_out || memcmp(&c.params, p, sizeof(*p)) != 0) continue;
      if (seeds_xy ? (c.n_seeds != n_seeds || (n_seeds > 0 && memcmp(c.seeds.data(), seeds_xy, sizeof(int) * 2 * n_seeds) != 0)) : c.n_seeds != -1) continue;
      pre = c.pre; seg = c.seg; hs = c.hs; k1_mode = c.k1_mode;
      if (i != 0) std::swap(wk->plans[0], wk->plans[i]);
      hit = true;
      break;
    }
  }
  if (!hit) {
    if ((rc = check_preprocess(p, &hs))) return rc;
    if ((rc = fill_pre(pre, p, 0, w, h, hs))) return rc;
    if ((rc = fill_seg_static(seg, p, w, h, seeds_xy, n_seeds, true))) return rc;
    if ((rc = se_halfwidths(p->dilate_size, p->se_shape, &seg.dil_n, seg.dil_hw))) return rc;
    if (erode_out) {
      if ((rc = se_halfwidths(p->erode_size, p->se_shape, &seg.ero_n, seg.ero_hw))) return rc;
    }
    k1_mode = exact ? nm03::K1_EXACT : fill_fast(pre, p, hs);
    Plan c;
    c.params = *p; c.w = w; c.h = h; c.exact = exact; c.erode_out = erode_out; c.hs = hs; c.k1_mode = k1_mode;
    c.n_seeds = seeds_xy ? n_seeds : -1;
    if (seeds_xy && n_seeds > 0) c.seeds.assign(seeds_xy, seeds_xy + 2 * n_seeds);
    c.pre = pre; c.seg = seg;
    std::lock_guard<std::mutex> lk(wk->plan_mu);
    wk->plans.insert(wk->plans.begin(), std::move(c));
    if (wk->plans.size() > 8) wk->plans.pop_back();
  }
  pre.n = n;
  const bool throughput = policy == SEG_THROUGHPUT || (policy == SEG_LOOK && others_in_flight(wk, s, st));
  if ((rc = seg_set_batch(seg, n, throughput))) return rc;
  const size_t pred_bytes = (size_t)n * h * pre.wpr * 4;
  if ((rc = ensure(s, (void**)&s->pred, &s->pred_bytes, pred_bytes))) return rc;
  if ((rc = order_after_last_use(s, st))) return rc;
  if ((rc = ensure_lut(wk, s, st, pre))) return rc;
  if (!s->sched) {
    if ((rc = ensure(s, (void**)&s->sched, &s->sched_bytes, 4 * sizeof(unsigned int)))) return rc;
    CK(cudaMemsetAsync(s->sched, 0, 4 * sizeof(unsigned int), st));
  }
  pre.sched = g_no_dynamic ? nullptr : s->sched;
  pre.in = d_slices;
  pre.median = d_out->median;
  pre.sharpened = d_out->sharpened;
  pre.predicate = s->pred;
  ProfRec rec;
  const bool prof = wk->profiling.load() != 0;
  if (prof) {
    std::lock_guard<std::mutex> lk(wk->prof_mu);
    if (!wk->prof_free.empty()) { rec = wk->prof_free.back(); wk->prof_free.pop_back(); }
    else for (auto& e : rec.e) CK(cudaEventCreate(&e));
  }
  if (prof) CK(cudaEventRecord(rec.e[0], st));
  if ((rc = launch_preprocess_any(s, pre, p->median_window, hs, k1_mode, st))) return rc;
  if (prof) CK(cudaEventRecord(rec.e[1], st));
  seg.pred = s->pred;
  seg.dilated = d_out->dilated;
  seg.dilated_bits = d_out->dilated_bits;
  seg.eroded = (p->erode_size > 0) ? d_out->eroded : nullptr;
  seg.rg_mask = d_out->rg_mask;
  seg.iterations = d_out->rg_iterations;
  if ((rc = run_segment(wk, s, st, seg))) return rc;
  if ((rc = enqueue_export(s, st, d_slices, n, w, h, pre.input_unsigned, d_out))) return rc;
  if (prof) {
    CK(cudaEventRecord(rec.e[2], st));
    std::lock_guard<std::mutex> lk(wk->prof_mu);
    wk->prof_pending.push_back(rec);
  }
  wk->launches += 1;
  return mark_used(s, st);
}

}  // namespace

// ---- scratch of the no-allocation entry point ------------------------------------------------
// The predicate bit-plane of nm03_process_batch_device_noalloc lives in the library image instead of
// in cudaMalloc'ed memory.  In host code `g_noalloc_pred` is only a handle for the symbol, not a
// device address: the address is asked for once per device, with that device current, and cached.
static __device__ __align__(256) uint32_t g_noalloc_pred[NM03_NOALLOC_MAX_PRED_WORDS];

namespace {

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_addr[kNoallocMaxDevices] = {};  // device address per CUDA ordinal, null = not looked up yet

// caller has made `device` current
int noalloc_scratch(int device, uint32_t** out) {
  if (device < 0 || device >= kNoallocMaxDevices)
    return fail(NM03_ERR_UNSUPPORTED, "device ordinal %d is beyond the %d the no-allocation path tracks", device,
                kNoallocMaxDevices);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_addr[device]) {
    uint32_t* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_pred);
    if (e != cudaSuccess || !ptr)
      return fail(NM03_ERR_CUDA, "cudaGetSymbolAddress(built-in scratch) failed on device %d: %s", device,
                  cudaGetErrorString(e));
    g_noalloc_addr[device] = ptr;
  }
  *out = g_noalloc_addr[device];
  return NM03_OK;
}

}  // namespace

extern "C" {

// See include/nm03.h for the contract.  No cudaMalloc/cudaFree, no stream or event creation, no slot
// lease (the slots' scratch is what this variant avoids); the only state touched is the cached symbol
// address and the launch counter.
int nm03_process_batch_device_noalloc(nm03_ctx* ctx, int gpu, const void* d_slices, int n, int w, int h,
                                      const nm03_params* params, const int* seeds_xy, int n_seeds,
                                      const nm03_outputs* d_out, void* stream) {
  Worker* wk;
  int rc, hs = 0;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!d_slices || !d_out || n < 0) return fail(NM03_ERR_INVALID, "null buffer or negative count");
  if (n == 0) return NM03_OK;
  if ((rc = check_params(params, w, h, true))) return rc;
  if ((rc = check_preprocess(params, &hs))) return rc;
  if (d_out->export_blocks || d_out->export_nz || d_out->export_rq)
    return fail(NM03_ERR_UNSUPPORTED, "export blocks need scratch memory and are only available through nm03_process_batch_device");
  if (params->median_window > nm03::K1_MAX_WINDOW)
    return fail(NM03_ERR_UNSUPPORTED, "median windows above %d need scratch memory for a second pass and are only available through "
                "nm03_process_batch_device, got %d", nm03::K1_MAX_WINDOW, params->median_window);
  const long long wpr = (w + 31) / 32;
  const long long words = (long long)n * h * wpr;
  if (words > (long long)NM03_NOALLOC_MAX_PRED_WORDS) {
    const long long per_slice = (long long)h * wpr;
    return fail(NM03_ERR_TOO_LARGE,
                "batch of %d slices %dx%d needs %lld words of predicate scratch but the built-in array holds %d; "
                "pass at most %lld slices of this size per call",
                n, w, h, words, (int)NM03_NOALLOC_MAX_PRED_WORDS,
                per_slice > 0 ? (long long)NM03_NOALLOC_MAX_PRED_WORDS / per_slice : 0LL);
  }
  nm03::PreParams pre;
  if ((rc = fill_pre(pre, params, n, w, h, hs))) return rc;
  nm03::SegParams seg;
  if ((rc = fill_seg(seg, params, n, w, h, seeds_xy, n_seeds, true))) return rc;
  if ((rc = se_halfwidths(params->dilate_size, params->se_shape, &seg.dil_n, seg.dil_hw))) return rc;
  if (params->erode_size > 0 && d_out->eroded) {
    if ((rc = se_halfwidths(params->erode_size, params->se_shape, &seg.ero_n, seg.ero_hw))) return rc;
  }
  DeviceGuard g(wk->device);
  if (seg.cluster == 0)
    return fail(NM03_ERR_TOO_LARGE, "slice %dx%d exceeds one cluster's shared memory; the fallback for such slices needs scratch "
                "memory and is only available through nm03_process_batch_device", w, h);
  uint32_t* scratch = nullptr;
  if ((rc = noalloc_scratch(wk->device, &scratch))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  pre.in = d_slices;
  pre.median = d_out->median;
  pre.sharpened = d_out->sharpened;
  pre.predicate = scratch;
  const int k1_mode = (d_out->sharpened || g_force_exact) ? nm03::K1_EXACT : fill_fast(pre, params, hs);
  CK(nm03::launch_preprocess(pre, params->median_window, hs, k1_mode, st));
  seg.pred = scratch;
  seg.dilated = d_out->dilated;
  seg.dilated_bits = d_out->dilated_bits;
  seg.eroded = (params->erode_size > 0) ? d_out->eroded : nullptr;
  seg.rg_mask = d_out->rg_mask;
  seg.iterations = d_out->rg_iterations;
  CK(nm03::launch_segment(seg, st));
  wk->launches += 2;
  return NM03_OK;
}

const char* nm03_last_error(void) { return g_err; }
int nm03_version(void) { return 102; }   // 101: nm03_outputs.dilated_bits, async tickets; 102: nm03_outputs.export_*

void nm03_default_params(nm03_params* p) {
  if (!p) return;
  p->norm_low = 0.5f; p->norm_high = 2.5f; p->norm_min = 0.0f; p->norm_max = 10000.0f;  // main_sequential.cpp:196
  p->clip_min = 0.68f; p->clip_max = 4000.0f;                                           // :200
  p->median_window = 7;                                                                 // :204
  p->sharpen_gain = 2.0f; p->sharpen_sigma = 0.5f; p->sharpen_mask = 9;                 // :208
  p->rg_min = 0.74f; p->rg_max = 0.91f;                                                 // :233
  p->dilate_size = 3;                                                                   // :250
  p->erode_size = 0;
  p->border_mode = NM03_BORDER_CLAMP;
  p->sharpen_fma = 0;
  p->rg_mode = NM03_RG_OPENCL;
  p->rg_connectivity = 4;
  p->rg_border_expand = 1;
  p->se_shape = NM03_SE_DISC;
  p->input_unsigned = 0;
}

int nm03_make_seeds(int w, int h, int* xy, int cap) {
  if (w < 10 || h < 10) return fail(NM03_ERR_TOO_SMALL, "image %dx%d too small for the seed grid", w, h);
  int n = 0;
  auto put = [&](int x, int y) {
    if (xy && n < cap) { xy[2 * n] = x; xy[2 * n + 1] = y; }
    ++n;
  };
  const int cx = w / 2, cy = h / 2, ox = w / 8, oy = h / 8;  // main_sequential.cpp:214-230
  put(cx, cy);
  put(cx + ox, cy);
  put(cx - ox, cy);
  put(cx, cy + oy);
  put(cx, cy - oy);
  for (int x = w / 4; x < w * 3 / 4; x += w / 10)            // main_sequential.cpp:237-241
    for (int y = h / 4; y < h * 3 / 4; y += h / 10) put(x, y);
  return n;
}

int nm03_init(const int* gpu_ids, int n_gpus, nm03_ctx** out) {
  if (!out) return fail(NM03_ERR_INVALID, "out is null");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(NM03_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  int def = 0;
  if (!gpu_ids || n_gpus <= 0) { gpu_ids = &def; n_gpus = 1; cudaGetDevice(&def); }
  nm03_ctx* ctx = new nm03_ctx();
  for (int i = 0; i < n_gpus; ++i) {
    if (gpu_ids[i] < 0 || gpu_ids[i] >= count) {
      nm03_shutdown(ctx);
      return fail(NM03_ERR_INVALID, "device ordinal %d out of range (%d devices)", gpu_ids[i], count);
    }
    Worker* wk = new Worker();
    wk->device = gpu_ids[i];
    ctx->workers.push_back(wk);
    DeviceGuard g(wk->device);
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, wk->device);
    if (e == cudaSuccess && prop.major < 10) {
      nm03_shutdown(ctx);
      return fail(NM03_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", wk->device,
                  prop.major, prop.minor);
    }
    if (e == cudaSuccess) e = nm03::configure_kernels();
    for (auto& s : wk->slots) {
      if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming | (g_spin_wait ? 0 : cudaEventBlockingSync));
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
      nm03_shutdown(ctx);
      return fail(NM03_ERR_CUDA, "device %d setup failed: %s", gpu_ids[i], cudaGetErrorString(e));
    }
  }
  *out = ctx;
  return NM03_OK;
}

void nm03_shutdown(nm03_ctx* ctx) {
  if (!ctx) return;
  for (Worker* wk : ctx->workers) {
    DeviceGuard g(wk->device);
    for (auto* v : {&wk->prof_pending, &wk->prof_free})
      for (auto& rec : *v)
        for (auto& e : rec.e) if (e) cudaEventDestroy(e);
    for (auto& s : wk->slots) {
      if (s.stream) cudaStreamSynchronize(s.stream);
      if (s.pred) cudaFree(s.pred);
      if (s.stage_in) cudaFree(s.stage_in);
      if (s.stage_out) cudaFree(s.stage_out);
      if (s.lut) cudaFree(s.lut);
      if (s.bigmed) cudaFree(s.bigmed);
      if (s.win) cudaFree(s.win);
      if (s.sched) cudaFree(s.sched);
      if (s.gmask) cudaFree(s.gmask);
      if (s.dflag) cudaFree(s.dflag);
      if (s.hflag) cudaFreeHost(s.hflag);
      if (s.done) cudaEventDestroy(s.done);
      if (s.fork) cudaEventDestroy(s.fork);
      if (s.stream) cudaStreamDestroy(s.stream);
    }
    delete wk;
  }
  delete ctx;
}

int nm03_gpu_count(const nm03_ctx* ctx) { return ctx ? (int)ctx->workers.size() : 0; }

int nm03_device_count(void) {
  int count = 0;
  return cudaGetDeviceCount(&count) == cudaSuccess ? count : 0;
}

long long nm03_kernel_launches(const nm03_ctx* ctx, int gpu) {
  if (!ctx || gpu < 0 || gpu >= (int)ctx->workers.size()) return -1;
  return ctx->workers[gpu]->launches.load();
}

int nm03_profile_enable(nm03_ctx* ctx, int gpu, int on) {
  Worker* wk;
  int rc;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  wk->profiling.store(on ? 1 : 0);
  return NM03_OK;
}

int nm03_profile_read(nm03_ctx* ctx, int gpu, double* preprocess_ms, double* segment_ms, long long* calls) {
  Worker* wk;
  int rc;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  DeviceGuard g(wk->device);
  std::lock_guard<std::mutex> lk(wk->prof_mu);
  double a = 0.0, b = 0.0;
  long long n = 0;
  for (auto& rec : wk->prof_pending) {
    CK(cudaEventSynchronize(rec.e[2]));
    float t1 = 0.f, t2 = 0.f;
    CK(cudaEventElapsedTime(&t1, rec.e[0], rec.e[1]));
    CK(cudaEventElapsedTime(&t2, rec.e[1], rec.e[2]));
    a += t1; b += t2; ++n;
    wk->prof_free.push_back(rec);
  }
  wk->prof_pending.clear();
  if (preprocess_ms) *preprocess_ms = a;
  if (segment_ms) *segment_ms = b;
  if (calls) *calls = n;
  return NM03_OK;
}

long long nm03_profile_work(nm03_ctx* ctx, int gpu) {
  Worker* wk;
  if (get_worker(ctx, gpu, &wk)) return -1;
  DeviceGuard g(wk->device);
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  unsigned long long total = 0;
  for (auto& s : wk->slots) {
    if (!s.sched) continue;
    unsigned long long v = 0;
    if (cudaMemcpy(&v, s.sched + 2, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    if (cudaMemset(s.sched + 2, 0, sizeof(v)) != cudaSuccess) return -1;
    total += v;
  }
  return (long long)total;
}

int nm03_median_network_ops(int median_window) { return nm03::median_network_ops(median_window); }

int nm03_host_alloc(void** p, size_t bytes) {
  if (!p) return fail(NM03_ERR_INVALID, "p is null");
  cudaError_t e = cudaHostAlloc(p, bytes, cudaHostAllocPortable);
  if (e != cudaSuccess) return fail(NM03_ERR_NOMEM, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
  return NM03_OK;
}

int nm03_host_free(void* p) {
  if (!p) return NM03_OK;
  CK(cudaFreeHost(p));
  return NM03_OK;
}

int nm03_process_batch_device(nm03_ctx* ctx, int gpu, const void* d_slices, int n, int w, int h,
                              const nm03_params* params, const int* seeds_xy, int n_seeds,
                              const nm03_outputs* d_out, void* stream) {
  Worker* wk;
  int rc;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!d_slices || !d_out || n < 0) return fail(NM03_ERR_INVALID, "null buffer or negative count");
  if (n == 0) return NM03_OK;
  DeviceGuard g(wk->device);
  SlotLease lease(wk, (cudaStream_t)stream);
  cudaStream_t st = stream ? (cudaStream_t)stream : lease.s->stream;
  // A caller that keeps several streams busy gets its batches overlapped by the GPU: the latency-bound fill of one runs
  // under the ALU-bound preprocess kernel of the next.  Opt-in experiment (NM03_SPLIT=1) for a caller with ONE stream:
  // when nothing else is in flight the batch is cut into parts, part 0 stays on the caller's stream, the others go to
  // library streams behind an event of the caller's stream (fork), and the caller's stream waits for their events
  // (join); everything is still complete where the caller's stream says it is.  Measured on B200, 24 slices 512x512
  // back to back on one stream: 79.5 us whole, 95 / 103 / 105 us in 2 / 3 / 4 parts (96 slices: 215 us whole,
  // 213 / 225 / 234 us in 2 / 4 / 6 parts) — stream order forces every call to end with an exposed fill of its last
  // part, the parts' fills run with the small clusters, and the events cost more than the overlap returns.  Hence off.
  int parts = 1;
  if (!g_no_split && n >= 12 && wk->profiling.load() == 0 && !others_in_flight(wk, lease.s, st)) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap == cudaStreamCaptureStatusNone) parts = n / 6 < g_split_parts ? n / 6 : g_split_parts;
    else (void)cudaGetLastError();
  }
  if (parts <= 1) return enqueue_pipeline(wk, lease.s, st, d_slices, n, w, h, params, seeds_xy, n_seeds, d_out, SEG_LOOK);
  if ((rc = check_params(params, w, h, true))) return rc;
  const size_t npx = (size_t)w * h, wpr = (size_t)(w + 31) / 32;
  CK(cudaEventRecord(lease.s->fork, st));
  // the slots of the parts are held until all parts are queued: a slot handed back early would be the next part's pick,
  // and the parts would line up on one stream
  struct Held {
    Worker* wk;
    std::vector<Slot*> v;
    ~Held() { for (Slot* s : v) release(wk, s); }
  } held{wk, {}};
  int start = 0;
  for (int i = 0; i < parts && start < n; ++i) {
    int cnt = (n - start) / (parts - i);
    Slot* s = lease.s;
    if (i > 0) {
      s = acquire(wk, false);
      if (!s) { s = lease.s; cnt = n - start; }          // no free slot: the rest follows part 0 on the caller's stream
      else held.v.push_back(s);
    }
    cudaStream_t ss = (s == lease.s) ? st : s->stream;
    nm03_outputs o = *d_out;
    if (o.dilated) o.dilated += start * npx;
    if (o.eroded) o.eroded += start * npx;
    if (o.rg_mask) o.rg_mask += start * npx;
    if (o.sharpened) o.sharpened += start * npx;
    if (o.median) o.median += start * npx;
    if (o.rg_iterations) o.rg_iterations += start;
    if (o.dilated_bits) o.dilated_bits += start * h * wpr;
    if (o.export_blocks) { o.export_blocks += (size_t)start * (w / 8) * (h / 8) * 64; o.export_nz += (size_t)start * (w / 8) * (h / 8); }
    if (o.export_seg_blocks) { o.export_seg_blocks += (size_t)start * (w / 8) * (h / 8) * 64; o.export_seg_nz += (size_t)start * (w / 8) * (h / 8); }
    rc = NM03_OK;
    if (ss != st) {
      cudaError_t e = cudaStreamWaitEvent(ss, lease.s->fork, 0);
      if (e != cudaSuccess) rc = fail(NM03_ERR_CUDA, "cudaStreamWaitEvent failed: %s", cudaGetErrorString(e));
    }
    const bool last = start + cnt >= n;
    if (!rc) rc = enqueue_pipeline(wk, s, ss, (const uint8_t*)d_slices + start * npx * 2, cnt, w, h, params, seeds_xy, n_seeds, &o,
                                   last ? SEG_LATENCY : SEG_THROUGHPUT);
    if (!rc && ss != st) {
      s->forked_from = st;
      cudaError_t e = cudaStreamWaitEvent(st, s->done, 0);   // join: recorded by enqueue_pipeline behind the part's last kernel
      if (e != cudaSuccess) rc = fail(NM03_ERR_CUDA, "cudaStreamWaitEvent failed: %s", cudaGetErrorString(e));
    }
    if (rc) return rc;
    start += cnt;
  }
  return NM03_OK;
}

// One chunk [start, start+cnt) of a host-buffer batch on slot s: H2D copy, K1 + K23, D2H copies of every requested
// output, and the slot's event recorded behind them.  n_total sizes the staging buffers (see below).
static int enqueue_host_chunk(Worker* wk, Slot* s, const void* slices, int start, int cnt, int n_total, int w, int h,
                              const nm03_params* params, const int* seeds_xy, int n_seeds, const nm03_outputs* out) {
  int rc;
  const size_t npx = (size_t)w * h, wpr = (size_t)(w + 31) / 32;
  // per-chunk device staging laid out as [inputs][dilated][eroded][rg_mask][sharpened][median][bits][iters]
  size_t out_bytes = 0;
  const size_t o_dil = out_bytes; if (out->dilated) out_bytes += cnt * npx;
  const size_t o_ero = out_bytes; if (out->eroded) out_bytes += cnt * npx;
  const size_t o_rg = out_bytes; if (out->rg_mask) out_bytes += cnt * npx;
  out_bytes = (out_bytes + 15) & ~(size_t)15;
  const size_t o_sh = out_bytes; if (out->sharpened) out_bytes += cnt * npx * 4;
  const size_t o_med = out_bytes; if (out->median) out_bytes += cnt * npx * 4;
  const size_t o_bits = out_bytes; if (out->dilated_bits) out_bytes += (size_t)cnt * h * wpr * 4;
  const size_t o_it = out_bytes; if (out->rg_iterations) out_bytes += (size_t)cnt * 4;
  const size_t nblk = (size_t)(w / 8) * (h / 8);   // export blocks: 128 bytes of coefficients + an 8-byte map each
  out_bytes = (out_bytes + 15) & ~(size_t)15;
  const size_t o_blk = out_bytes; if (out->export_blocks) out_bytes += (size_t)cnt * nblk * 128;
  const size_t o_nz = out_bytes; if (out->export_blocks) out_bytes += (size_t)cnt * nblk * 8;
  out_bytes = (out_bytes + 15) & ~(size_t)15;   // the block kernels store 16 bytes at a time
  const size_t o_sblk = out_bytes; if (out->export_seg_blocks) out_bytes += (size_t)cnt * nblk * 128;
  const size_t o_snz = out_bytes; if (out->export_seg_blocks) out_bytes += (size_t)cnt * nblk * 8;
  // staging is sized for the whole batch even when the call works in chunks: the next batch of this shape may
  // come unchunked, and growing a buffer then would put a cudaFree/cudaMalloc in its path
  const size_t per_slice = npx * ((out->dilated ? 1 : 0) + (out->eroded ? 1 : 0) + (out->rg_mask ? 1 : 0) +
                                  (out->sharpened ? 4 : 0) + (out->median ? 4 : 0)) + (out->dilated_bits ? h * wpr * 4 : 0) +
                           (out->rg_iterations ? 4 : 0) + (out->export_blocks ? nblk * 136 : 0) +
                           (out->export_seg_blocks ? nblk * 136 : 0);
  if ((rc = check_export(out, w, h))) return rc;
  if ((rc = ensure(s, &s->stage_in, &s->stage_in_bytes, (size_t)n_total * npx * 2))) return rc;
  if ((rc = ensure(s, &s->stage_out, &s->stage_out_bytes, std::max(out_bytes, (size_t)n_total * per_slice + 96)))) return rc;
  cudaStream_t st = s->stream;
  cudaError_t e = cudaMemcpyAsync(s->stage_in, (const uint8_t*)slices + (size_t)start * npx * 2, cnt * npx * 2,
                                  cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return fail(NM03_ERR_CUDA, "H2D copy failed: %s", cudaGetErrorString(e));
  uint8_t* so = (uint8_t*)s->stage_out;
  nm03_outputs d = {};
  if (out->dilated) d.dilated = so + o_dil;
  if (out->eroded) d.eroded = so + o_ero;
  if (out->rg_mask) d.rg_mask = so + o_rg;
  if (out->sharpened) d.sharpened = (float*)(so + o_sh);
  if (out->median) d.median = (float*)(so + o_med);
  if (out->dilated_bits) d.dilated_bits = (uint32_t*)(so + o_bits);
  if (out->rg_iterations) d.rg_iterations = (int32_t*)(so + o_it);
  if (out->export_blocks) { d.export_blocks = (int16_t*)(so + o_blk); d.export_nz = (uint64_t*)(so + o_nz); d.export_rq = out->export_rq; }
  if (out->export_seg_blocks) {
    d.export_seg_blocks = (int16_t*)(so + o_sblk); d.export_seg_nz = (uint64_t*)(so + o_snz); d.export_seg_style = out->export_seg_style;
  }
  // (host path: the other slots mostly hold copies in flight, and the chunk's latency is on the call's critical path)
  if ((rc = enqueue_pipeline(wk, s, st, s->stage_in, cnt, w, h, params, seeds_xy, n_seeds, &d, SEG_LATENCY))) return rc;
#define NM03_D2H(field, off, elem)                                                                          \
  if (out->field && e == cudaSuccess)                                                                       \
    e = cudaMemcpyAsync((uint8_t*)out->field + (size_t)start * npx * (elem), so + (off), cnt * npx * (elem), \
                        cudaMemcpyDeviceToHost, st);
  NM03_D2H(dilated, o_dil, 1)
  NM03_D2H(eroded, o_ero, 1)
  NM03_D2H(rg_mask, o_rg, 1)
  NM03_D2H(sharpened, o_sh, 4)
  NM03_D2H(median, o_med, 4)
#undef NM03_D2H
  if (out->dilated_bits && e == cudaSuccess)
    e = cudaMemcpyAsync(out->dilated_bits + (size_t)start * h * wpr, so + o_bits, (size_t)cnt * h * wpr * 4, cudaMemcpyDeviceToHost, st);
  if (out->rg_iterations && e == cudaSuccess)
    e = cudaMemcpyAsync(out->rg_iterations + start, so + o_it, (size_t)cnt * 4, cudaMemcpyDeviceToHost, st);
  if (out->export_blocks && e == cudaSuccess)
    e = cudaMemcpyAsync(out->export_blocks + (size_t)start * nblk * 64, so + o_blk, (size_t)cnt * nblk * 128, cudaMemcpyDeviceToHost, st);
  if (out->export_blocks && e == cudaSuccess)
    e = cudaMemcpyAsync(out->export_nz + (size_t)start * nblk, so + o_nz, (size_t)cnt * nblk * 8, cudaMemcpyDeviceToHost, st);
  if (out->export_seg_blocks && e == cudaSuccess)
    e = cudaMemcpyAsync(out->export_seg_blocks + (size_t)start * nblk * 64, so + o_sblk, (size_t)cnt * nblk * 128, cudaMemcpyDeviceToHost, st);
  if (out->export_seg_blocks && e == cudaSuccess)
    e = cudaMemcpyAsync(out->export_seg_nz + (size_t)start * nblk, so + o_snz, (size_t)cnt * nblk * 8, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaEventRecord(s->done, st);
  if (e != cudaSuccess) return fail(NM03_ERR_CUDA, "D2H copy failed: %s", cudaGetErrorString(e));
  return NM03_OK;
}

// wait for everything queued on the slot so far (its event is recorded behind the last D2H copy)
static cudaError_t wait_slot(Slot* s) { return cudaEventSynchronize(s->done); }

int nm03_process_batch(nm03_ctx* ctx, int gpu, const void* slices, int n, int w, int h, const nm03_params* params,
                       const int* seeds_xy, int n_seeds, const nm03_outputs* out, int* status) {
  Worker* wk;
  int rc;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!slices || !out || n < 0) return fail(NM03_ERR_INVALID, "null buffer or negative count");
  if (n == 0) return NM03_OK;
  if ((rc = check_params(params, w, h, true))) {
    if (status) for (int i = 0; i < n; ++i) status[i] = rc;
    return rc;
  }
  DeviceGuard g(wk->device);
  // A lone caller gets its batch pipelined in chunks over the stream slots, so that the H2D copy of one chunk, the
  // kernels of another and the D2H copy of a third overlap; when other host-buffer batches are in flight on this GPU they
  // already overlap each other's copies and kernels, and whole batches keep the H2D engine busier than interleaved
  // chunks do (3 feeder threads, 24 slices of 512x512: 96 000 slices/s unchunked, 84 000 in 4 chunks; one thread:
  // 63 500 in 4 chunks).
  struct InFlight {
    std::atomic<int>& c;
    int before;
    explicit InFlight(std::atomic<int>& x) : c(x), before(x.fetch_add(1)) {}
    ~InFlight() { c.fetch_sub(1); }
  } in_flight(wk->host_calls);
  const int n_chunks = (in_flight.before > 0 && !g_chunks_forced) ? 1 : kChunksPerCall;
  int chunk = (n + n_chunks - 1) / n_chunks;
  if (chunk < 4) chunk = 4;
  std::vector<Slot*> used;
  int result = NM03_OK;
  for (int start = 0; start < n; start += chunk) {
    const int cnt = (n - start < chunk) ? n - start : chunk;
    // never wait for a slot while holding one (two batches could starve each other): when none is
    // free, drain and reuse the oldest slot of this call instead
    Slot* s = acquire(wk, used.empty());
    if (!s) {
      s = used.front();
      used.erase(used.begin());
      wait_slot(s);
    }
    used.push_back(s);
    if ((rc = enqueue_host_chunk(wk, s, slices, start, cnt, n, w, h, params, seeds_xy, n_seeds, out))) { result = rc; break; }
    // keep at most kChunksPerCall chunks in flight: the oldest slot is drained and handed back
    if ((int)used.size() == n_chunks) {
      Slot* old = used.front();
      used.erase(used.begin());
      wait_slot(old);
      release(wk, old);
    }
  }
  for (Slot* s : used) {
    cudaError_t e = s->used ? wait_slot(s) : cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess && result == NM03_OK) result = fail(NM03_ERR_CUDA, "pipeline failed: %s", cudaGetErrorString(e));
    release(wk, s);
  }
  if (status) for (int i = 0; i < n; ++i) status[i] = result;
  return result;
}

struct nm03_ticket {
  Worker* wk;
  Slot* slot;      // held (busy) until nm03_wait
  int result;      // error found while queueing (the slot is idle then)
};

int nm03_process_batch_async(nm03_ctx* ctx, int gpu, const void* slices, int n, int w, int h, const nm03_params* params,
                             const int* seeds_xy, int n_seeds, const nm03_outputs* out, nm03_ticket** ticket) {
  Worker* wk;
  int rc;
  if (!ticket) return fail(NM03_ERR_INVALID, "ticket is null");
  *ticket = nullptr;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!slices || !out || n <= 0) return fail(NM03_ERR_INVALID, "null buffer or no slices");
  if ((rc = check_params(params, w, h, true))) return rc;
  DeviceGuard g(wk->device);
  Slot* s = acquire(wk, false);
  if (!s) return fail(NM03_ERR_BUSY, "all %d stream slots of this GPU hold unfinished tickets; call nm03_wait on one first", kSlots);
  wk->host_calls.fetch_add(1);
  rc = enqueue_host_chunk(wk, s, slices, 0, n, n, w, h, params, seeds_xy, n_seeds, out);
  if (rc) {   // nothing useful is in flight: drain what was queued and hand the slot back
    cudaStreamSynchronize(s->stream);
    wk->host_calls.fetch_sub(1);
    release(wk, s);
    return rc;
  }
  *ticket = new nm03_ticket{wk, s, NM03_OK};
  return NM03_OK;
}

int nm03_wait(nm03_ticket* t) {
  if (!t) return fail(NM03_ERR_INVALID, "ticket is null");
  int result = t->result;
  {
    DeviceGuard g(t->wk->device);
    cudaError_t e = wait_slot(t->slot);
    if (e != cudaSuccess && result == NM03_OK) result = fail(NM03_ERR_CUDA, "pipeline failed: %s", cudaGetErrorString(e));
  }
  t->wk->host_calls.fetch_sub(1);
  release(t->wk, t->slot);
  delete t;
  return result;
}

int nm03_preprocess(nm03_ctx* ctx, int gpu, const void* slice, int w, int h, const nm03_params* params,
                    float* median, float* sharpened, uint32_t* predicate) {
  Worker* wk;
  int rc, hs = 0;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!slice) return fail(NM03_ERR_INVALID, "slice is null");
  if ((rc = check_params(params, w, h, false))) return rc;
  if ((rc = check_preprocess(params, &hs))) return rc;
  nm03::PreParams pre;
  if ((rc = fill_pre(pre, params, 1, w, h, hs))) return rc;
  DeviceGuard g(wk->device);
  SlotLease lease(wk);
  Slot* s = lease.s;
  if ((rc = order_after_last_use(s, s->stream))) return rc;
  const size_t npx = (size_t)w * h, pw = (size_t)h * pre.wpr * 4;
  if ((rc = ensure(s, &s->stage_in, &s->stage_in_bytes, npx * 2))) return rc;
  if ((rc = ensure(s, &s->stage_out, &s->stage_out_bytes, npx * 8 + pw + 16))) return rc;
  uint8_t* so = (uint8_t*)s->stage_out;
  CK(cudaMemcpyAsync(s->stage_in, slice, npx * 2, cudaMemcpyHostToDevice, s->stream));
  pre.in = s->stage_in;
  pre.median = median ? (float*)so : nullptr;
  pre.sharpened = sharpened ? (float*)(so + npx * 4) : nullptr;
  pre.predicate = predicate ? (uint32_t*)(so + npx * 8) : nullptr;
  const int k1_mode = (sharpened || g_force_exact) ? nm03::K1_EXACT : fill_fast(pre, params, hs);
  if ((rc = launch_preprocess_any(s, pre, params->median_window, hs, k1_mode, s->stream))) return rc;
  wk->launches += 1;
  if (median) CK(cudaMemcpyAsync(median, pre.median, npx * 4, cudaMemcpyDeviceToHost, s->stream));
  if (sharpened) CK(cudaMemcpyAsync(sharpened, pre.sharpened, npx * 4, cudaMemcpyDeviceToHost, s->stream));
  if (predicate) CK(cudaMemcpyAsync(predicate, pre.predicate, pw, cudaMemcpyDeviceToHost, s->stream));
  if ((rc = mark_used(s, s->stream))) return rc;
  CK(cudaStreamSynchronize(s->stream));
  return NM03_OK;
}

int nm03_region_grow(nm03_ctx* ctx, int gpu, const float* image, int w, int h, const nm03_params* params,
                     const int* seeds_xy, int n_seeds, uint8_t* mask, int* iterations) {
  Worker* wk;
  int rc;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!image || !mask) return fail(NM03_ERR_INVALID, "null buffer");
  if ((rc = check_params(params, w, h, false))) return rc;
  if (!seeds_xy || n_seeds <= 0) return fail(NM03_ERR_NO_SEEDS, "no seed points");
  nm03::SegParams seg;
  if ((rc = fill_seg(seg, params, 1, w, h, seeds_xy, n_seeds, true))) return rc;
  DeviceGuard g(wk->device);
  SlotLease lease(wk);
  Slot* s = lease.s;
  if ((rc = order_after_last_use(s, s->stream))) return rc;
  const size_t npx = (size_t)w * h, pw = (size_t)h * seg.wpr * 4;
  if ((rc = ensure(s, &s->stage_in, &s->stage_in_bytes, npx * 4))) return rc;
  if ((rc = ensure(s, &s->stage_out, &s->stage_out_bytes, npx + 16))) return rc;
  if ((rc = ensure(s, (void**)&s->pred, &s->pred_bytes, pw))) return rc;
  CK(cudaMemcpyAsync(s->stage_in, image, npx * 4, cudaMemcpyHostToDevice, s->stream));
  CK(nm03::launch_threshold_pack((const float*)s->stage_in, 1, w, h, params->rg_min, params->rg_max, s->pred, s->stream));
  seg.pred = s->pred;
  seg.rg_mask = (uint8_t*)s->stage_out;
  seg.iterations = (int32_t*)((uint8_t*)s->stage_out + ((npx + 3) & ~(size_t)3));
  if ((rc = run_segment(wk, s, s->stream, seg))) return rc;
  wk->launches += 1;
  CK(cudaMemcpyAsync(mask, seg.rg_mask, npx, cudaMemcpyDeviceToHost, s->stream));
  int it = 0;
  CK(cudaMemcpyAsync(&it, seg.iterations, 4, cudaMemcpyDeviceToHost, s->stream));
  if ((rc = mark_used(s, s->stream))) return rc;
  CK(cudaStreamSynchronize(s->stream));
  if (iterations) *iterations = it;
  return NM03_OK;
}

int nm03_morphology(nm03_ctx* ctx, int gpu, const uint8_t* mask, int w, int h, int size, int se_shape, int op,
                    uint8_t* out) {
  Worker* wk;
  int rc;
  if ((rc = get_worker(ctx, gpu, &wk))) return rc;
  if (!mask || !out) return fail(NM03_ERR_INVALID, "null buffer");
  if (w <= 0 || h <= 0) return fail(NM03_ERR_INVALID, "bad image size %dx%d", w, h);
  if (op != NM03_MORPH_DILATE && op != NM03_MORPH_ERODE) return fail(NM03_ERR_INVALID, "bad morphology op %d", op);
  nm03_params dummy;
  nm03_default_params(&dummy);
  nm03::SegParams seg;
  if ((rc = fill_seg(seg, &dummy, 1, w, h, nullptr, 0, false))) return rc;
  if (op == NM03_MORPH_DILATE) rc = se_halfwidths(size, se_shape, &seg.dil_n, seg.dil_hw);
  else rc = se_halfwidths(size, se_shape, &seg.ero_n, seg.ero_hw);
  if (rc) return rc;
  DeviceGuard g(wk->device);
  SlotLease lease(wk);
  Slot* s = lease.s;
  if ((rc = order_after_last_use(s, s->stream))) return rc;
  const size_t npx = (size_t)w * h, pw = (size_t)h * seg.wpr * 4;
  if ((rc = ensure(s, &s->stage_in, &s->stage_in_bytes, npx))) return rc;
  if ((rc = ensure(s, &s->stage_out, &s->stage_out_bytes, npx))) return rc;
  if ((rc = ensure(s, (void**)&s->pred, &s->pred_bytes, pw))) return rc;
  CK(cudaMemcpyAsync(s->stage_in, mask, npx, cudaMemcpyHostToDevice, s->stream));
  CK(nm03::launch_pack_u8((const uint8_t*)s->stage_in, 1, w, h, s->pred, s->stream));
  seg.pred = s->pred;
  if (op == NM03_MORPH_DILATE) seg.dilated = (uint8_t*)s->stage_out;
  else seg.eroded = (uint8_t*)s->stage_out;
  if ((rc = run_segment(wk, s, s->stream, seg))) return rc;
  wk->launches += 1;
  CK(cudaMemcpyAsync(out, s->stage_out, npx, cudaMemcpyDeviceToHost, s->stream));
  if ((rc = mark_used(s, s->stream))) return rc;
  CK(cudaStreamSynchronize(s->stream));
  return NM03_OK;
}

}  // extern "C"
