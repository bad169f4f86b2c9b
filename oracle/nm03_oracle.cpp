// nm03_oracle.cpp — CPU ORACLE (test infrastructure only; PARITY UNPINNED — see nm03_oracle.h).
//
// One function per FAST operator the reference's processSingleImage() invokes
// (reference src/sequential/main_sequential.cpp:175-252).  Written for clarity, not speed: scalar,
// single-threaded per slice; the only parallelism is the OpenMP loop over slices that restates
// src/parallel/main_parallel.cpp:336-343.  Build with -ffp-contract=off so no multiply-add is fused
// unless the sharpen_fma switch asks for fmaf explicitly.
#include "nm03_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// sampler read: clamp-to-edge or zero outside the image
inline float sample(const float* img, int w, int h, int x, int y, int border_mode) {
  if (border_mode == NM03O_BORDER_ZERO) {
    if (x < 0 || y < 0 || x >= w || y >= h) return 0.0f;
    return img[(long)y * w + x];
  }
  return img[(long)clampi(y, 0, h - 1) * w + clampi(x, 0, w - 1)];
}

// structuring element offsets: disc a*a+b*b <= N*N (FAST: length((a,b)) > N is skipped) or square
std::vector<std::pair<int, int>> se_offsets(int size, int se_shape) {
  const int N = size / 2;
  std::vector<std::pair<int, int>> o;
  for (int b = -N; b <= N; ++b)
    for (int a = -N; a <= N; ++a)
      if (se_shape == NM03O_SE_SQUARE || a * a + b * b <= N * N) o.push_back({a, b});
  return o;
}

}  // namespace

extern "C" {

void nm03o_default_params(nm03o_params* p) {
  // the literals of main_sequential.cpp:196,200,204,208,233,250
  p->norm_low = 0.5f; p->norm_high = 2.5f; p->norm_min = 0.0f; p->norm_max = 10000.0f;
  p->clip_min = 0.68f; p->clip_max = 4000.0f;
  p->median_window = 7;
  p->sharpen_gain = 2.0f; p->sharpen_sigma = 0.5f; p->sharpen_mask = 9;
  p->rg_min = 0.74f; p->rg_max = 0.91f;
  p->dilate_size = 3; p->erode_size = 0;
  p->border_mode = NM03O_BORDER_CLAMP;
  p->sharpen_fma = 0;
  p->rg_mode = NM03O_RG_OPENCL;
  p->rg_connectivity = 4;
  p->rg_border_expand = 1;
  p->se_shape = NM03O_SE_DISC;
  p->input_unsigned = 0;
}

// main_sequential.cpp:214-241 — five centre-relative seeds, then the x-major grid over the
// central half, all in `int` arithmetic.
int nm03o_make_seeds(int w, int h, int* xy, int cap) {
  if (w < 10 || h < 10) return -1;  // step width/10 would be 0: the reference loop never ends
  int n = 0;
  auto put = [&](int x, int y) {
    if (n < cap && xy) { xy[2 * n] = x; xy[2 * n + 1] = y; }
    ++n;
  };
  const int cx = w / 2, cy = h / 2, ox = w / 8, oy = h / 8;
  put(cx, cy);
  put(cx + ox, cy);
  put(cx - ox, cy);
  put(cx, cy + oy);
  put(cx, cy - oy);
  for (int x = w / 4; x < w * 3 / 4; x += w / 10)
    for (int y = h / 4; y < h * 3 / 4; y += h / 10) put(x, y);
  return n;
}

// t = (float(v) - min) / (max - min); out = t * (high - low) + low.  No clamp.
int nm03o_normalize(const void* in, int is_unsigned, long n, float low, float high, float mn, float mx, float* out) {
  const float range = mx - mn;
  const float scale = high - low;
  for (long i = 0; i < n; ++i) {
    const float v = is_unsigned ? (float)((const uint16_t*)in)[i] : (float)((const int16_t*)in)[i];
    const float t = (v - mn) / range;
    out[i] = t * scale + low;
  }
  return 0;
}

// SURVEY §8c-2
int nm03o_clip(const float* in, long n, float lo, float hi, float* out) {
  for (long i = 0; i < n; ++i) out[i] = fminf(fmaxf(in[i], lo), hi);
  return 0;
}

// k odd, centred window, the (k*k+1)/2-th order statistic.  For one channel FAST's
// brute-force vector median (the sample minimising the summed distance to the rest) IS the scalar
// median value; an O(k^2) selection is used here instead of FAST's O(k^4) search.
int nm03o_median(const float* in, int w, int h, int k, int border_mode, float* out) {
  if (k < 1 || (k & 1) == 0) return -3;
  const int r = k / 2, cnt = k * k, mid = cnt / 2;
  std::vector<float> win(cnt);
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      int i = 0;
      for (int b = -r; b <= r; ++b)
        for (int a = -r; a <= r; ++a) win[i++] = sample(in, w, h, x + a, y + b, border_mode);
      std::nth_element(win.begin(), win.begin() + mid, win.end());
      out[(long)y * w + x] = win[mid];
    }
  return 0;
}

// normalised Gaussian, filled x-outer / y-inner, sum accumulated in the same order.
int nm03o_gauss_mask(float sigma, int m, float* mask) {
  if (m < 1 || (m & 1) == 0) return -3;
  const int hs = (m - 1) / 2;
  float sum = 0.0f;
  for (int x = -hs; x <= hs; ++x)
    for (int y = -hs; y <= hs; ++y) {
      const float v = expf(-(float)(x * x + y * y) / (2.0f * sigma * sigma));
      mask[x + hs + (y + hs) * m] = v;
      sum += v;
    }
  for (int i = 0; i < m * m; ++i) mask[i] /= sum;
  return 0;
}

// unsharp mask: out = v + gain * (v - blur); blur accumulated from 0.0f x-outer / y-inner.
int nm03o_sharpen(const float* in, int w, int h, float gain, float sigma, int m, int border_mode, int use_fma,
                  float* out) {
  if (m == 0) m = (int)ceilf(2.0f * sigma) * 2 + 1;
  if (m < 1 || (m & 1) == 0) return -3;
  const int hs = (m - 1) / 2;
  std::vector<float> mask((size_t)m * m);
  nm03o_gauss_mask(sigma, m, mask.data());
  for (int py = 0; py < h; ++py)
    for (int px = 0; px < w; ++px) {
      float blur = 0.0f;
      for (int x = -hs; x <= hs; ++x)
        for (int y = -hs; y <= hs; ++y) {
          const float wgt = mask[x + hs + (y + hs) * m];
          const float v = sample(in, w, h, px + x, py + y, border_mode);
          if (use_fma) blur = fmaf(wgt, v, blur);
          else { const float prod = wgt * v; blur = blur + prod; }
        }
      const float v = in[(long)py * w + px];
      const float d = v - blur;
      const float g = gain * d;
      out[(long)py * w + px] = v + g;
    }
  return 0;
}

// SURVEY §8c-5.
int nm03o_region_grow(const float* in, int w, int h, float lo, float hi, const int* seeds_xy, int nseeds,
                      int mode, int conn, int border_expand, uint8_t* out, int* iterations) {
  if (nseeds <= 0) return -4;                 // FAST: "no seed points" exception
  if (conn != 4 && conn != 8) return -3;
  for (int i = 0; i < nseeds; ++i) {
    const int x = seeds_xy[2 * i], y = seeds_xy[2 * i + 1];
    if (x < 0 || y < 0 || x >= w || y >= h) return -5;  // FAST: seed outside image
  }
  static const int dx[8] = {0, 1, -1, 0, 1, -1, 1, -1};
  static const int dy[8] = {1, 0, 0, -1, 1, 1, -1, -1};
  const long npx = (long)w * h;
  std::vector<uint8_t> label(npx, 0);
  int iters = 0;
  auto in_range = [&](long i) { return in[i] >= lo && in[i] <= hi; };

  if (mode == NM03O_RG_HOST) {
    // FAST executeOnHost: LIFO stack of seeds; a popped pixel's 4-neighbours are tested; the seed
    // itself is labelled only if it is reached from a neighbour.
    std::vector<long> stack;
    for (int i = 0; i < nseeds; ++i) stack.push_back((long)seeds_xy[2 * i + 1] * w + seeds_xy[2 * i]);
    while (!stack.empty()) {
      const long p = stack.back();
      stack.pop_back();
      const int x = (int)(p % w), y = (int)(p / w);
      for (int d = 0; d < conn; ++d) {
        const int nx = x + dx[d], ny = y + dy[d];
        if (nx < 0 || ny < 0 || nx >= w || ny >= h) continue;
        const long q = (long)ny * w + nx;
        if (label[q] != 1 && in_range(q)) { label[q] = 1; stack.push_back(q); }
      }
      ++iters;
    }
  } else {
    // FAST OpenCL path: labels 0 = background, 2 = queued, 1 = in region.  Each launch visits every
    // queued pixel; level-synchronous here (the fixed point does not depend on the visiting order).
    std::vector<long> frontier, next;
    for (int i = 0; i < nseeds; ++i) {
      const long p = (long)seeds_xy[2 * i + 1] * w + seeds_xy[2 * i];
      if (label[p] == 0) { label[p] = 2; frontier.push_back(p); }
    }
    while (!frontier.empty()) {
      ++iters;
      next.clear();
      for (long p : frontier) {
        if (label[p] != 2) continue;
        if (!in_range(p)) { label[p] = 0; continue; }
        label[p] = 1;
        const int x = (int)(p % w), y = (int)(p / w);
        if (!border_expand && (x == 0 || y == 0 || x == w - 1 || y == h - 1)) continue;
        for (int d = 0; d < conn; ++d) {
          const int nx = x + dx[d], ny = y + dy[d];
          if (nx < 0 || ny < 0 || nx >= w || ny >= h) continue;
          const long q = (long)ny * w + nx;
          if (label[q] == 0) {
            // a pixel that failed the range test goes back to 0 and may be queued again; it can
            // never pass, so skip it here — the fixed point is identical.
            if (in_range(q)) { label[q] = 2; next.push_back(q); }
          }
        }
      }
      frontier.swap(next);
    }
  }
  for (long i = 0; i < npx; ++i) out[i] = label[i] == 1 ? 1 : 0;
  if (iterations) *iterations = iters;
  return 0;
}

// every source pixel equal to 1 stamps the SE; writes outside the image are skipped.
int nm03o_dilate(const uint8_t* in, int w, int h, int size, int se_shape, uint8_t* out) {
  if (size <= 2 || (size & 1) == 0) return -3;  // FAST: size must be odd and > 2
  const auto se = se_offsets(size, se_shape);
  std::memset(out, 0, (size_t)w * h);
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      if (in[(long)y * w + x] == 1)
        for (auto [a, b] : se) {
          const int nx = x + a, ny = y + b;
          if (nx < 0 || ny < 0 || nx >= w || ny >= h) continue;
          out[(long)ny * w + nx] = 1;
        }
  return 0;
}

// keep a pixel iff it and every SE neighbour are 1; outside the image reads 0.
int nm03o_erode(const uint8_t* in, int w, int h, int size, int se_shape, uint8_t* out) {
  if (size <= 2 || (size & 1) == 0) return -3;
  const auto se = se_offsets(size, se_shape);
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      uint8_t keep = in[(long)y * w + x] == 1;
      if (keep)
        for (auto [a, b] : se) {
          const int nx = x + a, ny = y + b;
          if (nx < 0 || ny < 0 || nx >= w || ny >= h || in[(long)ny * w + nx] != 1) { keep = 0; break; }
        }
      out[(long)y * w + x] = keep;
    }
  return 0;
}

// processSingleImage(), main_sequential.cpp:170-252, stage by stage.
int nm03o_pipeline(const void* in, int w, int h, const nm03o_params* p, const nm03o_outputs* o) {
  if (w < 100 || h < 100) return -2;  // main_sequential.cpp:189-192
  const long n = (long)w * h;
  std::vector<float> a(n), b(n);
  int rc;
  if ((rc = nm03o_normalize(in, p->input_unsigned, n, p->norm_low, p->norm_high, p->norm_min, p->norm_max, a.data()))) return rc;
  if ((rc = nm03o_clip(a.data(), n, p->clip_min, p->clip_max, b.data()))) return rc;
  if ((rc = nm03o_median(b.data(), w, h, p->median_window, p->border_mode, a.data()))) return rc;
  if (o && o->median) std::memcpy(o->median, a.data(), n * sizeof(float));
  if ((rc = nm03o_sharpen(a.data(), w, h, p->sharpen_gain, p->sharpen_sigma, p->sharpen_mask, p->border_mode,
                          p->sharpen_fma, b.data()))) return rc;
  if (o && o->sharpened) std::memcpy(o->sharpened, b.data(), n * sizeof(float));
  int seeds[2 * 256];
  const int ns = nm03o_make_seeds(w, h, seeds, 256);
  if (ns < 0 || ns > 256) return -4;
  std::vector<uint8_t> m(n), d(n);
  int iters = 0;
  if ((rc = nm03o_region_grow(b.data(), w, h, p->rg_min, p->rg_max, seeds, ns, p->rg_mode, p->rg_connectivity,
                              p->rg_border_expand, m.data(), &iters))) return rc;
  if (o && o->rg_iterations) *o->rg_iterations = iters;
  if (o && o->rg_mask) std::memcpy(o->rg_mask, m.data(), n);
  // ImageCaster(TYPE_UINT8) is the identity on a {0,1} uint8 segmentation (main_sequential.cpp:246)
  if (p->erode_size > 0 && o && o->eroded) {
    if ((rc = nm03o_erode(m.data(), w, h, p->erode_size, p->se_shape, o->eroded))) return rc;
  }
  if ((rc = nm03o_dilate(m.data(), w, h, p->dilate_size, p->se_shape, d.data()))) return rc;
  if (o && o->dilated) std::memcpy(o->dilated, d.data(), n);
  return 0;
}

int nm03o_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// main_parallel.cpp:330-347 — batches of `batch` slices, one OpenMP loop per batch.
int nm03o_run_batch_omp(const void* slices, int n, int w, int h, const nm03o_params* p, int batch, int threads,
                        uint8_t* dilated) {
  int ok = 0;
  if (batch <= 0) batch = 25;  // DEFAULT_BATCH_SIZE, main_parallel.cpp:33
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
#else
  (void)threads;
#endif
  const long npx = (long)w * h;
  for (int start = 0; start < n; start += batch) {
    const int cur = std::min(batch, n - start);
#pragma omp parallel for schedule(auto) reduction(+ : ok)
    for (int i = 0; i < cur; ++i) {
      const long idx = start + i;
      nm03o_outputs o{};
      o.dilated = dilated ? dilated + idx * npx : nullptr;
      std::vector<uint8_t> tmp;
      if (!o.dilated) { tmp.resize(npx); o.dilated = tmp.data(); }
      const int rc = nm03o_pipeline((const uint8_t*)slices + idx * npx * 2, w, h, p, &o);
      if (rc == 0) ok++;
    }
  }
  return ok;
}

}  // extern "C"
