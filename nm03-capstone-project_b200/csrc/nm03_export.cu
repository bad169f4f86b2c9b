// nm03_export.cu — see nm03_export.h.  Two kernels per batch: the window of each slice (minimum and maximum sample),
// then one thread per 8x8 block: render, level shift, forward transform, quantise, zigzag.
//
// The arithmetic is the host writer's, operation for operation (host/nm03.hpp ImageRenderer::paint identity canvas,
// host/jpeg.hpp aan8 / quantise_sse2): IEEE fp32 add, subtract and multiply without contraction, so the coefficients —
// and the JPEG bytes entropy-coded from them — are those of the CPU path (tests/test_gpu_export.py).
#include "nm03_export.h"

namespace nm03 {

namespace {

__device__ __forceinline__ int sample_value(uint32_t u16, int is_unsigned) {
  return is_unsigned ? (int)u16 : (int)(short)u16;
}

// one CTA per slice
__global__ void __launch_bounds__(1024) k_slice_window(const uint16_t* __restrict__ in, int npx, int is_unsigned, int2* __restrict__ mm) {
  const uint4* __restrict__ src = reinterpret_cast<const uint4*>(in + (size_t)blockIdx.x * npx);
  int lo = 0x7fffffff, hi = -0x7fffffff - 1;
  for (int i = threadIdx.x; i < npx / 8; i += blockDim.x) {
    const uint4 q = __ldg(src + i);
    const uint32_t wds[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int a = sample_value(wds[k] & 0xffffu, is_unsigned), b = sample_value(wds[k] >> 16, is_unsigned);
      lo = min(lo, min(a, b));
      hi = max(hi, max(a, b));
    }
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, d));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, d));
  }
  __shared__ int s_lo[32], s_hi[32];
  if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = (blockDim.x + 31) >> 5;
    lo = threadIdx.x < nw ? s_lo[threadIdx.x] : 0x7fffffff;
    hi = threadIdx.x < nw ? s_hi[threadIdx.x] : -0x7fffffff - 1;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, d));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, d));
    }
    if (threadIdx.x == 0) mm[blockIdx.x] = make_int2(lo, hi);
  }
}

// 8-point forward transform after Arai, Agui and Nakajima: the statement of jpeg::detail::aan8, same operations, same order
__device__ __forceinline__ void aan8(float& d0, float& d1, float& d2, float& d3, float& d4, float& d5, float& d6, float& d7) {
  const float t0 = __fadd_rn(d0, d7), t7 = __fsub_rn(d0, d7), t1 = __fadd_rn(d1, d6), t6 = __fsub_rn(d1, d6);
  const float t2 = __fadd_rn(d2, d5), t5 = __fsub_rn(d2, d5), t3 = __fadd_rn(d3, d4), t4 = __fsub_rn(d3, d4);
  const float e0 = __fadd_rn(t0, t3), e3 = __fsub_rn(t0, t3), e1 = __fadd_rn(t1, t2), e2 = __fsub_rn(t1, t2);
  d0 = __fadd_rn(e0, e1);
  d4 = __fsub_rn(e0, e1);
  const float z1 = __fmul_rn(__fadd_rn(e2, e3), 0.707106781f);
  d2 = __fadd_rn(e3, z1);
  d6 = __fsub_rn(e3, z1);
  const float o0 = __fadd_rn(t4, t5), o1 = __fadd_rn(t5, t6), o2 = __fadd_rn(t6, t7);
  const float z5 = __fmul_rn(__fsub_rn(o0, o2), 0.382683433f);
  const float z2 = __fadd_rn(__fmul_rn(0.541196100f, o0), z5), z4 = __fadd_rn(__fmul_rn(1.306562965f, o2), z5);
  const float z3 = __fmul_rn(o1, 0.707106781f);
  const float z11 = __fadd_rn(t7, z3), z13 = __fsub_rn(t7, z3);
  d5 = __fadd_rn(z13, z2);
  d3 = __fsub_rn(z13, z2);
  d1 = __fadd_rn(z11, z4);
  d7 = __fsub_rn(z11, z4);
}

// zigzag position -> natural (row-major) index, T.81 figure A.6
__device__ constexpr unsigned char kZig[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                               41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                               30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// forward transform (rows, then columns), quantisation (round half away from zero) and zigzag of one level-shifted block;
// two coefficients per stored word
__device__ __forceinline__ void transform_quantise_store(float (&d)[8][8], const float* __restrict__ rq, int16_t* __restrict__ blocks,
                                                         unsigned long long* __restrict__ nzp) {
#pragma unroll
  for (int y = 0; y < 8; ++y) aan8(d[y][0], d[y][1], d[y][2], d[y][3], d[y][4], d[y][5], d[y][6], d[y][7]);   // rows
#pragma unroll
  for (int x = 0; x < 8; ++x) aan8(d[0][x], d[1][x], d[2][x], d[3][x], d[4][x], d[5][x], d[6][x], d[7][x]);   // columns
  unsigned long long nz = 0ull;
  uint32_t wout[32];
#pragma unroll
  for (int k = 0; k < 64; ++k) {
    const int i = kZig[k];
    const float qv = __fmul_rn(d[i >> 3][i & 7], rq[i]);
    const int c = __float2int_rz(__fadd_rn(qv, copysignf(0.5f, qv)));
    if (c != 0) nz |= 1ull << k;
    if (k & 1) wout[k >> 1] |= (uint32_t)(c & 0xffff) << 16;
    else wout[k >> 1] = (uint32_t)(c & 0xffff);
  }
  uint4* dst = reinterpret_cast<uint4*>(blocks);
#pragma unroll
  for (int k = 0; k < 8; ++k) dst[k] = make_uint4(wout[4 * k], wout[4 * k + 1], wout[4 * k + 2], wout[4 * k + 3]);
  *nzp = nz;
}

constexpr int kExportThreads = 128;

__global__ void __launch_bounds__(kExportThreads) k_render_dct(const __grid_constant__ ExportParams p) {
  const int bw = p.w >> 3, nb = bw * (p.h >> 3);
  const long long b = (long long)blockIdx.x * kExportThreads + threadIdx.x;
  if (b >= (long long)p.n * nb) return;
  const int slice = (int)(b / nb), bi = (int)(b - (long long)slice * nb);
  const int by = bi / bw, bx = bi - by * bw;
  // the renderer's window: the full intensity range of the slice (ImageRenderer::paint)
  const int2 m = p.mm[slice];
  const float lo = (float)m.x, hi = (float)m.y;
  const float inv = hi > lo ? __fdiv_rn(255.0f, __fsub_rn(hi, lo)) : 0.0f;
  const uint16_t* __restrict__ src = reinterpret_cast<const uint16_t*>(p.in) + ((size_t)slice * p.h + (size_t)by * 8) * p.w + (size_t)bx * 8;
  float d[8][8];
#pragma unroll
  for (int y = 0; y < 8; ++y) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(src + (size_t)y * p.w));
    const uint32_t wds[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int x = 0; x < 8; ++x) {
      const uint32_t u = (x & 1) ? (wds[x >> 1] >> 16) : (wds[x >> 1] & 0xffffu);
      const float v = (float)sample_value(u, p.input_unsigned);
      float g = __fmul_rn(__fsub_rn(v, lo), inv);
      g = fminf(fmaxf(g, 0.0f), 255.0f);
      const int grey = __float2int_rz(__fadd_rn(g, 0.5f));        // round half up, as the renderer's table does
      d[y][x] = __fsub_rn((float)grey, 128.0f);                   // level shift
    }
  }
  transform_quantise_store(d, p.rq, p.blocks + (size_t)b * 64, p.nz + b);
}

// 32 mask pixels of row y starting at pixel x0 - 8 (x0 a multiple of 8); pixels outside the image are 0
__device__ __forceinline__ uint32_t mask_window(const uint32_t* __restrict__ plane, int y, int h, int wpr, int x0) {
  if (y < 0 || y >= h) return 0u;
  const uint32_t* __restrict__ row = plane + (size_t)y * wpr;
  const int k = x0 >> 5, s = 24 + (x0 & 31);                    // start within the string (word k-1, word k, word k+1)
  const unsigned long long lo = k > 0 ? __ldg(row + k - 1) : 0u, mid = __ldg(row + k), hi = k + 1 < wpr ? __ldg(row + k + 1) : 0u;
  return s < 32 ? (uint32_t)((lo | (mid << 32)) >> s) : (uint32_t)((mid | (hi << 32)) >> (s - 32));
}

__global__ void __launch_bounds__(kExportThreads) k_seg_render_dct(const __grid_constant__ ExportSegParams p) {
  const int bw = p.w >> 3, nb = bw * (p.h >> 3);
  const long long b = (long long)blockIdx.x * kExportThreads + threadIdx.x;
  if (b >= (long long)p.n * nb) return;
  const int slice = (int)(b / nb), bi = (int)(b - (long long)slice * nb);
  const int by = bi / bw, bx = bi - by * bw;
  const uint32_t* __restrict__ plane = p.bits + (size_t)slice * p.h * p.wpr;
  const int R = p.radius, x0 = 8 * bx, y0 = 8 * by;
  // rows y0-R .. y0+7+R of the mask around the block; bit 8+j of a window = pixel x0+j
  uint32_t win[8 + 14];
#pragma unroll
  for (int r = 0; r < 8 + 14; ++r) win[r] = r < 8 + 2 * R ? mask_window(plane, y0 - R + r, p.h, p.wpr, x0) : 0u;
  float d[8][8];
#pragma unroll
  for (int y = 0; y < 8; ++y) {
    const uint32_t m = win[y + R];
    // interior = the mask eroded by the border disc: row offset b contributes the row eroded horizontally by hw[b]
    uint32_t interior = m;
    for (int bb = -R; bb <= R && interior; ++bb) {
      const uint32_t rw = win[y + R + bb];
      uint32_t e = rw;
      const int a = p.hw[bb + R];
      for (int sft = 1; sft <= a; ++sft) e &= (rw >> sft) & (rw << sft);
      interior &= e;
    }
#pragma unroll
    for (int x = 0; x < 8; ++x) {
      const bool lab = (m >> (8 + x)) & 1u, in = (interior >> (8 + x)) & 1u;
      const int grey = lab ? (in ? p.interior : p.border) : p.background;
      d[y][x] = __fsub_rn((float)grey, 128.0f);
    }
  }
  transform_quantise_store(d, p.rq, p.blocks + (size_t)b * 64, p.nz + b);
}

}  // namespace

cudaError_t launch_export_blocks(const ExportParams& p, cudaStream_t s) {
  if (p.n <= 0 || p.w <= 0 || p.h <= 0 || (p.w & 7) || (p.h & 7) || !p.in || !p.mm || !p.blocks || !p.nz) return cudaErrorInvalidValue;
  k_slice_window<<<p.n, 1024, 0, s>>>(reinterpret_cast<const uint16_t*>(p.in), p.w * p.h, p.input_unsigned, p.mm);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const long long total = (long long)p.n * (p.w >> 3) * (p.h >> 3);
  const long long grid = (total + kExportThreads - 1) / kExportThreads;
  if (grid > 0x7fffffffLL) return cudaErrorInvalidValue;
  k_render_dct<<<(unsigned)grid, kExportThreads, 0, s>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_export_seg_blocks(const ExportSegParams& p, cudaStream_t s) {
  if (p.n <= 0 || p.w <= 0 || p.h <= 0 || (p.w & 7) || (p.h & 7) || p.radius < 0 || p.radius > 7 || !p.bits || !p.blocks || !p.nz)
    return cudaErrorInvalidValue;
  const long long total = (long long)p.n * (p.w >> 3) * (p.h >> 3);
  const long long grid = (total + kExportThreads - 1) / kExportThreads;
  if (grid > 0x7fffffffLL) return cudaErrorInvalidValue;
  k_seg_render_dct<<<(unsigned)grid, kExportThreads, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace nm03
