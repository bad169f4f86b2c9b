// Microbenchmark: issue throughput of packed min/max flavours on sm_100a, alone and mixed.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_minmax ubench_minmax.cu
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed, int iters) {
  uint32_t a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = seed * (threadIdx.x + 1) + i * 0x01230123u; b[i] = a[i] ^ 0x00550055u; }
  float fa[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) fa[i] = (float)a[i];
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) {  // VIMNMX.S16x2 only: 2 ops
        uint32_t lo = __vmins2(a[i], b[i]), hi = __vmaxs2(a[i], b[i]); a[i] = lo; b[i] = hi ^ 1u;
      } else if (MODE == 1) {  // HMNMX2 only
        __half2 x = *reinterpret_cast<__half2*>(&a[i]), y = *reinterpret_cast<__half2*>(&b[i]);
        __half2 lo = __hmin2(x, y), hi = __hmax2(x, y);
        a[i] = *reinterpret_cast<uint32_t*>(&lo); b[i] = *reinterpret_cast<uint32_t*>(&hi) ^ 1u;
      } else if (MODE == 2) {  // half of the pairs on each
        if (i & 1) { uint32_t lo = __vmins2(a[i], b[i]), hi = __vmaxs2(a[i], b[i]); a[i] = lo; b[i] = hi ^ 1u; }
        else {
          __half2 x = *reinterpret_cast<__half2*>(&a[i]), y = *reinterpret_cast<__half2*>(&b[i]);
          __half2 lo = __hmin2(x, y), hi = __hmax2(x, y);
          a[i] = *reinterpret_cast<uint32_t*>(&lo); b[i] = *reinterpret_cast<uint32_t*>(&hi) ^ 1u;
        }
      } else if (MODE == 3) {  // VIMNMX + independent FFMA stream (does the fma pipe overlap?)
        uint32_t lo = __vmins2(a[i], b[i]), hi = __vmaxs2(a[i], b[i]); a[i] = lo; b[i] = hi ^ 1u;
        fa[i] = fmaf(fa[i], 1.0001f, 0.5f); fa[i] = fmaf(fa[i], 0.9999f, 0.25f);
      } else if (MODE == 4) {  // VIMNMX3
        uint32_t lo = __vimin3_s16x2(a[i], b[i], a[(i + 1) & 7]), hi = __vimax3_s16x2(a[i], b[i], b[(i + 1) & 7]);
        a[i] = lo; b[i] = hi ^ 1u;
      } else if (MODE == 5) {  // scalar 32-bit IMNMX
        int lo = min((int)a[i], (int)b[i]), hi = max((int)a[i], (int)b[i]); a[i] = lo; b[i] = hi ^ 1;
      } else if (MODE == 6) {  // HMNMX2 + FFMA
        __half2 x = *reinterpret_cast<__half2*>(&a[i]), y = *reinterpret_cast<__half2*>(&b[i]);
        __half2 lo = __hmin2(x, y), hi = __hmax2(x, y);
        a[i] = *reinterpret_cast<uint32_t*>(&lo); b[i] = *reinterpret_cast<uint32_t*>(&hi) ^ 1u;
        fa[i] = fmaf(fa[i], 1.0001f, 0.5f); fa[i] = fmaf(fa[i], 0.9999f, 0.25f);
      }
    }
  }
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) r ^= a[i] ^ b[i] ^ __float_as_uint(fa[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
void run(const char* name, uint32_t* d, double ops_per_iter_elem) {
  const int iters = 4096, blocks = 148 * 4, threads = 256;
  k<MODE><<<blocks, threads>>>(d, 3, 16);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(d, 7, iters);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double lane_ops = (double)blocks * threads * iters * 8 * ops_per_iter_elem;
  printf("%-28s %8.3f ms  %7.2f T lane-minmax/s  (%.1f lanes/clk/SM @1.9GHz)\n", name, ms, lane_ops / ms * 1e-9,
         lane_ops / (ms * 1e-3) / 148 / 1.9e9);
}

int main() {
  uint32_t* d; cudaMalloc(&d, 148 * 4 * 256 * 4);
  run<0>("VIMNMX.S16x2", d, 2);
  run<1>("HMNMX2", d, 2);
  run<2>("VIMNMX.S16x2 + HMNMX2 mixed", d, 2);
  run<3>("VIMNMX.S16x2 (+2 FFMA each)", d, 2);
  run<4>("VIMNMX3.S16x2", d, 2);
  run<5>("IMNMX 32-bit", d, 2);
  run<6>("HMNMX2 (+2 FFMA each)", d, 2);
  return 0;
}
