// Does write-combined pinned memory change the H2D rate in the bench's copy pattern (12 MiB in and 6 MiB out per step,
// on two streams at the same time)?   nvcc -O2 -o pcie_wc_probe pcie_wc_probe.cu && ./pcie_wc_probe
#include <chrono>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
int main() {
  const size_t n = 12u << 20;
  void *d, *d2, *h2;
  cudaMalloc(&d, n); cudaMalloc(&d2, n / 2); cudaHostAlloc(&h2, n / 2, cudaHostAllocPortable);
  cudaStream_t s1, s2; cudaStreamCreate(&s1); cudaStreamCreate(&s2);
  const unsigned flags[2] = {cudaHostAllocPortable, cudaHostAllocPortable | cudaHostAllocWriteCombined};
  const char* names[2] = {"portable", "portable + write-combined"};
  for (int k = 0; k < 2; ++k) {
    void* h; if (cudaHostAlloc(&h, n, flags[k]) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    memset(h, 1, n);
    for (int alone = 1; alone >= 0; --alone) {
      for (int i = 0; i < 5; ++i) cudaMemcpyAsync(d, h, n, cudaMemcpyHostToDevice, s1);
      cudaDeviceSynchronize();
      const int reps = 300;
      auto t0 = std::chrono::steady_clock::now();
      for (int i = 0; i < reps; ++i) {
        cudaMemcpyAsync(d, h, n, cudaMemcpyHostToDevice, s1);
        if (!alone) cudaMemcpyAsync(h2, d2, n / 2, cudaMemcpyDeviceToHost, s2);
      }
      cudaDeviceSynchronize();
      const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      printf("%-28s %s: H2D %.1f GB/s\n", names[k], alone ? "alone            " : "with D2H (half)  ", n * reps / dt / 1e9);
    }
    cudaFreeHost(h);
  }
  return 0;
}
