// Stand-alone check of the TMA bulk tensor load used by K1 (3-D uint16 tensor, 144 x 78 box, OOB zeros).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ CUtensorMap tmap, uint16_t* out, int gx0, int gy0, int z, int mode, int bw, int bh) {
  extern __shared__ __align__(128) uint8_t raw[];
  uint16_t* T = reinterpret_cast<uint16_t*>((reinterpret_cast<uintptr_t>(raw) + 127) & ~(uintptr_t)127);
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bw * bh * 2) : "memory");
    if (mode == 3)
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(T)), "l"(&tmap), "r"(smem_u32(&bar)), "r"(gx0), "r"(gy0), "r"(z) : "memory");
    else
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(T)), "l"(&tmap), "r"(smem_u32(&bar)), "r"(gx0), "r"(gy0) : "memory");
  }
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = T[i];
}
int main(int argc, char** argv) {
  const int BW = argc > 1 ? atoi(argv[1]) : 144, BH = argc > 2 ? atoi(argv[2]) : 78, RANK = argc > 3 ? atoi(argv[3]) : 3, L2P = argc > 4 ? atoi(argv[4]) : 1;
  const int w = 256, h = 256, n = 3;
  std::vector<uint16_t> img((size_t)w * h * n);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (uint16_t)(i * 7 + 1);
  uint16_t *d, *o; cudaMalloc(&d, img.size() * 2); cudaMalloc(&o, 256 * 256 * 2);
  cudaMemcpy(d, img.data(), img.size() * 2, cudaMemcpyHostToDevice);
  void* f = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
  typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap map; memset(&map, 0, sizeof(map));
  cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n}, strides[2] = {(cuuint64_t)w * 2, (cuuint64_t)w * h * 2};
  cuuint32_t box[3] = {(cuuint32_t)BW, (cuuint32_t)BH, 1}, es[3] = {1, 1, 1};
  CUresult r = ((Fn)f)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, RANK, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, L2P ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode rc=%d\n", (int)r);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int t = 0; t < 3; ++t) {
    int gx0 = t == 0 ? 20 : (t == 1 ? -7 : 200), gy0 = t == 0 ? 30 : (t == 1 ? -7 : 220), z = t;
    if (RANK == 2) z = 0;
    k<<<1, 256, BW * BH * 2 + 256>>>(map, o, gx0, gy0, z, RANK, BW, BH);
    cudaError_t e = cudaDeviceSynchronize();
    printf("case %d: %s\n", t, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<uint16_t> out(BW * BH); cudaMemcpy(out.data(), o, out.size() * 2, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int y = 0; y < BH; ++y) for (int x = 0; x < BW; ++x) {
      int gx = gx0 + x, gy = gy0 + y; uint16_t exp = (gx < 0 || gy < 0 || gx >= w || gy >= h) ? 0 : img[((size_t)z * h + gy) * w + gx];
      if (out[y * BW + x] != exp) ++bad;
    }
    printf("  mismatches: %d\n", bad);
  }
  return 0;
}
