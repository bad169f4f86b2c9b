// Canonical TMA example (CUDA programming guide) with libcu++ wrappers, for comparison.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
constexpr int BW = 64, BH = 32;
__global__ void k(const __grid_constant__ CUtensorMap tmap, uint16_t* out, int x, int y) {
  __shared__ alignas(128) uint16_t buf[BH][BW];
#pragma nv_diag_suppress static_var_with_dynamic_init
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token token;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_2d_global_to_shared(&buf, &tmap, x, y, bar);
    token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(buf));
  } else token = bar.arrive();
  bar.wait(std::move(token));
  for (int i = threadIdx.x; i < BW * BH; i += blockDim.x) out[i] = (&buf[0][0])[i];
}
int main() {
  const int w = 256, h = 256;
  std::vector<uint16_t> img((size_t)w * h);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (uint16_t)(i * 7 + 1);
  uint16_t *d, *o; cudaMalloc(&d, img.size() * 2); cudaMalloc(&o, BW * BH * 2);
  cudaMemcpy(d, img.data(), img.size() * 2, cudaMemcpyHostToDevice);
  void* f = nullptr; cudaDriverEntryPointQueryResult q;
  cudaError_t ge = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
  printf("entry point: %s q=%d f=%p\n", cudaGetErrorString(ge), (int)q, f);
  typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap map; memset(&map, 0, sizeof(map));
  cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h}, strides[1] = {(cuuint64_t)w * 2};
  cuuint32_t box[2] = {BW, BH}, es[2] = {1, 1};
  CUresult r = ((Fn)f)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode rc=%d\n", (int)r);
  const uint64_t* mw = (const uint64_t*)&map;
  for (int i = 0; i < 16; ++i) printf("%016llx%c", (unsigned long long)mw[i], i % 4 == 3 ? '\n' : ' ');
  k<<<1, 128>>>(map, o, 16, 8);
  cudaError_t e = cudaDeviceSynchronize();
  printf("run: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<uint16_t> out(BW * BH); cudaMemcpy(out.data(), o, out.size() * 2, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int y = 0; y < BH; ++y) for (int x = 0; x < BW; ++x) if (out[y * BW + x] != img[(size_t)(8 + y) * w + 16 + x]) ++bad;
  printf("mismatches %d\n", bad);
  return 0;
}
