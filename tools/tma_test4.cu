// Bisect: canonical libcu++ TMA example morphing into the raw-PTX form used by K1.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
constexpr int BW = 64, BH = 32;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int V>
__global__ void k(const __grid_constant__ CUtensorMap tmap, uint16_t* out, int x, int y) {
  __shared__ alignas(128) uint16_t sbuf[BH][BW];
  extern __shared__ __align__(128) uint8_t dyn[];
  uint16_t* buf = V >= 4 ? reinterpret_cast<uint16_t*>((reinterpret_cast<uintptr_t>(dyn) + 127) & ~(uintptr_t)127) : V >= 3 ? reinterpret_cast<uint16_t*>(dyn) : &sbuf[0][0];
  if (V == 0 || V == 1) {
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
      if (V == 0) cde::cp_async_bulk_tensor_2d_global_to_shared(buf, &tmap, x, y, bar);
      else asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                        ::"r"(smem_u32(buf)), "l"(&tmap), "r"(smem_u32(cuda::device::barrier_native_handle(bar))), "r"(x), "r"(y) : "memory");
      token = cuda::device::barrier_arrive_tx(bar, 1, BW * BH * 2);
    } else token = bar.arrive();
    bar.wait(std::move(token));
  } else {
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(BW * BH * 2) : "memory");
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                   ::"r"(smem_u32(buf)), "l"(&tmap), "r"(smem_u32(&bar)), "r"(x), "r"(y) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
  }
  for (int i = threadIdx.x; i < BW * BH; i += blockDim.x) out[i] = buf[i];
}
__global__ void k1(const __grid_constant__ CUtensorMap tmap, uint16_t* out, int gx0, int gy0, int z, int mode, int bw, int bh) {
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
  const int V = argc > 1 ? atoi(argv[1]) : 0;
  const int w = 256, h = 256;
  std::vector<uint16_t> img((size_t)w * h);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (uint16_t)(i * 7 + 1);
  uint16_t *d, *o; cudaMalloc(&d, img.size() * 2); cudaMalloc(&o, BW * BH * 2);
  cudaMemcpy(d, img.data(), img.size() * 2, cudaMemcpyHostToDevice);
  void* f = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
  typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap map; memset(&map, 0, sizeof(map));
  cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h}, strides[1] = {(cuuint64_t)w * 2};
  cuuint32_t box[2] = {BW, BH}, es[2] = {1, 1};
  ((Fn)f)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  const int X = argc > 2 ? atoi(argv[2]) : 16, Y = argc > 3 ? atoi(argv[3]) : 8;
  if (V == 10) k1<<<1, 256, BW * BH * 2 + 256>>>(map, o, X, Y, 0, 2, BW, BH);
  if (V == 0) k<0><<<1, 128>>>(map, o, 16, 8);
  if (V == 1) k<1><<<1, 128>>>(map, o, 16, 8);
  if (V == 2) k<2><<<1, 128>>>(map, o, 16, 8);
  if (V == 3) k<3><<<1, 128, BW * BH * 2>>>(map, o, 16, 8);
  if (V == 4) k<4><<<1, 128, BW * BH * 2 + 256>>>(map, o, 16, 8);
  if (V == 5) k<5><<<1, 256, BW * BH * 2 + 256>>>(map, o, 16, 8);
  if (V == 6) { cudaFuncSetAttribute(k<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024); k<6><<<1, 256, BW * BH * 2 + 256>>>(map, o, 16, 8); }
  cudaError_t e = cudaDeviceSynchronize();
  printf("variant %d run: %s\n", V, cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<uint16_t> out(BW * BH); cudaMemcpy(out.data(), o, out.size() * 2, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int y = 0; y < BH; ++y) for (int x = 0; x < BW; ++x) if (out[y * BW + x] != img[(size_t)(Y + y) * w + X + x]) ++bad;
  printf("  mismatches %d\n", bad);
  return 0;
}
