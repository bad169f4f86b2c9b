// nm03_k1_w.cu — the K1 instantiations of ONE median window (-DNM03_K1_WINDOW=k): every sharpen mask half-width and
// every sharpen mode.  One object per window keeps the build parallel (the 7x7 and 9x9 networks dominate compile time).
#include "nm03_k1.cuh"

#ifndef NM03_K1_WINDOW
#error "compile with -DNM03_K1_WINDOW=<1|3|5|7|9|11|13|15>"
#endif
#define NM03_CAT_(a, b) a##b
#define NM03_CAT(a, b) NM03_CAT_(a, b)

namespace nm03 {

cudaError_t NM03_CAT(launch_k1_w, NM03_K1_WINDOW)(const PreParams& p, int sharpen_half, int mode, cudaStream_t s) {
  switch (sharpen_half) {
    case 1: return launch_k1<NM03_K1_WINDOW, 1>(p, mode, s);
    case 2: return launch_k1<NM03_K1_WINDOW, 2>(p, mode, s);
    case 3: return launch_k1<NM03_K1_WINDOW, 3>(p, mode, s);
    case 4: return launch_k1<NM03_K1_WINDOW, 4>(p, mode, s);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t NM03_CAT(config_k1_w, NM03_K1_WINDOW)() {
  cudaError_t e = config_k1<NM03_K1_WINDOW, 1>();
  if (e == cudaSuccess) e = config_k1<NM03_K1_WINDOW, 2>();
  if (e == cudaSuccess) e = config_k1<NM03_K1_WINDOW, 3>();
  if (e == cudaSuccess) e = config_k1<NM03_K1_WINDOW, 4>();
  return e;
}

}  // namespace nm03
