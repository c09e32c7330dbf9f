// Temporary: entry points of the tcgen05 GEMMs until scaled_mm.cu / moe_gemm.cu land.
#include "common.cuh"
using namespace vb200;
extern "C" int vb200_moe_grouped_gemm_fp8(void*, const void*, const void*, const float*,
                                          const float*, const float*, const int32_t*,
                                          const int32_t*, const int32_t*, int, int, int, int, int,
                                          int, int, int, int, int, int, void*, size_t,
                                          vb200_stream_t) {
  VB_CHECK(false, "vb200_moe_grouped_gemm_fp8: not built yet");
  return 1;
}
extern "C" size_t vb200_moe_grouped_gemm_workspace_size(int, int) { return 0; }
