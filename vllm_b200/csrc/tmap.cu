// Host-side TMA tensor-map creation.  libvllm_b200 does not link libcuda: cuTensorMapEncodeTiled is
// resolved once through cudaGetDriverEntryPoint.
#include <mutex>

#include "tc_common.cuh"

namespace vb200 {
namespace tc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tmap_2d_u8(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                    uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  VB_CHECK(enc != nullptr, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  VB_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld & 15) == 0,
           "TMA operands need 16-byte aligned base and row pitch (base %p, pitch %llu)", base,
           (unsigned long long)ld);
  VB_CHECK(box_rows >= 1 && box_rows <= 256, "TMA box rows %u out of range", box_rows);
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {ld};
  const cuuint32_t box[2] = {128, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// 4-D byte tensor {inner bytes, dim1, dim2, dim3} with byte strides for dims 1..3 (any order of magnitudes: this is how
// one map covers both the NHD and the HND-permuted KV-cache views); box = {128 bytes, box1, 1, 1}, 128-byte swizzle.
int make_tmap_4d_u8(CUtensorMap* map, const void* base, const uint64_t dims[4], const uint64_t strides_bytes[3],
                    uint32_t box1) {
  EncodeTiledFn enc = get_encode();
  VB_CHECK(enc != nullptr, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  VB_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA operands need a 16-byte aligned base (%p)", base);
  for (int i = 0; i < 3; ++i)
    VB_CHECK((strides_bytes[i] & 15) == 0 && strides_bytes[i] > 0, "TMA strides must be positive multiples of 16 bytes (%llu)",
             (unsigned long long)strides_bytes[i]);
  VB_CHECK(dims[0] >= 128 && box1 >= 1 && box1 <= 256, "TMA box out of range");
  const cuuint64_t d[4] = {dims[0], dims[1], dims[2], dims[3]};
  const cuuint64_t st[3] = {strides_bytes[0], strides_bytes[1], strides_bytes[2]};
  const cuuint32_t box[4] = {128, box1, 1, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), d, st, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (4-D) failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace tc
}  // namespace vb200
