// Shared helpers for libvllm_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <type_traits>

#include "../../include/vllm_b200.h"

namespace vb200 {

// ---- error reporting (thread-local message, non-zero return) ---------------------------------
void set_error(const char* fmt, ...);

#define VB_CHECK(cond, ...)            \
  do {                                 \
    if (!(cond)) {                     \
      ::vb200::set_error(__VA_ARGS__); \
      return 1;                        \
    }                                  \
  } while (0)

#define VB_CUDA_CHECK(expr)                                                             \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      ::vb200::set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, \
                         __LINE__, cudaGetErrorString(_e));                             \
      return 2;                                                                         \
    }                                                                                   \
  } while (0)

#define VB_LAUNCH_CHECK() VB_CUDA_CHECK(cudaPeekAtLastError())

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// ---- programmatic dependent launch (PDL) --------------------------------------------------------
// Every kernel of the library starts with pdl_wait() (griddepcontrol.wait: the grids it depends on have completed
// and their memory is visible) before it touches global memory, then pdl_trigger() (the NEXT kernel on the stream
// may be scheduled; its CTAs park at their own pdl_wait()).  Host side, launch_k() adds
// cudaLaunchAttributeProgrammaticStreamSerialization, so the launch latency and the prologue (barrier init, TMEM
// allocation, tensor-map prefetch) of kernel i+1 overlap the tail of kernel i -- also inside captured CUDA graphs,
// where the attribute becomes a programmatic dependency edge.  Both instructions are no-ops for a kernel that was
// launched without the attribute.  VB200_PDL=0 turns the attribute off.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // runtime.cu

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                            Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Same, for a kernel whose CTAs form clusters of `cluster_x` along x (grid.x must be a multiple of it).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                    unsigned cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_x;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---- scalar conversions ----------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float to_float(T v);
template <>
__device__ __forceinline__ float to_float<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_float<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}

template <typename T>
__device__ __forceinline__ T from_float(float v);
template <>
__device__ __forceinline__ float from_float<float>(float v) { return v; }
template <>
__device__ __forceinline__ __half from_float<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// fp8 e4m3 quantisation exactly as the reference: __nv_cvt_float_to_fp8(v, SATFINITE, E4M3)
// (csrc/quantization/fp8/nvidia/quant_utils.cuh:466-479).
__device__ __forceinline__ uint8_t float_to_e4m3(float v) {
  return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
}
// Four clamped floats -> four e4m3 bytes (first argument in the low byte): the same clamp-then-RNE conversion as
// float_to_e4m3(fmaxf(-448, fminf(x, 448))) per element, two elements per cvt instruction.
__device__ __forceinline__ uint32_t e4m3x4_clamped(float a, float b, float c, float d) {
  a = fmaxf(-448.f, fminf(a, 448.f));
  b = fmaxf(-448.f, fminf(b, 448.f));
  c = fmaxf(-448.f, fminf(c, 448.f));
  d = fmaxf(-448.f, fminf(d, 448.f));
  uint16_t lo, hi;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(b), "f"(a));
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(d), "f"(c));
  return (uint32_t)lo | ((uint32_t)hi << 16);
}

// x / s for every element of one token's row, s fixed (the per-token quantisers: csrc/quantization/fp8/common.cu:74
// divides by the scale, it does not multiply by a reciprocal).  The IEEE quotient is reproduced bit for bit by
//     rc = RN(1 / s) (once per row);  q0 = RN(x * rc);  r = fma(-s, q0, x);  q = fma(r, rc, q0)
// whenever x carries at most 11 significant bits -- i.e. was rounded to f16 / bf16 before the division, which is the
// case at every call site -- three instructions instead of div.rn's dozen.  Proven by enumeration of all 2^23 scale
// significands against all x significands (oracle/checks/div_by_row_scale.c: 0 mismatches); rows whose maximum is
// not finite (s or x may be inf) take the plain division.
struct RowDivider {
  float s, rc;
  bool fast;
  __device__ __forceinline__ RowDivider(float scale, float row_absmax)
      : s(scale), rc(1.0f / scale), fast(row_absmax <= 3.402823466e38f) {}
  __device__ __forceinline__ float operator()(float x) const {
    if (fast) {
      const float q0 = __fmul_rn(x, rc);
      return __fmaf_rn(__fmaf_rn(-s, q0, x), rc, q0);
    }
    return x / s;
  }
};

// fp8 e4m3 -> float through half, as the reference (quant_utils.cuh:348-359).
__device__ __forceinline__ float e4m3_to_float(uint8_t v) {
  __half_raw h = __nv_cvt_fp8_to_halfraw(v, __NV_E4M3);
  return __half2float(*reinterpret_cast<__half*>(&h));
}

// two e4m3 (low byte first) -> f16x2
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint16_t v) {
  uint32_t r;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"(v));
  return r;
}

// ---- streaming (read-once / write-once) global accesses ---------------------------------------
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ldg_stream_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace vb200
