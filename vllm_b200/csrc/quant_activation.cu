// FP8 activation quantisation (static / dynamic per-tensor / dynamic per-token) and silu_and_mul
// for sm_100a.  Arithmetic follows /root/reference/csrc/quantization/fp8/common.cu:14-83,
// common.cuh:38-99 and /root/reference/csrc/activation_kernels.cu:26-63 operation for operation
// (reciprocal-multiply for the static path, true division for the per-token path, silu rounded to
// the tensor dtype before the multiply) so results are bit-identical; the kernels are HBM-bound
// streaming kernels: 16-byte loads, 8/16-byte stores, grids sized to the SM count.
#include "common.cuh"

namespace vb200 {

constexpr float kFp8Max = 448.f;

__device__ __forceinline__ uint8_t quant_clamped(float x) {
  // fmaxf/fminf clamp then RNE conversion == static_cast<c10::Float8_e4m3fn>(clamped)
  const float r = fmaxf(-kFp8Max, fminf(x, kFp8Max));
  return float_to_e4m3(r);
}

// INVERTED: x * s (s = 1 / scale, the static path).  Otherwise x / s, through RowDivider when T is a 16-bit type
// (`row_absmax` finite); fp32 inputs carry 24 significant bits and keep the plain division.
template <typename T, bool INVERTED>
__device__ __forceinline__ void quant_row(uint8_t* __restrict__ out, const T* __restrict__ in,
                                          int64_t n, float s, int tid, int nthreads, float row_absmax = 0.f) {
  // 16-byte vector path when both pointers are aligned
  constexpr int VE = 16 / sizeof(T);  // elements per 16-byte load (8 for 16-bit, 4 for fp32)
  const bool vec = ((reinterpret_cast<uintptr_t>(in) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(out) & (VE - 1)) == 0);
  const RowDivider div(s, (sizeof(T) == 2 && !INVERTED) ? row_absmax : __int_as_float(0x7f800000));
  int64_t done = 0;
  if (vec) {
    const int64_t nvec = n / VE;
    for (int64_t i = tid; i < nvec; i += nthreads) {
      union { uint4 u; T e[VE]; } v;
      v.u = *reinterpret_cast<const uint4*>(in + i * VE);
      float q[VE];
#pragma unroll
      for (int j = 0; j < VE; ++j) {
        const float x = to_float<T>(v.e[j]);
        q[j] = INVERTED ? x * s : div(x);
      }
      if (VE == 8) {
        uint2 o;
        o.x = e4m3x4_clamped(q[0], q[1], q[2], q[3]);
        o.y = e4m3x4_clamped(q[VE - 4], q[VE - 3], q[VE - 2], q[VE - 1]);
        *reinterpret_cast<uint2*>(out + i * VE) = o;
      } else {
        *reinterpret_cast<uint32_t*>(out + i * VE) = e4m3x4_clamped(q[0], q[1], q[2], q[3]);
      }
    }
    done = nvec * VE;
  }
  for (int64_t i = done + tid; i < n; i += nthreads) {
    const float x = to_float<T>(in[i]);
    out[i] = quant_clamped(INVERTED ? x * s : div(x));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) static_quant_kernel(uint8_t* __restrict__ out,
                                                           const T* __restrict__ in,
                                                           const float* __restrict__ scale,
                                                           int64_t n) {
  pdl_wait();
  pdl_trigger();
  const float inv = 1.0f / (*scale);  // common.cu:23
  quant_row<T, true>(out, in, n, inv, blockIdx.x * blockDim.x + threadIdx.x,
                     gridDim.x * blockDim.x);
}

template <typename T>
__device__ __forceinline__ float thread_absmax(const T* __restrict__ in, int64_t n, int tid,
                                               int nthreads) {
  constexpr int VE = 16 / sizeof(T);
  float m = 0.f;
  int64_t done = 0;
  if ((reinterpret_cast<uintptr_t>(in) & 15) == 0) {
    const int64_t nvec = n / VE;
    for (int64_t i = tid; i < nvec; i += nthreads) {
      union { uint4 u; T e[VE]; } v;
      v.u = *reinterpret_cast<const uint4*>(in + i * VE);
#pragma unroll
      for (int j = 0; j < VE; ++j) m = fmaxf(m, fabsf(to_float<T>(v.e[j])));
    }
    done = nvec * VE;
  }
  for (int64_t i = done + tid; i < n; i += nthreads) m = fmaxf(m, fabsf(to_float<T>(in[i])));
  return m;
}

__device__ __forceinline__ float block_max_256(float v, float* sm) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  // every warp reduces the per-warp maxima itself, so all threads return the block maximum
  float r = ((threadIdx.x & 31) < (blockDim.x >> 5)) ? sm[threadIdx.x & 31] : 0.f;
  return warp_max(r);
}

// per-tensor absmax / 448 -> atomic max into *scale (pre-zeroed by the caller), common.cuh:64-99
template <typename T>
__global__ void __launch_bounds__(256) absmax_scale_kernel(float* __restrict__ scale,
                                                           const T* __restrict__ in, int64_t n) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm[8];
  float m = thread_absmax<T>(in, n, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
  m = block_max_256(m, sm);
  if (threadIdx.x == 0) {
    // non-negative floats order like their int bit patterns (common.cuh:30-36)
    atomicMax(reinterpret_cast<int*>(scale), __float_as_int(m / kFp8Max));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) per_token_quant_kernel(uint8_t* __restrict__ out,
                                                              float* __restrict__ scales,
                                                              const T* __restrict__ in,
                                                              const float* __restrict__ scale_ub,
                                                              int hidden) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm[8];
  const int64_t row = blockIdx.x;
  const T* rin = in + row * hidden;
  uint8_t* rout = out + row * hidden;
  float m = thread_absmax<T>(rin, hidden, threadIdx.x, blockDim.x);
  m = block_max_256(m, sm);
  float s = scale_ub ? fminf(m, *scale_ub) : m;
  s = fmaxf(s / kFp8Max, 1.0f / (kFp8Max * 512.0f));  // common.cu:66-68, utils.cuh:49-51
  if (threadIdx.x == 0) scales[row] = s;
  quant_row<T, false>(rout, rin, hidden, s, threadIdx.x, blockDim.x, m);  // the quotient of a true division (:74)
}

// silu_and_mul: out = T(silu_f32(x)) *_T y
template <typename T>
__device__ __forceinline__ T silu_mul(T x, T y) {
  const float xf = to_float<T>(x);
  const T a = from_float<T>(xf / (1.0f + expf(-xf)));
  return from_float<T>(to_float<T>(a) * to_float<T>(y));  // T * T: exact product, one rounding
}

template <typename T>
__global__ void __launch_bounds__(256) silu_and_mul_kernel(T* __restrict__ out,
                                                           const T* __restrict__ in, int d,
                                                           bool vec) {
  pdl_wait();
  pdl_trigger();
  const int64_t row = blockIdx.x;
  const T* x = in + row * 2 * d;
  const T* y = x + d;
  T* o = out + row * d;
  constexpr int VE = 16 / sizeof(T);
  if (vec) {
    for (int i = threadIdx.x; i < d / VE; i += blockDim.x) {
      union { uint4 u; T e[VE]; } a, b, r;
      a.u = *reinterpret_cast<const uint4*>(x + i * VE);
      b.u = *reinterpret_cast<const uint4*>(y + i * VE);
#pragma unroll
      for (int j = 0; j < VE; ++j) r.e[j] = silu_mul<T>(a.e[j], b.e[j]);
      *reinterpret_cast<uint4*>(o + i * VE) = r.u;
    }
  } else {
    for (int i = threadIdx.x; i < d; i += blockDim.x) o[i] = silu_mul<T>(x[i], y[i]);
  }
}

// silu_and_mul that also leaves max|out| / 448 in *scale (atomic max; pre-zeroed by the caller): the first pass of the
// dynamic per-tensor quantisation that follows it in the fp8 MoE pipeline (fused_moe.py:1506-1527: activation, then
// moe_kernel_quantize_input), taken from the values as rounded to T -- what absmax_scale_kernel would read back.
template <typename T>
__global__ void __launch_bounds__(256) silu_and_mul_absmax_kernel(T* __restrict__ out, float* __restrict__ scale,
                                                                  const T* __restrict__ in, int d) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm[8];
  const int64_t row = blockIdx.x;
  const T* x = in + row * 2 * d;
  const T* y = x + d;
  T* o = out + row * d;
  constexpr int VE = 16 / sizeof(T);
  float m = 0.f;
  for (int i = threadIdx.x; i < d / VE; i += blockDim.x) {
    union { uint4 u; T e[VE]; } a, b, r;
    a.u = *reinterpret_cast<const uint4*>(x + i * VE);
    b.u = *reinterpret_cast<const uint4*>(y + i * VE);
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      r.e[j] = silu_mul<T>(a.e[j], b.e[j]);
      m = fmaxf(m, fabsf(to_float<T>(r.e[j])));
    }
    *reinterpret_cast<uint4*>(o + i * VE) = r.u;
  }
  m = block_max_256(m, sm);
  if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(scale), __float_as_int(m / kFp8Max));
}

static inline int stream_grid(int64_t n, int per_thread) {
  int64_t b = (n / per_thread + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace vb200

using namespace vb200;

#define DISPATCH_IN(dt, ...)                                                \
  switch (dt) {                                                             \
    case VB200_F32: { using scalar_t = float; __VA_ARGS__; break; }         \
    case VB200_F16: { using scalar_t = __half; __VA_ARGS__; break; }        \
    case VB200_BF16: { using scalar_t = __nv_bfloat16; __VA_ARGS__; break; } \
    default: VB_CHECK(false, "unsupported input dtype %d", dt);             \
  }

extern "C" int vb200_static_scaled_fp8_quant(void* out, const void* input, const float* scale,
                                             int64_t num_tokens, int hidden_size, int in_dtype,
                                             vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int64_t n = num_tokens * hidden_size;
  if (n <= 0) return 0;
  DISPATCH_IN(in_dtype, {
    VB_CUDA_CHECK(launch_k(static_quant_kernel<scalar_t>, dim3(stream_grid(n, 8)), dim3(256), 0, stream, (uint8_t*)out, (const scalar_t*)input, scale, n));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_dynamic_scaled_fp8_quant(void* out, const void* input, float* scale,
                                              int64_t num_tokens, int hidden_size, int in_dtype,
                                              vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int64_t n = num_tokens * hidden_size;
  if (n <= 0) return 0;
  DISPATCH_IN(in_dtype, {
    VB_CUDA_CHECK(launch_k(absmax_scale_kernel<scalar_t>, dim3(stream_grid(n, 8)), dim3(256), 0, stream, scale, (const scalar_t*)input, n));
    VB_CUDA_CHECK(launch_k(static_quant_kernel<scalar_t>, dim3(stream_grid(n, 8)), dim3(256), 0, stream, (uint8_t*)out, (const scalar_t*)input, scale, n));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_dynamic_per_token_scaled_fp8_quant(void* out, const void* input,
                                                        float* scales, const float* scale_ub,
                                                        int64_t num_tokens, int hidden_size,
                                                        int in_dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_tokens <= 0 || hidden_size <= 0) return 0;
  DISPATCH_IN(in_dtype, {
    VB_CUDA_CHECK(launch_k(per_token_quant_kernel<scalar_t>, dim3((unsigned)num_tokens), dim3(256), 0, stream, (uint8_t*)out, scales, (const scalar_t*)input, scale_ub, hidden_size));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d,
                                  int dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_tokens <= 0 || d <= 0) return 0;
  const int esz = dtype == VB200_F32 ? 4 : 2;
  const int ve = 16 / esz;
  const bool vec = (d % ve == 0) && ((reinterpret_cast<uintptr_t>(input) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
  const int threads = d / ve >= 256 ? 256 : (d / ve >= 128 ? 128 : 64);
  DISPATCH_IN(dtype, {
    VB_CUDA_CHECK(launch_k(silu_and_mul_kernel<scalar_t>, dim3((unsigned)num_tokens), dim3(threads), 0, stream, (scalar_t*)out, (const scalar_t*)input, d, vec));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

// First pass of dynamic_scaled_fp8_quant on its own: *scale = max(*scale, max|input| / 448) (scale pre-zeroed by the
// caller).  The fp8 MoE pipeline quantises while it gathers the rows into expert order, so it needs only the scale.
extern "C" int vb200_absmax_scale(float* scale, const void* input, int64_t numel, int in_dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (numel <= 0) return 0;
  DISPATCH_IN(in_dtype, {
    VB_CUDA_CHECK(launch_k(absmax_scale_kernel<scalar_t>, dim3(stream_grid(numel, 8)), dim3(256), 0, stream, scale, (const scalar_t*)input, numel));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_silu_and_mul_absmax(void* out, float* scale, const void* input, int64_t num_tokens, int d, int dtype,
                                         vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_tokens <= 0 || d <= 0) return 0;
  VB_CHECK(dtype == VB200_F16 || dtype == VB200_BF16, "silu_and_mul_absmax: dtype must be f16 or bf16");
  VB_CHECK(d % 8 == 0 && (reinterpret_cast<uintptr_t>(input) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
           "silu_and_mul_absmax: d (%d) must be a multiple of 8 and the tensors 16-byte aligned", d);
  DISPATCH_IN(dtype, {
    VB_CUDA_CHECK(launch_k(silu_and_mul_absmax_kernel<scalar_t>, dim3((unsigned)num_tokens), dim3(256), 0, stream, (scalar_t*)out, scale, (const scalar_t*)input, d));
  });
  VB_LAUNCH_CHECK();
  return 0;
}
