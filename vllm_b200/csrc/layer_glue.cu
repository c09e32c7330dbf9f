// The small per-token kernels that sit between the contractions of one decode layer: RMSNorm
// (plain / fused residual add / fused with the dynamic per-token fp8 quant that feeds the next
// scaled_mm), rotary embedding, and silu_and_mul fused with the per-token fp8 quant.
//
// Arithmetic follows the reference operation for operation, including every rounding point:
//   rms_norm, fused_add_rms_norm        /root/reference/csrc/layernorm_kernels.cu:16-160
//   rms_norm_dynamic_per_token_quant    /root/reference/csrc/quantization/fused_kernels/
//                                       layernorm_utils.cuh:19-125 + fused_layernorm_dynamic_per_token_quant.cu:9-43
//   rotary_embedding                    /root/reference/csrc/pos_encoding_kernels.cu:10-98
//   silu_and_mul (+ per-token quant)    /root/reference/csrc/activation_kernels.cu:26-63 followed by
//                                       /root/reference/csrc/quantization/fp8/common.cu:28-83
// One CTA per token; the row is read with 16-byte loads and kept in registers between the
// reduction pass and the write pass (one HBM read per element instead of the reference's two or
// three), which is what makes these launch-latency-bound rather than bandwidth-bound at decode
// batch sizes.
#include <cstdlib>

#include "common.cuh"
#include "token_ops.cuh"

namespace vb200 {

constexpr float kFp8MaxG = 448.f;
constexpr int kGlueThreads = 256;
constexpr int kMaxVecPerThread = 8;  // rows up to 256 * 8 * 8 = 16384 16-bit elements stay in registers

template <bool IS_MAX>
__device__ __forceinline__ float block_reduce(float v, float* sm) {
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  __syncthreads();  // sm may still be read from a previous reduction
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = ((threadIdx.x & 31) < (blockDim.x >> 5)) ? sm[threadIdx.x & 31] : 0.f;
  return IS_MAX ? warp_max(r) : warp_sum(r);
}

template <typename T>
union Vec8 {
  uint4 u;
  T e[8];
};

// MODE 0: rms_norm            out = T(T(x * inv) * w)
// MODE 1: fused_add_rms_norm  z = T(x + r); residual = z; input = T(T(z * inv) * w)   (in place)
// MODE 2: rms_norm_dynamic_per_token_quant (optional residual): z = float(x) + float(r) unrounded,
//         residual = T(z); y = T(T(z * inv) * w); s = max(min(amax|y|, ub) / 448, 1/(448*512));
//         out = fp8(clamp(y / s)); scales[token] = s
// KV: 16-byte vectors per thread, sized to the row by the launcher (hidden <= KV * 256 * 8) so that the registers
// holding the row between the passes do not cap the CTAs per SM.
template <typename T, int MODE, int KV>
__global__ void __launch_bounds__(kGlueThreads) rms_norm_kernel(
    void* __restrict__ out_, T* __restrict__ input, T* __restrict__ residual,
    const T* __restrict__ weight, float* __restrict__ scales, const float* __restrict__ scale_ub,
    float eps, int hidden, int64_t input_stride) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm[8];
  const int64_t row = blockIdx.x;
  T* in = input + row * input_stride;
  T* res = residual ? residual + row * hidden : nullptr;
  const int nvec = hidden >> 3;
  float zf[KV][8];  // the (residual-added) row, fp32
  Vec8<T> wv[KV];   // the weight: fetched with the row, not after the reduction that needs no weight
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int i = threadIdx.x + k * kGlueThreads;
    if (i < nvec) wv[k].u = *reinterpret_cast<const uint4*>(weight + i * 8);
  }
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int i = threadIdx.x + k * kGlueThreads;
    if (i < nvec) {
      Vec8<T> a;
      a.u = *reinterpret_cast<const uint4*>(in + i * 8);
      if (res != nullptr) {
        Vec8<T> b, z;
        b.u = *reinterpret_cast<const uint4*>(res + i * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float s = to_float<T>(a.e[j]) + to_float<T>(b.e[j]);
          z.e[j] = from_float<T>(s);
          zf[k][j] = (MODE == 1) ? to_float<T>(z.e[j]) : s;
        }
        *reinterpret_cast<uint4*>(res + i * 8) = z.u;
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) zf[k][j] = to_float<T>(a.e[j]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss = fmaf(zf[k][j], zf[k][j], ss);
    }
  }
  ss = block_reduce<false>(ss, sm);
  const float inv = rsqrtf(ss / (float)hidden + eps);
  float amax = 0.f;
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int i = threadIdx.x + k * kGlueThreads;
    if (i < nvec) {
      Vec8<T> y;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const T n = from_float<T>(zf[k][j] * inv);
        y.e[j] = from_float<T>(to_float<T>(n) * to_float<T>(wv[k].e[j]));
        if (MODE == 2) {
          zf[k][j] = to_float<T>(y.e[j]);
          amax = fmaxf(amax, fabsf(zf[k][j]));
        }
      }
      if (MODE == 0) *reinterpret_cast<uint4*>(reinterpret_cast<T*>(out_) + row * hidden + i * 8) = y.u;
      if (MODE == 1) *reinterpret_cast<uint4*>(in + i * 8) = y.u;
    }
  }
  if (MODE == 2) {
    amax = block_reduce<true>(amax, sm);
    float s = scale_ub ? fminf(amax, *scale_ub) : amax;
    s = fmaxf(s / kFp8MaxG, 1.0f / (kFp8MaxG * 512.0f));
    if (threadIdx.x == 0) scales[row] = s;
    uint8_t* o = reinterpret_cast<uint8_t*>(out_) + row * hidden;
    const RowDivider div(s, amax);
#pragma unroll
    for (int k = 0; k < KV; ++k) {
      const int i = threadIdx.x + k * kGlueThreads;
      if (i < nvec) {
        uint2 q;
        q.x = e4m3x4_clamped(div(zf[k][0]), div(zf[k][1]), div(zf[k][2]), div(zf[k][3]));
        q.y = e4m3x4_clamped(div(zf[k][4]), div(zf[k][5]), div(zf[k][6]), div(zf[k][7]));
        *reinterpret_cast<uint2*>(o + i * 8) = q;
      }
    }
  }
}

// rotary_embedding: one CTA per token, q and k heads rotated in place.  T arithmetic with a
// rounding after every operation (x*cos, y*sin and the sum are each rounded to T), as the
// reference's scalar_t expression does.
template <typename T, bool NEOX>
__global__ void __launch_bounds__(256) rotary_kernel(
    const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
    const T* __restrict__ cos_sin_cache, int rot_dim, int64_t query_stride, int64_t key_stride,
    int64_t head_stride, int num_heads, int num_kv_heads) {
  pdl_wait();
  pdl_trigger();
  rotate_token<T, NEOX>(blockIdx.x, positions, query, key, cos_sin_cache, rot_dim, query_stride, key_stride,
                        head_stride, num_heads, num_kv_heads);
}

// silu_and_mul followed by dynamic per-token fp8 quant in ONE pass: act = T(T(silu(x)) * y) is
// kept in registers, amax reduced, then quantised with the true division of common.cu:74.
// Bit-identical to silu_and_mul -> dynamic_per_token_scaled_fp8_quant run back to back.
// The kernel is INSTRUCTION-bound, not bandwidth-bound (precise expf + two IEEE divisions per element, ~50
// instructions each; round 1 measured 1.3 TB/s with one CTA per row: 256 rows on 148 SMs leave 40 SMs with twice the
// work).  A row is therefore cut into SPLIT column slices, one CTA each, the SPLIT CTAs of a row forming a thread-
// block cluster that exchanges its slice maxima through distributed shared memory (one st.shared::cluster per peer +
// one cluster barrier) -- rows x SPLIT CTAs spread evenly over all SMs.
// Two floats rounded to T (round-to-nearest-even each, as from_float<T>) by ONE pack instruction, returned as floats.
template <typename T>
__device__ __forceinline__ void round_pair(float& a, float& b);
template <>
__device__ __forceinline__ void round_pair<__half>(float& a, float& b) {
  const float2 f = __half22float2(__floats2half2_rn(a, b));
  a = f.x;
  b = f.y;
}
template <>
__device__ __forceinline__ void round_pair<__nv_bfloat16>(float& a, float& b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  const uint32_t u = *reinterpret_cast<const uint32_t*>(&h);
  a = __uint_as_float(u << 16);
  b = __uint_as_float(u & 0xffff0000u);
}

// KV: 16-byte vectors per thread (the slice is at most KV * 256 vectors) -- sized to the slice so that the registers
// holding the activations between the two passes do not cap the CTAs per SM (rows x SPLIT CTAs should be ONE wave).
template <typename T, int SPLIT, int KV>
__global__ void __launch_bounds__(kGlueThreads) silu_mul_quant_kernel(
    uint8_t* __restrict__ out, float* __restrict__ scales, const T* __restrict__ in,
    const float* __restrict__ scale_ub, int d) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm[8];
  __shared__ float peer_max[8];
  // "I have started" (non-blocking): a peer may write my shared memory only once I am running
  if (SPLIT > 1) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  const int64_t row = blockIdx.x / SPLIT;
  const int part = blockIdx.x % SPLIT;     // == %cluster_ctarank (clusters are SPLIT consecutive CTAs along x)
  const int dpart = d / SPLIT;             // host guarantees d % (8 * SPLIT) == 0
  const T* x = in + row * 2 * d + part * dpart;
  const T* y = x + d;
  const int nvec = dpart >> 3;
  float act[KV][8];
  float amax = 0.f;
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int i = threadIdx.x + k * kGlueThreads;
    if (i < nvec) {
      Vec8<T> a, b;
      a.u = ldg_stream_v4(x + i * 8);
      b.u = ldg_stream_v4(y + i * 8);
#pragma unroll
      for (int j = 0; j < 8; j += 2) {
        const float x0 = to_float<T>(a.e[j]), x1 = to_float<T>(a.e[j + 1]);
        float s0 = x0 / (1.0f + expf(-x0)), s1 = x1 / (1.0f + expf(-x1));
        round_pair<T>(s0, s1);                                   // T(silu(x))
        float p0 = s0 * to_float<T>(b.e[j]), p1 = s1 * to_float<T>(b.e[j + 1]);
        round_pair<T>(p0, p1);                                   // T(T(silu(x)) * y)
        act[k][j] = p0;
        act[k][j + 1] = p1;
        amax = fmaxf(amax, fmaxf(fabsf(p0), fabsf(p1)));
      }
    }
  }
  amax = block_reduce<true>(amax, sm);
  if (SPLIT > 1) {
    // every CTA of the cluster posts its slice maximum into slot `part` of every CTA's peer_max[]
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");   // all peers have started
    if (threadIdx.x < SPLIT) {
      uint32_t local = static_cast<uint32_t>(__cvta_generic_to_shared(&peer_max[part])), remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(threadIdx.x));
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(amax) : "memory");
    }
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
#pragma unroll
    for (int r = 0; r < SPLIT; ++r) amax = fmaxf(amax, peer_max[r]);
  }
  float s = scale_ub ? fminf(amax, *scale_ub) : amax;
  s = fmaxf(s / kFp8MaxG, 1.0f / (kFp8MaxG * 512.0f));
  if (threadIdx.x == 0 && part == 0) scales[row] = s;
  uint8_t* o = out + row * d + part * dpart;
  const RowDivider div(s, amax);
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int i = threadIdx.x + k * kGlueThreads;
    if (i < nvec) {
      uint2 q;
      q.x = e4m3x4_clamped(div(act[k][0]), div(act[k][1]), div(act[k][2]), div(act[k][3]));
      q.y = e4m3x4_clamped(div(act[k][4]), div(act[k][5]), div(act[k][6]), div(act[k][7]));
      *reinterpret_cast<uint2*>(o + i * 8) = q;
    }
  }
}

static int check_row(const char* op, int hidden, int dtype, const void* p0, const void* p1) {
  VB_CHECK(dtype == VB200_F16 || dtype == VB200_BF16, "%s: dtype must be f16 or bf16", op);
  VB_CHECK(hidden > 0 && hidden % 8 == 0 && hidden <= kGlueThreads * kMaxVecPerThread * 8,
           "%s: hidden size %d must be a multiple of 8 and <= %d", op, hidden,
           kGlueThreads * kMaxVecPerThread * 8);
  VB_CHECK(((reinterpret_cast<uintptr_t>(p0) | reinterpret_cast<uintptr_t>(p1)) & 15) == 0,
           "%s: tensors must be 16-byte aligned", op);
  return 0;
}

}  // namespace vb200

using namespace vb200;

#define DISPATCH_16(dt, ...)                                     \
  if ((dt) == VB200_F16) {                                       \
    using scalar_t = __half;                                     \
    __VA_ARGS__;                                                 \
  } else {                                                       \
    using scalar_t = __nv_bfloat16;                              \
    __VA_ARGS__;                                                 \
  }

// vectors per thread for a row of `hidden` 16-bit elements: 1, 2, 4 or 8
#define VB_NORM_KV(hidden, ...)                                    \
  do {                                                             \
    const int vpt_ = ((hidden) / 8 + kGlueThreads - 1) / kGlueThreads; \
    if (vpt_ <= 1) {                                               \
      constexpr int KV = 1;                                        \
      __VA_ARGS__;                                                 \
    } else if (vpt_ <= 2) {                                        \
      constexpr int KV = 2;                                        \
      __VA_ARGS__;                                                 \
    } else if (vpt_ <= 4) {                                        \
      constexpr int KV = 4;                                        \
      __VA_ARGS__;                                                 \
    } else {                                                       \
      constexpr int KV = 8;                                        \
      __VA_ARGS__;                                                 \
    }                                                              \
  } while (0)

extern "C" int vb200_rms_norm(void* out, const void* input, const void* weight, float epsilon,
                              int num_tokens, int hidden_size, int64_t input_stride, int dtype,
                              vb200_stream_t stream) {
  if (num_tokens <= 0) return 0;
  if (check_row("rms_norm", hidden_size, dtype, out, input)) return 1;
  VB_CHECK(input_stride % 8 == 0, "rms_norm: input stride must be a multiple of 8");
  VB_NORM_KV(hidden_size, DISPATCH_16(dtype, VB_CUDA_CHECK(launch_k(rms_norm_kernel<scalar_t, 0, KV>, dim3(num_tokens), dim3(kGlueThreads), 0, (cudaStream_t)stream, out, (scalar_t*)input, nullptr, (const scalar_t*)weight, nullptr, nullptr,
                         epsilon, hidden_size, input_stride))));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_fused_add_rms_norm(void* input, void* residual, const void* weight,
                                        float epsilon, int num_tokens, int hidden_size,
                                        int64_t input_stride, int dtype, vb200_stream_t stream) {
  if (num_tokens <= 0) return 0;
  if (check_row("fused_add_rms_norm", hidden_size, dtype, input, residual)) return 1;
  VB_CHECK(input_stride % 8 == 0, "fused_add_rms_norm: input stride must be a multiple of 8");
  VB_NORM_KV(hidden_size, DISPATCH_16(dtype, VB_CUDA_CHECK(launch_k(rms_norm_kernel<scalar_t, 1, KV>, dim3(num_tokens), dim3(kGlueThreads), 0, (cudaStream_t)stream, nullptr, (scalar_t*)input, (scalar_t*)residual, (const scalar_t*)weight,
                         nullptr, nullptr, epsilon, hidden_size, input_stride))));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_rms_norm_dynamic_per_token_quant(void* out, const void* input,
                                                      const void* weight, float* scales,
                                                      float epsilon, const float* scale_ub,
                                                      void* residual, int num_tokens,
                                                      int hidden_size, int dtype,
                                                      vb200_stream_t stream) {
  if (num_tokens <= 0) return 0;
  if (check_row("rms_norm_dynamic_per_token_quant", hidden_size, dtype, input, residual)) return 1;
  VB_CHECK((reinterpret_cast<uintptr_t>(out) & 7) == 0, "rms_norm_dynamic_per_token_quant: out must be 8-byte aligned");
  VB_NORM_KV(hidden_size, DISPATCH_16(dtype, VB_CUDA_CHECK(launch_k(rms_norm_kernel<scalar_t, 2, KV>, dim3(num_tokens), dim3(kGlueThreads), 0, (cudaStream_t)stream, out, (scalar_t*)input, (scalar_t*)residual, (const scalar_t*)weight, scales,
                         scale_ub, epsilon, hidden_size, (int64_t)hidden_size))));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_rotary_embedding(const int64_t* positions, void* query, void* key,
                                      int num_tokens, int num_heads, int num_kv_heads,
                                      int head_size, int rot_dim, const void* cos_sin_cache,
                                      int is_neox, int64_t query_stride, int64_t key_stride,
                                      int64_t head_stride, int dtype, vb200_stream_t stream) {
  if (num_tokens <= 0) return 0;
  VB_CHECK(dtype == VB200_F16 || dtype == VB200_BF16, "rotary_embedding: dtype must be f16 or bf16");
  VB_CHECK(rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size,
           "rotary_embedding: rot_dim %d must be even and <= head_size %d", rot_dim, head_size);
  const int work = (num_heads + (key ? num_kv_heads : 0)) * rot_dim / 2;
  const int threads = work >= 256 ? 256 : (work >= 128 ? 128 : 64);
  DISPATCH_16(dtype, {
    if (is_neox)
      VB_CUDA_CHECK(launch_k(rotary_kernel<scalar_t, true>, dim3(num_tokens), dim3(threads), 0, (cudaStream_t)stream, positions, (scalar_t*)query, (scalar_t*)key, (const scalar_t*)cos_sin_cache, rot_dim,
          query_stride, key_stride, head_stride, num_heads, num_kv_heads));
    else
      VB_CUDA_CHECK(launch_k(rotary_kernel<scalar_t, false>, dim3(num_tokens), dim3(threads), 0, (cudaStream_t)stream, positions, (scalar_t*)query, (scalar_t*)key, (const scalar_t*)cos_sin_cache, rot_dim,
          query_stride, key_stride, head_stride, num_heads, num_kv_heads));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_silu_and_mul_dynamic_per_token_quant(void* out, float* scales,
                                                          const void* input,
                                                          const float* scale_ub, int num_tokens,
                                                          int d, int dtype,
                                                          vb200_stream_t stream) {
  if (num_tokens <= 0) return 0;
  if (check_row("silu_and_mul_dynamic_per_token_quant", d, dtype, input, input)) return 1;
  VB_CHECK((reinterpret_cast<uintptr_t>(out) & 7) == 0, "silu_and_mul_dynamic_per_token_quant: out must be 8-byte aligned");
  // column slices per row: enough CTAs for about two waves of light CTAs (measured at 256 x 14336: 1 / 2 / 4 / 8 slices
  // 13.2 / 11.4 / 11.0 / 9.5 us -- a second wave's loads overlap the first wave's arithmetic); whole 16-byte vectors
  int split = 1;
  const int want = (14 * num_sms() + num_tokens - 1) / num_tokens;
  for (int s = 2; s <= 8; s *= 2)
    if (s <= want && d % (8 * s) == 0 && d / s >= 1024) split = s;
  static const char* force = getenv("VB200_SILU_SPLIT");   // experiments (benchmarks/bench_layer_glue.py)
  if (force != nullptr) {
    const int f = atoi(force);
    if ((f == 1 || f == 2 || f == 4 || f == 8) && d % (8 * f) == 0) split = f;
  }
  const dim3 grid((unsigned)num_tokens * split), block(kGlueThreads);
  const int vec_per_thread = (d / split / 8 + kGlueThreads - 1) / kGlueThreads;
#define VB_SILU_Q(S, KV)                                                                                           \
  do {                                                                                                             \
    if (S > 1) {                                                                                                   \
      DISPATCH_16(dtype, VB_CUDA_CHECK(launch_k_cluster(silu_mul_quant_kernel<scalar_t, S, KV>, grid, block, 0,    \
                                                        (cudaStream_t)stream, S, (uint8_t*)out, scales,            \
                                                        (const scalar_t*)input, scale_ub, d)));                    \
    } else {                                                                                                       \
      DISPATCH_16(dtype, VB_CUDA_CHECK(launch_k(silu_mul_quant_kernel<scalar_t, S, KV>, grid, block, 0,            \
                                                (cudaStream_t)stream, (uint8_t*)out, scales,                       \
                                                (const scalar_t*)input, scale_ub, d)));                            \
    }                                                                                                              \
  } while (0)
#define VB_SILU_Q_KV(S)                                   \
  do {                                                    \
    if (vec_per_thread <= 1) VB_SILU_Q(S, 1);             \
    else if (vec_per_thread <= 2) VB_SILU_Q(S, 2);        \
    else if (vec_per_thread <= 4) VB_SILU_Q(S, 4);        \
    else VB_SILU_Q(S, 8);                                 \
  } while (0)
  if (split == 8) VB_SILU_Q_KV(8);
  else if (split == 4) VB_SILU_Q_KV(4);
  else if (split == 2) VB_SILU_Q_KV(2);
  else VB_SILU_Q_KV(1);
#undef VB_SILU_Q_KV
#undef VB_SILU_Q
  VB_LAUNCH_CHECK();
  return 0;
}
