// Per-token device bodies shared by the stand-alone kernels (rotary_embedding in layer_glue.cu, reshape_and_cache in
// cache_kernels.cu) and by the fused RoPE + KV-cache-write kernel (cache_kernels.cu): one CTA works on one token.
#pragma once
#include "common.cuh"

namespace vb200 {

// rotary_embedding of one token's q and k heads, in place.  T arithmetic with a rounding after every operation
// (x*cos, y*sin and the sum are each rounded to T), as the reference's scalar_t expression does
// (/root/reference/csrc/pos_encoding_kernels.cu:10-98).
template <typename T, bool NEOX>
__device__ __forceinline__ void rotate_token(int token, const int64_t* __restrict__ positions, T* query, T* key,
                                             const T* __restrict__ cos_sin_cache, int rot_dim, int64_t query_stride,
                                             int64_t key_stride, int64_t head_stride, int num_heads,
                                             int num_kv_heads) {
  const int64_t pos = positions[token];
  const T* cache = cos_sin_cache + pos * rot_dim;
  const int embed = rot_dim >> 1;
  const int total = (num_heads + (key ? num_kv_heads : 0)) * embed;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int head = i / embed;
    const int off = i - head * embed;
    T* arr = head < num_heads ? query + token * query_stride + head * head_stride
                              : key + token * key_stride + (head - num_heads) * head_stride;
    const int xi = NEOX ? off : 2 * off;
    const int yi = NEOX ? embed + off : 2 * off + 1;
    const float c = to_float<T>(cache[off]);
    const float s = to_float<T>(cache[embed + off]);
    const float x = to_float<T>(arr[xi]);
    const float y = to_float<T>(arr[yi]);
    const float xc = to_float<T>(from_float<T>(x * c)), ys = to_float<T>(from_float<T>(y * s));
    const float yc = to_float<T>(from_float<T>(y * c)), xs = to_float<T>(from_float<T>(x * s));
    arr[xi] = from_float<T>(xc - ys);
    arr[yi] = from_float<T>(yc + xs);
  }
}

// reshape_and_cache of one token into the paged layout  K: [blk, h, d/x, off, x]   V: [blk, h, d, off]
// (/root/reference/csrc/cache_kernels.cu:210-262).  ksrc / vsrc: this token's [num_heads * head_size] rows.  The
// pointers carry no __restrict__ / const promise: the fused kernel reads K right after rotating it in place.
template <typename scalar_t, bool FP8>
__device__ __forceinline__ void scatter_token_paged(const scalar_t* ksrc, const scalar_t* vsrc, void* key_cache_,
                                                    void* value_cache_, int64_t slot_idx, int num_heads,
                                                    int head_size, int block_size, int x, float k_scale,
                                                    float v_scale, bool vec_ok) {
  using cache_t = typename std::conditional<FP8, uint8_t, scalar_t>::type;
  cache_t* key_cache = reinterpret_cast<cache_t*>(key_cache_);
  cache_t* value_cache = reinterpret_cast<cache_t*>(value_cache_);
  const int64_t block_idx = slot_idx / block_size;
  const int64_t block_offset = slot_idx % block_size;
  const int n = num_heads * head_size;
  cache_t* kdst = key_cache + block_idx * (int64_t)n * block_size;
  cache_t* vdst = value_cache + block_idx * (int64_t)n * block_size + block_offset;

  if (vec_ok) {
    // K: one x-run (x * sizeof(cache_t) == 16 bytes) per thread-iteration.
    const int runs = n / x;
    const int runs_per_head = head_size / x;
    for (int r = threadIdx.x; r < runs; r += blockDim.x) {
      const int h = r / runs_per_head;
      const int xi = r - h * runs_per_head;
      const scalar_t* s = ksrc + (int64_t)r * x;
      cache_t* d = kdst + ((int64_t)(h * runs_per_head + xi) * block_size + block_offset) * x;
      if (!FP8) {
        *reinterpret_cast<uint4*>(d) = *reinterpret_cast<const uint4*>(s);
      } else {
        // x == 16 source elements -> 16 fp8 bytes
        union { uint4 u; uint8_t b[16]; } o;
#pragma unroll
        for (int j = 0; j < 16; ++j) o.b[j] = float_to_e4m3(to_float<scalar_t>(s[j]) / k_scale);
        *reinterpret_cast<uint4*>(d) = o.u;
      }
    }
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int h = i / head_size;
      const int ho = i - h * head_size;
      const int xi = ho / x, xo = ho - xi * x;
      const int64_t t =
          ((int64_t)(h * (head_size / x) + xi) * block_size + block_offset) * x + xo;
      if (!FP8) {
        kdst[t] = *reinterpret_cast<const cache_t*>(&ksrc[i]);
      } else {
        reinterpret_cast<uint8_t*>(kdst)[t] = float_to_e4m3(to_float<scalar_t>(ksrc[i]) / k_scale);
      }
    }
  }
  // V: element (h, d) lands at stride block_size -- inherent to the layout.
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (!FP8) {
      vdst[(int64_t)i * block_size] = *reinterpret_cast<const cache_t*>(&vsrc[i]);
    } else {
      reinterpret_cast<uint8_t*>(vdst)[(int64_t)i * block_size] =
          float_to_e4m3(to_float<scalar_t>(vsrc[i]) / v_scale);
    }
  }
}

}  // namespace vb200
