// KV-cache scatter kernels for sm_100a: reshape_and_cache (paged layout), reshape_and_cache_flash
// (flash layout through runtime strides) and the convert_fp8 test helper.
// Semantics follow /root/reference/csrc/cache_kernels.cu:210-306,369-450,526-571; the kernels are
// a fresh design: one CTA per token, 16-byte vector loads of the source row, 16-byte stores of the
// x-element K runs / flash rows, bit-exact copy (AUTO) or reference-exact fp8 quantisation.
#include "common.cuh"
#include "token_ops.cuh"

namespace vb200 {

template <typename T>
struct Vec16 {  // 16 bytes of T
  static constexpr int N = 16 / sizeof(T);
  T v[N];
};

// ---------------------------------------------------------------------------------------------
// Paged layout.  K: [blk, h, d/x, off, x]   V: [blk, h, d, off]
// ---------------------------------------------------------------------------------------------
template <typename scalar_t, bool FP8>
__global__ void __launch_bounds__(256) reshape_and_cache_kernel(
    const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
    void* __restrict__ key_cache_, void* __restrict__ value_cache_,
    const int64_t* __restrict__ slot_mapping, int64_t key_stride, int64_t value_stride,
    int num_heads, int head_size, int block_size, int x, const float* __restrict__ k_scale_p,
    const float* __restrict__ v_scale_p, bool vec_ok) {
  pdl_wait();
  pdl_trigger();
  const int64_t token_idx = blockIdx.x;
  const int64_t slot_idx = slot_mapping[token_idx];
  if (slot_idx < 0) return;  // padding token
  float k_scale = 1.f, v_scale = 1.f;
  if (FP8) {
    k_scale = *k_scale_p;
    v_scale = *v_scale_p;
  }
  scatter_token_paged<scalar_t, FP8>(key + token_idx * key_stride, value + token_idx * value_stride, key_cache_,
                                     value_cache_, slot_idx, num_heads, head_size, block_size, x, k_scale, v_scale,
                                     vec_ok);
}

// RoPE + KV-cache write in ONE launch: rotary_embedding(positions, q, k) in place, then
// reshape_and_cache(k, v) of the same token.  Reference: /root/reference/csrc/pos_encoding_kernels.cu:77-101 followed
// by /root/reference/csrc/cache_kernels.cu:210-262, two launches; the arithmetic (and so every bit of q, k and both
// caches) is that of the two kernels run back to back -- the CTA that rotated a token's K re-reads it (bar.sync makes
// its own global writes visible to all its threads) and scatters it.
template <typename scalar_t, bool NEOX, bool FP8>
__global__ void __launch_bounds__(256) rope_reshape_and_cache_kernel(
    const int64_t* __restrict__ positions, scalar_t* query, scalar_t* key, const scalar_t* value,
    const scalar_t* __restrict__ cos_sin_cache, void* __restrict__ key_cache_, void* __restrict__ value_cache_,
    const int64_t* __restrict__ slot_mapping, int rot_dim, int64_t query_stride, int64_t key_stride,
    int64_t value_stride, int num_heads, int num_kv_heads, int head_size, int block_size, int x,
    const float* __restrict__ k_scale_p, const float* __restrict__ v_scale_p, bool vec_ok) {
  pdl_wait();
  pdl_trigger();
  const int token = blockIdx.x;
  rotate_token<scalar_t, NEOX>(token, positions, query, key, cos_sin_cache, rot_dim, query_stride, key_stride,
                               (int64_t)head_size, num_heads, num_kv_heads);
  const int64_t slot_idx = slot_mapping[token];
  if (slot_idx < 0) return;  // padding token: rotated, not cached (what the two kernels would do)
  __syncthreads();
  float k_scale = 1.f, v_scale = 1.f;
  if (FP8) {
    k_scale = *k_scale_p;
    v_scale = *v_scale_p;
  }
  scatter_token_paged<scalar_t, FP8>(key + token * key_stride, value + token * value_stride, key_cache_, value_cache_,
                                     slot_idx, num_kv_heads, head_size, block_size, x, k_scale, v_scale, vec_ok);
}

// The same operation for the shapes a decode step has (full rotary: rot_dim == head_size, 16-byte-aligned rows,
// K runs of x * sizeof(cache_t) == 16 bytes): every thread owns ONE 16-byte piece of the token -- 8 (x, y) pairs of a
// q / k head, or 8 elements of a v head -- so all of a token's loads are in flight at once (the per-element kernel
// above does ten dependent rounds of 2-byte loads and stores per thread), and the rotated K goes to the cache from
// registers instead of being re-read.  Same arithmetic, element for element: every product and sum rounded to T, the
// cache quantised from the rounded K.
template <typename scalar_t, bool FP8>
__device__ __forceinline__ void store_cache8(void* dst, const scalar_t (&v)[8], float scale) {
  if (!FP8) {
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
  } else {
    union { uint2 u; uint8_t b[8]; } o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.b[j] = float_to_e4m3(to_float<scalar_t>(v[j]) / scale);
    *reinterpret_cast<uint2*>(dst) = o.u;
  }
}

template <typename scalar_t, bool NEOX, bool FP8>
__global__ void __launch_bounds__(512) rope_reshape_and_cache_vec_kernel(
    const int64_t* __restrict__ positions, scalar_t* query, scalar_t* key, const scalar_t* __restrict__ value,
    const scalar_t* __restrict__ cos_sin_cache, void* __restrict__ key_cache_, void* __restrict__ value_cache_,
    const int64_t* __restrict__ slot_mapping, int64_t query_stride, int64_t key_stride, int64_t value_stride,
    int num_heads, int num_kv_heads, int head_size, int block_size, int x, const float* __restrict__ k_scale_p,
    const float* __restrict__ v_scale_p) {
  pdl_wait();
  pdl_trigger();
  using cache_t = typename std::conditional<FP8, uint8_t, scalar_t>::type;
  const int64_t token = blockIdx.x;
  const int64_t pos = positions[token];
  const int64_t slot_idx = slot_mapping[token];
  const bool cached = slot_idx >= 0;           // padding token: rotated, not cached
  float k_scale = 1.f, v_scale = 1.f;
  if (FP8) {
    k_scale = *k_scale_p;
    v_scale = *v_scale_p;
  }
  const int embed = head_size >> 1;
  const int rot_per_head = NEOX ? embed >> 3 : head_size >> 3;
  const int nrot = (num_heads + num_kv_heads) * rot_per_head;
  const int v_per_head = head_size >> 3;
  const int nv = cached ? num_kv_heads * v_per_head : 0;
  const scalar_t* cache = cos_sin_cache + pos * head_size;
  const int64_t block_idx = cached ? slot_idx / block_size : 0;
  const int block_offset = cached ? (int)(slot_idx % block_size) : 0;
  const int64_t n = (int64_t)num_kv_heads * head_size;
  cache_t* kdst = reinterpret_cast<cache_t*>(key_cache_) + block_idx * n * block_size;
  cache_t* vdst = reinterpret_cast<cache_t*>(value_cache_) + block_idx * n * block_size + block_offset;
  const int runs_per_head = head_size / x;

  for (int t = threadIdx.x; t < nrot + nv; t += blockDim.x) {
    if (t < nrot) {
      const int head = t / rot_per_head, r = t - head * rot_per_head;
      const bool is_k = head >= num_heads;
      const int hk = head - num_heads;
      scalar_t* arr = is_k ? key + token * key_stride + (int64_t)hk * head_size
                           : query + token * query_stride + (int64_t)head * head_size;
      Vec16<scalar_t> a, b, c, sn;
      if (NEOX) {   // x = arr[off .. off+8), y = arr[embed + off ..), cos = cache[off ..), sin = cache[embed + off ..)
        const int off = r * 8;
        *reinterpret_cast<uint4*>(&a) = *reinterpret_cast<const uint4*>(arr + off);
        *reinterpret_cast<uint4*>(&b) = *reinterpret_cast<const uint4*>(arr + embed + off);
        *reinterpret_cast<uint4*>(&c) = *reinterpret_cast<const uint4*>(cache + off);
        *reinterpret_cast<uint4*>(&sn) = *reinterpret_cast<const uint4*>(cache + embed + off);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float cf = to_float<scalar_t>(c.v[j]), sf = to_float<scalar_t>(sn.v[j]);
          const float xf = to_float<scalar_t>(a.v[j]), yf = to_float<scalar_t>(b.v[j]);
          const float xc = to_float<scalar_t>(from_float<scalar_t>(xf * cf));
          const float ys = to_float<scalar_t>(from_float<scalar_t>(yf * sf));
          const float yc = to_float<scalar_t>(from_float<scalar_t>(yf * cf));
          const float xs = to_float<scalar_t>(from_float<scalar_t>(xf * sf));
          a.v[j] = from_float<scalar_t>(xc - ys);
          b.v[j] = from_float<scalar_t>(yc + xs);
        }
        *reinterpret_cast<uint4*>(arr + off) = *reinterpret_cast<const uint4*>(&a);
        *reinterpret_cast<uint4*>(arr + embed + off) = *reinterpret_cast<const uint4*>(&b);
        if (is_k && cached) {
          const int d0 = off, d1 = embed + off;
          store_cache8<scalar_t, FP8>(
              kdst + ((int64_t)(hk * runs_per_head + d0 / x) * block_size + block_offset) * x + d0 % x, a.v, k_scale);
          store_cache8<scalar_t, FP8>(
              kdst + ((int64_t)(hk * runs_per_head + d1 / x) * block_size + block_offset) * x + d1 % x, b.v, k_scale);
        }
      } else {      // pairs (arr[2i], arr[2i+1]) for i in [r*4, r*4+4): cos = cache[i], sin = cache[embed + i]
        const int off = r * 4;
        *reinterpret_cast<uint4*>(&a) = *reinterpret_cast<const uint4*>(arr + 2 * off);
        *reinterpret_cast<uint2*>(&c) = *reinterpret_cast<const uint2*>(cache + off);
        *reinterpret_cast<uint2*>(&sn) = *reinterpret_cast<const uint2*>(cache + embed + off);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float cf = to_float<scalar_t>(c.v[j]), sf = to_float<scalar_t>(sn.v[j]);
          const float xf = to_float<scalar_t>(a.v[2 * j]), yf = to_float<scalar_t>(a.v[2 * j + 1]);
          const float xc = to_float<scalar_t>(from_float<scalar_t>(xf * cf));
          const float ys = to_float<scalar_t>(from_float<scalar_t>(yf * sf));
          const float yc = to_float<scalar_t>(from_float<scalar_t>(yf * cf));
          const float xs = to_float<scalar_t>(from_float<scalar_t>(xf * sf));
          a.v[2 * j] = from_float<scalar_t>(xc - ys);
          a.v[2 * j + 1] = from_float<scalar_t>(yc + xs);
        }
        *reinterpret_cast<uint4*>(arr + 2 * off) = *reinterpret_cast<const uint4*>(&a);
        if (is_k && cached) {
          const int d0 = 2 * off;
          store_cache8<scalar_t, FP8>(
              kdst + ((int64_t)(hk * runs_per_head + d0 / x) * block_size + block_offset) * x + d0 % x, a.v, k_scale);
        }
      }
    } else {        // 8 elements of a v head: element (h, d) lands at stride block_size -- inherent to the layout
      const int i = (t - nrot) * 8;   // h * head_size + d
      Vec16<scalar_t> a;
      *reinterpret_cast<uint4*>(&a) = *reinterpret_cast<const uint4*>(value + token * value_stride + i);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (!FP8) {
          vdst[(int64_t)(i + j) * block_size] = *reinterpret_cast<const cache_t*>(&a.v[j]);
        } else {
          reinterpret_cast<uint8_t*>(vdst)[(int64_t)(i + j) * block_size] =
              float_to_e4m3(to_float<scalar_t>(a.v[j]) / v_scale);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Flash layout through runtime strides: cache[blk*block_stride + off*page_stride + h*head_stride + d]
// ---------------------------------------------------------------------------------------------
template <typename scalar_t, bool FP8>
__global__ void __launch_bounds__(256) reshape_and_cache_flash_kernel(
    const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
    void* __restrict__ key_cache_, void* __restrict__ value_cache_,
    const int64_t* __restrict__ slot_mapping, int64_t block_stride, int64_t page_stride,
    int64_t head_stride, int64_t key_stride, int64_t value_stride, int num_heads, int head_size,
    int block_size, const float* __restrict__ k_scale_p, const float* __restrict__ v_scale_p,
    bool vec_ok) {
  pdl_wait();
  pdl_trigger();
  using cache_t = typename std::conditional<FP8, uint8_t, scalar_t>::type;
  cache_t* key_cache = reinterpret_cast<cache_t*>(key_cache_);
  cache_t* value_cache = reinterpret_cast<cache_t*>(value_cache_);
  const int64_t token_idx = blockIdx.x;
  const int64_t slot_idx = slot_mapping[token_idx];
  if (slot_idx < 0) return;
  const int64_t block_idx = slot_idx / block_size;
  const int64_t block_offset = slot_idx % block_size;
  float k_scale = 1.f, v_scale = 1.f;
  if (FP8) {
    k_scale = *k_scale_p;
    v_scale = *v_scale_p;
  }
  const scalar_t* ksrc = key + token_idx * key_stride;
  const scalar_t* vsrc = value + token_idx * value_stride;
  const int64_t base = block_idx * block_stride + block_offset * page_stride;
  const int n = num_heads * head_size;
  if (vec_ok) {
    constexpr int VE = FP8 ? 16 : (int)(16 / sizeof(scalar_t));  // elements per thread-iteration
    const int vec_per_head = head_size / VE;
    const int nvec = num_heads * vec_per_head;
    for (int r = threadIdx.x; r < 2 * nvec; r += blockDim.x) {
      const bool is_v = r >= nvec;
      const int rr = is_v ? r - nvec : r;
      const int h = rr / vec_per_head;
      const int ho = (rr - h * vec_per_head) * VE;
      const scalar_t* s = (is_v ? vsrc : ksrc) + h * head_size + ho;
      cache_t* d = (is_v ? value_cache : key_cache) + base + h * head_stride + ho;
      if (!FP8) {
        *reinterpret_cast<uint4*>(d) = *reinterpret_cast<const uint4*>(s);
      } else {
        const float sc = is_v ? v_scale : k_scale;
        union { uint4 u; uint8_t b[16]; } o;
#pragma unroll
        for (int j = 0; j < 16; ++j) o.b[j] = float_to_e4m3(to_float<scalar_t>(s[j]) / sc);
        *reinterpret_cast<uint4*>(d) = o.u;
      }
    }
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int h = i / head_size;
      const int ho = i - h * head_size;
      const int64_t t = base + h * head_stride + ho;
      if (!FP8) {
        key_cache[t] = *reinterpret_cast<const cache_t*>(&ksrc[i]);
        value_cache[t] = *reinterpret_cast<const cache_t*>(&vsrc[i]);
      } else {
        reinterpret_cast<uint8_t*>(key_cache)[t] =
            float_to_e4m3(to_float<scalar_t>(ksrc[i]) / k_scale);
        reinterpret_cast<uint8_t*>(value_cache)[t] =
            float_to_e4m3(to_float<scalar_t>(vsrc[i]) / v_scale);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// convert_fp8: flat conversion, both directions
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void to_fp8_kernel(uint8_t* __restrict__ dst, const T* __restrict__ src, int64_t n,
                              float scale) {
  pdl_wait();
  pdl_trigger();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = float_to_e4m3(to_float<T>(src[i]) / scale);
}
template <typename T>
__global__ void from_fp8_kernel(T* __restrict__ dst, const uint8_t* __restrict__ src, int64_t n,
                                float scale) {
  pdl_wait();
  pdl_trigger();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = from_float<T>(e4m3_to_float(src[i]) * scale);
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace vb200

using namespace vb200;

#define DISPATCH_SRC(dt, ...)                                   \
  switch (dt) {                                                 \
    case VB200_F32: { using scalar_t = float; __VA_ARGS__; break; }         \
    case VB200_F16: { using scalar_t = __half; __VA_ARGS__; break; }        \
    case VB200_BF16: { using scalar_t = __nv_bfloat16; __VA_ARGS__; break; } \
    default: VB_CHECK(false, "unsupported source dtype %d", dt);  \
  }

extern "C" int vb200_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                       void* value_cache, const int64_t* slot_mapping,
                                       int num_tokens, int num_heads, int head_size,
                                       int block_size, int x, int64_t key_stride,
                                       int64_t value_stride, int src_dtype, int kv_dtype,
                                       const float* k_scale, const float* v_scale,
                                       vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(kv_dtype == VB200_KV_AUTO || kv_dtype == VB200_KV_FP8_E4M3,
           "Unsupported data type of kv cache: %d", kv_dtype);
  VB_CHECK(x > 0 && head_size % x == 0, "head_size %d must be a multiple of x %d", head_size, x);
  if (num_tokens <= 0) return 0;
  const bool fp8 = kv_dtype == VB200_KV_FP8_E4M3;
  VB_CHECK(!fp8 || (k_scale && v_scale), "fp8 kv cache needs k_scale and v_scale");
  const int src_size = src_dtype == VB200_F32 ? 4 : 2;
  const int cache_elt = fp8 ? 1 : src_size;
  // vector path: an x-run is exactly 16 bytes of cache, source rows 16-byte aligned.
  bool vec_ok = (x * cache_elt == 16) && aligned16(key) && aligned16(key_cache) &&
                ((key_stride * src_size) % 16 == 0);
  if (fp8) vec_ok = vec_ok && (x == 16);
  const int n = num_heads * head_size;
  const int threads = n >= 1024 ? 256 : (n >= 256 ? 128 : 64);
  DISPATCH_SRC(src_dtype, {
    if (fp8)
      VB_CUDA_CHECK(launch_k(reshape_and_cache_kernel<scalar_t, true>, dim3(num_tokens), dim3(threads), 0, stream, (const scalar_t*)key, (const scalar_t*)value, key_cache, value_cache, slot_mapping,
          key_stride, value_stride, num_heads, head_size, block_size, x, k_scale, v_scale, vec_ok));
    else
      VB_CUDA_CHECK(launch_k(reshape_and_cache_kernel<scalar_t, false>, dim3(num_tokens), dim3(threads), 0, stream, (const scalar_t*)key, (const scalar_t*)value, key_cache, value_cache, slot_mapping,
          key_stride, value_stride, num_heads, head_size, block_size, x, k_scale, v_scale, vec_ok));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_rotary_embedding_reshape_and_cache(
    const int64_t* positions, void* query, void* key, const void* value, const void* cos_sin_cache, void* key_cache,
    void* value_cache, const int64_t* slot_mapping, int num_tokens, int num_heads, int num_kv_heads, int head_size,
    int rot_dim, int is_neox, int block_size, int x, int64_t query_stride, int64_t key_stride, int64_t value_stride,
    int src_dtype, int kv_dtype, const float* k_scale, const float* v_scale, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(src_dtype == VB200_F16 || src_dtype == VB200_BF16,
           "rotary_embedding_reshape_and_cache: dtype must be f16 or bf16");
  VB_CHECK(kv_dtype == VB200_KV_AUTO || kv_dtype == VB200_KV_FP8_E4M3,
           "Unsupported data type of kv cache: %d", kv_dtype);
  VB_CHECK(rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size,
           "rotary_embedding_reshape_and_cache: rot_dim %d must be even and <= head_size %d", rot_dim, head_size);
  VB_CHECK(x > 0 && head_size % x == 0, "head_size %d must be a multiple of x %d", head_size, x);
  VB_CHECK(key != nullptr && value != nullptr, "rotary_embedding_reshape_and_cache: key and value are required");
  if (num_tokens <= 0) return 0;
  const bool fp8 = kv_dtype == VB200_KV_FP8_E4M3;
  VB_CHECK(!fp8 || (k_scale && v_scale), "fp8 kv cache needs k_scale and v_scale");
  const int cache_elt = fp8 ? 1 : 2;
  bool vec_ok = (x * cache_elt == 16) && aligned16(key) && aligned16(key_cache) && ((key_stride * 2) % 16 == 0);
  if (fp8) vec_ok = vec_ok && (x == 16);
  // decode shapes: one 16-byte piece of the token per thread (rope_reshape_and_cache_vec_kernel)
  const bool pieces = vec_ok && rot_dim == head_size && head_size % 16 == 0 && x % 8 == 0 && aligned16(query) &&
                      aligned16(value) && aligned16(cos_sin_cache) && query_stride % 8 == 0 && value_stride % 8 == 0;
  if (pieces) {
    const int tasks = (num_heads + num_kv_heads) * (is_neox ? head_size / 16 : head_size / 8) + num_kv_heads * head_size / 8;
    const int threads = tasks >= 512 ? 512 : (tasks + 31) / 32 * 32;
#define VB_ROPE_VEC(T, NEOX, FP8)                                                                                     \
  VB_CUDA_CHECK(launch_k(rope_reshape_and_cache_vec_kernel<T, NEOX, FP8>, dim3(num_tokens), dim3(threads), 0, stream, \
                         positions, (T*)query, (T*)key, (const T*)value, (const T*)cos_sin_cache, key_cache,          \
                         value_cache, slot_mapping, query_stride, key_stride, value_stride, num_heads, num_kv_heads,  \
                         head_size, block_size, x, k_scale, v_scale))
#define VB_ROPE_VEC_T(T)                          \
  if (is_neox) {                                  \
    if (fp8) VB_ROPE_VEC(T, true, true);          \
    else VB_ROPE_VEC(T, true, false);             \
  } else {                                        \
    if (fp8) VB_ROPE_VEC(T, false, true);         \
    else VB_ROPE_VEC(T, false, false);            \
  }
    if (src_dtype == VB200_F16) {
      VB_ROPE_VEC_T(__half)
    } else {
      VB_ROPE_VEC_T(__nv_bfloat16)
    }
#undef VB_ROPE_VEC_T
#undef VB_ROPE_VEC
    VB_LAUNCH_CHECK();
    return 0;
  }
  const int work = (num_heads + num_kv_heads) * rot_dim / 2;
  const int threads = work >= 256 ? 256 : (work >= 128 ? 128 : 64);
#define VB_ROPE_CACHE(T, NEOX, FP8)                                                                                  \
  VB_CUDA_CHECK(launch_k(rope_reshape_and_cache_kernel<T, NEOX, FP8>, dim3(num_tokens), dim3(threads), 0, stream,     \
                         positions, (T*)query, (T*)key, (const T*)value, (const T*)cos_sin_cache, key_cache,          \
                         value_cache, slot_mapping, rot_dim, query_stride, key_stride, value_stride, num_heads,       \
                         num_kv_heads, head_size, block_size, x, k_scale, v_scale, vec_ok))
#define VB_ROPE_CACHE_T(T)                          \
  if (is_neox) {                                    \
    if (fp8) VB_ROPE_CACHE(T, true, true);          \
    else VB_ROPE_CACHE(T, true, false);             \
  } else {                                          \
    if (fp8) VB_ROPE_CACHE(T, false, true);         \
    else VB_ROPE_CACHE(T, false, false);            \
  }
  if (src_dtype == VB200_F16) {
    VB_ROPE_CACHE_T(__half)
  } else {
    VB_ROPE_CACHE_T(__nv_bfloat16)
  }
#undef VB_ROPE_CACHE_T
#undef VB_ROPE_CACHE
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                             void* value_cache, const int64_t* slot_mapping,
                                             int num_tokens, int num_heads, int head_size,
                                             int block_size, int64_t block_stride,
                                             int64_t page_stride, int64_t head_stride,
                                             int64_t key_stride, int64_t value_stride,
                                             int src_dtype, int kv_dtype, const float* k_scale,
                                             const float* v_scale, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(kv_dtype == VB200_KV_AUTO || kv_dtype == VB200_KV_FP8_E4M3,
           "Unsupported data type of kv cache: %d", kv_dtype);
  if (num_tokens <= 0) return 0;
  const bool fp8 = kv_dtype == VB200_KV_FP8_E4M3;
  VB_CHECK(!fp8 || (k_scale && v_scale), "fp8 kv cache needs k_scale and v_scale");
  const int src_size = src_dtype == VB200_F32 ? 4 : 2;
  const int cache_elt = fp8 ? 1 : src_size;
  const int ve = fp8 ? 16 : 16 / src_size;
  const bool vec_ok = (head_size % ve == 0) && aligned16(key) && aligned16(value) &&
                      aligned16(key_cache) && aligned16(value_cache) &&
                      ((key_stride * src_size) % 16 == 0) && ((value_stride * src_size) % 16 == 0) &&
                      ((block_stride * cache_elt) % 16 == 0) &&
                      ((page_stride * cache_elt) % 16 == 0) && ((head_stride * cache_elt) % 16 == 0);
  const int n = num_heads * head_size;
  const int threads = n >= 1024 ? 256 : (n >= 256 ? 128 : 64);
  DISPATCH_SRC(src_dtype, {
    if (fp8)
      VB_CUDA_CHECK(launch_k(reshape_and_cache_flash_kernel<scalar_t, true>, dim3(num_tokens), dim3(threads), 0, stream, (const scalar_t*)key, (const scalar_t*)value, key_cache, value_cache, slot_mapping,
          block_stride, page_stride, head_stride, key_stride, value_stride, num_heads, head_size,
          block_size, k_scale, v_scale, vec_ok));
    else
      VB_CUDA_CHECK(launch_k(reshape_and_cache_flash_kernel<scalar_t, false>, dim3(num_tokens), dim3(threads), 0, stream, (const scalar_t*)key, (const scalar_t*)value, key_cache, value_cache, slot_mapping,
          block_stride, page_stride, head_stride, key_stride, value_stride, num_heads, head_size,
          block_size, k_scale, v_scale, vec_ok));
  });
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_convert_fp8(void* dst, const void* src, int64_t numel, int dst_dtype,
                                 int src_dtype, float scale, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (numel <= 0) return 0;
  const int threads = 256;
  const int blocks = (int)((numel + threads - 1) / threads < 148 * 16
                               ? (numel + threads - 1) / threads
                               : 148 * 16);
  if (dst_dtype == VB200_FP8_E4M3) {
    DISPATCH_SRC(src_dtype, {
      VB_CUDA_CHECK(launch_k(to_fp8_kernel<scalar_t>, dim3(blocks), dim3(threads), 0, stream, (uint8_t*)dst, (const scalar_t*)src, numel, scale));
    });
  } else {
    VB_CHECK(src_dtype == VB200_FP8_E4M3, "convert_fp8: one side must be fp8 (src %d dst %d)",
             src_dtype, dst_dtype);
    DISPATCH_SRC(dst_dtype, {
      VB_CUDA_CHECK(launch_k(from_fp8_kernel<scalar_t>, dim3(blocks), dim3(threads), 0, stream, (scalar_t*)dst, (const uint8_t*)src, numel, scale));
    });
  }
  VB_LAUNCH_CHECK();
  return 0;
}
