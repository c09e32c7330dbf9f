// Routing metadata and row permutation of the CUTLASS-style fused-MoE API:
//   get_cutlass_moe_mm_data      /root/reference/csrc/quantization/cutlass_w8a8/moe/moe_data.cu:11-159
//   get_cutlass_pplx_moe_mm_data /root/reference/csrc/quantization/cutlass_w8a8/moe/moe_data.cu:161-195
//   shuffle_rows                 /root/reference/vllm/_custom_ops.py:912-922 (torch.ops._moe_C.shuffle_rows)
// The reference runs five small kernels (zero-fill, per-expert count with atomics, serial offsets, per-expert
// arg-sort with atomics -> order inside an expert depends on thread scheduling).  Here ONE CTA does a stable
// counting sort in shared memory: same expert_offsets / problem sizes, and the permutation is the deterministic
// member (ascending flat index inside every expert) of the set of permutations the reference may return.
#include "common.cuh"

namespace vb200 {

constexpr int kDataThreads = 1024;
constexpr int kDataWarps = kDataThreads / 32;

// smem: hist[E] | start[E+1] | run[E] | warp_cnt[NW][E]
__global__ void __launch_bounds__(kDataThreads) moe_mm_data_kernel(
    const int* __restrict__ topk_ids, int numel, int topk, int num_experts, int n, int k, int swap_ab,
    int* __restrict__ expert_offsets, int* __restrict__ problem_sizes1, int* __restrict__ problem_sizes2,
    int* __restrict__ input_permutation, int* __restrict__ output_permutation,
    int* __restrict__ blockscale_offsets) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ int sm[];
  int* hist = sm;
  int* start = hist + num_experts;
  int* run = start + num_experts + 1;
  int* warp_cnt = run + num_experts;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < num_experts; i += kDataThreads) {
    hist[i] = 0;
    run[i] = 0;
  }
  for (int i = tid; i < kDataWarps * num_experts; i += kDataThreads) warp_cnt[i] = 0;
  __syncthreads();
  for (int i = tid; i < numel; i += kDataThreads) {
    const int e = topk_ids[i];
    if (e >= 0 && e < num_experts) atomicAdd(&hist[e], 1);
  }
  __syncthreads();
  // problem sizes (moe_data.cu:28-44): (m_e, 2n, k) / (m_e, k, n), first two entries swapped for swap_ab
  for (int e = tid; e < num_experts; e += kDataThreads) {
    const int m = hist[e];
    problem_sizes1[e * 3 + (swap_ab ? 1 : 0)] = m;
    problem_sizes1[e * 3 + (swap_ab ? 0 : 1)] = 2 * n;
    problem_sizes1[e * 3 + 2] = k;
    problem_sizes2[e * 3 + (swap_ab ? 1 : 0)] = m;
    problem_sizes2[e * 3 + (swap_ab ? 0 : 1)] = k;
    problem_sizes2[e * 3 + 2] = n;
  }
  // exclusive prefix over experts (serial over <= 1024 shared-memory words: moe_data.cu:48-59)
  if (tid == 0) {
    int tot = 0, tot_round = 0;
    if (blockscale_offsets) blockscale_offsets[0] = 0;
    for (int e = 0; e < num_experts; ++e) {
      start[e] = tot;
      expert_offsets[e] = tot;
      tot += hist[e];
      if (blockscale_offsets) {
        tot_round += (hist[e] + 127) / 128 * 128;   // moe_data.cu:75
        blockscale_offsets[e + 1] = tot_round;
      }
    }
    start[num_experts] = tot;
    expert_offsets[num_experts] = tot;
  }
  __syncthreads();
  const int total = start[num_experts];
  // stable scatter, 1024 flat indices per pass
  for (int base = 0; base < numel; base += kDataThreads) {
    const int i = base + tid;
    int e = -1;
    bool invalid = false;
    if (i < numel) {
      e = topk_ids[i];
      invalid = e == -1;
      if (e < 0 || e >= num_experts) e = -1;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, e);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const bool leader = rank == 0 && e >= 0;
    if (leader) warp_cnt[warp * num_experts + e] = __popc(peers);
    __syncthreads();
    int pos = -1;
    if (e >= 0) {
      int before = 0;
      for (int w = 0; w < warp; ++w) before += warp_cnt[w * num_experts + e];
      pos = start[e] + run[e] + before + rank;
    }
    __syncthreads();
    if (leader) {
      atomicAdd(&run[e], __popc(peers));
      warp_cnt[warp * num_experts + e] = 0;
    }
    if (pos >= 0) {
      input_permutation[pos] = i / topk;
      output_permutation[i] = pos;
    } else if (invalid) {
      output_permutation[i] = total;  // moe_data.cu:92-99: row `total` of the zero-padded output
    }
    __syncthreads();
  }
}

__global__ void moe_pplx_data_kernel(int* __restrict__ expert_offsets, int* __restrict__ problem_sizes1,
                                     int* __restrict__ problem_sizes2,
                                     const int* __restrict__ expert_num_tokens, int num_local_experts,
                                     int padded_m, int n, int k) {
  pdl_wait();
  pdl_trigger();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= num_local_experts) return;
  expert_offsets[e] = e * padded_m;
  const int m = expert_num_tokens[e];
  problem_sizes1[e * 3] = m;
  problem_sizes1[e * 3 + 1] = 2 * n;
  problem_sizes1[e * 3 + 2] = k;
  problem_sizes2[e * 3] = m;
  problem_sizes2[e * 3 + 1] = k;
  problem_sizes2[e * 3 + 2] = n;
}

// out[r, :] = in[dst2src[r], :], rows of row_bytes bytes (16-byte vectors when aligned, bytes otherwise)
template <typename V>
__global__ void __launch_bounds__(256) shuffle_rows_kernel(uint8_t* __restrict__ out,
                                                           const uint8_t* __restrict__ in,
                                                           const int* __restrict__ dst2src, int num_dst,
                                                           int num_src, int64_t row_bytes) {
  pdl_wait();
  pdl_trigger();
  const int64_t nvec = row_bytes / (int64_t)sizeof(V);
  for (int r = blockIdx.x; r < num_dst; r += gridDim.x) {
    const int s = dst2src[r];
    if (s < 0 || s >= num_src) continue;
    const V* src = reinterpret_cast<const V*>(in + (int64_t)s * row_bytes);
    V* dst = reinterpret_cast<V*>(out + (int64_t)r * row_bytes);
    for (int64_t i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = src[i];
  }
}

// ---------------------------------------------------------------------------------------------
// The two ends of the permuted pipeline as single launches (what the B200 path of cutlass_moe_fp8 uses instead of
// "quantise, then shuffle_rows" and "shuffle_rows, multiply by the routing weights, moe_sum"):
//   gather + quantise   a_q[r, :] = fp8(a[src_row[r], :] / s_r)   s_r = the row's dynamic per-token scale
//                       (/root/reference/csrc/quantization/fp8/common.cu:28-83) or the static *scale (:14-26);
//                       per-token quantisation is row-wise, so quantising the gathered row gives the same bytes as
//                       gathering the quantised row (cutlass_moe.py:142-150 `_fp8_perm`).
//   un-permute + weight + reduce   out[t, :] = sum_j T(float(c[dst_row[t*k+j], :]) * w[t, j]), accumulated like
//                       moe_sum (in T, j order for k in 2..4; fp32 otherwise), rows past *num_routed count as zero
//                       (TopKWeightAndReduceContiguous, topk_weight_and_reduce.py:60-100 + moe_data.cu:92-99).
// ---------------------------------------------------------------------------------------------
constexpr float kE4m3Max = 448.f;

template <typename T>
__global__ void __launch_bounds__(256) moe_gather_quant_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                                               const T* __restrict__ in,
                                                               const int* __restrict__ src_row,
                                                               const int* __restrict__ num_routed,
                                                               const float* __restrict__ static_scale,
                                                               const float* __restrict__ scale_ub, int num_src,
                                                               int hidden, int src_div) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm[8];
  const int r = blockIdx.x;
  if (num_routed != nullptr && r >= *num_routed) return;   // rows no local expert owns: never read downstream
  // src_row holds source rows (src_div = 1) or (token, slot) pair ids whose source row is id / top_k (the
  // sorted_token_ids of moe_align_block_size; its padding ids are >= num_src * src_div)
  int s_row = src_row[r];
  if (s_row < 0 || s_row >= num_src * src_div) return;
  s_row /= src_div;
  constexpr int VE = 16 / sizeof(T);
  const T* rin = in + (int64_t)s_row * hidden;
  uint8_t* rout = out + (int64_t)r * hidden;
  const int nvec = hidden / VE;   // host guarantees hidden % VE == 0 and 16-byte aligned rows
  float s;
  if (static_scale != nullptr) {
    s = 1.0f / *static_scale;     // common.cu:23: multiply by the reciprocal
  } else {
    float m = 0.f;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      union { uint4 u; T e[VE]; } v;
      v.u = *reinterpret_cast<const uint4*>(rin + (int64_t)i * VE);
#pragma unroll
      for (int j = 0; j < VE; ++j) m = fmaxf(m, fabsf(to_float<T>(v.e[j])));
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    m = warp_max((threadIdx.x & 31) < (blockDim.x >> 5) ? sm[threadIdx.x & 31] : 0.f);
    s = scale_ub ? fminf(m, *scale_ub) : m;
    s = fmaxf(s / kE4m3Max, 1.0f / (kE4m3Max * 512.0f));   // common.cu:66-68
    if (threadIdx.x == 0) scales[r] = s;
  }
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    union { uint4 u; T e[VE]; } v;
    v.u = *reinterpret_cast<const uint4*>(rin + (int64_t)i * VE);   // second pass hits L1/L2
    union { uint2 u2; uint8_t b[8]; } o;
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      const float x = to_float<T>(v.e[j]);
      const float q = static_scale != nullptr ? x * s : x / s;      // true division on the per-token path (:74)
      o.b[j] = float_to_e4m3(fmaxf(-kE4m3Max, fminf(q, kE4m3Max)));
    }
    *reinterpret_cast<uint2*>(rout + (int64_t)i * VE) = o.u2;
  }
}

template <typename T, int TOPK>
__global__ void __launch_bounds__(256) moe_unpermute_reduce_kernel(T* __restrict__ out, const T* __restrict__ c,
                                                                   const int* __restrict__ dst_row,
                                                                   const float* __restrict__ weights,
                                                                   const int* __restrict__ num_routed, int topk,
                                                                   int hidden) {
  pdl_wait();
  pdl_trigger();
  constexpr int VE = 16 / sizeof(T);
  constexpr int KMAX = TOPK > 0 ? TOPK : 1;
  const int64_t t = blockIdx.x;
  const int limit = *num_routed;
  const int nvec = hidden / VE;
  if (TOPK > 0) {
    int rows[KMAX];
    float w[KMAX];
#pragma unroll
    for (int j = 0; j < KMAX; ++j) {
      rows[j] = dst_row[t * TOPK + j];
      w[j] = weights ? weights[t * TOPK + j] : 1.f;
    }
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      union { uint4 u; T e[VE]; } acc, v;
#pragma unroll
      for (int e = 0; e < VE; ++e) acc.e[e] = from_float<T>(0.f);
#pragma unroll
      for (int j = 0; j < KMAX; ++j) {
        if (rows[j] >= 0 && rows[j] < limit) v.u = ldg_stream_v4(c + (int64_t)rows[j] * hidden + (int64_t)i * VE);
        else v.u = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int e = 0; e < VE; ++e) {
          // fused_out.mul_(w): fp32 product rounded once to T; moe_sum: adds in T, j order, from 0
          const float p = weights ? to_float<T>(from_float<T>(to_float<T>(v.e[e]) * w[j])) : to_float<T>(v.e[e]);
          acc.e[e] = from_float<T>(to_float<T>(acc.e[e]) + p);
        }
      }
      *reinterpret_cast<uint4*>(out + t * hidden + (int64_t)i * VE) = acc.u;
    }
  } else {
    // other top-k: at::sum_out semantics (fp32 accumulation of the T-rounded products, one final rounding)
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      float acc[VE];
#pragma unroll
      for (int e = 0; e < VE; ++e) acc[e] = 0.f;
      for (int j = 0; j < topk; ++j) {
        const int row = dst_row[t * topk + j];
        if (row < 0 || row >= limit) continue;
        const float wj = weights ? weights[t * topk + j] : 1.f;
        union { uint4 u; T e[VE]; } v;
        v.u = ldg_stream_v4(c + (int64_t)row * hidden + (int64_t)i * VE);
#pragma unroll
        for (int e = 0; e < VE; ++e)
          acc[e] += weights ? to_float<T>(from_float<T>(to_float<T>(v.e[e]) * wj)) : to_float<T>(v.e[e]);
      }
      union { uint4 u; T e[VE]; } o;
#pragma unroll
      for (int e = 0; e < VE; ++e) o.e[e] = from_float<T>(acc[e]);
      *reinterpret_cast<uint4*>(out + t * hidden + (int64_t)i * VE) = o.u;
    }
  }
}

}  // namespace vb200

using namespace vb200;

extern "C" int vb200_get_cutlass_moe_mm_data(const int32_t* topk_ids, int num_tokens, int topk,
                                             int32_t* expert_offsets, int32_t* problem_sizes1,
                                             int32_t* problem_sizes2, int32_t* input_permutation,
                                             int32_t* output_permutation, int num_experts, int n, int k,
                                             int32_t* blockscale_offsets, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(num_experts >= 1 && num_experts <= 1024,
           "get_cutlass_moe_mm_data: num_experts %d must be in [1, 1024]", num_experts);
  VB_CHECK(topk >= 1 && num_tokens >= 0, "get_cutlass_moe_mm_data: bad topk_ids shape");
  const int numel = num_tokens * topk;
  // moe_data.cu:9,121-122: problem sizes are written M/N-swapped for <= 64 routed rows (never for the fp4 path)
  const int swap_ab = (blockscale_offsets == nullptr && numel <= 64) ? 1 : 0;
  const size_t smem = ((size_t)num_experts * 3 + 1 + (size_t)kDataWarps * num_experts) * 4;
  if (smem > 48 * 1024)
    VB_CUDA_CHECK(cudaFuncSetAttribute(moe_mm_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
  VB_CUDA_CHECK(launch_k(moe_mm_data_kernel, dim3(1), dim3(kDataThreads), smem, stream, topk_ids, numel, topk, num_experts, n, k, swap_ab,
                                                        expert_offsets, problem_sizes1, problem_sizes2,
                                                        input_permutation, output_permutation,
                                                        blockscale_offsets));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_get_cutlass_pplx_moe_mm_data(int32_t* expert_offsets, int32_t* problem_sizes1,
                                                  int32_t* problem_sizes2,
                                                  const int32_t* expert_num_tokens, int num_local_experts,
                                                  int padded_m, int n, int k, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(num_local_experts >= 1, "get_cutlass_pplx_moe_mm_data: no experts");
  VB_CUDA_CHECK(launch_k(moe_pplx_data_kernel, dim3((num_local_experts + 127) / 128), dim3(128), 0, stream, expert_offsets, problem_sizes1, problem_sizes2, expert_num_tokens, num_local_experts, padded_m, n, k));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_shuffle_rows(void* out, const void* in, const int32_t* dst2src_map, int num_dst_rows,
                                  int num_src_rows, int64_t row_bytes, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(row_bytes > 0, "shuffle_rows: empty rows");
  if (num_dst_rows <= 0) return 0;
  const int grid = num_dst_rows < num_sms() * 8 ? num_dst_rows : num_sms() * 8;
  const bool vec = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(in) & 15) == 0;
  uint8_t* o = reinterpret_cast<uint8_t*>(out);
  const uint8_t* i = reinterpret_cast<const uint8_t*>(in);
  if (vec)
    VB_CUDA_CHECK(launch_k(shuffle_rows_kernel<uint4>, dim3(grid), dim3(row_bytes >= 4096 ? 256 : 128), 0, stream, o, i, dst2src_map, num_dst_rows, num_src_rows, row_bytes));
  else
    VB_CUDA_CHECK(launch_k(shuffle_rows_kernel<uint8_t>, dim3(grid), dim3(256), 0, stream, o, i, dst2src_map, num_dst_rows, num_src_rows,
                                                           row_bytes));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_moe_gather_quant_fp8(void* out, float* scales, const void* input, const int32_t* src_row,
                                          const int32_t* num_routed, const float* static_scale,
                                          const float* scale_ub, int num_dst_rows, int num_src_rows, int hidden,
                                          int in_dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_dst_rows <= 0 || hidden <= 0) return 0;
  VB_CHECK(in_dtype == VB200_BF16 || in_dtype == VB200_F16, "moe_gather_quant_fp8: input must be bf16 or f16");
  VB_CHECK(hidden % 8 == 0 && (reinterpret_cast<uintptr_t>(input) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0,
           "moe_gather_quant_fp8: hidden (%d) must be a multiple of 8 and the rows 16-byte aligned", hidden);
  VB_CHECK(static_scale != nullptr || scales != nullptr, "moe_gather_quant_fp8: dynamic quantisation needs `scales`");
  const int threads = hidden >= 2048 ? 256 : 128;
  if (in_dtype == VB200_BF16)
    VB_CUDA_CHECK(launch_k(moe_gather_quant_kernel<__nv_bfloat16>, dim3(num_dst_rows), dim3(threads), 0, stream,
                           (uint8_t*)out, scales, (const __nv_bfloat16*)input, src_row, num_routed, static_scale,
                           scale_ub, num_src_rows, hidden, 1));
  else
    VB_CUDA_CHECK(launch_k(moe_gather_quant_kernel<__half>, dim3(num_dst_rows), dim3(threads), 0, stream,
                           (uint8_t*)out, scales, (const __half*)input, src_row, num_routed, static_scale, scale_ub,
                           num_src_rows, hidden, 1));
  VB_LAUNCH_CHECK();
  return 0;
}

// The same gather + static quantisation driven by moe_align_block_size's sorted_token_ids (pair ids, source row =
// id / top_k): used by vb200_moe_grouped_gemm_fp8_q16 (moe_gemm.cu) in place of "quantise the tensor, then gather".
namespace vb200 {
int launch_moe_gather_quant_sorted(uint8_t* out, const void* input, const int32_t* sorted_token_ids,
                                   const int32_t* num_tokens_post_pad, const float* static_scale, int sorted_capacity,
                                   int num_src_rows, int top_k, int hidden, int in_dtype, cudaStream_t stream) {
  const int threads = hidden >= 2048 ? 256 : 128;
  if (in_dtype == VB200_BF16)
    VB_CUDA_CHECK(launch_k(moe_gather_quant_kernel<__nv_bfloat16>, dim3(sorted_capacity), dim3(threads), 0, stream, out,
                           (float*)nullptr, (const __nv_bfloat16*)input, sorted_token_ids, num_tokens_post_pad,
                           static_scale, (const float*)nullptr, num_src_rows, hidden, top_k));
  else
    VB_CUDA_CHECK(launch_k(moe_gather_quant_kernel<__half>, dim3(sorted_capacity), dim3(threads), 0, stream, out,
                           (float*)nullptr, (const __half*)input, sorted_token_ids, num_tokens_post_pad, static_scale,
                           (const float*)nullptr, num_src_rows, hidden, top_k));
  VB_LAUNCH_CHECK();
  return 0;
}
}  // namespace vb200

template <typename T>
static int launch_unpermute_reduce(void* out, const void* c, const int32_t* dst_row, const float* weights,
                                   const int32_t* num_routed, int num_tokens, int topk, int hidden,
                                   cudaStream_t stream) {
  const int threads = hidden >= 2048 ? 256 : 128;
  T* o = (T*)out;
  const T* ci = (const T*)c;
  switch (topk) {
    case 1: VB_CUDA_CHECK(launch_k(moe_unpermute_reduce_kernel<T, 1>, dim3(num_tokens), dim3(threads), 0, stream, o, ci, dst_row, weights, num_routed, topk, hidden)); break;
    case 2: VB_CUDA_CHECK(launch_k(moe_unpermute_reduce_kernel<T, 2>, dim3(num_tokens), dim3(threads), 0, stream, o, ci, dst_row, weights, num_routed, topk, hidden)); break;
    case 3: VB_CUDA_CHECK(launch_k(moe_unpermute_reduce_kernel<T, 3>, dim3(num_tokens), dim3(threads), 0, stream, o, ci, dst_row, weights, num_routed, topk, hidden)); break;
    case 4: VB_CUDA_CHECK(launch_k(moe_unpermute_reduce_kernel<T, 4>, dim3(num_tokens), dim3(threads), 0, stream, o, ci, dst_row, weights, num_routed, topk, hidden)); break;
    default: VB_CUDA_CHECK(launch_k(moe_unpermute_reduce_kernel<T, 0>, dim3(num_tokens), dim3(threads), 0, stream, o, ci, dst_row, weights, num_routed, topk, hidden)); break;
  }
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_moe_unpermute_reduce(void* out, const void* c, const int32_t* dst_row, const float* topk_weights,
                                          const int32_t* num_routed, int num_tokens, int topk, int hidden, int dtype,
                                          vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_tokens <= 0 || hidden <= 0) return 0;
  VB_CHECK(dtype == VB200_BF16 || dtype == VB200_F16, "moe_unpermute_reduce: dtype must be bf16 or f16");
  VB_CHECK(topk >= 1 && num_routed != nullptr, "moe_unpermute_reduce: bad arguments");
  VB_CHECK(hidden % 8 == 0 && (reinterpret_cast<uintptr_t>(c) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
           "moe_unpermute_reduce: hidden (%d) must be a multiple of 8 and the rows 16-byte aligned", hidden);
  if (dtype == VB200_BF16)
    return launch_unpermute_reduce<__nv_bfloat16>(out, c, dst_row, topk_weights, num_routed, num_tokens, topk, hidden, stream);
  return launch_unpermute_reduce<__half>(out, c, dst_row, topk_weights, num_routed, num_tokens, topk, hidden, stream);
}
