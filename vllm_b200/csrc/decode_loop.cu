// Kernels either side of the decode hot path:
//   advance_step_flashattn  /root/reference/csrc/prepare_inputs/advance_step.cu:13-65,176-232
//       on-device update of the next step's inputs (token ids, positions, seq_lens, slot_mapping) so a decode loop
//       replays one CUDA graph with no host round trip;
//   merge_attn_states       /root/reference/csrc/attention/merge_attn_states.cu:12-89,121-178
//       log-sum-exp merge of two partial attention results (split-KV / cascade attention).
//   greedy_argmax           the greedy branch of /root/reference/vllm/v1/sample/sampler.py:87-96 (`logits.argmax(dim=-1)`)
//       on the logits in the dtype the lm_head GEMM wrote them (no fp32 copy of a [tokens, vocab] matrix): arg-max
//       and max per row, first index on ties, NaN counts as the largest value (torch.argmax's rules).
// All are HBM-trivial; one launch each, 16-byte accesses.
#include <cmath>
#include <limits>

#include "common.cuh"

namespace vb200 {

__global__ void __launch_bounds__(256) advance_step_kernel(
    int num_seqs, int num_queries, int block_size, int64_t* __restrict__ input_tokens,
    const int64_t* __restrict__ sampled_token_ids, int64_t* __restrict__ input_positions,
    int* __restrict__ seq_lens, int64_t* __restrict__ slot_mapping, const int* __restrict__ block_tables,
    int64_t block_tables_stride) {
  pdl_wait();
  pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= num_seqs) return;
  if (i >= num_queries) {  // CUDA-graph padding rows (advance_step.cu:21-30)
    input_tokens[i] = 0;
    input_positions[i] = 0;
    slot_mapping[i] = -1;
    return;
  }
  input_tokens[i] = sampled_token_ids[i];
  const int next_len = seq_lens[i] + 1;
  const int pos = next_len - 1;
  seq_lens[i] = next_len;
  input_positions[i] = pos;
  const int blk = block_tables[block_tables_stride * i + pos / block_size];
  slot_mapping[i] = (int64_t)(blk * block_size + pos % block_size);  // int arithmetic as the reference (:62)
}

template <typename T>
__global__ void __launch_bounds__(128) merge_attn_states_kernel(
    T* __restrict__ output, float* __restrict__ output_lse, const T* __restrict__ prefix_output,
    const float* __restrict__ prefix_lse, const T* __restrict__ suffix_output,
    const float* __restrict__ suffix_lse, unsigned num_tokens, unsigned num_heads, unsigned head_size) {
  pdl_wait();
  pdl_trigger();
  constexpr unsigned kPack = 16 / sizeof(T);
  const unsigned threads_per_head = head_size / kPack;
  const unsigned gid = blockIdx.x * 128u + threadIdx.x;
  if (gid >= num_tokens * num_heads * threads_per_head) return;
  const unsigned th = gid / threads_per_head, pack = gid % threads_per_head;
  const unsigned token = th / num_heads, head = th % num_heads;
  const size_t off = ((size_t)token * num_heads + head) * head_size + (size_t)pack * kPack;
  float p_lse = prefix_lse[(size_t)head * num_tokens + token];
  float s_lse = suffix_lse[(size_t)head * num_tokens + token];
  const float ninf = -std::numeric_limits<float>::infinity();
  p_lse = isinf(p_lse) ? ninf : p_lse;   // +inf is treated as "no contribution" (:45-46)
  s_lse = isinf(s_lse) ? ninf : s_lse;
  const float max_lse = fmaxf(p_lse, s_lse);
  const float p_se = expf(p_lse - max_lse), s_se = expf(s_lse - max_lse);
  const float out_se = p_se + s_se;
  const float p_scale = p_se / out_se, s_scale = s_se / out_se;
  const uint4 pv = *reinterpret_cast<const uint4*>(prefix_output + off);
  const uint4 sv = *reinterpret_cast<const uint4*>(suffix_output + off);
  uint4 ov;
  const T* pp = reinterpret_cast<const T*>(&pv);
  const T* sp = reinterpret_cast<const T*>(&sv);
  T* op = reinterpret_cast<T*>(&ov);
#pragma unroll
  for (unsigned i = 0; i < kPack; ++i) {
    // p * p_scale + (s * s_scale): one fused multiply-add around the rounded product, as the reference compiles (:73)
    op[i] = from_float<T>(fmaf(to_float(pp[i]), p_scale, to_float(sp[i]) * s_scale));
  }
  *reinterpret_cast<uint4*>(output + off) = ov;
  if (output_lse != nullptr && pack == 0) output_lse[(size_t)head * num_tokens + token] = logf(out_se) + max_lse;
}

// One CTA per row of logits.  (value, index) pairs are compared as torch.argmax does: larger value wins, NaN beats
// every number, the lower index wins a tie.
__device__ __forceinline__ bool argmax_better(float v, int i, float bv, int bi) {
  const bool vn = v != v, bn = bv != bv;
  if (vn || bn) return vn && (!bn || i < bi);
  return v > bv || (v == bv && i < bi);
}

template <typename T>
__global__ void __launch_bounds__(512) greedy_argmax_kernel(int64_t* __restrict__ out_idx, float* __restrict__ out_val,
                                                            const T* __restrict__ logits, int vocab,
                                                            int64_t row_stride, int vec_ok) {
  pdl_wait();
  pdl_trigger();
  constexpr int kPack = 16 / sizeof(T);
  const T* row = logits + (int64_t)blockIdx.x * row_stride;
  float bv = -std::numeric_limits<float>::infinity();
  int bi = 0x7fffffff;
  if (vec_ok) {
    const int nvec = vocab / kPack;
#pragma unroll 4
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      const uint4 u = ldg_stream_v4(row + (int64_t)i * kPack);
      const T* e = reinterpret_cast<const T*>(&u);
#pragma unroll
      for (int j = 0; j < kPack; ++j) {
        const float v = to_float<T>(e[j]);
        if (argmax_better(v, i * kPack + j, bv, bi)) { bv = v; bi = i * kPack + j; }
      }
    }
    for (int i = nvec * kPack + threadIdx.x; i < vocab; i += blockDim.x) {
      const float v = to_float<T>(row[i]);
      if (argmax_better(v, i, bv, bi)) { bv = v; bi = i; }
    }
  } else {
    for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
      const float v = to_float<T>(row[i]);
      if (argmax_better(v, i, bv, bi)) { bv = v; bi = i; }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (argmax_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
  }
  __shared__ float sv[16];
  __shared__ int si[16];
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = blockDim.x >> 5;
    bv = threadIdx.x < nw ? sv[threadIdx.x] : -std::numeric_limits<float>::infinity();
    bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (argmax_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (threadIdx.x == 0) {
      out_idx[blockIdx.x] = bi;
      if (out_val != nullptr) out_val[blockIdx.x] = bv;
    }
  }
}

}  // namespace vb200

using namespace vb200;

extern "C" int vb200_greedy_argmax(int64_t* out_idx, float* out_val, const void* logits, int num_rows, int vocab,
                                   int64_t row_stride, int dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(dtype == VB200_F32 || dtype == VB200_F16 || dtype == VB200_BF16, "greedy_argmax: unsupported logits dtype %d", dtype);
  VB_CHECK(vocab > 0 && row_stride >= vocab, "greedy_argmax: need vocab > 0 and row_stride >= vocab");
  if (num_rows <= 0) return 0;
  const int esz = dtype == VB200_F32 ? 4 : 2;
  const int vec_ok = ((reinterpret_cast<uintptr_t>(logits) & 15) == 0 && (row_stride * esz) % 16 == 0) ? 1 : 0;
#define VB_ARGMAX(T)                                                                                               \
  VB_CUDA_CHECK(launch_k(greedy_argmax_kernel<T>, dim3(num_rows), dim3(512), 0, stream, out_idx, out_val,          \
                         (const T*)logits, vocab, row_stride, vec_ok))
  if (dtype == VB200_F32) VB_ARGMAX(float);
  else if (dtype == VB200_F16) VB_ARGMAX(__half);
  else VB_ARGMAX(__nv_bfloat16);
#undef VB_ARGMAX
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_advance_step_flashattn(int num_seqs, int num_queries, int block_size,
                                            int64_t* input_tokens, const int64_t* sampled_token_ids,
                                            int64_t* input_positions, int32_t* seq_lens,
                                            int64_t* slot_mapping, const int32_t* block_tables,
                                            int64_t block_tables_stride, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(num_queries >= 0 && num_seqs >= num_queries, "advance_step_flashattn: need 0 <= num_queries <= num_seqs");
  VB_CHECK(block_size > 0, "advance_step_flashattn: bad block_size %d", block_size);
  if (num_seqs == 0) return 0;
  VB_CUDA_CHECK(launch_k(advance_step_kernel, dim3((num_seqs + 255) / 256), dim3(256), 0, stream, num_seqs, num_queries, block_size, input_tokens, sampled_token_ids, input_positions, seq_lens, slot_mapping,
      block_tables, block_tables_stride));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                                       const float* prefix_lse, const void* suffix_output,
                                       const float* suffix_lse, int num_tokens, int num_heads, int head_size,
                                       int dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int esz = dtype == VB200_F32 ? 4 : 2;
  VB_CHECK(dtype == VB200_F32 || dtype == VB200_F16 || dtype == VB200_BF16, "Unsupported data type of O: %d", dtype);
  const int pack = 16 / esz;
  VB_CHECK(head_size > 0 && head_size % pack == 0, "headsize must be multiple of pack_size:%d", pack);
  VB_CHECK(((reinterpret_cast<uintptr_t>(output) | reinterpret_cast<uintptr_t>(prefix_output) |
             reinterpret_cast<uintptr_t>(suffix_output)) & 15) == 0,
           "merge_attn_states: outputs must be 16-byte aligned");
  if (num_tokens <= 0 || num_heads <= 0) return 0;
  const long total = (long)num_tokens * num_heads * (head_size / pack);
  const unsigned grid = (unsigned)((total + 127) / 128);
#define VB_MERGE(T)                                                                                       \
  VB_CUDA_CHECK(launch_k(merge_attn_states_kernel<T>, dim3(grid), dim3(128), 0, stream, (T*)output, output_lse, (const T*)prefix_output, \
                                                        prefix_lse, (const T*)suffix_output, suffix_lse, \
                                                        num_tokens, num_heads, head_size))
  if (dtype == VB200_F32) VB_MERGE(float);
  else if (dtype == VB200_F16) VB_MERGE(__half);
  else VB_MERGE(__nv_bfloat16);
#undef VB_MERGE
  VB_LAUNCH_CHECK();
  return 0;
}
