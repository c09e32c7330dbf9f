// torch dispatcher shim over the C ABI of libvllm_b200.so (include/vllm_b200.h): registers the decode
// hot-path ops under the reference's own schema strings
//   /root/reference/csrc/torch_bindings.cpp:45-69 (paged attention), :111 (silu_and_mul), :171-219 (norms, rope),
//   :422-441 (cutlass_scaled_mm), :575-591 (fp8 quant), :676-707 (cache ops),
//   /root/reference/csrc/moe/torch_bindings.cpp:6-22 (MoE routing)
// so `torch.ops.<lib>.<op>` calls written against vLLM run unchanged.  The library names are build-time macros
// like the reference's TORCH_EXTENSION_NAME (csrc/core/registration.h): the default build registers
// _b200_C / _b200_C_cache_ops / _b200_moe_C for side-by-side tests; a drop-in build passes
// -DVB200_LIB_C=_C -DVB200_LIB_CACHE=_C_cache_ops -DVB200_LIB_MOE=_moe_C.
// This file is the binding a reference maintainer would add; it contains no kernels.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/all.h>
#include <torch/library.h>

#include <optional>
#include <string>

#include "../include/vllm_b200.h"

#ifndef VB200_LIB_C
#define VB200_LIB_C _b200_C
#endif
#ifndef VB200_LIB_CACHE
#define VB200_LIB_CACHE _b200_C_cache_ops
#endif
#ifndef VB200_LIB_MOE
#define VB200_LIB_MOE _b200_moe_C
#endif

namespace {

using torch::Tensor;
using OptTensor = std::optional<Tensor>;

inline void vb_check(int rc) { TORCH_CHECK(rc == 0, vb200_last_error()); }

inline int dt(const Tensor& t) {
  switch (t.scalar_type()) {
    case at::kFloat: return VB200_F32;
    case at::kHalf: return VB200_F16;
    case at::kBFloat16: return VB200_BF16;
    case at::kFloat8_e4m3fn:
    case at::kByte: return VB200_FP8_E4M3;
    case at::kInt: return VB200_I32;
    case at::kLong: return VB200_I64;
    case at::kUInt32: return VB200_U32;
    default: TORCH_CHECK(false, "vllm_b200: unsupported dtype ", t.scalar_type());
  }
  return -1;
}

inline int kvd(const std::string& s) {
  if (s == "auto") return VB200_KV_AUTO;
  if (s == "fp8" || s == "fp8_e4m3") return VB200_KV_FP8_E4M3;
  TORCH_CHECK(false, "Unsupported data type of kv cache: ", s);
  return -1;
}

inline void* stream_of(const Tensor& t) {
  return at::cuda::getCurrentCUDAStream(t.get_device()).stream();
}
inline const void* optptr(const OptTensor& t) { return t.has_value() ? t->const_data_ptr() : nullptr; }

// ---- attention ---------------------------------------------------------------------------------------
void paged_attention_v1(Tensor& out, Tensor& query, Tensor& key_cache, Tensor& value_cache,
                        int64_t num_kv_heads, double scale, Tensor& block_tables, Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const OptTensor& alibi_slopes,
                        const std::string& kv_cache_dtype, Tensor& k_scale, Tensor& v_scale, int64_t tp_rank,
                        int64_t blocksparse_local_blocks, int64_t blocksparse_vert_stride,
                        int64_t blocksparse_block_size, int64_t blocksparse_head_sliding_step) {
  TORCH_CHECK(blocksparse_vert_stride <= 1, "vllm_b200: block-sparse paged attention is not supported");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int S = query.size(0), H = query.size(1), D = query.size(2);
  const int kv = kvd(kv_cache_dtype);
  const int part = vb200_paged_attention_v1_partition_size(S, H, num_kv_heads, D, max_seq_len, block_size, kv);
  const size_t ws_bytes = vb200_paged_attention_v1_workspace_size(S, H, D, max_seq_len, dt(query), part);
  Tensor ws;
  if (ws_bytes) ws = torch::empty({(int64_t)ws_bytes}, query.options().dtype(at::kByte));
  vb_check(vb200_paged_attention_v1(
      out.data_ptr(), query.const_data_ptr(), key_cache.const_data_ptr(), value_cache.const_data_ptr(), S, H, D,
      num_kv_heads, (float)scale, block_tables.const_data_ptr<int32_t>(), seq_lens.const_data_ptr<int32_t>(),
      block_size, max_seq_len, block_tables.stride(0), (const float*)optptr(alibi_slopes), dt(query), kv,
      k_scale.const_data_ptr<float>(), v_scale.const_data_ptr<float>(), query.stride(0), key_cache.stride(0),
      key_cache.stride(1), part, ws_bytes ? ws.data_ptr() : nullptr, ws_bytes, stream_of(query)));
}

void paged_attention_v2(Tensor& out, Tensor& exp_sums, Tensor& max_logits, Tensor& tmp_out, Tensor& query,
                        Tensor& key_cache, Tensor& value_cache, int64_t num_kv_heads, double scale,
                        Tensor& block_tables, Tensor& seq_lens, int64_t block_size, int64_t max_seq_len,
                        const OptTensor& alibi_slopes, const std::string& kv_cache_dtype, Tensor& k_scale,
                        Tensor& v_scale, int64_t tp_rank, int64_t blocksparse_local_blocks,
                        int64_t blocksparse_vert_stride, int64_t blocksparse_block_size,
                        int64_t blocksparse_head_sliding_step) {
  TORCH_CHECK(blocksparse_vert_stride <= 1, "vllm_b200: block-sparse paged attention is not supported");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  TORCH_CHECK(exp_sums.size(2) == (max_seq_len + 511) / 512,
              "paged_attention_v2: partial buffers must have ceil(max_seq_len/512) partitions");
  vb_check(vb200_paged_attention_v2(
      out.data_ptr(), exp_sums.data_ptr<float>(), max_logits.data_ptr<float>(), tmp_out.data_ptr(),
      query.const_data_ptr(), key_cache.const_data_ptr(), value_cache.const_data_ptr(), query.size(0), query.size(1),
      query.size(2), num_kv_heads, (float)scale, block_tables.const_data_ptr<int32_t>(),
      seq_lens.const_data_ptr<int32_t>(), block_size, max_seq_len, block_tables.stride(0),
      (const float*)optptr(alibi_slopes), dt(query), kvd(kv_cache_dtype), k_scale.const_data_ptr<float>(),
      v_scale.const_data_ptr<float>(), query.stride(0), key_cache.stride(0), key_cache.stride(1), 512,
      stream_of(query)));
}

// ---- cache ops -----------------------------------------------------------------------------------------
void reshape_and_cache(Tensor& key, Tensor& value, Tensor& key_cache, Tensor& value_cache, Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, Tensor& k_scale, Tensor& v_scale) {
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  vb_check(vb200_reshape_and_cache(
      key.const_data_ptr(), value.const_data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      slot_mapping.const_data_ptr<int64_t>(), slot_mapping.size(0), key.size(1), key.size(2), key_cache.size(3),
      key_cache.size(4), key.stride(0), value.stride(0), dt(key), kvd(kv_cache_dtype),
      k_scale.const_data_ptr<float>(), v_scale.const_data_ptr<float>(), stream_of(key)));
}

void reshape_and_cache_flash(Tensor& key, Tensor& value, Tensor& key_cache, Tensor& value_cache,
                             Tensor& slot_mapping, const std::string& kv_cache_dtype, Tensor& k_scale,
                             Tensor& v_scale) {
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0));
  vb_check(vb200_reshape_and_cache_flash(
      key.const_data_ptr(), value.const_data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      slot_mapping.const_data_ptr<int64_t>(), slot_mapping.size(0), key.size(1), key.size(2), key_cache.size(1),
      key_cache.stride(0), key_cache.stride(1), key_cache.stride(2), key.stride(0), value.stride(0), dt(key),
      kvd(kv_cache_dtype), k_scale.const_data_ptr<float>(), v_scale.const_data_ptr<float>(), stream_of(key)));
}

void convert_fp8(Tensor& dst_cache, Tensor& src_cache, double scale, const std::string& kv_cache_dtype) {
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  kvd(kv_cache_dtype);
  vb_check(vb200_convert_fp8(dst_cache.data_ptr(), src_cache.const_data_ptr(), src_cache.numel(), dt(dst_cache),
                             dt(src_cache), (float)scale, stream_of(src_cache)));
}

// ---- fp8 linear -----------------------------------------------------------------------------------------
at::Tensor& gemm_workspace(const Tensor& like, size_t bytes) {
  // zero-initialised once per device (split-K tickets must start at zero; every launch leaves them zero)
  static at::Tensor ws[64];
  const int dev = like.get_device();
  const size_t want = bytes < (64u << 20) ? (64u << 20) : bytes;
  if (!ws[dev].defined() || (size_t)ws[dev].numel() < want)
    ws[dev] = torch::zeros({(int64_t)want}, like.options().dtype(at::kByte));
  return ws[dev];
}

void cutlass_scaled_mm(Tensor& out, const Tensor& a, const Tensor& b, const Tensor& a_scales,
                       const Tensor& b_scales, const OptTensor& bias) {
  // checks of csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:169-187
  TORCH_CHECK(a.dim() == 2 && b.dim() == 2 && out.dim() == 2);
  TORCH_CHECK(out.size(0) == a.size(0) && a.size(1) == b.size(0) && b.size(1) == out.size(1));
  TORCH_CHECK(a.stride(1) == 1 && out.stride(1) == 1);  // row-major
  TORCH_CHECK(b.stride(0) == 1);                        // column-major
  TORCH_CHECK(out.stride(0) % 16 == 0 && b.stride(1) % 16 == 0);
  TORCH_CHECK(a.scalar_type() == at::kFloat8_e4m3fn && b.scalar_type() == at::kFloat8_e4m3fn,
              "vllm_b200 cutlass_scaled_mm: only the fp8 e4m3 path is built");
  if (bias) TORCH_CHECK(bias->numel() == b.size(1) && bias->is_contiguous() && bias->dim() == 1);
  const at::cuda::OptionalCUDAGuard guard(device_of(a));
  const int M = a.size(0), N = b.size(1), K = a.size(1);
  const size_t ws_bytes = vb200_cutlass_scaled_mm_workspace_size(M, N, K);  // 0: split-K reduces through DSMEM
  void* ws_ptr = nullptr;
  size_t ws_len = 0;
  if (ws_bytes) {
    at::Tensor& ws = gemm_workspace(a, ws_bytes);
    ws_ptr = ws.data_ptr();
    ws_len = ws.numel();
  }
  vb_check(vb200_cutlass_scaled_mm(out.data_ptr(), a.const_data_ptr(), b.const_data_ptr(),
                                   a_scales.const_data_ptr<float>(), b_scales.const_data_ptr<float>(), optptr(bias), M,
                                   N, K, a.stride(0), b.stride(1), out.stride(0), a_scales.numel(), b_scales.numel(),
                                   dt(out), ws_ptr, ws_len, stream_of(a)));
}

bool cutlass_scaled_mm_supports_fp8(int64_t cuda_device_capability) {
  return vb200_cutlass_scaled_mm_supports_fp8((int)cuda_device_capability) != 0;
}

void static_scaled_fp8_quant(Tensor& out, const Tensor& input, const Tensor& scale) {
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int hidden = input.size(-1);
  vb_check(vb200_static_scaled_fp8_quant(out.data_ptr(), input.const_data_ptr(), scale.const_data_ptr<float>(),
                                         input.numel() / hidden, hidden, dt(input), stream_of(input)));
}
void dynamic_scaled_fp8_quant(Tensor& out, const Tensor& input, Tensor& scale) {
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int hidden = input.size(-1);
  vb_check(vb200_dynamic_scaled_fp8_quant(out.data_ptr(), input.const_data_ptr(), scale.data_ptr<float>(),
                                          input.numel() / hidden, hidden, dt(input), stream_of(input)));
}
void dynamic_per_token_scaled_fp8_quant(Tensor& out, const Tensor& input, Tensor& scales, const OptTensor& scale_ub) {
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int hidden = input.size(-1);
  vb_check(vb200_dynamic_per_token_scaled_fp8_quant(out.data_ptr(), input.const_data_ptr(), scales.data_ptr<float>(),
                                                    (const float*)optptr(scale_ub), input.numel() / hidden, hidden,
                                                    dt(input), stream_of(input)));
}
void silu_and_mul(Tensor& out, Tensor& input) {
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int d = input.size(-1) / 2;
  vb_check(vb200_silu_and_mul(out.data_ptr(), input.const_data_ptr(), input.numel() / input.size(-1), d, dt(input),
                              stream_of(input)));
}

// ---- norms / rope ----------------------------------------------------------------------------------------
void rms_norm(Tensor& out, Tensor& input, Tensor& weight, double epsilon) {
  TORCH_CHECK(out.is_contiguous() && input.stride(-1) == 1 && weight.is_contiguous());
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int hidden = input.size(-1);
  vb_check(vb200_rms_norm(out.data_ptr(), input.const_data_ptr(), weight.const_data_ptr(), (float)epsilon,
                          input.numel() / hidden, hidden, input.dim() >= 2 ? input.stride(-2) : hidden, dt(input),
                          stream_of(input)));
}
void fused_add_rms_norm(Tensor& input, Tensor& residual, Tensor& weight, double epsilon) {
  TORCH_CHECK(residual.is_contiguous() && weight.is_contiguous() && input.stride(-1) == 1);
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int hidden = input.size(-1);
  vb_check(vb200_fused_add_rms_norm(input.data_ptr(), residual.data_ptr(), weight.const_data_ptr(), (float)epsilon,
                                    input.numel() / hidden, hidden, input.dim() >= 2 ? input.stride(-2) : hidden,
                                    dt(input), stream_of(input)));
}
void rms_norm_dynamic_per_token_quant(Tensor& out, const Tensor& input, const Tensor& weight, Tensor& scales,
                                      double epsilon, OptTensor scale_ub, OptTensor residual) {
  TORCH_CHECK(out.is_contiguous() && input.is_contiguous());
  TORCH_CHECK(out.scalar_type() == at::kFloat8_e4m3fn, "vllm_b200: only the fp8 e4m3 output is built");
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int hidden = input.size(-1);
  vb_check(vb200_rms_norm_dynamic_per_token_quant(
      out.data_ptr(), input.const_data_ptr(), weight.const_data_ptr(), scales.data_ptr<float>(), (float)epsilon,
      (const float*)optptr(scale_ub), residual.has_value() ? residual->data_ptr() : nullptr, input.numel() / hidden,
      hidden, dt(input), stream_of(input)));
}
void rotary_embedding(Tensor& positions, Tensor& query, OptTensor key, int64_t head_size, Tensor& cos_sin_cache,
                      bool is_neox) {
  // shape rules of csrc/pos_encoding_kernels.cu:133-185 ([num_tokens] positions)
  TORCH_CHECK(positions.dim() == 1, "vllm_b200 rotary_embedding: positions must be [num_tokens]");
  const int64_t T = positions.numel();
  TORCH_CHECK(query.size(0) == T && (!key.has_value() || key->size(0) == T),
              "query, key and positions must have the same number of tokens");
  const int qh = query.numel() / T, kh = key.has_value() ? key->numel() / T : 0;
  TORCH_CHECK(qh % head_size == 0 && kh % head_size == 0);
  const int nh = qh / head_size, nkv = key.has_value() ? kh / head_size : nh;
  TORCH_CHECK(nh % nkv == 0);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  vb_check(vb200_rotary_embedding(positions.const_data_ptr<int64_t>(), query.data_ptr(),
                                  key.has_value() ? key->data_ptr() : nullptr, T, nh, nkv, head_size,
                                  cos_sin_cache.size(1), cos_sin_cache.const_data_ptr(), is_neox, query.stride(0),
                                  key.has_value() ? key->stride(0) : 0,
                                  query.dim() == 3 ? query.stride(1) : head_size, dt(query), stream_of(query)));
}

// ---- MoE routing ------------------------------------------------------------------------------------------
void topk_softmax(Tensor& topk_weights, Tensor& topk_indices, Tensor& token_expert_indices, Tensor& gating_output) {
  const at::cuda::OptionalCUDAGuard guard(device_of(gating_output));
  const int E = gating_output.size(-1);
  vb_check(vb200_topk_softmax(topk_weights.data_ptr<float>(), topk_indices.data_ptr(),
                              token_expert_indices.data_ptr<int32_t>(), gating_output.const_data_ptr<float>(),
                              gating_output.numel() / E, E, topk_weights.size(-1), dt(topk_indices),
                              stream_of(gating_output)));
}
void moe_sum(Tensor& input, Tensor& output) {
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  vb_check(vb200_moe_sum(output.data_ptr(), input.const_data_ptr(), input.size(0), input.size(1), input.size(2),
                         dt(input), stream_of(input)));
}
void moe_align_block_size(Tensor topk_ids, int64_t num_experts, int64_t block_size, Tensor sorted_token_ids,
                          Tensor experts_ids, Tensor num_tokens_post_pad) {
  const at::cuda::OptionalCUDAGuard guard(device_of(topk_ids));
  vb_check(vb200_moe_align_block_size(topk_ids.const_data_ptr<int32_t>(), topk_ids.numel(), num_experts, block_size,
                                      sorted_token_ids.data_ptr<int32_t>(), sorted_token_ids.numel(),
                                      experts_ids.data_ptr<int32_t>(), experts_ids.numel(),
                                      num_tokens_post_pad.data_ptr<int32_t>(), nullptr, 0, stream_of(topk_ids)));
}

// ---- CUTLASS-style fused-MoE API (csrc/ops.h:237-262, csrc/moe/moe_ops.h shuffle_rows) ----
bool cutlass_group_gemm_supported(int64_t cuda_device_capability) { return cuda_device_capability == 100; }
void cutlass_moe_mm(Tensor& out_tensors, const Tensor& a_tensors, const Tensor& b_tensors, const Tensor& a_scales,
                    const Tensor& b_scales, const Tensor& expert_offsets, const Tensor& problem_sizes,
                    const Tensor& a_strides, const Tensor& b_strides, const Tensor& c_strides, bool per_act_token,
                    bool per_out_ch) {
  (void)a_strides; (void)b_strides; (void)c_strides; (void)per_act_token; (void)per_out_ch;
  TORCH_CHECK(a_tensors.size(0) > 0, "No input A tensors provided.");
  TORCH_CHECK(b_tensors.size(0) > 0, "No input B tensors provided.");
  TORCH_CHECK(out_tensors.size(0) > 0, "No output tensors provided.");
  TORCH_CHECK(a_tensors.dtype() == at::kFloat8_e4m3fn, "A tensors must be of type float8_e4m3fn.");
  TORCH_CHECK(b_tensors.dtype() == at::kFloat8_e4m3fn, "B tensors must be of type float8_e4m3fn.");
  TORCH_CHECK(a_tensors.is_contiguous() && b_tensors.is_contiguous() && out_tensors.is_contiguous());
  const int64_t E = b_tensors.size(0), N = b_tensors.size(1), K = b_tensors.size(2);
  TORCH_CHECK(a_tensors.size(1) == K && out_tensors.size(1) == N && out_tensors.size(0) == a_tensors.size(0));
  const at::cuda::OptionalCUDAGuard guard(device_of(a_tensors));
  // scale granularity from the element counts (moe/get_group_starts.cuh:58-59)
  vb_check(vb200_cutlass_moe_mm(out_tensors.data_ptr(), a_tensors.const_data_ptr(), b_tensors.const_data_ptr(),
                                a_scales.const_data_ptr<float>(), b_scales.const_data_ptr<float>(),
                                expert_offsets.const_data_ptr<int32_t>(), problem_sizes.const_data_ptr<int32_t>(),
                                a_tensors.size(0), E, N, K, a_scales.numel() != 1, b_scales.numel() != E,
                                dt(out_tensors), stream_of(a_tensors)));
}
void get_cutlass_moe_mm_data(const Tensor& topk_ids, Tensor& expert_offsets, Tensor& problem_sizes1,
                             Tensor& problem_sizes2, Tensor& input_permutation, Tensor& output_permutation,
                             int64_t num_experts, int64_t n, int64_t k, const OptTensor& blockscale_offsets) {
  TORCH_CHECK(topk_ids.dtype() == at::kInt && topk_ids.dim() == 2 && topk_ids.is_contiguous());
  const at::cuda::OptionalCUDAGuard guard(device_of(topk_ids));
  vb_check(vb200_get_cutlass_moe_mm_data(
      topk_ids.const_data_ptr<int32_t>(), topk_ids.size(0), topk_ids.size(1), expert_offsets.data_ptr<int32_t>(),
      problem_sizes1.data_ptr<int32_t>(), problem_sizes2.data_ptr<int32_t>(), input_permutation.data_ptr<int32_t>(),
      output_permutation.data_ptr<int32_t>(), num_experts, n, k,
      blockscale_offsets.has_value() ? blockscale_offsets->data_ptr<int32_t>() : nullptr, stream_of(topk_ids)));
}
void get_cutlass_pplx_moe_mm_data(Tensor& expert_offsets, Tensor& problem_sizes1, Tensor& problem_sizes2,
                                  const Tensor& expert_num_tokens, int64_t num_local_experts, int64_t padded_m,
                                  int64_t n, int64_t k) {
  const at::cuda::OptionalCUDAGuard guard(device_of(expert_offsets));
  vb_check(vb200_get_cutlass_pplx_moe_mm_data(expert_offsets.data_ptr<int32_t>(), problem_sizes1.data_ptr<int32_t>(),
                                              problem_sizes2.data_ptr<int32_t>(),
                                              expert_num_tokens.const_data_ptr<int32_t>(), num_local_experts,
                                              padded_m, n, k, stream_of(expert_offsets)));
}
void shuffle_rows(const Tensor& input_tensor, const Tensor& dst2src_map, Tensor& output_tensor) {
  TORCH_CHECK(input_tensor.dim() == 2 && input_tensor.is_contiguous() && output_tensor.is_contiguous());
  TORCH_CHECK(dst2src_map.dtype() == at::kInt && dst2src_map.is_contiguous());
  const at::cuda::OptionalCUDAGuard guard(device_of(input_tensor));
  vb_check(vb200_shuffle_rows(output_tensor.data_ptr(), input_tensor.const_data_ptr(),
                              dst2src_map.const_data_ptr<int32_t>(), dst2src_map.size(0), input_tensor.size(0),
                              input_tensor.size(1) * input_tensor.element_size(), stream_of(input_tensor)));
}

// ---- callers either side of the path (csrc/ops.h advance_step_flashattn, merge_attn_states) ----
void advance_step_flashattn(int64_t num_seqs, int64_t num_queries, int64_t block_size, Tensor& input_tokens,
                            Tensor& sampled_token_ids, Tensor& input_positions, Tensor& seq_lens,
                            Tensor& slot_mapping, Tensor& block_tables) {
  TORCH_CHECK(input_tokens.dtype() == at::kLong && sampled_token_ids.dtype() == at::kLong &&
              input_positions.dtype() == at::kLong && slot_mapping.dtype() == at::kLong &&
              seq_lens.dtype() == at::kInt && block_tables.dtype() == at::kInt);
  TORCH_CHECK(input_tokens.size(0) == num_seqs && sampled_token_ids.size(0) == num_queries &&
              seq_lens.size(0) == num_seqs && slot_mapping.size(0) == num_seqs && block_tables.size(0) == num_seqs);
  const at::cuda::OptionalCUDAGuard guard(device_of(sampled_token_ids));
  vb_check(vb200_advance_step_flashattn(num_seqs, num_queries, block_size, input_tokens.data_ptr<int64_t>(),
                                        sampled_token_ids.const_data_ptr<int64_t>(),
                                        input_positions.data_ptr<int64_t>(), seq_lens.data_ptr<int32_t>(),
                                        slot_mapping.data_ptr<int64_t>(), block_tables.const_data_ptr<int32_t>(),
                                        block_tables.stride(0), stream_of(sampled_token_ids)));
}
void merge_attn_states(Tensor& output, OptTensor output_lse, const Tensor& prefix_output, const Tensor& prefix_lse,
                       const Tensor& suffix_output, const Tensor& suffix_lse) {
  TORCH_CHECK(output.dim() == 3 && output.is_contiguous() && prefix_output.is_contiguous() &&
              suffix_output.is_contiguous());
  const at::cuda::OptionalCUDAGuard guard(device_of(output));
  vb_check(vb200_merge_attn_states(output.data_ptr(), output_lse.has_value() ? output_lse->data_ptr<float>() : nullptr,
                                   prefix_output.const_data_ptr(), prefix_lse.const_data_ptr<float>(),
                                   suffix_output.const_data_ptr(), suffix_lse.const_data_ptr<float>(), output.size(0),
                                   output.size(1), output.size(2), dt(output), stream_of(output)));
}

}  // namespace

#define VB_CAT_(a, b) a##b
#define VB_CAT(a, b) VB_CAT_(a, b)

TORCH_LIBRARY(VB200_LIB_C, ops) {
  ops.def(
      "paged_attention_v1(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads,"
      " float scale, Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len, Tensor? alibi_slopes,"
      " str kv_cache_dtype, Tensor k_scale, Tensor v_scale, int tp_rank, int blocksparse_local_blocks,"
      " int blocksparse_vert_stride, int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def(
      "paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits, Tensor! tmp_out, Tensor query,"
      " Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens,"
      " int block_size, int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
      " int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size,"
      " int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);
  ops.def("silu_and_mul(Tensor! result, Tensor input) -> ()");
  ops.impl("silu_and_mul", torch::kCUDA, &silu_and_mul);
  ops.def("rms_norm(Tensor! result, Tensor input, Tensor weight, float epsilon) -> ()");
  ops.impl("rms_norm", torch::kCUDA, &rms_norm);
  ops.def("fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor weight, float epsilon) -> ()");
  ops.impl("fused_add_rms_norm", torch::kCUDA, &fused_add_rms_norm);
  ops.def(
      "rms_norm_dynamic_per_token_quant(Tensor! result, Tensor input, Tensor weight, Tensor! scale, float epsilon,"
      " Tensor? scale_ub, Tensor!? residual) -> ()");
  ops.impl("rms_norm_dynamic_per_token_quant", torch::kCUDA, &rms_norm_dynamic_per_token_quant);
  ops.def(
      "rotary_embedding(Tensor positions, Tensor! query, Tensor!? key, int head_size, Tensor cos_sin_cache,"
      " bool is_neox) -> ()");
  ops.impl("rotary_embedding", torch::kCUDA, &rotary_embedding);
  ops.def(
      "cutlass_scaled_mm(Tensor! out, Tensor a, Tensor b, Tensor a_scales, Tensor b_scales, Tensor? bias) -> ()");
  ops.impl("cutlass_scaled_mm", torch::kCUDA, &cutlass_scaled_mm);
  ops.def("cutlass_scaled_mm_supports_fp8(int cuda_device_capability) -> bool");
  ops.impl("cutlass_scaled_mm_supports_fp8", &cutlass_scaled_mm_supports_fp8);
  ops.def("static_scaled_fp8_quant(Tensor! result, Tensor input, Tensor scale) -> ()");
  ops.impl("static_scaled_fp8_quant", torch::kCUDA, &static_scaled_fp8_quant);
  ops.def("dynamic_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale) -> ()");
  ops.impl("dynamic_scaled_fp8_quant", torch::kCUDA, &dynamic_scaled_fp8_quant);
  ops.def("dynamic_per_token_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale, Tensor? scale_ub) -> ()");
  ops.impl("dynamic_per_token_scaled_fp8_quant", torch::kCUDA, &dynamic_per_token_scaled_fp8_quant);
  ops.def(
      "advance_step_flashattn(int num_seqs, int num_queries, int block_size, Tensor! input_tokens,"
      " Tensor sampled_token_ids, Tensor! input_positions, Tensor! seq_lens, Tensor! slot_mapping,"
      " Tensor block_tables) -> ()");
  ops.impl("advance_step_flashattn", torch::kCUDA, &advance_step_flashattn);
  ops.def(
      "merge_attn_states(Tensor! output, Tensor!? output_lse, Tensor prefix_output, Tensor prefix_lse,"
      " Tensor suffix_output, Tensor suffix_lse) -> ()");
  ops.impl("merge_attn_states", torch::kCUDA, &merge_attn_states);
  ops.def("cutlass_group_gemm_supported(int cuda_device_capability) -> bool");
  ops.impl("cutlass_group_gemm_supported", &cutlass_group_gemm_supported);
  ops.def(
      "cutlass_moe_mm(Tensor! out_tensors, Tensor a_tensors, Tensor b_tensors, Tensor a_scales, Tensor b_scales,"
      " Tensor expert_offsets, Tensor problem_sizes, Tensor a_strides, Tensor b_strides, Tensor c_strides,"
      " bool per_act_token, bool per_out_ch) -> ()");
  ops.impl("cutlass_moe_mm", torch::kCUDA, &cutlass_moe_mm);
  ops.def(
      "get_cutlass_moe_mm_data(Tensor topk_ids, Tensor! expert_offsets, Tensor! problem_sizes1,"
      " Tensor! problem_sizes2, Tensor! input_permutation, Tensor! output_permutation, int num_experts, int n,"
      " int k, Tensor? blockscale_offsets) -> ()");
  ops.impl("get_cutlass_moe_mm_data", torch::kCUDA, &get_cutlass_moe_mm_data);
  ops.def(
      "get_cutlass_pplx_moe_mm_data(Tensor! expert_offsets, Tensor! problem_sizes1, Tensor! problem_sizes2,"
      " Tensor expert_num_tokens, int num_local_experts, int padded_m, int n, int k) -> ()");
  ops.impl("get_cutlass_pplx_moe_mm_data", torch::kCUDA, &get_cutlass_pplx_moe_mm_data);
}

TORCH_LIBRARY(VB200_LIB_CACHE, ops) {
  ops.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping,"
      " str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache,"
      " Tensor slot_mapping, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  ops.def("convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()");
  ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
}

TORCH_LIBRARY(VB200_LIB_MOE, ops) {
  ops.def(
      "topk_softmax(Tensor! topk_weights, Tensor! topk_indices, Tensor! token_expert_indices,"
      " Tensor gating_output) -> ()");
  ops.impl("topk_softmax", torch::kCUDA, &topk_softmax);
  ops.def("moe_sum(Tensor input, Tensor! output) -> ()");
  ops.impl("moe_sum", torch::kCUDA, &moe_sum);
  ops.def(
      "moe_align_block_size(Tensor topk_ids, int num_experts, int block_size, Tensor! sorted_token_ids,"
      " Tensor! experts_ids, Tensor! num_tokens_post_pad) -> ()");
  ops.impl("moe_align_block_size", torch::kCUDA, &moe_align_block_size);
  ops.def("shuffle_rows(Tensor input_tensor, Tensor dst2src_map, Tensor! output_tensor) -> ()");
  ops.impl("shuffle_rows", torch::kCUDA, &shuffle_rows);
}
