// TEST INFRASTRUCTURE ONLY (oracle).  Binding shim that exposes the reference's own CUDA
// kernels (compiled for sm_100a from the sources where they lie under /root/reference/csrc
// -- nothing is copied) under the private torch library name `_ref_C`, so that the GPU
// parity tests can call THE reference op next to ours on the same inputs.
// Schemas: /root/reference/csrc/torch_bindings.cpp:45-69 (attention), :111 (silu_and_mul),
// :171-219 (norms, rope), :575-591 (fp8 quant), :676-707 (cache ops); /root/reference/csrc/moe/torch_bindings.cpp:6-22.
#include <torch/all.h>
#include <torch/library.h>
#include "ops.h"
#include "cache.h"
#include "moe/moe_ops.h"

TORCH_LIBRARY(_ref_C, m) {
  m.def(
      "paged_attention_v1(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache,"
      " int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size,"
      " int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
      " int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,"
      " int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()");
  m.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  m.def(
      "paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits, Tensor! tmp_out,"
      " Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale,"
      " Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len,"
      " Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale, int tp_rank,"
      " int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size,"
      " int blocksparse_head_sliding_step) -> ()");
  m.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);
  m.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache,"
      " Tensor slot_mapping, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  m.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  m.def(
      "reshape_and_cache_flash(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache,"
      " Tensor slot_mapping, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  m.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  m.def("convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()");
  m.impl("convert_fp8", torch::kCUDA, &convert_fp8);
  m.def("silu_and_mul(Tensor! result, Tensor input) -> ()");
  m.impl("silu_and_mul", torch::kCUDA, &silu_and_mul);
  m.def("static_scaled_fp8_quant(Tensor! result, Tensor input, Tensor scale) -> ()");
  m.impl("static_scaled_fp8_quant", torch::kCUDA, &static_scaled_fp8_quant);
  m.def("dynamic_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale) -> ()");
  m.impl("dynamic_scaled_fp8_quant", torch::kCUDA, &dynamic_scaled_fp8_quant);
  m.def(
      "dynamic_per_token_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale,"
      " Tensor? scale_ub) -> ()");
  m.impl("dynamic_per_token_scaled_fp8_quant", torch::kCUDA,
         &dynamic_per_token_scaled_fp8_quant);
  m.def("rms_norm(Tensor! result, Tensor input, Tensor weight, float epsilon) -> ()");
  m.impl("rms_norm", torch::kCUDA, &rms_norm);
  m.def("fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor weight, float epsilon) -> ()");
  m.impl("fused_add_rms_norm", torch::kCUDA, &fused_add_rms_norm);
  m.def(
      "rms_norm_dynamic_per_token_quant(Tensor! result, Tensor input, Tensor weight, Tensor! scale,"
      " float epsilon, Tensor? scale_ub, Tensor!? residual) -> ()");
  m.impl("rms_norm_dynamic_per_token_quant", torch::kCUDA, &rms_norm_dynamic_per_token_quant);
  m.def(
      "rotary_embedding(Tensor positions, Tensor! query, Tensor!? key, int head_size,"
      " Tensor cos_sin_cache, bool is_neox) -> ()");
  m.impl("rotary_embedding", torch::kCUDA, &rotary_embedding);
  m.def(
      "topk_softmax(Tensor! topk_weights, Tensor! topk_indices, Tensor! token_expert_indices,"
      " Tensor gating_output) -> ()");
  m.impl("topk_softmax", torch::kCUDA, &topk_softmax);
  m.def("moe_sum(Tensor input, Tensor! output) -> ()");
  m.impl("moe_sum", torch::kCUDA, &moe_sum);
  m.def(
      "moe_align_block_size(Tensor topk_ids, int num_experts, int block_size,"
      " Tensor! sorted_token_ids, Tensor! experts_ids, Tensor! num_tokens_post_pad) -> ()");
  m.impl("moe_align_block_size", torch::kCUDA, &moe_align_block_size);
}
