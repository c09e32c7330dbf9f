// TEST INFRASTRUCTURE ONLY (oracle).  Binding shim that exposes the reference's own
// CPU kernels (compiled from /root/reference/csrc/cpu/{attention,cache}.cpp where they
// lie -- no reference source is copied into this repo) under the private torch library
// name `_ref_cpu`, with the schema strings of /root/reference/csrc/cpu/torch_bindings.cpp:80-104
// (paged attention) and :277-283 (reshape_and_cache).  Function declarations come from the
// reference headers csrc/ops.h:32-54 and csrc/cache.h:21-25 via -I.
#include <torch/all.h>
#include <torch/library.h>
#include "ops.h"
#include "cache.h"

TORCH_LIBRARY(_ref_cpu, m) {
  m.def(
      "paged_attention_v1(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache,"
      " int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size,"
      " int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
      " int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,"
      " int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()");
  m.impl("paged_attention_v1", torch::kCPU, &paged_attention_v1);
  m.def(
      "paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits, Tensor! tmp_out,"
      " Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale,"
      " Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len,"
      " Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale, int tp_rank,"
      " int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size,"
      " int blocksparse_head_sliding_step) -> ()");
  m.impl("paged_attention_v2", torch::kCPU, &paged_attention_v2);
  m.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache,"
      " Tensor slot_mapping, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  m.impl("reshape_and_cache", torch::kCPU, &reshape_and_cache);
}
