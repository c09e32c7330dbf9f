/*
 * vllm_b200.h -- C ABI of libvllm_b200.so: the B200-native (sm_100a) implementation of vLLM's
 * decode hot path.  Plain pointers and sizes only; no torch types.  Every entry point is the
 * device-side body of ONE reference operator; the reference-side binding a maintainer would add
 * (a TORCH_LIBRARY shim that unwraps at::Tensor into these arguments) is in INTEGRATION.md, and
 * vllm_b200/torch_shim.cpp is that shim.
 *
 * Conventions (mirror the reference operator contract):
 *   - all data pointers are DEVICE pointers on the current CUDA device; the caller owns every
 *     buffer (outputs, workspaces); nothing is retained after the call returns;
 *   - `stream` is a cudaStream_t (passed as void*); kernels are launched on it, no host sync,
 *     CUDA-graph-capturable;
 *   - scales (k_scale, v_scale, a_scales, ...) are DEVICE fp32 pointers dereferenced in-kernel
 *     (reference: csrc/attention/attention_kernels.cuh:283);
 *   - return value: 0 on success, non-zero on error; vb200_last_error() gives the message of the
 *     last failure on this thread (the Python host raises RuntimeError with it, matching the
 *     reference's TORCH_CHECK -> c10::Error -> RuntimeError convention).
 *
 * All citations are relative to /root/reference.
 */
#ifndef VLLM_B200_H_
#define VLLM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* element types */
enum {
  VB200_F32 = 0,
  VB200_F16 = 1,
  VB200_BF16 = 2,
  VB200_FP8_E4M3 = 3,
  VB200_I32 = 4,
  VB200_I64 = 5,
  VB200_U32 = 6
};

/* kv_cache_dtype strings of the reference ("auto", "fp8"/"fp8_e4m3";
 * csrc/quantization/fp8/nvidia/quant_utils.cuh:532-569) */
enum { VB200_KV_AUTO = 0, VB200_KV_FP8_E4M3 = 1 };

typedef void* vb200_stream_t;

const char* vb200_last_error(void);
/* library / build identification: "vllm_b200 <ver> sm_100a" */
const char* vb200_version(void);

/* ------------------------------------------------------------------------------------------ */
/* KV-cache scatter                                                                            */
/* ------------------------------------------------------------------------------------------ */

/* reshape_and_cache: csrc/cache_kernels.cu:369-395 (kernel :210-262), schema
 * csrc/torch_bindings.cpp:676-681.
 *   key/value  [num_tokens, num_heads, head_size] of `src_dtype`, row strides key_stride /
 *              value_stride (elements);
 *   key_cache  [num_blocks, num_heads, head_size/x, block_size, x]; value_cache
 *              [num_blocks, num_heads, head_size, block_size]; contiguous; element type is
 *              src_dtype (kv_dtype AUTO) or fp8 e4m3 bytes (kv_dtype FP8_E4M3, quantised as
 *              fp8_satfinite(float(v) / *scale), quant_utils.cuh:466-479);
 *   slot_mapping int64 [num_tokens]; negative slots are skipped. */
int vb200_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                            const int64_t* slot_mapping, int num_tokens, int num_heads,
                            int head_size, int block_size, int x, int64_t key_stride,
                            int64_t value_stride, int src_dtype, int kv_dtype,
                            const float* k_scale, const float* v_scale, vb200_stream_t stream);

/* reshape_and_cache_flash: csrc/cache_kernels.cu:412-450 (kernel :264-306), schema
 * csrc/torch_bindings.cpp:685-691.  Cache [num_blocks, block_size, num_heads, head_size] addressed
 * through runtime strides (elements) so NHD and HND-permuted views both work; the key and value
 * caches share the same strides (the reference passes key_cache's).  num_tokens =
 * slot_mapping.size(0) may be smaller than key.size(0) (:421-431). */
int vb200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                  void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                  int num_heads, int head_size, int block_size,
                                  int64_t block_stride, int64_t page_stride, int64_t head_stride,
                                  int64_t key_stride, int64_t value_stride, int src_dtype,
                                  int kv_dtype, const float* k_scale, const float* v_scale,
                                  vb200_stream_t stream);

/* RoPE + KV-cache write in one launch: exactly rotary_embedding(positions, query, key, head_size,
 * cos_sin_cache, is_neox) (csrc/pos_encoding_kernels.cu:77-101,133-185) followed by reshape_and_cache(key, value, ...)
 * (csrc/cache_kernels.cu:210-262) -- same bits in query, key and both caches as the two ops run back to back.
 * query [T, num_heads*head_size], key / value [T, num_kv_heads*head_size] with the given token strides (elements),
 * heads contiguous; paged cache layout as vb200_reshape_and_cache. */
int vb200_rotary_embedding_reshape_and_cache(
    const int64_t* positions, void* query, void* key, const void* value, const void* cos_sin_cache, void* key_cache,
    void* value_cache, const int64_t* slot_mapping, int num_tokens, int num_heads, int num_kv_heads, int head_size,
    int rot_dim, int is_neox, int block_size, int x, int64_t query_stride, int64_t key_stride, int64_t value_stride,
    int src_dtype, int kv_dtype, const float* k_scale, const float* v_scale, vb200_stream_t stream);

/* convert_fp8 (test helper of the reference): csrc/cache_kernels.cu:526-571, schema
 * csrc/torch_bindings.cpp:705-707.  Flat element-wise conversion of `numel` elements:
 *   src 16/32-bit -> dst fp8:   fp8_satfinite(float(src) / scale)
 *   src fp8 -> dst 16/32-bit:   dst_t(float(half(fp8)) * scale)        (quant_utils.cuh:348-359) */
int vb200_convert_fp8(void* dst, const void* src, int64_t numel, int dst_dtype, int src_dtype,
                      float scale, vb200_stream_t stream);

/* ------------------------------------------------------------------------------------------ */
/* Paged-attention decode                                                                      */
/* ------------------------------------------------------------------------------------------ */

/* paged_attention_v2: csrc/attention/paged_attention_v2.cu:167-192, kernels
 * csrc/attention/attention_kernels.cuh:524-664, schema csrc/torch_bindings.cpp:58-68.
 *   out        [num_seqs, num_heads, head_size]  q_dtype
 *   exp_sums, max_logits  fp32 [num_seqs, num_heads, max_num_partitions]
 *   tmp_out    [num_seqs, num_heads, max_num_partitions, head_size]  q_dtype
 *   query      [num_seqs, num_heads, head_size], row stride q_stride (elements)
 *   key_cache / value_cache in the paged layout above, strides kv_block_stride / kv_head_stride
 *   block_tables int32 [num_seqs, max_num_blocks_per_seq]; seq_lens int32 [num_seqs]
 *   partition_size: tokens per split-K partition (the reference fixes 512,
 *   paged_attention_v2.cu:29); max_num_partitions = ceil(max_seq_len / partition_size).
 * Same per-partition outputs as the reference: tmp_out holds the partition's NORMALISED output,
 * exp_sums the un-normalised sum at the partition max, max_logits the partition max. */
int vb200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                             const void* query, const void* key_cache, const void* value_cache,
                             int num_seqs, int num_heads, int head_size, int num_kv_heads,
                             float scale, const int32_t* block_tables, const int32_t* seq_lens,
                             int block_size, int max_seq_len, int max_num_blocks_per_seq,
                             const float* alibi_slopes, int q_dtype, int kv_dtype,
                             const float* k_scale, const float* v_scale, int64_t q_stride,
                             int64_t kv_block_stride, int64_t kv_head_stride, int partition_size,
                             vb200_stream_t stream);

/* paged_attention_v1: csrc/attention/paged_attention_v1.cu:160-182, kernel
 * attention_kernels.cuh:494-522, schema csrc/torch_bindings.cpp:45-54.  Same math without
 * caller-visible partials.  The B200 kernel is split-K by design: when `workspace` is non-NULL it
 * is used for the partials ( vb200_paged_attention_v1_workspace_size() bytes, any alignment of
 * 16 ) and `partition_size` (0 = library heuristic) selects the split; with workspace == NULL the
 * whole sequence is processed by one CTA per (sequence, kv-head group). */
size_t vb200_paged_attention_v1_workspace_size(int num_seqs, int num_heads, int head_size,
                                               int max_seq_len, int q_dtype, int partition_size);
int vb200_paged_attention_v1_partition_size(int num_seqs, int num_heads, int num_kv_heads,
                                            int head_size, int max_seq_len, int block_size,
                                            int kv_dtype);
int vb200_paged_attention_v1(void* out, const void* query, const void* key_cache,
                             const void* value_cache, int num_seqs, int num_heads, int head_size,
                             int num_kv_heads, float scale, const int32_t* block_tables,
                             const int32_t* seq_lens, int block_size, int max_seq_len,
                             int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype,
                             int kv_dtype, const float* k_scale, const float* v_scale,
                             int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
                             int partition_size, void* workspace, size_t workspace_bytes,
                             vb200_stream_t stream);

/* paged_attention_v1 that allocates nothing and takes no workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, partition plan, kernels and results as
 * vb200_paged_attention_v1 given a workspace of vb200_paged_attention_v1_workspace_size() bytes; the split-K partials
 * live in a library-owned static device array of vb200_paged_attention_v1_static_max_bytes() = 16 MiB, resident on
 * every GPU that loads the library until the process exits.  A call whose partials need more (num_seqs * num_heads *
 * partitions * (8 + 2 * head_size) + 64 bytes) is rejected with an error; use the workspace form for it.  No
 * cudaMalloc / cudaFree or allocating library routine is called.  All calls on a device share the array and it keeps
 * the last call's contents (nothing in it has to start at zero: every slot read is written first by the same call);
 * calls that overlap on the GPU -- different streams, or a graph replay beside eager work -- race on it. */
size_t vb200_paged_attention_v1_static_max_bytes(void);
int vb200_paged_attention_v1_static(void* out, const void* query, const void* key_cache, const void* value_cache,
                                    int num_seqs, int num_heads, int head_size, int num_kv_heads, float scale,
                                    const int32_t* block_tables, const int32_t* seq_lens, int block_size,
                                    int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype,
                                    int kv_dtype, const float* k_scale, const float* v_scale, int64_t q_stride,
                                    int64_t kv_block_stride, int64_t kv_head_stride, int partition_size,
                                    vb200_stream_t stream);

/* paged_attention_v1 with the dynamic per-token fp8 quantisation of its output fused into the split-K reduce kernel
 * (SURVEY.md 8f row 1): out_q fp8 [num_seqs, num_heads*head_size] and out_scales fp32 [num_seqs] are exactly what
 * vb200_paged_attention_v1 followed by vb200_dynamic_per_token_scaled_fp8_quant (fp8/common.cu:28-83) produce on the
 * same partition plan -- the activation the FP8 o_proj consumes -- in one launch fewer and without the bf16 round trip
 * through HBM.  Needs block_size 16, head_size % 32 == 0, num_heads <= 128 and a workspace of
 * vb200_paged_attention_v1_quant_workspace_size() bytes (the partials are always materialised). */
size_t vb200_paged_attention_v1_quant_workspace_size(int num_seqs, int num_heads, int head_size, int max_seq_len,
                                                     int q_dtype, int partition_size);
int vb200_paged_attention_v1_quant(void* out_q, float* out_scales, const float* scale_ub, const void* query,
                                   const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
                                   int head_size, int num_kv_heads, float scale, const int32_t* block_tables,
                                   const int32_t* seq_lens, int block_size, int max_seq_len, int max_num_blocks_per_seq,
                                   const float* alibi_slopes, int q_dtype, int kv_dtype, const float* k_scale,
                                   const float* v_scale, int64_t q_stride, int64_t kv_block_stride,
                                   int64_t kv_head_stride, int partition_size, void* workspace, size_t workspace_bytes,
                                   vb200_stream_t stream);

/* Decode attention over the FLASH cache layout: the cache vLLM's V1 backends allocate,
 * kv_cache [2, num_blocks, block_size, num_kv_heads, head_size] (vllm/v1/attention/backends/flash_attn.py:79-102),
 * in NHD order or as the HND-permuted view Blackwell uses (vllm/platforms/cuda.py:264-290) -- the layout
 * vb200_reshape_and_cache_flash writes.  key_cache / value_cache are the two halves (kv_cache.unbind(0)); element
 * (block, token, head, d) lives at block*block_stride + token*token_stride + head*head_stride + d (elements; both
 * halves share the strides).  Same math, query/out shapes, scales, alibi and split-K workspace contract as
 * vb200_paged_attention_v1 (one query token per sequence, seq_lens = context length including it); it is the decode
 * case of the reference's flash_attn_varlen_func(q, k=key_cache, v=value_cache, block_table=..., seqused_k=...)
 * call (flash_attn.py:497-519) with max_seqlen_q == 1.  block_size must be 16; head_size * sizeof(cache element) must
 * be a multiple of 128 (f16/bf16: 64, 128, 192, 256; fp8: 128, 256); strides * element size must be multiples of 16
 * bytes.  The kernel builds two TMA tensor maps on the host per call (no allocation, graph-capturable). */
size_t vb200_paged_attention_flash_workspace_size(int num_seqs, int num_heads, int head_size, int max_seq_len,
                                                  int q_dtype, int partition_size);
int vb200_paged_attention_flash(void* out, const void* query, const void* key_cache, const void* value_cache,
                                int num_seqs, int num_heads, int head_size, int num_kv_heads, float scale,
                                const int32_t* block_tables, const int32_t* seq_lens, int block_size, int num_blocks,
                                int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype,
                                int kv_dtype, const float* k_scale, const float* v_scale, int64_t q_stride,
                                int64_t block_stride, int64_t token_stride, int64_t head_stride, int partition_size,
                                void* workspace, size_t workspace_bytes, vb200_stream_t stream);
/* The same decode with a sliding window (flash-attn's window_size = (window_left, 0) = (sliding_window - 1, 0),
 * flash_attn.py:366-369,507-510): the query at position seq_len - 1 attends keys seq_len - 1 - window_left ..
 * seq_len - 1.  Blocks left of the window are never loaded (partitions are counted from the block holding the first
 * visible key; the split-K plan is made for min(max_seq_len, window_left + 16) tokens -- pass that to the workspace /
 * partition helpers).  window_left < 0: no window.  softcap > 0: logits soft-capping, softcap * tanh(logit / softcap)
 * before ALiBi and the mask (flash_attn.py:371-374); 0: off. */
int vb200_paged_attention_flash_window(void* out, const void* query, const void* key_cache, const void* value_cache,
                                       int num_seqs, int num_heads, int head_size, int num_kv_heads, float scale,
                                       const int32_t* block_tables, const int32_t* seq_lens, int block_size,
                                       int num_blocks, int max_seq_len, int max_num_blocks_per_seq,
                                       const float* alibi_slopes, int q_dtype, int kv_dtype, const float* k_scale,
                                       const float* v_scale, int64_t q_stride, int64_t block_stride, int64_t token_stride,
                                       int64_t head_stride, int partition_size, void* workspace, size_t workspace_bytes,
                                       int window_left, float softcap, vb200_stream_t stream);

/* Variable-length causal attention over the FLASH cache layout (prefill with paged KV): the
 * general case of flash_attn_varlen_func(q, k=key_cache, v=value_cache, cu_seqlens_q=..., seqused_k=...,
 * block_table=..., causal=True) as vllm/v1/attention/backends/flash_attn.py:497-519 calls it -- prefill, chunked
 * prefill and mixed prefill / decode batches.  query / out [total_q, num_heads, head_size] (f16 / bf16; token rows
 * q_stride / out_stride elements apart), sequence s owning rows cu_seqlens_q[s] .. cu_seqlens_q[s+1] (int32
 * [num_seqs + 1]); the cache (layout and strides as vb200_paged_attention_flash; kv_dtype AUTO = the query's dtype,
 * FP8_E4M3 = bytes with per-tensor k_scale / v_scale as there; any power-of-two block_size) already holds all
 * seqused_k[s] tokens of the sequence, the new ones last; row i of a sequence with
 * q_len new tokens attends keys 0 .. seqused_k[s] - q_len + i when `causal`, all seqused_k[s] keys otherwise.
 * max_seqlen_q: the longest q_len (the reference call's argument; picks 64- or 128-row tiles, <= 0 = unknown).
 * head_size 64 / 128 / 192 / 256; strides multiples of 16 bytes; no workspace, no host read of the device arrays. */
int vb200_flash_attn_varlen_paged(void* out, const void* query, const void* key_cache, const void* value_cache,
                                  const int32_t* cu_seqlens_q, const int32_t* seqused_k, const int32_t* block_table,
                                  int num_seqs, int total_q, int num_heads, int num_kv_heads, int head_size,
                                  int block_size, int max_num_blocks_per_seq, int max_seqlen_q, float scale, int causal,
                                  int dtype, int kv_dtype, const float* k_scale, const float* v_scale,
                                  int64_t q_stride, int64_t out_stride, int64_t block_stride, int64_t token_stride,
                                  int64_t head_stride, vb200_stream_t stream);
/* The same call with the two remaining arguments the reference backend passes (flash_attn.py:342-350,507-510):
 *   window_left: sliding window, flash-attn's window_size = (window_left, 0) = (sliding_window - 1, 0): a query at
 *                absolute position p attends keys p - window_left .. p only (needs causal); < 0 = no window;
 *   softcap:     logits soft-capping, logits = softcap * tanh(logits / softcap) before the mask; 0 = off.
 * Golden: ref_paged_attn, tests/kernels/attention/test_flash_attn.py:25-77 (sliding_window :62-67, soft_cap :68-69);
 * the reference test runs sliding_window in {None, 256} and soft_cap in {None, 10, 50} (:86-88,:201-203).
 *   softmax_lse: optional fp32 [num_heads, lse_stride >= total_q]: the natural-log sum of exponentials of every
 *                query row's (scaled, capped, masked) logits -- flash_attn_varlen_func(return_softmax_lse=True), what
 *                cascade_attention (flash_attn.py:674-764) feeds to merge_attn_states; -inf for a row without a
 *                visible key; NULL = not wanted.
 *   alibi_slopes: optional fp32 [num_heads]: -slope * |key position - query position| is added to the scaled
 *                (and capped) logit -- flash-attn's -slope * |i + seqlen_k - seqlen_q - j|, which for one query token
 *                is paged_attention_v1's slope * (token - seq_len + 1) (attention_kernels.cuh:283-285); NULL = none.
 * K / V tiles that lie left of every row's window are never loaded.  Decode batches (q_len 1) are accepted. */
int vb200_flash_attn_varlen_paged_opts(void* out, const void* query, const void* key_cache, const void* value_cache,
                                       const int32_t* cu_seqlens_q, const int32_t* seqused_k,
                                       const int32_t* block_table, int num_seqs, int total_q, int num_heads,
                                       int num_kv_heads, int head_size, int block_size, int max_num_blocks_per_seq,
                                       int max_seqlen_q, float scale, int causal, int dtype, int kv_dtype,
                                       const float* k_scale, const float* v_scale, int64_t q_stride, int64_t out_stride,
                                       int64_t block_stride, int64_t token_stride, int64_t head_stride, int window_left,
                                       float softcap, float* softmax_lse, int64_t lse_stride,
                                       const float* alibi_slopes, vb200_stream_t stream);

/* ------------------------------------------------------------------------------------------ */
/* FP8 activation quantisation + activation                                                    */
/* ------------------------------------------------------------------------------------------ */

/* static_scaled_fp8_quant: csrc/quantization/fp8/common.cu:87-111 (kernel :14-26):
 * out = fp8(clamp(x * (1 / *scale), +-448)).  input [num_tokens, hidden] contiguous. */
int vb200_static_scaled_fp8_quant(void* out, const void* input, const float* scale,
                                  int64_t num_tokens, int hidden_size, int in_dtype,
                                  vb200_stream_t stream);
/* dynamic_scaled_fp8_quant: common.cu:113-141: *scale = max(*scale, amax(|x|)/448) through an
 * atomic max (scale must be pre-zeroed, vllm/_custom_ops.py:1297), then the static kernel. */
int vb200_dynamic_scaled_fp8_quant(void* out, const void* input, float* scale, int64_t num_tokens,
                                   int hidden_size, int in_dtype, vb200_stream_t stream);
/* dynamic_per_token_scaled_fp8_quant: common.cu:143-171 (kernel :28-83): per row
 * s = max(min(amax, *scale_ub) / 448, 1/(448*512)); out = fp8(clamp(x / s, +-448)); scales[row]=s */
int vb200_dynamic_per_token_scaled_fp8_quant(void* out, const void* input, float* scales,
                                             const float* scale_ub, int64_t num_tokens,
                                             int hidden_size, int in_dtype, vb200_stream_t stream);
/* silu_and_mul: csrc/activation_kernels.cu:129-136 (kernel :26-39):
 * out[t,i] = T(silu_f32(x[t,i])) *_T x[t,d+i]; input [num_tokens, 2d], out [num_tokens, d]. */
int vb200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d, int dtype,
                       vb200_stream_t stream);
/* The two halves of the dynamic per-tensor quantisation around the MoE activation (fused_moe.py:1506-1527: activation,
 * then moe_kernel_quantize_input -> scaled_fp8_quant(x, None), fp8/common.cu:113-137) for a pipeline that quantises
 * while it gathers (vb200_moe_grouped_gemm_fp8_q16): vb200_absmax_scale is the first pass of
 * vb200_dynamic_scaled_fp8_quant alone (*scale = max(*scale, max|x| / 448), scale pre-zeroed by the caller);
 * vb200_silu_and_mul_absmax is vb200_silu_and_mul (f16 / bf16, d % 8 == 0) that leaves the same quantity of its
 * OUTPUT (as rounded to the tensor dtype) in *scale. */
int vb200_absmax_scale(float* scale, const void* input, int64_t numel, int in_dtype, vb200_stream_t stream);
int vb200_silu_and_mul_absmax(void* out, float* scale, const void* input, int64_t num_tokens, int d, int dtype,
                              vb200_stream_t stream);

/* ------------------------------------------------------------------------------------------ */
/* Per-token glue between the contractions                         */
/* ------------------------------------------------------------------------------------------ */

/* rms_norm: csrc/layernorm_kernels.cu:16-100, schema csrc/torch_bindings.cpp:171-173.
 * out[t,:] = T(T(x[t,:] * rsqrt(mean(x^2) + eps)) * weight); input rows input_stride apart. */
int vb200_rms_norm(void* out, const void* input, const void* weight, float epsilon,
                   int num_tokens, int hidden_size, int64_t input_stride, int dtype,
                   vb200_stream_t stream);
/* fused_add_rms_norm: csrc/layernorm_kernels.cu:102-200, schema torch_bindings.cpp:177-179.
 * In place: residual = T(input + residual); input = rms_norm(residual). */
int vb200_fused_add_rms_norm(void* input, void* residual, const void* weight, float epsilon,
                             int num_tokens, int hidden_size, int64_t input_stride, int dtype,
                             vb200_stream_t stream);
/* rms_norm_dynamic_per_token_quant: csrc/quantization/fused_kernels/
 * fused_layernorm_dynamic_per_token_quant.cu:9-43,100-170, schema torch_bindings.cpp:207-211.
 * z = input (+ residual, fp32 sum; residual <- T(z) when given); y = T(T(z*inv_rms)*weight);
 * scales[t] = max(min(amax|y|, *scale_ub)/448, 1/(448*512)); out = fp8(clamp(y / scales[t])). */
int vb200_rms_norm_dynamic_per_token_quant(void* out, const void* input, const void* weight,
                                           float* scales, float epsilon, const float* scale_ub,
                                           void* residual, int num_tokens, int hidden_size,
                                           int dtype, vb200_stream_t stream);
/* rotary_embedding: csrc/pos_encoding_kernels.cu:10-98,130-178, schema torch_bindings.cpp:216-219.
 * In place on query [num_tokens, num_heads, head_size] and key (may be NULL); positions int64;
 * cos_sin_cache [max_position, rot_dim] = cos | sin halves; strides in elements. */
int vb200_rotary_embedding(const int64_t* positions, void* query, void* key, int num_tokens,
                           int num_heads, int num_kv_heads, int head_size, int rot_dim,
                           const void* cos_sin_cache, int is_neox, int64_t query_stride,
                           int64_t key_stride, int64_t head_stride, int dtype,
                           vb200_stream_t stream);
/* silu_and_mul followed by dynamic_per_token_scaled_fp8_quant in one pass (bit-identical to the
 * two reference ops run back to back: csrc/activation_kernels.cu:26-63 then
 * csrc/quantization/fp8/common.cu:28-83).  input [num_tokens, 2d] -> out fp8 [num_tokens, d]. */
int vb200_silu_and_mul_dynamic_per_token_quant(void* out, float* scales, const void* input,
                                               const float* scale_ub, int num_tokens, int d,
                                               int dtype, vb200_stream_t stream);

/* ------------------------------------------------------------------------------------------ */
/* FP8 w8a8 linear                                                                             */
/* ------------------------------------------------------------------------------------------ */

/* cutlass_scaled_mm (fp8 e4m3 path): csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:169-240,
 * epilogue csrc/cutlass_extensions/epilogue/scaled_mm_epilogues_c3x.hpp:137-216, schema
 * csrc/torch_bindings.cpp:422-426.
 *   out[m,n] = OutT( a_scales[m|0] * ( b_scales[n|0] * sum_k a[m,k]*b[k,n] ) (+ bias[n]) )
 *   a  fp8 [M,K] row-major, row stride lda (elements);  b fp8 [K,N] COLUMN-major, i.e. the
 *   [N,K] weight, row stride ldb;  out [M,N] of out_dtype (bf16/f16), row stride ldc;
 *   a_scales fp32, numel 1 or M;  b_scales fp32, numel 1 or N;  bias NULL or [N] of out_dtype.
 * Requires K % 16 == 0 and N % 16 == 0 (vllm/_custom_ops.py:732-736).
 * The tcgen05/TMA kernel builds its TMA descriptors on the host per call (driver entry point
 * resolved at load time; no allocation, graph-capturable).  Split-K problems reduce their partial sums through
 * distributed shared memory inside a thread-block cluster: the kernel needs NO scratch.  `workspace` /
 * `workspace_bytes` are kept in the signature for ABI stability and ignored (pass NULL, 0);
 * vb200_cutlass_scaled_mm_workspace_size() returns 0. */
size_t vb200_cutlass_scaled_mm_workspace_size(int M, int N, int K);
int vb200_cutlass_scaled_mm(void* out, const void* a, const void* b, const float* a_scales,
                            const float* b_scales, const void* bias, int M, int N, int K,
                            int64_t lda, int64_t ldb, int64_t ldc, int a_scales_numel,
                            int b_scales_numel, int out_dtype, void* workspace,
                            size_t workspace_bytes, vb200_stream_t stream);
/* cutlass_scaled_mm_supports_fp8(cuda_device_capability): torch_bindings.cpp:441 */
int vb200_cutlass_scaled_mm_supports_fp8(int cuda_device_capability);

/* ------------------------------------------------------------------------------------------ */
/* Fused MoE                                                                                   */
/* ------------------------------------------------------------------------------------------ */

/* topk_softmax: csrc/moe/topk_softmax_kernels.cu:499-556, schema csrc/moe/torch_bindings.cpp:6-8.
 * gating_output fp32 [num_tokens, num_experts]; topk_weights fp32 [num_tokens, topk];
 * topk_indices [num_tokens, topk] of indices_dtype (I32 / U32 / I64); token_expert_indices int32
 * [num_tokens, topk] = k_idx * num_tokens + row (:373).  Lowest-index tie-break (:336-358). */
int vb200_topk_softmax(float* topk_weights, void* topk_indices, int32_t* token_expert_indices,
                       const float* gating_output, int num_tokens, int num_experts, int topk,
                       int indices_dtype, vb200_stream_t stream);

/* fused_topk (vllm/model_executor/layers/fused_moe/fused_moe.py:882-915) in ONE launch: the reference casts the router
 * logits to fp32 on the host, calls topk_softmax and renormalises with two more torch ops (:869-870, :908); here the
 * kernel reads the logits in their own dtype (F32 / F16 / BF16; widening is exact), and divides the k winners by their
 * sum (accumulated in k order) when `renormalize` is set.  topk_indices / token_expert_indices int32. */
int vb200_fused_topk(float* topk_weights, int32_t* topk_indices, int32_t* token_expert_indices,
                     const void* gating_output, int num_tokens, int num_experts, int topk, int renormalize,
                     int in_dtype, vb200_stream_t stream);

/* moe_align_block_size: csrc/moe/moe_align_sum_kernels.cu:204-275, schema
 * csrc/moe/torch_bindings.cpp:18-22.  topk_ids int32 [numel]; sorted_token_ids int32
 * [numel + num_experts*(block_size-1)] (caller may pre-fill; this call writes every entry:
 * padding = numel); expert_ids int32 [ceil(max_padded/block_size)] (trailing blocks 0);
 * num_tokens_post_pad int32 [1].  Order inside an expert is the STABLE order (ascending flat
 * index) -- deterministic, equal to the reference test's golden.
 * workspace: int32 scratch of vb200_moe_align_workspace_size() bytes. */
size_t vb200_moe_align_workspace_size(int numel, int num_experts);
int vb200_moe_align_block_size(const int32_t* topk_ids, int numel, int num_experts, int block_size,
                               int32_t* sorted_token_ids, int sorted_capacity, int32_t* expert_ids,
                               int expert_ids_capacity, int32_t* num_tokens_post_pad,
                               void* workspace, size_t workspace_bytes, vb200_stream_t stream);

/* The router of a MoE layer in one launch: logits = hidden @ w_gate^T (the replicated, unquantised gate Linear,
 * `router_logits, _ = self.gate(hidden_states)`, vllm/model_executor/models/mixtral.py:105-112), rounded to the
 * activation dtype as a GEMM output is, then fused_topk (fp32 softmax, top-k with lowest-index ties, optional
 * renormalisation: fused_moe.py:866-915; identical to vb200_fused_topk on those logits).  hidden [num_tokens,
 * hidden_size] and w_gate [num_experts, hidden_size] of `dtype` (F16 / BF16), contiguous, 16-byte aligned; num_experts
 * <= 32; hidden_size % 8 == 0, <= 8192.  topk_weights fp32 / topk_indices int32 [num_tokens, topk];
 * token_expert_indices int32 [num_tokens, topk] or NULL; logits_out [num_tokens, num_experts] of `dtype` or NULL;
 * absmax_scale: NULL, or a pre-zeroed fp32 scalar that receives max|hidden| / 448 (atomic max) -- the dynamic
 * per-tensor fp8 scale of the experts' first GEMM (what vb200_absmax_scale computes), for free while the row is in
 * registers. */
int vb200_moe_router_topk(float* topk_weights, int32_t* topk_indices, int32_t* token_expert_indices, void* logits_out,
                          float* absmax_scale, const void* hidden, const void* w_gate, int num_tokens, int hidden_size,
                          int num_experts, int topk, int renormalize, int dtype, vb200_stream_t stream);

/* moe_sum: csrc/moe/moe_align_sum_kernels.cu:277-318 (kernel :109-123):
 * out[t,:] = sum_j in[t,j,:], accumulated in the tensor dtype in j order. */
int vb200_moe_sum(void* out, const void* input, int num_tokens, int topk, int hidden, int dtype,
                  vb200_stream_t stream);

/* Grouped FP8 GEMM of fused_experts (the work of invoke_fused_moe_kernel,
 * vllm/model_executor/layers/fused_moe/fused_moe.py:478-647, Triton kernel :267-475), driven by
 * moe_align_block_size's outputs computed with block_size == block_m:
 *   for M-block p: expert e = expert_ids[p] (e < 0: the block's rows of C are zeroed, :370-377);
 *   rows = sorted_token_ids[p*block_m : (p+1)*block_m] (entries >= num_valid_tokens are padding);
 *   A row = token_id / top_k;
 *   C[token_id, n] = OutT( ((sum_k A[row,k] * B[e,n,k]) * [topk_weights[token_id]]) * a_scale * b_scale )
 *   (routed weight multiplied BEFORE the scales, in fp32: fused_moe.py:455-468).
 *   a  fp8 [num_a_rows, K] contiguous;  b fp8 [E, N, K] contiguous;  c [num_valid_tokens, N] out_dtype.
 *   a_scales: fp32, 1 (per-tensor) or num_a_rows (per-token, a_scales_per_token);
 *   b_scales: fp32 [E] (per-tensor per expert) or [E, N] (per-channel, b_scales_per_channel).
 *   block_m: multiple of 16 in [16, 256]; K % 16 == 0.
 *   workspace: vb200_moe_grouped_gemm_workspace_size(sorted_capacity, K) bytes of device scratch
 *   (the activation rows gathered in sorted order); no initialisation needed. */
size_t vb200_moe_grouped_gemm_workspace_size(int sorted_capacity, int K);
int vb200_moe_grouped_gemm_fp8(void* c, const void* a, const void* b, const float* a_scales,
                               const float* b_scales, const float* topk_weights,
                               const int32_t* sorted_token_ids, const int32_t* expert_ids,
                               const int32_t* num_tokens_post_pad, int num_valid_tokens,
                               int num_a_rows, int top_k, int E, int N, int K, int block_m,
                               int sorted_capacity, int a_scales_per_token,
                               int b_scales_per_channel, int mul_routed_weight, int out_dtype,
                               void* workspace, size_t workspace_bytes, vb200_stream_t stream);

/* vb200_moe_grouped_gemm_fp8 fed with the UNQUANTISED activations a [num_a_rows, K] (a_dtype F16 / BF16) and their
 * per-tensor scale a_scale (device scalar): rows are quantised (fp8(x * (1 / scale)), fp8/common.cu:14-26) while they
 * are gathered into expert order -- one launch instead of static_scaled_fp8_quant + gather.  Bit-identical to
 * vb200_static_scaled_fp8_quant followed by vb200_moe_grouped_gemm_fp8 with a per-tensor a_scales. */
int vb200_moe_grouped_gemm_fp8_q16(void* c, const void* a, int a_dtype, const void* b, const float* a_scale,
                                   const float* b_scales, const float* topk_weights, const int32_t* sorted_token_ids,
                                   const int32_t* expert_ids, const int32_t* num_tokens_post_pad, int num_valid_tokens,
                                   int num_a_rows, int top_k, int E, int N, int K, int block_m, int sorted_capacity,
                                   int b_scales_per_channel, int mul_routed_weight, int out_dtype, void* workspace,
                                   size_t workspace_bytes, vb200_stream_t stream);

/* Unquantised (f16 / bf16) experts: the same grouped GEMM for fused_experts with use_fp8_w8a8 = False
 * (fused_moe.py:1349-1563; the bf16-weights case of tests/kernels/moe/test_moe.py:117-233):
 *   C[token_id, n] = OutT( (sum_k A[token_id / top_k, k] * B[e, n, k]) * [topk_weights[token_id]] ), fp32 accumulate.
 * a [num_a_rows, K], b [E, N, K], c [num_valid_tokens, N] all of `dtype` (F16 / BF16), contiguous; K % 8 == 0.
 * workspace: vb200_moe_grouped_gemm_workspace_size(sorted_capacity, 2 * K) bytes. */
int vb200_moe_grouped_gemm_16(void* c, const void* a, const void* b, const float* topk_weights,
                              const int32_t* sorted_token_ids, const int32_t* expert_ids,
                              const int32_t* num_tokens_post_pad, int num_valid_tokens, int num_a_rows, int top_k,
                              int E, int N, int K, int block_m, int sorted_capacity, int mul_routed_weight, int dtype,
                              void* workspace, size_t workspace_bytes, vb200_stream_t stream);

/* ---- CUTLASS-style fused-MoE API ----------------------------------------- */

/* get_cutlass_moe_mm_data: csrc/quantization/cutlass_w8a8/moe/moe_data.cu:103-159, schema
 * csrc/torch_bindings.cpp:465-471.  topk_ids int32 [num_tokens, topk] (-1 = expert not on this rank);
 * expert_offsets int32 [num_experts + 1]; problem_sizes1/2 int32 [num_experts, 3] = (m_e, 2n, k) / (m_e, k, n),
 * first two entries swapped when num_tokens*topk <= 64 and blockscale_offsets == NULL (:9,121-122);
 * input_permutation int32 [num_tokens*topk]: sorted row -> source token (flat index / topk);
 * output_permutation int32 [num_tokens*topk]: flat index -> sorted row (ids == -1 -> total routed rows);
 * blockscale_offsets: NULL or int32 [num_experts + 1], rows rounded up to 128 per expert (:61-79).
 * Order inside an expert is the stable one (the reference's is scheduling-dependent). */
int vb200_get_cutlass_moe_mm_data(const int32_t* topk_ids, int num_tokens, int topk,
                                  int32_t* expert_offsets, int32_t* problem_sizes1,
                                  int32_t* problem_sizes2, int32_t* input_permutation,
                                  int32_t* output_permutation, int num_experts, int n, int k,
                                  int32_t* blockscale_offsets, vb200_stream_t stream);

/* get_cutlass_pplx_moe_mm_data: moe_data.cu:161-195 (batched-experts format: expert e starts at e*padded_m). */
int vb200_get_cutlass_pplx_moe_mm_data(int32_t* expert_offsets, int32_t* problem_sizes1,
                                       int32_t* problem_sizes2, const int32_t* expert_num_tokens,
                                       int num_local_experts, int padded_m, int n, int k,
                                       vb200_stream_t stream);

/* shuffle_rows: vllm/_custom_ops.py:912-922.  out[r, :] = in[dst2src_map[r], :]; rows are row_bytes bytes. */
int vb200_shuffle_rows(void* out, const void* in, const int32_t* dst2src_map, int num_dst_rows,
                       int num_src_rows, int64_t row_bytes, vb200_stream_t stream);

/* cutlass_moe_mm: csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:242-270, schema
 * csrc/torch_bindings.cpp:450-456.  For every expert e with m_e rows (problem_sizes[e] = (m_e, N, K),
 * or (N, m_e, K) as get_cutlass_moe_mm_data writes it for <= 64 routed rows; the kernel accepts either):
 *   out[off_e : off_e + m_e, :] = OutT( (A[off_e : off_e + m_e, :] @ B[e]^T) * a_scale * b_scale ),
 *   off_e = expert_offsets[e] (int32 [E]).
 * a fp8 [num_rows, K] contiguous; b fp8 [E, N, K] contiguous (the reference's column-major B per expert, stride K);
 * out [num_rows, N] bf16/f16 contiguous; a_scales fp32 [num_rows] (per_act_token) or [1]; b_scales fp32 [E, N]
 * (per_out_ch) or [E].  Rows owned by no expert are left untouched.  The a/b/c stride tensors of the reference
 * schema must describe these contiguous layouts (its only caller passes K, K, N: cutlass_moe.py:153-168).
 */
int vb200_cutlass_moe_mm(void* out, const void* a, const void* b, const float* a_scales,
                         const float* b_scales, const int32_t* expert_offsets,
                         const int32_t* problem_sizes, int num_rows, int E, int N, int K,
                         int per_act_token, int per_out_ch, int out_dtype, vb200_stream_t stream);

/* The two ends of the permuted MoE pipeline as single launches (B200 path of cutlass_moe_fp8; the reference runs
 * scaled_fp8_quant + shuffle_rows (+ shuffle_rows of the scales) before GEMM 1 and shuffle_rows + mul_ + moe_sum after
 * GEMM 2: fused_moe/cutlass_moe.py:142-150,205-214, topk_weight_and_reduce.py:60-100).
 * moe_gather_quant_fp8: for sorted row r < *num_routed (all rows when num_routed == NULL):
 *   out[r, :] = fp8(input[src_row[r], :] quantised per row) -- dynamic per-token (static_scale == NULL; scales[r]
 *   receives the row's scale, fp8/common.cu:28-83 arithmetic) or static per-tensor (*static_scale, common.cu:14-26).
 *   Same bytes as quantising `input` first and gathering rows afterwards.  input bf16/f16 [num_src_rows, hidden],
 *   hidden % 8 == 0. */
int vb200_moe_gather_quant_fp8(void* out, float* scales, const void* input, const int32_t* src_row,
                               const int32_t* num_routed, const float* static_scale, const float* scale_ub,
                               int num_dst_rows, int num_src_rows, int hidden, int in_dtype, vb200_stream_t stream);
/* moe_unpermute_reduce: out[t, :] = sum_j T(float(c[dst_row[t*topk + j], :]) * topk_weights[t, j]); rows with
 * dst_row >= *num_routed (ids no local expert owns, moe_data.cu:92-99) contribute zero; topk_weights == NULL skips
 * the multiply.  Rounding points = `fused_out.mul_(topk_weights)` followed by moe_sum (accumulation in T in j order
 * for topk 2..4, fp32 otherwise: moe_align_sum_kernels.cu:109-123,277-318).  c [*, hidden], out [num_tokens, hidden]
 * bf16/f16, hidden % 8 == 0. */
int vb200_moe_unpermute_reduce(void* out, const void* c, const int32_t* dst_row, const float* topk_weights,
                               const int32_t* num_routed, int num_tokens, int topk, int hidden, int dtype,
                               vb200_stream_t stream);

/* ---- callers either side of the path --------------------------------- */

/* advance_step_flashattn: csrc/prepare_inputs/advance_step.cu:176-232 (kernel :13-65), schema
 * csrc/torch_bindings.cpp:150-155.  For i < num_queries: input_tokens[i] = sampled_token_ids[i];
 * seq_lens[i] += 1; input_positions[i] = seq_lens[i] - 1; slot_mapping[i] = block_tables[i][pos / block_size] *
 * block_size + pos % block_size.  Rows num_queries..num_seqs-1 (CUDA-graph padding): tokens/positions 0, slot -1.
 * input_tokens / input_positions / slot_mapping / sampled_token_ids int64; seq_lens / block_tables int32;
 * block_tables_stride in elements. */
int vb200_advance_step_flashattn(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                                 const int64_t* sampled_token_ids, int64_t* input_positions,
                                 int32_t* seq_lens, int64_t* slot_mapping, const int32_t* block_tables,
                                 int64_t block_tables_stride, vb200_stream_t stream);

/* greedy_argmax: the greedy branch of vllm/v1/sample/sampler.py:87-96 (`logits.argmax(dim=-1)`), additive (the
 * reference runs torch's reduction on an fp32 copy of the logits).  logits [num_rows, vocab] of `dtype`
 * (F32/F16/BF16), rows `row_stride` elements apart; out_idx int64 [num_rows]; out_val fp32 [num_rows] or NULL (the
 * row maximum, for the vocab-parallel merge).  First index wins a tie; NaN counts as the largest value. */
int vb200_greedy_argmax(int64_t* out_idx, float* out_val, const void* logits, int num_rows, int vocab,
                        int64_t row_stride, int dtype, vb200_stream_t stream);

/* merge_attn_states: csrc/attention/merge_attn_states.cu:121-178 (kernel :12-89), schema
 * csrc/torch_bindings.cpp:75-83.  output/prefix_output/suffix_output [num_tokens, num_heads, head_size] of
 * `dtype` (F32/F16/BF16), contiguous; *_lse fp32 [num_heads, num_tokens]; output_lse may be NULL.
 * out = p * e^(p_lse-m)/(e^(p_lse-m)+e^(s_lse-m)) + s * (...), infinities in an lse count as -inf. */
int vb200_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                            const float* prefix_lse, const void* suffix_output, const float* suffix_lse,
                            int num_tokens, int num_heads, int head_size, int dtype, vb200_stream_t stream);

/* ---- row-parallel all-reduce over NVLink peer memory, fused with its consumer ---- */

/* The reference's analogues: csrc/custom_all_reduce.cuh:297-314 (one-shot P2P all-reduce, flag barriers :184-260)
 * and vllm/compilation/collective_fusion.py (all-reduce + residual add + RMSNorm fusion).
 * peer_data / peer_signals are HOST arrays of `world` device pointers: this process's mappings (CUDA IPC) of every
 * rank's partial-sum buffer ([rows, hidden] contiguous, 16-byte aligned) and signal block
 * (vb200_peer_allreduce_signal_bytes() bytes, zero-initialised once); entry `rank` is this rank's own.
 *   mode 0: out = T(sum_r data_r)                                  (out: T [rows, hidden])
 *   mode 1: + fused_add_rms_norm: residual = T(sum + residual); out = rmsnorm(residual) * weight
 *   mode 2: + rms_norm_dynamic_per_token_quant: residual = T(sum + residual); out = fp8 [rows, hidden],
 *           scales[rows] fp32 (scale_ub optional)                  (layernorm_utils.cuh:19-125 arithmetic)
 * The sum is taken in rank order in fp32 on every rank (bitwise identical everywhere).  world in {2, 4, 8}; every
 * rank must issue the same sequence of calls with the same shapes (the barrier is per CTA index). */
size_t vb200_peer_allreduce_signal_bytes(void);
/* Peer-visible device buffers (csrc/custom_all_reduce.cu allocate_shared_buffer_and_handle / open_mem_handle):
 * alloc = cudaMalloc + zero fill + cudaIpcGetMemHandle (64-byte handle to send to the peers); open =
 * cudaIpcOpenMemHandle with lazy peer access on the CALLER's current device; close / free undo them. */
int vb200_peer_buffer_alloc(size_t bytes, void** ptr, void* handle64);
int vb200_peer_buffer_open(const void* handle64, void** ptr);
int vb200_peer_buffer_close(void* ptr);
int vb200_peer_buffer_free(void* ptr);
int vb200_peer_allreduce(int mode, void* out, const void* const* peer_data, void* const* peer_signals,
                         int rank, int world, int rows, int hidden, int dtype, void* residual,
                         const void* weight, float epsilon, float* scales, const float* scale_ub,
                         vb200_stream_t stream);
/* Two-shot variant (reference: cross_device_reduce_2stage, csrc/custom_all_reduce.cuh:320-360, chosen above
 * 256 KB at 8 ranks / 512 KB at 4 ranks, :560-572): rank r reduces only its shard of rows into its GATHER buffer
 * (peer_gather: HOST array of `world` device pointers, each at least rows*hidden elements, 16-byte aligned,
 * peer-visible like peer_data), then every rank reads each row once from its owner and applies the same fused
 * consumer.  Per GPU 2(N-1)/N of the tensor crosses NVLink instead of (N-1).  Same results bit for bit as
 * vb200_peer_allreduce (the sum of a row is taken once, in rank order, by its owner). */
int vb200_peer_allreduce_2shot(int mode, void* out, const void* const* peer_data, void* const* peer_gather,
                               void* const* peer_signals, int rank, int world, int rows, int hidden, int dtype,
                               void* residual, const void* weight, float epsilon, float* scales,
                               const float* scale_ub, vb200_stream_t stream);

/* Split variant: reduce-scatter fused with the consumer (each rank reduces, norms and quantises ONLY its rows) followed
 * by an all-gather of the finished rows, as two launches -- the single-kernel two-shot pays a fenced grid-wide barrier in
 * the middle that costs more than the NVLink time it saves (profiles/r02_multi_gpu_tp8_s44.txt); here that barrier is a
 * kernel boundary and both kernels start with the cheap unfenced flag barrier.  peer_gather: HOST array of `world`
 * device pointers, each at least 2 * rows * hidden * sizeof(T) + 4 * rows bytes, peer-visible.  Results (out, residual,
 * scales on every rank) are bit-identical to vb200_peer_allreduce. */
int vb200_peer_allreduce_split(int mode, void* out, const void* const* peer_data, void* const* peer_gather,
                               void* const* peer_signals, int rank, int world, int rows, int hidden, int dtype,
                               void* residual, const void* weight, float epsilon, float* scales,
                               const float* scale_ub, vb200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* VLLM_B200_H_ */
