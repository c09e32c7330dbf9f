"""Decode over the V1 ("flash") KV-cache layout on B200.

Mirrors the pieces of /root/reference/vllm/v1/attention/backends/flash_attn.py a decode step touches:
  * `B200FlashAttentionBackend`      FlashAttentionBackend (:38-101): cache shape `(2, num_blocks, block_size,
                                     num_kv_heads, head_size)` and the NHD / HND stride order
                                     (`get_kv_cache_layout`, vllm/v1/attention/backends/utils.py:135-151; HND is what
                                     vllm/platforms/cuda.py:264-290 selects on Blackwell -- the default here);
  * `B200FlashAttentionMetadata`     FlashAttentionMetadata (:103-133), the fields the decode path reads;
  * `B200FlashAttentionImpl.forward` FlashAttentionImpl.forward (:372-520): reshape_and_cache_flash, then attention
                                     over the paged cache -- the reference calls `flash_attn_varlen_func(q,
                                     k=key_cache, v=value_cache, block_table=..., seqused_k=...)`;
  * `flash_attn_varlen_func`         that call's signature for the decode case (max_seqlen_q == 1), on this package's
                                     vb200_paged_attention_flash kernel (TMA tensor-map loads of the cache as it lies,
                                     csrc/paged_attention_flash.cuh).
Prefill / chunked prefill / mixed batches (max_seqlen_q > 1) run on the variable-length kernel of
csrc/paged_prefill.cu (16-bit and fp8 caches), with a sliding window (`window_size = (left, 0)`: tiles left of the
window are skipped) and logits soft-capping (in front of the mask); pure decode batches keep the decode kernel with
both (blocks left of the window are never read).  Cascade attention (`cascade_attention`, flash_attn.py:674-764: shared prefix non-causal + per-query
suffix, both returning the log-sum-exp, merged by merge_attn_states) runs on the same kernel.  fp8 caches: the reference also quantises the QUERY to fp8 (FA3,
:478-486); this kernel keeps the query in
bf16/f16 and folds the per-tensor k / v descales into the fp32 logits / output, which is the dequantised result the
parity bar is stated against.
"""
from dataclasses import dataclass
from typing import Optional

import torch

from .. import _custom_ops as ops

_KV_CACHE_LAYOUT = "HND"   # vllm/platforms/cuda.py:281-290 sets HND on SM 10.0


def set_kv_cache_layout(cache_layout: str) -> None:
    """vllm/v1/attention/backends/utils.py:149-151."""
    global _KV_CACHE_LAYOUT
    if cache_layout not in ("NHD", "HND"):
        raise ValueError(f"Unknown cache layout format {cache_layout}.")
    _KV_CACHE_LAYOUT = cache_layout


def get_kv_cache_layout() -> str:
    return _KV_CACHE_LAYOUT


def flash_attn_varlen_func(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, max_seqlen_q: int,
                           cu_seqlens_q: torch.Tensor, max_seqlen_k: int, seqused_k: torch.Tensor,
                           softmax_scale: Optional[float] = None, causal: bool = True, window_size=None,
                           softcap: float = 0.0, alibi_slopes: Optional[torch.Tensor] = None,
                           block_table: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                           q_descale=None, k_descale: Optional[torch.Tensor] = None,
                           v_descale: Optional[torch.Tensor] = None, scheduler_metadata=None, fa_version: int = 3,
                           num_splits: int = 0, cu_seqlens_k=None, return_softmax_lse: bool = False):
    """vllm.vllm_flash_attn.flash_attn_varlen_func as flash_attn.py:497-519 calls it.  Decode (max_seqlen_q == 1): q
    [num_seqs, num_heads, head_size] (one token per sequence, cu_seqlens_q = 0..num_seqs) on the HBM-roofline decode
    kernel; anything else (prefill, chunked prefill, mixed batches: q [total_q, ...] split by cu_seqlens_q, bottom-right
    aligned causal mask against seqused_k) on the variable-length tensor-core kernel.  k / v the two halves of the
    paged cache `[num_blocks, block_size, num_kv_heads, head_size]` (any stride order with a contiguous head
    dimension), block_table int32 [num_seqs, max_blocks], seqused_k int32 [num_seqs].  k_descale / v_descale: fp32
    per-tensor scales of an fp8 cache (the reference passes them expanded to [num_seqs, num_kv_heads]; element 0 is
    used).  return_softmax_lse: also return the fp32 [num_heads, total_q] log-sum-exp (variable-length kernel)."""
    if block_table is None:
        raise NotImplementedError("flash_attn_varlen_func: only the paged-KV (block_table) form is built")
    softcap = float(softcap or 0.0)
    window_left = -1
    if window_size is not None and tuple(window_size) not in ((-1, -1), (-1, 0)):
        # (left, 0) is what the reference backend passes: (sliding_window - 1, 0), flash_attn.py:366-369; a causal call
        # ignores the right side (keys right of the diagonal are masked anyway)
        left, right = (int(x) for x in window_size)
        if not causal or left < 0:
            raise NotImplementedError("flash_attn_varlen_func: only causal (left, 0) sliding windows are built")
        window_left = left
    if cu_seqlens_q is not None and cu_seqlens_q.numel() != seqused_k.shape[0] + 1:
        raise RuntimeError("cu_seqlens_q must hold num_seqs + 1 offsets")
    fp8 = k.element_size() == 1
    if fp8 and (k_descale is None or v_descale is None):
        raise RuntimeError("an fp8 cache needs k_descale and v_descale")
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** -0.5
    if out is None:
        out = torch.empty_like(q)
    decode = max_seqlen_q == 1 and q.shape[0] == seqused_k.shape[0]
    head_bytes = q.shape[-1] * k.element_size()
    # a decode batch -- window and soft-cap included -- stays on the HBM-roofline decode kernel when that kernel takes
    # the shape
    decode_kernel_ok = decode and not return_softmax_lse and k.shape[1] == 16 and head_bytes % 128 == 0
    if not decode_kernel_ok:
        # prefill / chunked prefill / mixed batch, or an LSE output: the variable-length kernel (csrc/paged_prefill.cu)
        if cu_seqlens_q is None:
            if max_seqlen_q != 1:
                raise RuntimeError("flash_attn_varlen_func: cu_seqlens_q is required")
            cu_seqlens_q = torch.arange(q.shape[0] + 1, dtype=torch.int32, device=q.device)
        ks = k_descale.reshape(-1)[:1].float().contiguous() if fp8 else None
        vs = v_descale.reshape(-1)[:1].float().contiguous() if fp8 else None
        lse = torch.empty((q.shape[1], q.shape[0]), dtype=torch.float32, device=q.device) if return_softmax_lse else None
        ops.flash_attn_varlen_paged(out, q, k, v, cu_seqlens_q, seqused_k, block_table, softmax_scale, causal,
                                    int(max_seqlen_q), "fp8" if fp8 else "auto", ks, vs, window_left, softcap, lse,
                                    alibi_slopes)
        return (out, lse) if return_softmax_lse else out
    ks = k_descale.reshape(-1)[:1].float().contiguous() if k_descale is not None else None
    vs = v_descale.reshape(-1)[:1].float().contiguous() if v_descale is not None else None
    ops.paged_attention_flash(out, q, k, v, softmax_scale, block_table, seqused_k, int(max_seqlen_k), alibi_slopes,
                              "fp8" if fp8 else "auto", ks, vs, window_left=window_left, softcap=softcap)
    return out


@dataclass
class B200FlashAttentionMetadata:
    """FlashAttentionMetadata (flash_attn.py:103-133): the fields a decode step reads."""
    num_actual_tokens: int
    max_query_len: int
    query_start_loc: torch.Tensor     # int32 [num_seqs + 1]
    max_seq_len: int
    seq_lens: torch.Tensor            # int32 [num_seqs]
    block_table: torch.Tensor         # int32 [num_seqs, max_blocks]
    slot_mapping: torch.Tensor        # int64 [num_tokens]
    use_cascade: bool = False
    common_prefix_len: int = 0
    cu_prefix_query_lens: Optional[torch.Tensor] = None   # int32 [2] = [0, num_actual_tokens]   (:121-123)
    prefix_kv_lens: Optional[torch.Tensor] = None         # int32 [1] = [common_prefix_len]
    suffix_kv_lens: Optional[torch.Tensor] = None         # int32 [num_seqs] = seq_lens - common_prefix_len
    causal: bool = True


def cascade_attention(output: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                      cu_query_lens: torch.Tensor, max_query_len: int, cu_prefix_query_lens: torch.Tensor,
                      prefix_kv_lens: torch.Tensor, suffix_kv_lens: torch.Tensor, max_kv_len: int, softmax_scale: float,
                      alibi_slopes: Optional[torch.Tensor], sliding_window: tuple, logits_soft_cap: float,
                      block_table: torch.Tensor, common_prefix_len: int, fa_version: int = 3,
                      prefix_scheduler_metadata=None, suffix_scheduler_metadata=None, q_descale=None,
                      k_descale: Optional[torch.Tensor] = None, v_descale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """cascade_attention (flash_attn.py:674-764): every query attends the batch's SHARED prefix (the first
    common_prefix_len tokens, i.e. the first blocks of sequence 0's table) as one non-causal problem over all tokens, then
    its own suffix causally through the remaining block-table columns; the two partial results are merged through their
    log-sum-exps.  Same restrictions as the reference (:698-707): no ALiBi, no sliding window, the prefix a whole number
    of blocks."""
    if alibi_slopes is not None:
        raise AssertionError("Cascade attention does not support ALiBi.")
    if tuple(sliding_window) != (-1, -1):
        raise AssertionError("Cascade attention does not support sliding window.")
    num_tokens = query.shape[0]
    block_size = key_cache.shape[-3]
    if common_prefix_len % block_size != 0 or common_prefix_len < block_size:
        raise AssertionError("common_prefix_len must be a positive multiple of the block size")
    num_common_kv_blocks = common_prefix_len // block_size
    prefix_output, prefix_lse = flash_attn_varlen_func(
        q=query, k=key_cache, v=value_cache, cu_seqlens_q=cu_prefix_query_lens, seqused_k=prefix_kv_lens,
        max_seqlen_q=num_tokens, max_seqlen_k=common_prefix_len, softmax_scale=softmax_scale, causal=False,
        window_size=sliding_window, block_table=block_table[:1], softcap=logits_soft_cap, return_softmax_lse=True,
        k_descale=k_descale, v_descale=v_descale)
    suffix_output, suffix_lse = flash_attn_varlen_func(
        q=query, k=key_cache, v=value_cache, cu_seqlens_q=cu_query_lens, seqused_k=suffix_kv_lens,
        max_seqlen_q=max_query_len, max_seqlen_k=max_kv_len - common_prefix_len, softmax_scale=softmax_scale, causal=True,
        window_size=sliding_window, block_table=block_table[:, num_common_kv_blocks:], softcap=logits_soft_cap,
        return_softmax_lse=True, k_descale=k_descale, v_descale=v_descale)
    ops.merge_attn_states(output, prefix_output, prefix_lse, suffix_output, suffix_lse)
    return output


class B200FlashAttentionImpl:
    """FlashAttentionImpl (flash_attn.py:318-520), decode branch."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[list] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: Optional[float] = None,
                 attn_type: str = "decoder", kv_sharing_target_layer_name: Optional[str] = None) -> None:
        # flash_attn.py:366-374: flash-attn's window is (sliding_window - 1, 0); a soft cap of 0 means none
        self.sliding_window = (-1, -1) if sliding_window is None else (sliding_window - 1, 0)
        self.logits_soft_cap = float(logits_soft_cap or 0.0)
        if attn_type != "decoder":
            raise NotImplementedError("B200FlashAttentionImpl: decoder self-attention only")
        B200FlashAttentionBackend.validate_head_size(head_size)
        self.num_heads, self.head_size, self.scale = num_heads, head_size, float(scale)
        self.num_kv_heads = num_kv_heads or num_heads
        self.alibi_slopes = None if alibi_slopes is None else torch.tensor(alibi_slopes, dtype=torch.float32)
        self.kv_cache_dtype = kv_cache_dtype
        self.kv_sharing_target_layer_name = kv_sharing_target_layer_name

    def forward(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                attn_metadata: B200FlashAttentionMetadata, output: Optional[torch.Tensor] = None,
                output_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """query [T, num_heads, head_size], key / value [T, num_kv_heads, head_size], kv_cache the
        `[2, num_blocks, block_size, num_kv_heads, head_size]` tensor (or its HND-permuted view)."""
        if output is None:
            raise AssertionError("Output tensor must be provided.")
        if output_scale is not None:
            raise NotImplementedError("fused output quantization is not yet supported for B200FlashAttentionImpl")
        if attn_metadata is None:   # profiling run
            return output
        n = attn_metadata.num_actual_tokens
        key_cache, value_cache = kv_cache.unbind(0)
        if self.kv_sharing_target_layer_name is None:
            ops.reshape_and_cache_flash(key, value, key_cache, value_cache, attn_metadata.slot_mapping,
                                        self.kv_cache_dtype, layer._k_scale, layer._v_scale)
        if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
            self.alibi_slopes = self.alibi_slopes.to(query.device)
        fp8 = self.kv_cache_dtype.startswith("fp8")
        if attn_metadata.use_cascade:   # flash_attn.py:520-545
            if not output[:n].is_contiguous():
                raise RuntimeError("cascade attention needs a contiguous output")
            cascade_attention(output[:n], query[:n], key_cache, value_cache, cu_query_lens=attn_metadata.query_start_loc,
                              max_query_len=attn_metadata.max_query_len,
                              cu_prefix_query_lens=attn_metadata.cu_prefix_query_lens,
                              prefix_kv_lens=attn_metadata.prefix_kv_lens, suffix_kv_lens=attn_metadata.suffix_kv_lens,
                              max_kv_len=attn_metadata.max_seq_len, softmax_scale=self.scale,
                              alibi_slopes=self.alibi_slopes, sliding_window=self.sliding_window,
                              logits_soft_cap=self.logits_soft_cap, block_table=attn_metadata.block_table,
                              common_prefix_len=attn_metadata.common_prefix_len,
                              k_descale=layer._k_scale if fp8 else None, v_descale=layer._v_scale if fp8 else None)
            return output
        flash_attn_varlen_func(q=query[:n], k=key_cache, v=value_cache, out=output[:n],
                               cu_seqlens_q=attn_metadata.query_start_loc, max_seqlen_q=attn_metadata.max_query_len,
                               seqused_k=attn_metadata.seq_lens, max_seqlen_k=attn_metadata.max_seq_len,
                               softmax_scale=self.scale, causal=attn_metadata.causal, alibi_slopes=self.alibi_slopes,
                               window_size=self.sliding_window, softcap=self.logits_soft_cap,
                               block_table=attn_metadata.block_table,
                               k_descale=layer._k_scale if self.kv_cache_dtype.startswith("fp8") else None,
                               v_descale=layer._v_scale if self.kv_cache_dtype.startswith("fp8") else None)
        return output


class B200FlashAttentionBackend:
    """FlashAttentionBackend (flash_attn.py:38-101)."""
    accept_output_buffer: bool = True

    @classmethod
    def get_supported_dtypes(cls) -> list:
        return [torch.float16, torch.bfloat16]

    @classmethod
    def get_supported_head_sizes(cls) -> list:
        return [64, 128, 192, 256]

    @classmethod
    def validate_head_size(cls, head_size: int) -> None:
        if head_size not in cls.get_supported_head_sizes():
            raise ValueError(f"Head size {head_size} is not supported by {cls.__name__}. Supported head sizes are: "
                             f"{cls.get_supported_head_sizes()}.")

    @staticmethod
    def get_name() -> str:
        return "B200_FLASH_ATTN_V1"

    @staticmethod
    def get_impl_cls():
        return B200FlashAttentionImpl

    @staticmethod
    def get_metadata_cls():
        return B200FlashAttentionMetadata

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> tuple:
        if block_size != 16:
            raise ValueError("Block size must be 16 for the B200 flash-layout decode kernel.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order() -> tuple:
        """Permutation from `get_kv_cache_shape` to the memory layout (flash_attn.py:88-101)."""
        layout = get_kv_cache_layout()
        if layout == "NHD":
            return (0, 1, 2, 3, 4)
        if layout == "HND":
            return (0, 1, 3, 2, 4)
        raise ValueError(f"Unknown cache layout format {layout}.")

    @staticmethod
    def allocate_kv_cache(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int, dtype: torch.dtype,
                          device) -> torch.Tensor:
        """What vllm/v1/worker/gpu_model_runner.py does with the two methods above: allocate in the PHYSICAL order and
        hand the backend the logical `[2, nb, bs, H, D]` view."""
        shape = B200FlashAttentionBackend.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)
        order = B200FlashAttentionBackend.get_kv_cache_stride_order()
        phys = torch.zeros(tuple(shape[i] for i in order), dtype=dtype, device=device)
        inv = [order.index(i) for i in range(len(order))]
        return phys.permute(*inv)
