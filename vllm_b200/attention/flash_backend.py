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
csrc/paged_prefill.cu (16-bit and fp8 caches).  Sliding windows, soft-capping and cascade attention raise.  fp8 caches: the reference also quantises the QUERY to fp8 (FA3, :478-486); this kernel keeps the query in
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
                           num_splits: int = 0, cu_seqlens_k=None) -> torch.Tensor:
    """vllm.vllm_flash_attn.flash_attn_varlen_func as flash_attn.py:497-519 calls it.  Decode (max_seqlen_q == 1): q
    [num_seqs, num_heads, head_size] (one token per sequence, cu_seqlens_q = 0..num_seqs) on the HBM-roofline decode
    kernel; anything else (prefill, chunked prefill, mixed batches: q [total_q, ...] split by cu_seqlens_q, bottom-right
    aligned causal mask against seqused_k) on the variable-length tensor-core kernel.  k / v the two halves of the
    paged cache `[num_blocks, block_size, num_kv_heads, head_size]` (any stride order with a contiguous head
    dimension), block_table int32 [num_seqs, max_blocks], seqused_k int32 [num_seqs].  k_descale / v_descale: fp32
    per-tensor scales of an fp8 cache (the reference passes them expanded to [num_seqs, num_kv_heads]; element 0 is
    used)."""
    if block_table is None:
        raise NotImplementedError("flash_attn_varlen_func: only the paged-KV (block_table) form is built")
    if softcap:
        raise NotImplementedError("flash_attn_varlen_func: logits soft-capping is not built")
    if window_size is not None and tuple(window_size) not in ((-1, -1), (-1, 0)):
        raise NotImplementedError("flash_attn_varlen_func: sliding windows are not built")
    if cu_seqlens_q is not None and cu_seqlens_q.numel() != seqused_k.shape[0] + 1:
        raise RuntimeError("cu_seqlens_q must hold num_seqs + 1 offsets")
    fp8 = k.element_size() == 1
    if fp8 and (k_descale is None or v_descale is None):
        raise RuntimeError("an fp8 cache needs k_descale and v_descale")
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** -0.5
    if out is None:
        out = torch.empty_like(q)
    if max_seqlen_q != 1 or q.shape[0] != seqused_k.shape[0]:
        # prefill / chunked prefill / mixed batch: the variable-length kernel (csrc/paged_prefill.cu)
        if alibi_slopes is not None:
            raise NotImplementedError("flash_attn_varlen_func: ALiBi is built for the decode case only")
        if cu_seqlens_q is None:
            raise RuntimeError("flash_attn_varlen_func: cu_seqlens_q is required")
        ks = k_descale.reshape(-1)[:1].float().contiguous() if fp8 else None
        vs = v_descale.reshape(-1)[:1].float().contiguous() if fp8 else None
        ops.flash_attn_varlen_paged(out, q, k, v, cu_seqlens_q, seqused_k, block_table, softmax_scale, causal,
                                    int(max_seqlen_q), "fp8" if fp8 else "auto", ks, vs)
        return out
    ks = k_descale.reshape(-1)[:1].float().contiguous() if k_descale is not None else None
    vs = v_descale.reshape(-1)[:1].float().contiguous() if v_descale is not None else None
    ops.paged_attention_flash(out, q, k, v, softmax_scale, block_table, seqused_k, int(max_seqlen_k), alibi_slopes,
                              "fp8" if fp8 else "auto", ks, vs)
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
    causal: bool = True


class B200FlashAttentionImpl:
    """FlashAttentionImpl (flash_attn.py:318-520), decode branch."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[list] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: Optional[float] = None,
                 attn_type: str = "decoder", kv_sharing_target_layer_name: Optional[str] = None) -> None:
        if sliding_window is not None or logits_soft_cap:
            raise NotImplementedError("B200FlashAttentionImpl: sliding window / soft cap are not built")
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
        if attn_metadata.use_cascade:
            raise NotImplementedError("cascade attention is outside the decode hot path")
        n = attn_metadata.num_actual_tokens
        key_cache, value_cache = kv_cache.unbind(0)
        if self.kv_sharing_target_layer_name is None:
            ops.reshape_and_cache_flash(key, value, key_cache, value_cache, attn_metadata.slot_mapping,
                                        self.kv_cache_dtype, layer._k_scale, layer._v_scale)
        if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
            self.alibi_slopes = self.alibi_slopes.to(query.device)
        flash_attn_varlen_func(q=query[:n], k=key_cache, v=value_cache, out=output[:n],
                               cu_seqlens_q=attn_metadata.query_start_loc, max_seqlen_q=attn_metadata.max_query_len,
                               seqused_k=attn_metadata.seq_lens, max_seqlen_k=attn_metadata.max_seq_len,
                               softmax_scale=self.scale, causal=attn_metadata.causal, alibi_slopes=self.alibi_slopes,
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
