"""B200 paged-attention backend implementing the vllm/attention interface
(/root/reference/vllm/attention/backends/abstract.py:37-107 AttentionBackend, :110-164 AttentionMetadata,
:220-237 AttentionMetadataBuilder, :260-290 AttentionImpl), decode branch modelled on the reference's
decode-capable implementation XFormersImpl.forward (/root/reference/vllm/attention/backends/xformers.py:518-547,
601-629).  Prefill is outside the hot path (north_star: decode): forward() raises for prefill tokens.

Registration from outside the reference tree: a platform plugin whose get_attn_backend_cls returns
"vllm_b200.attention.backend.B200PagedAttentionBackend" (vllm/attention/selector.py:195-201); see INTEGRATION.md.
"""
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple, Type

import torch

from .paged_attn import PagedAttention, PagedAttentionMetadata


@dataclass
class B200PagedAttentionMetadata(PagedAttentionMetadata):
    """Fields of abstract.AttentionMetadata (:110-137) + PagedAttentionMetadata (ops/paged_attn.py:19-33).
    Dataclass field order: the three PagedAttentionMetadata fields come first, as in XFormersMetadata."""
    num_prefills: int = 0
    num_prefill_tokens: int = 0
    num_decode_tokens: int = 0
    slot_mapping: Optional[torch.Tensor] = None          # int64 [num_tokens]
    multi_modal_placeholder_index_maps: Optional[Dict[str, Any]] = None
    enable_kv_scales_calculation: bool = False
    use_cuda_graph: bool = False

    @property
    def prefill_metadata(self) -> Optional["B200PagedAttentionMetadata"]:
        return None if self.num_prefills == 0 else self

    @property
    def decode_metadata(self) -> Optional["B200PagedAttentionMetadata"]:
        return None if self.num_decode_tokens == 0 else self

    def asdict_zerocopy(self, skip_fields=None) -> Dict[str, Any]:
        from dataclasses import fields
        skip_fields = skip_fields or set()
        return {f.name: getattr(self, f.name) for f in fields(self) if f.name not in skip_fields}


class B200PagedAttentionMetadataBuilder:
    """abstract.AttentionMetadataBuilder (:220-237) for decode-only batches: turns the scheduler's per-sequence
    block lists into the device tensors the kernels read."""

    def __init__(self, block_size: int, device: torch.device):
        self.block_size = block_size
        self.device = device
        self.prepare()

    def prepare(self) -> None:
        self._seq_lens: List[int] = []
        self._block_tables: List[List[int]] = []

    def add_decode_seq(self, seq_len: int, block_table: List[int]) -> None:
        """seq_len counts the token being decoded (its K/V are written to slot seq_len-1 first)."""
        self._seq_lens.append(seq_len)
        self._block_tables.append(block_table)

    def build(self, seq_lens: Optional[List[int]] = None, query_lens: Optional[List[int]] = None,
              cuda_graph_pad_size: int = -1, batch_size: int = 0) -> B200PagedAttentionMetadata:
        seq_lens = self._seq_lens if seq_lens is None else seq_lens
        n = len(seq_lens)
        max_blocks = max((len(b) for b in self._block_tables), default=0)
        bt = torch.zeros((n, max(max_blocks, 1)), dtype=torch.int32)
        slots = torch.empty(n, dtype=torch.int64)
        for i, (L, blocks) in enumerate(zip(seq_lens, self._block_tables)):
            bt[i, :len(blocks)] = torch.tensor(blocks, dtype=torch.int32)
            slots[i] = blocks[(L - 1) // self.block_size] * self.block_size + (L - 1) % self.block_size
        return B200PagedAttentionMetadata(
            seq_lens_tensor=torch.tensor(seq_lens, dtype=torch.int32).to(self.device, non_blocking=True),
            max_decode_seq_len=max(seq_lens, default=0),
            block_tables=bt.to(self.device, non_blocking=True),
            num_prefills=0, num_prefill_tokens=0, num_decode_tokens=n,
            slot_mapping=slots.to(self.device, non_blocking=True))


class B200PagedAttentionImpl:
    """abstract.AttentionImpl (:260-290)."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: Optional[float] = None,
                 attn_type: str = "decoder", kv_sharing_target_layer_name: Optional[str] = None) -> None:
        if kv_sharing_target_layer_name is not None:
            raise NotImplementedError("KV sharing is not supported by the B200 paged-attention backend")
        if logits_soft_cap is not None:
            raise ValueError("the B200 paged-attention backend does not support logits soft cap")
        if sliding_window is not None:
            raise ValueError("the B200 paged-attention backend does not support sliding window")
        if attn_type != "decoder":
            raise NotImplementedError("only decoder self-attention is implemented")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.alibi_slopes = None if alibi_slopes is None else torch.tensor(alibi_slopes, dtype=torch.float32)
        self.kv_cache_dtype = kv_cache_dtype
        assert self.num_heads % self.num_kv_heads == 0
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        supported = PagedAttention.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(f"Head size {head_size} is not supported by PagedAttention. Supported head sizes are: "
                             f"{supported}.")

    def forward(self, layer, query: torch.Tensor, key: Optional[torch.Tensor], value: Optional[torch.Tensor],
                kv_cache: torch.Tensor, attn_metadata: B200PagedAttentionMetadata,
                output: Optional[torch.Tensor] = None, output_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """query [T, H*D]; key/value [T, Hkv*D]; kv_cache [2, num_blocks, block_size*Hkv*D]; `layer` supplies
        _k_scale / _v_scale fp32 tensors (abstract.py:240-247).  Returns [T, H*D]."""
        if output_scale is not None:
            raise NotImplementedError("fused output quantization is not supported")
        if attn_metadata.prefill_metadata is not None:
            raise NotImplementedError("the B200 paged-attention backend implements the decode path only")
        query = query.view(-1, self.num_heads, self.head_size)
        key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads, self.head_size)
        if key is not None and value is not None:
            key = key.view(-1, self.num_kv_heads, self.head_size)
            value = value.view(-1, self.num_kv_heads, self.head_size)
            PagedAttention.write_to_paged_cache(key, value, key_cache, value_cache, attn_metadata.slot_mapping,
                                                self.kv_cache_dtype, layer._k_scale, layer._v_scale)
        if output is None:
            output = torch.empty((query.shape[0], self.num_heads, self.head_size), dtype=query.dtype,
                                 device=query.device)
        out3 = output.view(-1, self.num_heads, self.head_size)
        decode_meta = attn_metadata.decode_metadata
        if decode_meta is not None:
            if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
                self.alibi_slopes = self.alibi_slopes.to(query.device)
            PagedAttention.forward_decode(query, key_cache, value_cache, decode_meta.block_tables,
                                          decode_meta.seq_lens_tensor, decode_meta.max_decode_seq_len,
                                          self.kv_cache_dtype, self.num_kv_heads, self.scale, self.alibi_slopes,
                                          layer._k_scale, layer._v_scale, output=out3)
        return output.view(-1, self.num_heads * self.head_size)

    def supports_fused_output_quant(self, block_size: int) -> bool:
        return block_size == 16 and self.head_size % 32 == 0 and self.num_heads <= 128

    def forward_quant(self, layer, query: torch.Tensor, kv_cache: torch.Tensor,
                      attn_metadata: B200PagedAttentionMetadata, scale_ub: Optional[torch.Tensor] = None):
        """Decode attention whose output leaves as the dynamically per-token quantised fp8 activation of the FP8
        o_proj: (fp8 [T, H*D], scales [T, 1]).  The quantisation runs inside the split-K reduce kernel
        (ops.paged_attention_v1_quant); results are bit-identical to forward() followed by
        scaled_fp8_quant(use_per_token_if_dynamic=True).  The new token's K / V must already be in the cache."""
        from .. import _custom_ops as ops
        if attn_metadata.prefill_metadata is not None:
            raise NotImplementedError("the B200 paged-attention backend implements the decode path only")
        query = query.view(-1, self.num_heads, self.head_size)
        key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads, self.head_size)
        meta = attn_metadata.decode_metadata
        if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
            self.alibi_slopes = self.alibi_slopes.to(query.device)
        return ops.paged_attention_v1_quant(query, key_cache, value_cache, self.num_kv_heads, self.scale,
                                            meta.block_tables, meta.seq_lens_tensor, value_cache.shape[3],
                                            meta.max_decode_seq_len, self.alibi_slopes, self.kv_cache_dtype,
                                            layer._k_scale, layer._v_scale, scale_ub)


class B200PagedAttentionBackend:
    """abstract.AttentionBackend (:37-107)."""
    accept_output_buffer: bool = True

    @staticmethod
    def get_name() -> str:
        return "B200_PAGED"

    @staticmethod
    def get_impl_cls() -> Type[B200PagedAttentionImpl]:
        return B200PagedAttentionImpl

    @staticmethod
    def get_metadata_cls() -> Type[B200PagedAttentionMetadata]:
        return B200PagedAttentionMetadata

    @classmethod
    def make_metadata(cls, *args, **kwargs) -> B200PagedAttentionMetadata:
        return cls.get_metadata_cls()(*args, **kwargs)

    @staticmethod
    def get_builder_cls() -> Type[B200PagedAttentionMetadataBuilder]:
        return B200PagedAttentionMetadataBuilder

    @staticmethod
    def get_state_cls():
        raise NotImplementedError("AttentionState (V0 model-runner graph state) lives outside the hot path")

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return PagedAttention.get_supported_head_sizes()

    @classmethod
    def validate_head_size(cls, head_size: int) -> None:
        if head_size not in cls.get_supported_head_sizes():
            raise ValueError(f"Head size {head_size} is not supported by {cls.get_name()}.")

    @staticmethod
    def get_supported_dtypes() -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> Tuple[int, ...]:
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order() -> Tuple[int, ...]:
        raise NotImplementedError  # as the reference's paged backends: identity order

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        PagedAttention.swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        PagedAttention.copy_blocks(kv_caches, src_to_dists)
