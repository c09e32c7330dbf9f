"""Mirror of /root/reference/vllm/attention/ops/paged_attn.py:19-193 (the decode half; prefix prefill is a Triton
kernel outside the hot path).  Same class, method names, arguments and cache views, so the KV cache stays
byte-compatible with paged_attention_v1/v2."""
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from .. import _custom_ops as ops

# Should be the same as PARTITION_SIZE of paged_attention_v2 (csrc/attention/paged_attention_v2.cu:29).
_PARTITION_SIZE = 512


@dataclass
class PagedAttentionMetadata:
    """ops/paged_attn.py:19-33."""
    seq_lens_tensor: Optional[torch.Tensor]   # (batch_size,) int32
    max_decode_seq_len: int                   # 0 for a prefill-only batch
    block_tables: Optional[torch.Tensor]      # (batch_size, max_blocks_per_seq) int32


class PagedAttention:

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [32, 64, 80, 96, 112, 120, 128, 192, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> Tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int, head_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        x = 16 // kv_cache.element_size()
        num_blocks = kv_cache.shape[1]
        key_cache = kv_cache[0].view(num_blocks, num_kv_heads, head_size // x, -1, x)
        value_cache = kv_cache[1].view(num_blocks, num_kv_heads, head_size, -1)
        return key_cache, value_cache

    @staticmethod
    def write_to_paged_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                             value_cache: torch.Tensor, slot_mapping: torch.Tensor, kv_cache_dtype: str,
                             k_scale: torch.Tensor, v_scale: torch.Tensor) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping.flatten(), kv_cache_dtype, k_scale,
                              v_scale)

    @staticmethod
    def forward_decode(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int, kv_cache_dtype: str,
                       num_kv_heads: int, scale: float, alibi_slopes: Optional[torch.Tensor], k_scale: torch.Tensor,
                       v_scale: torch.Tensor, tp_rank: int = 0, blocksparse_local_blocks: int = 0,
                       blocksparse_vert_stride: int = 0, blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0, output: Optional[torch.Tensor] = None) -> torch.Tensor:
        """ops/paged_attn.py:90-193.  The reference picks v1 or v2 by a heuristic (:129-130) because its v1 is
        unsplit; here v1 is split-K internally (the library picks the partition from the SM count), so the v1 entry
        is always used -- same result, caller-visible partial buffers are only needed through paged_attention_v2."""
        if output is None:
            output = torch.empty_like(query)
        block_size = value_cache.shape[3]
        ops.paged_attention_v1(output, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                               block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
                               blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                               blocksparse_head_sliding_step)
        return output

    @staticmethod
    def forward_decode_v2(query, key_cache, value_cache, block_tables, seq_lens, max_seq_len, kv_cache_dtype,
                          num_kv_heads, scale, alibi_slopes, k_scale, v_scale) -> torch.Tensor:
        """The reference's v2 branch (:156-192) verbatim: caller-allocated partials, PARTITION_SIZE 512."""
        output = torch.empty_like(query)
        block_size = value_cache.shape[3]
        num_seqs, num_heads, head_size = query.shape
        max_num_partitions = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
        assert _PARTITION_SIZE % block_size == 0
        tmp_output = torch.empty((num_seqs, num_heads, max_num_partitions, head_size), dtype=output.dtype,
                                 device=output.device)
        exp_sums = torch.empty((num_seqs, num_heads, max_num_partitions), dtype=torch.float32, device=output.device)
        max_logits = torch.empty_like(exp_sums)
        ops.paged_attention_v2(output, exp_sums, max_logits, tmp_output, query, key_cache, value_cache, num_kv_heads,
                               scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
                               k_scale, v_scale)
        return output

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        """ops/paged_attn.py:236-247 (csrc/cache_kernels.cu:24-66): block copies between two caches; host-driven
        cudaMemcpyAsync in the reference, tensor copies here (not on the decode path)."""
        for src, dst in src_to_dst.tolist():
            dst_kv_cache[:, dst].copy_(src_kv_cache[:, src], non_blocking=True)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        """ops/paged_attn.py:249-257 (csrc/cache_kernels.cu:68-140): copy-on-write of whole blocks."""
        for src, dst in src_to_dists.tolist():
            for kv in kv_caches:
                kv[:, dst].copy_(kv[:, src])
