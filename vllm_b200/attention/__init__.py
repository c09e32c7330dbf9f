"""vllm/attention interface of the B200 decode path: PagedAttention helper + AttentionBackend."""
from .backend import (B200PagedAttentionBackend, B200PagedAttentionImpl, B200PagedAttentionMetadata,
                      B200PagedAttentionMetadataBuilder)
from .flash_backend import (B200FlashAttentionBackend, B200FlashAttentionImpl, B200FlashAttentionMetadata,
                            flash_attn_varlen_func)
from .paged_attn import PagedAttention, PagedAttentionMetadata

__all__ = ["PagedAttention", "PagedAttentionMetadata", "B200PagedAttentionBackend", "B200PagedAttentionImpl",
           "B200PagedAttentionMetadata", "B200PagedAttentionMetadataBuilder", "B200FlashAttentionBackend",
           "B200FlashAttentionImpl", "B200FlashAttentionMetadata", "flash_attn_varlen_func"]
