"""vllm/attention interface of the B200 decode path: PagedAttention helper + AttentionBackend."""
from .backend import (B200PagedAttentionBackend, B200PagedAttentionImpl, B200PagedAttentionMetadata,
                      B200PagedAttentionMetadataBuilder)
from .paged_attn import PagedAttention, PagedAttentionMetadata

__all__ = ["PagedAttention", "PagedAttentionMetadata", "B200PagedAttentionBackend", "B200PagedAttentionImpl",
           "B200PagedAttentionMetadata", "B200PagedAttentionMetadataBuilder"]
