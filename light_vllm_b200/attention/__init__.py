from light_vllm_b200.attention.abstract import (AttentionBackend, AttentionImpl, AttentionMetadata,
                                                AttentionMetadataBuilder, AttentionType)
from light_vllm_b200.attention.backend import (PagedB200AttentionBackend, PagedB200AttentionImpl,
                                               PagedB200AttentionMetadata,
                                               PagedB200AttentionMetadataBuilder)
from light_vllm_b200.attention.paged_attn import PagedAttention, PagedAttentionMetadata

__all__ = [
    "AttentionBackend", "AttentionImpl", "AttentionMetadata", "AttentionMetadataBuilder",
    "AttentionType", "PagedAttention", "PagedAttentionMetadata", "PagedB200AttentionBackend",
    "PagedB200AttentionImpl", "PagedB200AttentionMetadata", "PagedB200AttentionMetadataBuilder",
]
