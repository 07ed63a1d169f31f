"""``PagedAttention``: the call conventions of the paged KV cache, on top of the B200 ops.

Behavioural spec: light_vllm/decoding/backends/attention/ops/paged_attn.py:34-248 (orphaned in
the reference) -- same static-method names, argument meaning, KV pool shape
``(2, num_blocks, block_size * num_kv_heads * head_size)``, the v1/v2 choice, caller-owned v2
scratch, K-then-V swap, one copy call for all layers.
"""
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from light_vllm_b200 import _custom_ops as ops
from light_vllm_b200.attention.prefix_prefill import context_attention_fwd

# Fixed by the op contract: partition size of paged_attention_v2 (paged_attn.py:14).
_PARTITION_SIZE = 512


@dataclass
class PagedAttentionMetadata:
    """Decode-side metadata (paged_attn.py:17-31)."""
    seq_lens_tensor: Optional[torch.Tensor]
    max_decode_seq_len: int
    block_tables: Optional[torch.Tensor]


class PagedAttention:
    # v2 scratch is reused across layers/steps of one shape instead of three torch.empty per call
    # (reference :156-166); keyed by (device, stream, dtype, B, H, P, D) -- per stream because the
    # executor may keep two steps in flight on different streams (core/executor.py:62-185).

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [64, 80, 96, 112, 120, 128, 192, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int,
                       head_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """Views K [nb, kh, d/x, bs, x] and V [nb, kh, d, bs] of one layer's pool."""
        x = 16 // kv_cache.element_size()
        nb = kv_cache.shape[1]
        key_cache = kv_cache[0].view(nb, num_kv_heads, head_size // x, -1, x)
        value_cache = kv_cache[1].view(nb, num_kv_heads, head_size, -1)
        return key_cache, value_cache

    @staticmethod
    def write_to_paged_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                             value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                             kv_cache_dtype: str, k_scale: float, v_scale: float) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping.flatten(),
                              kv_cache_dtype, k_scale, v_scale)

    @staticmethod
    def use_v1(max_seq_len: int, num_seqs: int, num_heads: int) -> bool:
        """The reference's heuristic (paged_attn.py:126-128), kept verbatim in behaviour."""
        max_num_partitions = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
        return max_seq_len <= 8192 and (max_num_partitions == 1 or num_seqs * num_heads > 512)

    @staticmethod
    def _v2_scratch(query: torch.Tensor, max_num_partitions: int):
        """Per-call ``torch.empty`` scratch, exactly as the reference allocates it
        (paged_attn.py:156-166): the caching allocator makes it cheap, and a CUDA graph that captured
        the call owns the memory through its graph pool -- nothing cached here can be freed under it."""
        B, H, D = query.shape
        tmp = torch.empty((B, H, max_num_partitions, D), dtype=query.dtype, device=query.device)
        exp_sums = torch.empty((B, H, max_num_partitions), dtype=torch.float32, device=query.device)
        return exp_sums, torch.empty_like(exp_sums), tmp

    @staticmethod
    def forward_decode(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int,
                       kv_cache_dtype: str, num_kv_heads: int, scale: float,
                       alibi_slopes: Optional[torch.Tensor], k_scale: float, v_scale: float,
                       tp_rank: int = 0, blocksparse_local_blocks: int = 0,
                       blocksparse_vert_stride: int = 0, blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> torch.Tensor:
        block_size = value_cache.shape[3]
        if blocksparse_vert_stride is not None and blocksparse_vert_stride > 1:
            assert blocksparse_block_size > 0 and blocksparse_block_size % block_size == 0, (
                f"{blocksparse_block_size=} needs to be a multiple of {block_size=} used in "
                "block_tables.")
        output = torch.empty_like(query)
        num_seqs, num_heads, _ = query.shape
        tail = (tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                blocksparse_head_sliding_step)
        if PagedAttention.use_v1(max_seq_len, num_seqs, num_heads):
            ops.paged_attention_v1(output, query, key_cache, value_cache, num_kv_heads, scale,
                                   block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                                   kv_cache_dtype, k_scale, v_scale, *tail)
        else:
            assert _PARTITION_SIZE % block_size == 0
            max_num_partitions = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
            exp_sums, max_logits, tmp_output = PagedAttention._v2_scratch(query, max_num_partitions)
            ops.paged_attention_v2(output, exp_sums, max_logits, tmp_output, query, key_cache,
                                   value_cache, num_kv_heads, scale, block_tables, seq_lens,
                                   block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
                                   v_scale, *tail)
        return output

    @staticmethod
    def forward_prefix(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                       key_cache: torch.Tensor, value_cache: torch.Tensor,
                       block_tables: torch.Tensor, query_start_loc: torch.Tensor,
                       seq_lens_tensor: torch.Tensor, context_lens: torch.Tensor,
                       max_query_len: int, alibi_slopes: Optional[torch.Tensor],
                       sliding_window: Optional[int]) -> torch.Tensor:
        """Prefill of new tokens over a cached prefix (reference: paged_attn.py:193-225)."""
        output = torch.empty_like(query)
        # query_start_loc is (batch_size + 1,)
        context_attention_fwd(query, key, value, output, key_cache, value_cache, block_tables,
                              query_start_loc[:-1], seq_lens_tensor, context_lens, max_query_len,
                              alibi_slopes, sliding_window)
        return output

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                    src_to_dst: torch.Tensor) -> None:
        ops.swap_blocks(src_kv_cache[0], dst_kv_cache[0], src_to_dst)
        ops.swap_blocks(src_kv_cache[1], dst_kv_cache[1], src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        ops.copy_blocks([kv[0] for kv in kv_caches], [kv[1] for kv in kv_caches], src_to_dists)
