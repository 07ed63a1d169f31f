"""Python op surface of the B200 paged-KV hot path.

Same function names, argument order, defaults and error behaviour (``RuntimeError`` from the
native layer) as the attention / cache subset of the reference's
``light_vllm/backends/_custom_ops.py`` (:72-130 paged attention, :424-480 cache ops and device
attributes), so a caller written against that module can ``import light_vllm_b200._custom_ops as
ops`` instead.  Everything dispatches to ``torch.ops._C*`` registered by ``_C.so`` -> C ABI of
``libb200_paged.so`` -> hand-written sm_100a kernels.  There is no fallback path.
"""
from typing import List, Optional

import torch

from light_vllm_b200 import _lib  # noqa: F401  (loads libb200_paged.so and registers torch.ops._C*)

_attn = torch.ops._C
_cache = torch.ops._C_cache_ops
_cuda_utils = torch.ops._C_cuda_utils


def is_custom_op_supported(op_name: str) -> bool:
    """``op_name`` like ``"_C::paged_attention_v1"`` (reference :21-23)."""
    op, _ = torch._C._jit_get_operation(op_name)
    return op is not None


# ------------------------------------------------------------------ paged attention ----
def paged_attention_v1(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, block_size: int,
                       max_seq_len: int, alibi_slopes: Optional[torch.Tensor],
                       kv_cache_dtype: str, k_scale: float, v_scale: float, tp_rank: int = 0,
                       blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> None:
    """Single-pass decode attention over the paged cache; writes ``out[B, H, D]`` in place."""
    _attn.paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale,
                             block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                             kv_cache_dtype, k_scale, v_scale, tp_rank,
                             blocksparse_local_blocks, blocksparse_vert_stride,
                             blocksparse_block_size, blocksparse_head_sliding_step)


def paged_attention_v2(out: torch.Tensor, exp_sum: torch.Tensor, max_logits: torch.Tensor,
                       tmp_out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, block_size: int,
                       max_seq_len: int, alibi_slopes: Optional[torch.Tensor],
                       kv_cache_dtype: str, k_scale: float, v_scale: float, tp_rank: int = 0,
                       blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> None:
    """Split-context decode attention (512-token partitions + merge) using caller scratch."""
    _attn.paged_attention_v2(out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache,
                             num_kv_heads, scale, block_tables, seq_lens, block_size,
                             max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale,
                             tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
                             blocksparse_block_size, blocksparse_head_sliding_step)


# ------------------------------------------------------------------------ cache ops ----
def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                      value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                      kv_cache_dtype: str, k_scale: float, v_scale: float) -> None:
    """Scatter new K/V rows into the paged layout (optionally quantising to fp8)."""
    _cache.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype,
                             k_scale, v_scale)


def reshape_and_cache_flash(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                            value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                            kv_cache_dtype: str, k_scale: float, v_scale: float) -> None:
    """Same for the ``[num_blocks, block_size, heads, head_size]`` layout."""
    _cache.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                                   kv_cache_dtype, k_scale, v_scale)


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: torch.Tensor) -> None:
    """For every layer and every ``(src, dst)`` row: copy the K block and the V block."""
    _cache.copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """Move blocks between pools (H2D / D2H / same-GPU D2D); ``block_mapping`` is a CPU tensor."""
    _cache.swap_blocks(src, dst, block_mapping)


def convert_fp8(output: torch.Tensor, input: torch.Tensor, scale: float = 1.0,
                kv_dtype: str = "fp8") -> None:
    """Elementwise scaled fp8-e4m3 conversion; direction follows the dtypes."""
    _cache.convert_fp8(output, input, scale, kv_dtype)


def rotary_embedding_reshape_and_cache(positions: torch.Tensor, query: torch.Tensor,
                                       key: torch.Tensor, value: torch.Tensor, head_size: int,
                                       cos_sin_cache: torch.Tensor, is_neox: bool,
                                       key_cache: torch.Tensor, value_cache: torch.Tensor,
                                       slot_mapping: torch.Tensor, kv_cache_dtype: str,
                                       k_scale: float, v_scale: float) -> None:
    """B200 extension: ``rotary_embedding`` then ``reshape_and_cache`` in ONE launch -- q and k are
    rotated in place, the rotated k and v land in the paged cache (fp8 quantisation included).
    Bit-identical to calling the two reference ops back to back."""
    _cache.rotary_embedding_reshape_and_cache(positions, query, key, value, head_size,
                                              cos_sin_cache, is_neox, key_cache, value_cache,
                                              slot_mapping, kv_cache_dtype, k_scale, v_scale)


# ------------------------------------------------------------------- per-step ops ----
# Same names / arguments as light_vllm/backends/_custom_ops.py:134-153,167-176.
def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                     head_size: int, cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    _attn.rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def batched_rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                             head_size: int, cos_sin_cache: torch.Tensor, is_neox: bool,
                             rot_dim: int, cos_sin_cache_offsets: torch.Tensor) -> None:
    _attn.batched_rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox,
                                   rot_dim, cos_sin_cache_offsets)


def advance_step(num_seqs: int, num_queries: int, block_size: int, input_tokens: torch.Tensor,
                 sampled_token_ids: torch.Tensor, input_positions: torch.Tensor,
                 seq_lens: torch.Tensor, slot_mapping: torch.Tensor,
                 block_tables: torch.Tensor) -> None:
    """Advance a step on GPU for existing inputs for a multi-step runner"""
    return _attn.advance_step(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids,
                              input_positions, seq_lens, slot_mapping, block_tables)


# ---------------------------------------------------------------- device attributes ----
def get_device_attribute(attribute: int, device: int) -> int:
    return _cuda_utils.get_device_attribute(attribute, device)


def get_max_shared_memory_per_block_device_attribute(device: int) -> int:
    return _cuda_utils.get_max_shared_memory_per_block_device_attribute(device)
