"""Host-side helpers of the attention plugin: slot mapping and padded tensors.

Behaviour follows light_vllm/decoding/backends/attention/backends/utils.py:16-75
(``PAD_SLOT_ID``, ``compute_slot_mapping*``) and light_vllm/utils.py ``make_tensor_with_pad`` /
``is_pin_memory_available``; implemented with numpy so a decode batch costs O(1) Python calls.
"""
from typing import Dict, List, Optional, Union

import numpy as np
import torch

PAD_SLOT_ID = -1


def pin_memory_available() -> bool:
    return torch.cuda.is_available()


def is_block_tables_empty(block_tables: Union[None, Dict]) -> bool:
    """None, or a dict whose values are all None (profiling run / embedding models)."""
    if block_tables is None:
        return True
    return isinstance(block_tables, dict) and all(v is None for v in block_tables.values())


def compute_slot_mapping_start_idx(is_prompt: bool, query_len: int, context_len: int,
                                   sliding_window: Optional[int],
                                   use_v2_block_manager: bool) -> int:
    """First prompt position that is written to the cache (sliding window drops older ones)."""
    if is_prompt and sliding_window is not None:
        assert use_v2_block_manager or context_len == 0, (
            "Prefix caching is currently not supported with sliding window attention in V1 "
            "block manager")
        return max(0, query_len - sliding_window)
    return 0


def compute_slot_mapping(is_profile_run: bool, slot_mapping: List[int], seq_id: int, seq_len: int,
                         context_len: int, start_idx: int, block_size: int,
                         block_tables: Dict[int, List[int]]) -> None:
    """Append the slots of tokens [context_len, seq_len) of ``seq_id`` (-1 = not cached)."""
    if is_profile_run:
        slot_mapping.extend([PAD_SLOT_ID] * seq_len)
        return
    first = max(start_idx, context_len)
    slot_mapping.extend([PAD_SLOT_ID] * max(0, start_idx - context_len))
    if first >= seq_len:
        return
    table = np.asarray(block_tables[seq_id], dtype=np.int64)
    pos = np.arange(first, seq_len, dtype=np.int64)
    slot_mapping.extend((table[pos // block_size] * block_size + pos % block_size).tolist())


def make_tensor_with_pad(rows: List[List[int]], pad: int, dtype: torch.dtype,
                         max_len: Optional[int] = None, pin_memory: bool = False) -> torch.Tensor:
    """[len(rows), max_len] CPU tensor, rows right-padded with ``pad``."""
    width = max((len(r) for r in rows), default=0) if max_len is None else max_len
    arr = np.full((len(rows), width), pad, dtype=np.int64)
    for i, r in enumerate(rows):
        assert len(r) <= width
        arr[i, :len(r)] = r
    t = torch.from_numpy(arr).to(dtype)
    return t.pin_memory() if pin_memory else t
