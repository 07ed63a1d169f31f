"""Backend selection: the branch a light-vllm maintainer adds to
``light_vllm/decoding/backends/attention/selector.py:37-75`` (whose ``which_attn_to_use`` always
answers FLASH_ATTN).  Same class / method names, so ``AttnBackend.from_engine(engine)`` keeps
working; ``VLLM_ATTENTION_BACKEND=PAGED_B200`` (or the default here) picks the paged B200 backend.
"""
import enum
import os
from typing import Optional

import torch

from light_vllm_b200.attention.abstract import AttentionType
from light_vllm_b200.attention.paged_attn import PagedAttention


class _Backend(enum.Enum):
    PAGED_B200 = enum.auto()

    @staticmethod
    def backend_name_to_enum(backend_name: str) -> "_Backend":
        assert backend_name is not None
        members = _Backend.__members__
        if backend_name not in members:
            raise ValueError(f"Invalid attention backend '{backend_name}'. "
                             f"Available backends: {', '.join(members)} (case-sensitive).")
        return _Backend[backend_name]


class AttnBackend:

    @classmethod
    def from_engine(cls, engine):
        model_config = engine.engine_config.model_config
        backend = cls.which_attn_to_use(model_config.get_num_attention_heads(),
                                        model_config.get_head_size(),
                                        model_config.get_num_kv_heads(),
                                        model_config.get_sliding_window(), model_config.dtype)
        backend_cls = cls.get_backend_cls(backend)
        attn_type = AttentionType.attn_type_name_to_enum(engine.workflow.attn_type)
        return backend_cls(attn_type)

    @staticmethod
    def get_backend_cls(backend):
        if backend == _Backend.PAGED_B200:
            from light_vllm_b200.attention.backend import PagedB200AttentionBackend
            return PagedB200AttentionBackend
        raise ValueError("Invalid attention backend.")

    @classmethod
    def which_attn_to_use(cls, num_heads: int, head_size: int, num_kv_heads: int,
                          sliding_window: Optional[int], dtype: torch.dtype):
        name = os.environ.get("VLLM_ATTENTION_BACKEND") or "PAGED_B200"
        selected = _Backend.backend_name_to_enum(name)
        if head_size not in PagedAttention.get_supported_head_sizes():
            raise ValueError(f"Head size {head_size} is not supported by PagedAttention. "
                             f"Supported head sizes are: {PagedAttention.get_supported_head_sizes()}.")
        if dtype not in (torch.float16, torch.bfloat16, torch.float32):
            raise ValueError(f"Unsupported data type: {dtype}")
        if num_heads % num_kv_heads != 0:
            raise ValueError(f"num_heads ({num_heads}) must be a multiple of num_kv_heads ({num_kv_heads})")
        return selected
