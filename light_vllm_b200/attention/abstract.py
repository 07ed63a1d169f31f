"""The attention plugin interface the backend satisfies.

When light-vllm itself is importable the reference's own abstract classes are used, so
``PagedB200AttentionBackend`` *is a* ``DecodeOnlyAttentionBackend`` and can be returned by
``AttnBackend.from_engine`` (light_vllm/decoding/backends/attention/selector.py:40-75).  Otherwise
(stand-alone use, tests, the GPU box) structurally identical stand-ins are defined here with the
same names, members and signatures:
  light_vllm/backends/attention/abstract.py:9-124          AttentionType/Backend/Metadata/Impl
  light_vllm/decoding/backends/attention/backends/abstract.py:15-166   DecodeOnly* variants
"""
from abc import ABC, abstractmethod
from dataclasses import dataclass, fields
from enum import Enum, auto
from typing import Any, Dict, Generic, List, Optional, Set, Tuple, Type, TypeVar

import torch

try:  # pragma: no cover - exercised only inside a light-vllm checkout
    from light_vllm.backends.attention.abstract import (AttentionBackend, AttentionImpl,
                                                        AttentionMetadata,
                                                        AttentionMetadataBuilder, AttentionType)
    from light_vllm.decoding.backends.attention.backends.abstract import (
        DecodeOnlyAttentionBackend, DecodeOnlyAttentionImpl, DecodeOnlyAttentionMetadata,
        DecodeOnlyAttentionMetadataBuilder)
    HAVE_LIGHT_VLLM = True
except Exception:  # light_vllm absent (or its heavy imports unavailable)
    HAVE_LIGHT_VLLM = False

    class AttentionType(Enum):
        DECODER = auto()
        ENCODER = auto()
        ENCODER_DECODER = auto()

        @staticmethod
        def attn_type_name_to_enum(attn_type: str) -> "AttentionType":
            if attn_type not in AttentionType.__members__:
                raise ValueError(f"Invalid attn_type '{attn_type}'. Available backends: "
                                 f"{', '.join(AttentionType.__members__)} (case-sensitive).")
            return AttentionType[attn_type]

    class AttentionBackend(ABC):

        def __init__(self, attn_type: AttentionType):
            self._attn_type = attn_type

        @property
        def attn_type(self) -> AttentionType:
            return self._attn_type

        @staticmethod
        @abstractmethod
        def get_name() -> str:
            ...

        @staticmethod
        @abstractmethod
        def get_impl_cls() -> Type["AttentionImpl"]:
            ...

        @staticmethod
        @abstractmethod
        def get_metadata_cls() -> Type["AttentionMetadata"]:
            ...

        @classmethod
        def make_metadata(cls, *args, **kwargs):
            return cls.get_metadata_cls()(*args, **kwargs)

        @staticmethod
        @abstractmethod
        def get_builder_cls() -> Type["AttentionMetadataBuilder"]:
            ...

        @classmethod
        def make_metadata_builder(cls, *args, **kwargs):
            return cls.get_builder_cls()(*args, **kwargs)

    @dataclass
    class AttentionMetadata:

        def to(self, device, non_blocking=True):
            for k, v in self.__dict__.items():
                if isinstance(v, torch.Tensor):
                    self.__dict__[k] = v.to(device, non_blocking=non_blocking)
            return self

    _T = TypeVar("_T", bound=AttentionMetadata)

    class AttentionMetadataBuilder(ABC, Generic[_T]):

        @abstractmethod
        def __init__(self) -> None:
            ...

        @abstractmethod
        def __call__(self, *args, **kwargs) -> _T:
            ...

    class AttentionImpl(ABC, Generic[_T]):

        @abstractmethod
        def __init__(self, num_heads: int, head_size: int, scale: float,
                     num_kv_heads: Optional[int] = None,
                     alibi_slopes: Optional[List[float]] = None,
                     sliding_window: Optional[int] = None, kv_cache_dtype: str = "auto",
                     blocksparse_params: Optional[Dict[str, Any]] = None,
                     logits_soft_cap: Optional[float] = None) -> None:
            ...

        @abstractmethod
        def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                    kv_cache: torch.Tensor, attn_metadata: _T, k_scale: float = 1.0,
                    v_scale: float = 1.0,
                    attn_type: AttentionType = AttentionType.DECODER) -> torch.Tensor:
            ...

    class DecodeOnlyAttentionBackend(AttentionBackend, ABC):

        @staticmethod
        @abstractmethod
        def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                               head_size: int) -> Tuple[int, ...]:
            ...

        @staticmethod
        @abstractmethod
        def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                        src_to_dst: torch.Tensor) -> None:
            ...

        @staticmethod
        @abstractmethod
        def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
            ...

    @dataclass
    class DecodeOnlyAttentionMetadata(AttentionMetadata):
        num_prefills: int
        num_prefill_tokens: int
        num_decode_tokens: int
        slot_mapping: torch.Tensor

        @property
        @abstractmethod
        def prefill_metadata(self):
            ...

        @property
        @abstractmethod
        def decode_metadata(self):
            ...

        def asdict_zerocopy(self, skip_fields: Optional[Set[str]] = None) -> Dict[str, Any]:
            skip = skip_fields or set()
            return {f.name: getattr(self, f.name) for f in fields(self) if f.name not in skip}

    class DecodeOnlyAttentionMetadataBuilder(AttentionMetadataBuilder, ABC, Generic[_T]):

        @abstractmethod
        def __init__(self, input_builder) -> None:
            ...

        @abstractmethod
        def build(self, seq_lens: List[int], query_lens: List[int], cuda_graph_pad_size: int,
                  batch_size: int) -> _T:
            ...

    class DecodeOnlyAttentionImpl(AttentionImpl, ABC, Generic[_T]):
        pass
