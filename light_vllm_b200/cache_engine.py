"""KV pool owner: per-layer GPU pools, a pinned-CPU swap pool, and the swap / copy entry points.

Behavioural spec: light_vllm/decoding/worker/cache_engine.py:15-103 (``CacheEngine``) and the
caller ``Worker.non_blocking_h2d`` (decoding/worker/gpu_worker.py:206-221).  The constructor takes
plain numbers instead of light-vllm's Config objects so that it is usable stand-alone;
``from_configs`` adapts the reference's ``CacheConfig`` / ``ModelConfig`` / ``DeviceConfig``.

B200 sizing: one layer pool is ``2 * num_blocks * block_size * num_kv_heads * head_size`` elements;
with 180 GB of HBM3e a Llama-3-8B-shaped model (32 layers, 8 kv heads x 128) holds ~1.2 M tokens
of bf16 KV next to its weights, so pools are plain ``torch.zeros`` tensors (no paging of the pool
itself) and block ids are GPU-local.
"""
from typing import List, Optional

import torch

from light_vllm_b200.attention.backend import PagedB200AttentionBackend

STR_DTYPE_TO_TORCH_DTYPE = {
    "half": torch.half, "bfloat16": torch.bfloat16, "float": torch.float,
    "fp8": torch.uint8, "fp8_e4m3": torch.uint8, "fp8_e5m2": torch.uint8,
}  # light_vllm/utils.py:37-44


class CacheEngine:

    def __init__(self, num_layers: int, num_kv_heads: int, head_size: int, block_size: int,
                 num_gpu_blocks: int, num_cpu_blocks: int, dtype: torch.dtype,
                 cache_dtype: str = "auto", device: str = "cuda",
                 attn_backend=PagedB200AttentionBackend) -> None:
        self.num_attention_layers = num_layers
        self.num_kv_heads = num_kv_heads
        self.head_size = head_size
        self.block_size = block_size
        self.num_gpu_blocks = num_gpu_blocks
        self.num_cpu_blocks = num_cpu_blocks
        self.cache_dtype = cache_dtype
        self.dtype = dtype if cache_dtype == "auto" else STR_DTYPE_TO_TORCH_DTYPE[cache_dtype]
        self.attn_backend = attn_backend
        self.device = device
        self.gpu_cache = self._allocate_kv_cache(num_gpu_blocks, device)
        self.cpu_cache = self._allocate_kv_cache(num_cpu_blocks, "cpu")

    @classmethod
    def from_configs(cls, cache_config, model_config, device_config, attn_backend) -> "CacheEngine":
        """Adapter for the reference's constructor signature (cache_engine.py:23-49)."""
        return cls(model_config.get_num_attention_layers(), model_config.get_num_kv_heads(),
                   model_config.get_head_size(), cache_config.block_size,
                   cache_config.num_gpu_blocks, cache_config.num_cpu_blocks, model_config.dtype,
                   cache_config.cache_dtype, device_config.device_type, attn_backend)

    def _allocate_kv_cache(self, num_blocks: int, device: str) -> List[torch.Tensor]:
        shape = self.attn_backend.get_kv_cache_shape(num_blocks, self.block_size, self.num_kv_heads,
                                                     self.head_size)
        pin = device == "cpu" and torch.cuda.is_available()
        # zero-filled like the reference (the allocator's null block must read as zeros)
        return [torch.zeros(shape, dtype=self.dtype, pin_memory=pin, device=device)
                for _ in range(self.num_attention_layers)]

    # -- the three operations the worker issues before each model step ---------------------
    def swap_in(self, src_to_dst: torch.Tensor) -> None:
        for cpu, gpu in zip(self.cpu_cache, self.gpu_cache):
            self.attn_backend.swap_blocks(cpu, gpu, src_to_dst)

    def swap_out(self, src_to_dst: torch.Tensor) -> None:
        for cpu, gpu in zip(self.cpu_cache, self.gpu_cache):
            self.attn_backend.swap_blocks(gpu, cpu, src_to_dst)

    def copy(self, src_to_dsts: torch.Tensor) -> None:
        self.attn_backend.copy_blocks(self.gpu_cache, src_to_dsts)

    def apply(self, blocks_to_swap_in: Optional[torch.Tensor] = None,
              blocks_to_swap_out: Optional[torch.Tensor] = None,
              blocks_to_copy: Optional[torch.Tensor] = None) -> None:
        """What ``Worker.non_blocking_h2d`` does with a step's worker input: swap in, swap out,
        copy -- in that order, all on the current stream, nothing synchronises."""
        if blocks_to_swap_in is not None and blocks_to_swap_in.numel() > 0:
            self.swap_in(blocks_to_swap_in)
        if blocks_to_swap_out is not None and blocks_to_swap_out.numel() > 0:
            self.swap_out(blocks_to_swap_out)
        if blocks_to_copy is not None and blocks_to_copy.numel() > 0:
            self.copy(blocks_to_copy)

    @staticmethod
    def get_cache_block_size(block_size: int, num_kv_heads: int, head_size: int, num_layers: int,
                             dtype: torch.dtype) -> int:
        """Bytes of one block across all layers, K and V (cache_engine.py:86-103)."""
        per_layer = 2 * block_size * num_kv_heads * head_size
        return num_layers * per_layer * torch.empty((), dtype=dtype).element_size()
