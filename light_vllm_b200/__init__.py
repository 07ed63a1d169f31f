"""light_vllm_b200: B200-native (sm_100a) drop-in for light-vllm's decode hot path.

Paged attention over the block-table KV cache (``paged_attention_v1/v2``) plus the cache ops
that feed it (``reshape_and_cache[_flash]``, ``copy_blocks``, ``swap_blocks``, ``convert_fp8``),
behind the reference's own operator surface and attention plugin.  See DESIGN.md.
"""
__version__ = "0.1.0"
