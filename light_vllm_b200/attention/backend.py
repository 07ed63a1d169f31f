"""``PagedB200AttentionBackend``: the decode attention plugin over the B200 paged-KV ops.

This is the backend the reference lost: light-vllm's only live decode
backend is FlashAttention (decoding/backends/attention/backends/flash_attn.py:20-571); its
``PagedAttention`` helper is orphaned.  The classes here satisfy the same plugin ABI
(decoding/backends/attention/backends/abstract.py:15-166) and route
  * cache writes  -> reshape_and_cache   (paged layout, optional fused fp8-e4m3 quantisation)
  * decode        -> paged_attention_v1 / v2
  * swap / copy   -> swap_blocks / copy_blocks
through ``light_vllm_b200._custom_ops``.  Prompt (prefill) tokens -- with or without a cached prefix,
with the model's sliding window or ALiBi slopes -- run the package's own sm_100a
``context_attention_fwd`` kernel (the reference calls flash_attn_varlen_func /
PagedAttention.forward_prefix there).
"""
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple, Type

import torch

from light_vllm_b200.attention.abstract import (AttentionType, DecodeOnlyAttentionBackend,
                                                DecodeOnlyAttentionImpl,
                                                DecodeOnlyAttentionMetadata,
                                                DecodeOnlyAttentionMetadataBuilder)
from light_vllm_b200.attention.paged_attn import PagedAttention
from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
from light_vllm_b200.attention.utils import (compute_slot_mapping,
                                             compute_slot_mapping_start_idx,
                                             is_block_tables_empty, make_tensor_with_pad,
                                             pin_memory_available)


class PagedB200AttentionBackend(DecodeOnlyAttentionBackend):

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return PagedAttention.get_supported_head_sizes()

    @staticmethod
    def get_name() -> str:
        return "paged-b200"

    @staticmethod
    def get_impl_cls() -> Type["PagedB200AttentionImpl"]:
        return PagedB200AttentionImpl

    @staticmethod
    def get_metadata_cls() -> Type["PagedB200AttentionMetadata"]:
        return PagedB200AttentionMetadata

    @staticmethod
    def get_builder_cls() -> Type["PagedB200AttentionMetadataBuilder"]:
        return PagedB200AttentionMetadataBuilder

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        if block_size not in (8, 16, 32):
            raise ValueError("Block size must be 8, 16 or 32 for the paged KV layout.")
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                    src_to_dst: torch.Tensor) -> None:
        PagedAttention.swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        PagedAttention.copy_blocks(kv_caches, src_to_dists)


@dataclass
class PagedB200AttentionMetadata(DecodeOnlyAttentionMetadata):
    """Same fields as the reference's decode metadata (flash_attn.py:77-137) so the input builder
    and model runner need no change."""
    seq_lens: Optional[List[int]]
    seq_lens_tensor: Optional[torch.Tensor]
    max_query_len: Optional[int]
    max_prefill_seq_len: int
    max_decode_seq_len: int
    query_start_loc: Optional[torch.Tensor]
    seq_start_loc: Optional[torch.Tensor]
    context_lens_tensor: Optional[torch.Tensor]
    block_tables: Optional[torch.Tensor]
    use_cuda_graph: bool
    _cached_prefill_metadata: Optional["PagedB200AttentionMetadata"] = None
    _cached_decode_metadata: Optional["PagedB200AttentionMetadata"] = None

    @property
    def prefill_metadata(self) -> Optional["PagedB200AttentionMetadata"]:
        if self.num_prefills == 0:
            return None
        if self._cached_prefill_metadata is None:
            n = self.num_prefills
            self._cached_prefill_metadata = PagedB200AttentionMetadata(
                num_prefills=n, num_prefill_tokens=self.num_prefill_tokens, num_decode_tokens=0,
                slot_mapping=self.slot_mapping[:self.num_prefill_tokens],
                seq_lens=self.seq_lens[:n], seq_lens_tensor=self.seq_lens_tensor[:n],
                max_query_len=self.max_query_len, max_prefill_seq_len=self.max_prefill_seq_len,
                max_decode_seq_len=0, query_start_loc=self.query_start_loc[:n + 1],
                seq_start_loc=self.seq_start_loc[:n + 1],
                context_lens_tensor=self.context_lens_tensor[:n],
                block_tables=self.block_tables[:n], use_cuda_graph=False)
        return self._cached_prefill_metadata

    @property
    def decode_metadata(self) -> Optional["PagedB200AttentionMetadata"]:
        if self.num_decode_tokens == 0:
            return None
        if self._cached_decode_metadata is None:
            self._cached_decode_metadata = PagedB200AttentionMetadata(
                num_prefills=0, num_prefill_tokens=0, num_decode_tokens=self.num_decode_tokens,
                slot_mapping=self.slot_mapping[self.num_prefill_tokens:], seq_lens=None,
                seq_lens_tensor=self.seq_lens_tensor[self.num_prefills:], max_query_len=None,
                max_prefill_seq_len=0, max_decode_seq_len=self.max_decode_seq_len,
                query_start_loc=None, seq_start_loc=None, context_lens_tensor=None,
                block_tables=self.block_tables[self.num_prefills:],
                use_cuda_graph=self.use_cuda_graph)
        return self._cached_decode_metadata

    def to(self, device, non_blocking=True):
        for k, v in self.__dict__.items():
            if isinstance(v, torch.Tensor):
                self.__dict__[k] = v.to(device=device, non_blocking=non_blocking)
        self._cached_prefill_metadata = None
        self._cached_decode_metadata = None
        return self


class PagedB200AttentionMetadataBuilder(
        DecodeOnlyAttentionMetadataBuilder[PagedB200AttentionMetadata]):
    """Collects per-sequence-group data from the model-input builder and emits pinned CPU tensors
    (int64 slot_mapping, int32 block_tables padded with 0, int32 seq_lens), as
    flash_attn.py:209-365 does; the worker moves them to the GPU with ``.to(non_blocking=True)``.
    """

    def __init__(self, input_builder):
        self.slot_mapping: List[int] = []
        self.prefill_seq_lens: List[int] = []
        self.context_lens: List[int] = []
        self.block_tables: List[List[int]] = []
        self.curr_seq_lens: List[int] = []
        self.num_prefills = 0
        self.num_prefill_tokens = 0
        self.num_decode_tokens = 0
        self.input_builder = input_builder
        self.sliding_window = input_builder.sliding_window
        self.block_size = input_builder.block_size
        self.use_v2_block_manager = input_builder.scheduler_config.use_v2_block_manager

    def _add_seq_group(self, inter_data, chunked_prefill_enabled: bool, prefix_cache_hit: bool):
        is_prompt = inter_data.is_prompt
        block_tables = inter_data.block_tables
        for (seq_id, token_len, seq_len, curr_seq_len, query_len, context_len,
             curr_sliding_window_block) in zip(
                 inter_data.seq_ids, [len(t) for t in inter_data.input_tokens],
                 inter_data.orig_seq_lens, inter_data.seq_lens, inter_data.query_lens,
                 inter_data.context_lens, inter_data.curr_sliding_window_blocks):
            self.context_lens.append(context_len)
            if is_prompt:
                self.num_prefills += 1
                self.num_prefill_tokens += token_len
                self.prefill_seq_lens.append(seq_len)
            else:
                assert query_len == 1, f"seq_len: {seq_len}, context_len: {context_len}, " \
                                       f"query_len: {query_len}"
                self.num_decode_tokens += query_len
                self.curr_seq_lens.append(curr_seq_len)
            table: List[int] = []
            if prefix_cache_hit:
                table = block_tables[seq_id]
            elif (chunked_prefill_enabled or not is_prompt) and block_tables is not None:
                table = block_tables[seq_id][-curr_sliding_window_block:]
            self.block_tables.append(table)
            start_idx = compute_slot_mapping_start_idx(is_prompt, query_len, context_len,
                                                       self.sliding_window,
                                                       self.use_v2_block_manager)
            compute_slot_mapping(is_block_tables_empty(block_tables), self.slot_mapping, seq_id,
                                 seq_len, context_len, start_idx, self.block_size, block_tables)

    def build(self, seq_lens: List[int], query_lens: List[int], cuda_graph_pad_size: int,
              batch_size: int) -> PagedB200AttentionMetadata:
        groups = self.input_builder.inter_data_list
        prefix_cache_hit = any(g.prefix_cache_hit for g in groups)
        for g in groups:
            self._add_seq_group(g, self.input_builder.chunked_prefill_enabled, prefix_cache_hit)
        max_query_len = max(query_lens)
        assert max_query_len > 0, f"query_lens: {query_lens}"
        pin = pin_memory_available()

        def cpu(data, dtype):
            t = torch.tensor(data, dtype=dtype, device="cpu")
            return t.pin_memory() if pin else t

        seq_lens_tensor = cpu(seq_lens, torch.int)
        query_lens_tensor = cpu(query_lens, torch.long)
        query_start_loc = torch.zeros(len(query_lens) + 1, dtype=torch.int32, pin_memory=pin)
        seq_start_loc = torch.zeros(len(seq_lens) + 1, dtype=torch.int32, pin_memory=pin)
        torch.cumsum(seq_lens_tensor, dim=0, dtype=torch.int32, out=seq_start_loc[1:])
        torch.cumsum(query_lens_tensor, dim=0, dtype=torch.int32, out=query_start_loc[1:])
        return PagedB200AttentionMetadata(
            num_prefills=self.num_prefills, slot_mapping=cpu(self.slot_mapping, torch.long),
            num_prefill_tokens=self.num_prefill_tokens, num_decode_tokens=self.num_decode_tokens,
            seq_lens=seq_lens, seq_lens_tensor=seq_lens_tensor, max_query_len=max_query_len,
            max_prefill_seq_len=max(self.prefill_seq_lens, default=0),
            max_decode_seq_len=max(self.curr_seq_lens, default=0), query_start_loc=query_start_loc,
            seq_start_loc=seq_start_loc, context_lens_tensor=cpu(self.context_lens, torch.int),
            block_tables=make_tensor_with_pad(self.block_tables, pad=0, dtype=torch.int,
                                              pin_memory=pin),
            use_cuda_graph=False)

    def __call__(self, *args, **kwargs):
        pass


class PagedB200AttentionImpl(DecodeOnlyAttentionImpl):
    """Token layout of the flattened batch (as in the reference, flash_attn.py:372-396):
    ``|<-- num_prefill_tokens -->|<-- num_decode_tokens -->|``; decode tokens have length 1."""

    def __init__(self, num_heads: int, head_size: int, scale: float,
                 num_kv_heads: Optional[int] = None, alibi_slopes: Optional[List[float]] = None,
                 sliding_window: Optional[int] = None, kv_cache_dtype: str = "auto",
                 blocksparse_params: Optional[Dict[str, Any]] = None,
                 logits_soft_cap: Optional[float] = None) -> None:
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.alibi_slopes = (torch.tensor(alibi_slopes, dtype=torch.float32)
                             if alibi_slopes is not None else None)
        # Sliding window is realised by the input builder truncating the block table to the last
        # N blocks (model_input_builder.py:281-308); the kernel needs nothing extra.
        self.sliding_window = sliding_window
        self.kv_cache_dtype = kv_cache_dtype
        if logits_soft_cap:
            raise ValueError("logits_soft_cap is not supported by the paged-KV decode kernels.")
        bs = blocksparse_params or {}
        self.blocksparse = (int(bs.get("tp_rank", 0)), int(bs.get("local_blocks", 0)),
                            int(bs.get("vert_stride", 0)), int(bs.get("block_size", 64)),
                            int(bs.get("head_sliding_step", 0)))
        assert self.num_heads % self.num_kv_heads == 0
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        supported = PagedAttention.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(f"Head size {head_size} is not supported by PagedAttention. "
                             f"Supported head sizes are: {supported}.")

    def _prefill(self, query, key, value, meta) -> torch.Tensor:
        """Causal self-attention over freshly computed prompt K/V (no cached prefix).

        The reference calls ``flash_attn_varlen_func(causal=True, window_size=self.sliding_window,
        alibi_slopes=...)`` here (flash_attn.py:522-535) with ``window_size = (W, W)``
        (flash_attn.py:419-420), i.e. query i sees keys ``i - W <= j <= i``.  Served by the package's
        own sm_100a prefix-prefill kernel with an empty context (``ctx_len = 0``): its window keeps
        ``i - j < w`` (prefix_prefill.py:130-143), so ``w = W + 1`` reproduces the flash semantics.
        ALiBi *and* a window together (no model of the reference's zoo uses both; the Triton kernel
        drops the window in that case, prefix_prefill.py:712-760) go through a banded-mask SDPA."""
        n_seq = meta.num_prefills
        W = self.sliding_window
        if self.alibi_slopes is not None and W:
            return self._prefill_masked_sdpa(query, key, value, meta)
        dev = query.device
        if meta.seq_lens_tensor is not None:
            lens = meta.seq_lens_tensor[:n_seq]
        else:
            lens = torch.tensor(meta.seq_lens[:n_seq], dtype=torch.int32, device=dev)
        lens = lens.to(device=dev, dtype=torch.int32)
        if meta.query_start_loc is not None:
            starts = meta.query_start_loc[:n_seq].to(device=dev, dtype=torch.int32)
        else:
            starts = (torch.cumsum(lens, 0, dtype=torch.int32) - lens)
        max_len = meta.max_prefill_seq_len or meta.max_query_len or int(lens.max())
        alibi = self.alibi_slopes
        if alibi is not None and alibi.device != dev:
            alibi = self.alibi_slopes = alibi.to(dev)
        # no cached prefix: one dummy block, never dereferenced (context length 0 for every sequence)
        x = 16 // query.element_size()
        kc = torch.empty(1, self.num_kv_heads, self.head_size // x, 16, x, dtype=query.dtype, device=dev)
        vc = torch.empty(1, self.num_kv_heads, self.head_size, 16, dtype=query.dtype, device=dev)
        out = torch.empty_like(query)
        context_attention_fwd(query, key, value, out, kc, vc,
                              torch.zeros(n_seq, 1, dtype=torch.int32, device=dev), starts, lens,
                              torch.zeros(n_seq, dtype=torch.int32, device=dev), int(max_len), alibi,
                              (W + 1) if W else None)
        return out

    def _prefill_masked_sdpa(self, query, key, value, meta) -> torch.Tensor:
        import torch.nn.functional as F
        out = torch.empty_like(query)
        starts = meta.seq_start_loc.tolist() if meta.seq_lens is None else None
        slopes = self.alibi_slopes.to(query.device)
        W = self.sliding_window
        pos = 0
        for i in range(meta.num_prefills):
            n = meta.seq_lens[i] if meta.seq_lens is not None else starts[i + 1] - starts[i]
            q = query[pos:pos + n].transpose(0, 1).unsqueeze(0)          # [1, H, n, D]
            k = key[pos:pos + n].transpose(0, 1).unsqueeze(0)            # [1, KH, n, D]
            v = value[pos:pos + n].transpose(0, 1).unsqueeze(0)
            idx = torch.arange(n, device=query.device)
            rel = (idx[None, :] - idx[:, None]).to(torch.float32)         # j - i
            bias = slopes[:, None, None] * rel[None]
            bias = bias.masked_fill((rel[None] > 0) | (rel[None] < -W), float("-inf")).to(query.dtype)
            o = F.scaled_dot_product_attention(q, k, v, attn_mask=bias.unsqueeze(0), scale=self.scale,
                                               enable_gqa=self.num_queries_per_kv > 1)
            out[pos:pos + n] = o.squeeze(0).transpose(0, 1)
            pos += n
        return out

    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                kv_cache: Optional[torch.Tensor], attn_metadata: PagedB200AttentionMetadata,
                k_scale: float = 1.0, v_scale: float = 1.0,
                attn_type: AttentionType = AttentionType.DECODER) -> torch.Tensor:
        """query [T, H*D], key/value [T, KH*D] (slices of the fused qkv output are fine),
        kv_cache [2, num_blocks, block_size*KH*D] or None (memory-profiling run) -> [T, H*D]."""
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("Encoder self-attention and encoder/decoder cross-attention "
                                      "are not implemented for PagedB200AttentionImpl")
        num_tokens, hidden_size = query.shape
        query = query.view(-1, self.num_heads, self.head_size)
        key = key.view(-1, self.num_kv_heads, self.head_size)
        value = value.view(-1, self.num_kv_heads, self.head_size)
        key_cache = value_cache = None
        if kv_cache is not None:
            key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads,
                                                                   self.head_size)
            # write-then-read on one stream is part of the contract
            PagedAttention.write_to_paged_cache(key, value, key_cache, value_cache,
                                                attn_metadata.slot_mapping, self.kv_cache_dtype,
                                                k_scale, v_scale)
        n_prefill = attn_metadata.num_prefill_tokens
        n_decode = attn_metadata.num_decode_tokens
        assert key.shape[0] == n_prefill + n_decode and value.shape[0] == n_prefill + n_decode
        decode_meta = attn_metadata.decode_metadata
        if n_prefill == 0 and decode_meta is not None:  # pure decode step: the hot path
            return self._decode(query, key_cache, value_cache, decode_meta, k_scale,
                                v_scale).view(num_tokens, hidden_size)
        output = torch.empty_like(query)
        if prefill_meta := attn_metadata.prefill_metadata:
            if (kv_cache is None or prefill_meta.block_tables is None
                    or prefill_meta.block_tables.numel() == 0):
                output[:n_prefill] = self._prefill(query[:n_prefill], key[:n_prefill],
                                                   value[:n_prefill], prefill_meta)
            else:
                # prefix-cache hit / chunked prefill: new tokens attend the cached context through
                # the block table plus themselves causally (reference: xformers.py / rocm / ipex
                # backends -> PagedAttention.forward_prefix, paged_attn.py:193-225)
                if self.kv_cache_dtype != "auto":
                    raise NotImplementedError(
                        "prefix-cached prefill reads the cache in the query dtype; the reference "
                        "kernel (prefix_prefill.py) has no fp8 path either")
                alibi = self.alibi_slopes
                if alibi is not None and alibi.device != query.device:
                    alibi = self.alibi_slopes = alibi.to(query.device)
                output[:n_prefill] = PagedAttention.forward_prefix(
                    query[:n_prefill], key[:n_prefill], value[:n_prefill], key_cache, value_cache,
                    prefill_meta.block_tables, prefill_meta.query_start_loc,
                    prefill_meta.seq_lens_tensor, prefill_meta.context_lens_tensor,
                    prefill_meta.max_query_len, alibi, self.sliding_window)
        if decode_meta is not None:
            output[n_prefill:] = self._decode(query[n_prefill:], key_cache, value_cache,
                                              decode_meta, k_scale, v_scale)
        return output.view(num_tokens, hidden_size)

    def _decode(self, q, key_cache, value_cache, meta, k_scale, v_scale) -> torch.Tensor:
        alibi = self.alibi_slopes
        if alibi is not None and alibi.device != q.device:
            alibi = self.alibi_slopes = alibi.to(q.device)
        return PagedAttention.forward_decode(q, key_cache, value_cache, meta.block_tables,
                                             meta.seq_lens_tensor, meta.max_decode_seq_len,
                                             self.kv_cache_dtype, self.num_kv_heads, self.scale,
                                             alibi, k_scale, v_scale, *self.blocksparse)
