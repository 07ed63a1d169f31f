"""CPU: host-side logic of the attention plugin (no kernels run): KV pool shapes / views, the
v1-v2 heuristic, slot mapping, the metadata builder, metadata views, the cache engine's sizes."""
import types

import pytest
import torch

from light_vllm_b200.attention import (AttentionType, PagedAttention, PagedB200AttentionBackend,
                                       PagedB200AttentionImpl, PagedB200AttentionMetadata,
                                       PagedB200AttentionMetadataBuilder)
from light_vllm_b200.attention.utils import (PAD_SLOT_ID, compute_slot_mapping,
                                             compute_slot_mapping_start_idx, make_tensor_with_pad)


def test_backend_statics():
    B = PagedB200AttentionBackend
    assert B.get_name() == "paged-b200"
    assert B.get_impl_cls() is PagedB200AttentionImpl
    assert B.get_metadata_cls() is PagedB200AttentionMetadata
    assert B.get_builder_cls() is PagedB200AttentionMetadataBuilder
    assert B.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16 * 8 * 128)   # paged_attn.py:41-47
    assert B.get_supported_head_sizes() == [64, 80, 96, 112, 120, 128, 192, 256]
    with pytest.raises(ValueError):
        B.get_kv_cache_shape(10, 24, 8, 128)
    assert B(AttentionType.DECODER).attn_type == AttentionType.DECODER


@pytest.mark.parametrize("dtype,x", [(torch.float16, 8), (torch.uint8, 16), (torch.float32, 4)])
def test_split_kv_cache_views(dtype, x):
    kv = torch.zeros(PagedAttention.get_kv_cache_shape(5, 16, 2, 64), dtype=dtype)
    k, v = PagedAttention.split_kv_cache(kv, 2, 64)
    assert k.shape == (5, 2, 64 // x, 16, x) and v.shape == (5, 2, 64, 16)
    assert k.data_ptr() == kv[0].data_ptr() and v.data_ptr() == kv[1].data_ptr()  # views, no copy
    assert k.stride(0) == 16 * 2 * 64 and k.stride(1) == 16 * 64


def test_v1_v2_heuristic_matches_reference():
    # paged_attn.py:126-128
    f = PagedAttention.use_v1
    assert f(512, 1, 1)             # single partition
    assert not f(513, 1, 32)        # 2 partitions, little parallelism -> v2
    assert f(4096, 256, 32)         # B*H > 512 -> v1
    assert not f(8193, 512, 32)     # > 8192 always v2
    assert f(8192, 17, 32) and not f(8192, 16, 32)


def test_slot_mapping():
    tables = {7: [5, 2, 9]}
    out = []
    compute_slot_mapping(False, out, 7, seq_len=35, context_len=30, start_idx=0, block_size=16,
                         block_tables=tables)
    assert out == [2 * 16 + 14, 2 * 16 + 15, 9 * 16 + 0, 9 * 16 + 1, 9 * 16 + 2]
    out = []
    compute_slot_mapping(True, out, 7, 4, 0, 0, 16, tables)
    assert out == [PAD_SLOT_ID] * 4
    # sliding window 8 over a 10-token prompt, block 4: first two tokens are not cached
    out = []
    start = compute_slot_mapping_start_idx(True, 10, 0, 8, False)
    compute_slot_mapping(False, out, 0, 10, 0, start, 4, {0: [0, 1, 2]})
    assert out == [-1, -1, 2, 3, 4, 5, 6, 7, 8, 9]
    t = make_tensor_with_pad([[1, 2, 3], [], [4]], pad=0, dtype=torch.int)
    assert t.tolist() == [[1, 2, 3], [0, 0, 0], [4, 0, 0]] and t.dtype == torch.int32


def _fake_input_builder(groups, block_size=16):
    return types.SimpleNamespace(sliding_window=None, block_size=block_size,
                                 scheduler_config=types.SimpleNamespace(use_v2_block_manager=False),
                                 chunked_prefill_enabled=False, inter_data_list=groups)


def _group(is_prompt, seq_id, tokens, seq_len, context_len, table):
    return types.SimpleNamespace(
        is_prompt=is_prompt, block_tables={seq_id: table} if table is not None else None,
        seq_ids=[seq_id], input_tokens=[tokens], orig_seq_lens=[seq_len], seq_lens=[seq_len],
        query_lens=[len(tokens)], context_lens=[context_len],
        curr_sliding_window_blocks=[len(table) if table else 0], prefix_cache_hit=False)


def test_metadata_builder_decode_batch():
    groups = [_group(False, 0, [11], 18, 17, [3, 8]), _group(False, 1, [12], 5, 4, [6])]
    b = PagedB200AttentionMetadataBuilder(_fake_input_builder(groups))
    m = b.build(seq_lens=[18, 5], query_lens=[1, 1], cuda_graph_pad_size=-1, batch_size=2)
    assert (m.num_prefills, m.num_prefill_tokens, m.num_decode_tokens) == (0, 0, 2)
    assert m.slot_mapping.tolist() == [8 * 16 + 1, 6 * 16 + 4] and m.slot_mapping.dtype == torch.int64
    assert m.block_tables.tolist() == [[3, 8], [6, 0]] and m.block_tables.dtype == torch.int32
    assert m.seq_lens_tensor.tolist() == [18, 5] and m.seq_lens_tensor.dtype == torch.int32
    assert m.max_decode_seq_len == 18 and m.max_prefill_seq_len == 0
    d = m.decode_metadata
    assert d is m.decode_metadata and m.prefill_metadata is None
    assert d.block_tables.shape == (2, 2) and d.max_decode_seq_len == 18


def test_metadata_builder_mixed_batch_and_views():
    groups = [_group(True, 0, [1, 2, 3], 3, 0, None), _group(False, 1, [9], 33, 32, [4, 5, 7])]
    groups[0].block_tables = {0: [2]}
    b = PagedB200AttentionMetadataBuilder(_fake_input_builder(groups))
    m = b.build(seq_lens=[3, 33], query_lens=[3, 1], cuda_graph_pad_size=-1, batch_size=2)
    assert (m.num_prefills, m.num_prefill_tokens, m.num_decode_tokens) == (1, 3, 1)
    assert m.slot_mapping.tolist() == [32, 33, 34, 7 * 16 + 0]
    assert m.query_start_loc.tolist() == [0, 3, 4] and m.seq_start_loc.tolist() == [0, 3, 36]
    p, d = m.prefill_metadata, m.decode_metadata
    assert p.slot_mapping.tolist() == [32, 33, 34] and p.seq_lens == [3]
    assert d.slot_mapping.tolist() == [7 * 16] and d.seq_lens_tensor.tolist() == [33]
    assert d.block_tables.tolist() == [[4, 5, 7]]
    assert "slot_mapping" in m.asdict_zerocopy()


def test_impl_constructor_contract():
    impl = PagedB200AttentionImpl(32, 128, 0.088, 8, None, None, "fp8",
                                  {"vert_stride": 4, "local_blocks": 2, "block_size": 64})
    assert impl.num_queries_per_kv == 4 and impl.blocksparse == (0, 2, 4, 64, 0)
    with pytest.raises(ValueError, match="Head size"):
        PagedB200AttentionImpl(8, 100, 1.0, 8, None, None, "auto")
    with pytest.raises(ValueError, match="logits_soft_cap"):
        PagedB200AttentionImpl(8, 128, 1.0, 8, None, None, "auto", None, 30.0)
    q = torch.zeros(2, 32 * 128)
    with pytest.raises(NotImplementedError):
        impl.forward(q, q, q, None, None, attn_type=AttentionType.ENCODER)


def test_cache_engine_block_size():
    from light_vllm_b200.cache_engine import CacheEngine
    # Llama-3-8B-shaped: 32 layers x (K+V) x 16 tokens x 8 heads x 128 x 2 B = 2 MiB per block
    assert CacheEngine.get_cache_block_size(16, 8, 128, 32, torch.bfloat16) == 2 * 1024 * 1024
    eng = CacheEngine(2, 2, 64, 16, 4, 3, torch.float16, "fp8", device="cpu")
    assert eng.dtype == torch.uint8 and len(eng.gpu_cache) == 2
    assert eng.gpu_cache[0].shape == (2, 4, 16 * 2 * 64) and eng.cpu_cache[1].shape == (2, 3, 2048)
    eng.apply(None, torch.empty(0, 2, dtype=torch.int64), None)  # empty worker input: no op calls


def test_selector_mirrors_reference(monkeypatch):
    """selector.py:37-75: AttnBackend.from_engine -> backend instance; unknown names raise."""
    import types
    from light_vllm_b200.attention.selector import AttnBackend, _Backend
    mc = types.SimpleNamespace(get_num_attention_heads=lambda: 32, get_head_size=lambda: 128,
                               get_num_kv_heads=lambda: 8, get_sliding_window=lambda: None,
                               dtype=torch.bfloat16)
    engine = types.SimpleNamespace(engine_config=types.SimpleNamespace(model_config=mc),
                                   workflow=types.SimpleNamespace(attn_type="DECODER"))
    monkeypatch.delenv("VLLM_ATTENTION_BACKEND", raising=False)
    backend = AttnBackend.from_engine(engine)
    assert backend.get_name() == PagedB200AttentionBackend.get_name()
    assert AttnBackend.which_attn_to_use(32, 128, 8, None, torch.float16) == _Backend.PAGED_B200
    monkeypatch.setenv("VLLM_ATTENTION_BACKEND", "FLASH_ATTN")
    with pytest.raises(ValueError, match="Invalid attention backend"):
        AttnBackend.which_attn_to_use(32, 128, 8, None, torch.float16)
    monkeypatch.setenv("VLLM_ATTENTION_BACKEND", "PAGED_B200")
    with pytest.raises(ValueError, match="Head size"):
        AttnBackend.which_attn_to_use(32, 100, 8, None, torch.float16)
