"""TEST INFRASTRUCTURE ONLY -- ctypes front-end of oracle/paged_oracle.c.

The functions take CPU torch tensors and have the argument order of the reference ops
(csrc/torch_bindings.cpp:24-48,227-265) so parity tests can call oracle and product
interchangeably.  Built by ``make -C oracle oracle`` (or ``__graft_entry__.build()``) into
``oracle/_build/liboracle.so``.
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_KV = {"auto": 0, "fp8": 1, "fp8_e4m3": 1, "fp8_e5m2": 2}
PARTITION_SIZE = 512


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "paged_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-std=c11", "-O2", "-fPIC", "-shared", "-fopenmp",
                               "-ffp-contract=off", "-o", _SO, src, "-lm"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_fp8_encode.argtypes = [ctypes.c_float, ctypes.c_int]
        _lib.oracle_fp8_encode.restype = ctypes.c_uint8
        _lib.oracle_fp8_decode.argtypes = [ctypes.c_uint8, ctypes.c_int]
        _lib.oracle_fp8_decode.restype = ctypes.c_float
    return _lib


def _p(t):
    if t is None:
        return ctypes.c_void_p(0)
    assert t.device.type == "cpu", "the oracle runs on CPU tensors"
    return ctypes.c_void_p(t.data_ptr())


def _i64(v):
    return ctypes.c_int64(int(v))


def _f(v):
    return ctypes.c_float(float(v))


def _bs5(tp_rank, local_blocks, vert_stride, block_size, head_sliding_step):
    return (ctypes.c_int32 * 5)(tp_rank, local_blocks, vert_stride, block_size, head_sliding_step)


def _attn_common(query, key_cache, block_tables):
    assert query.stride(2) == 1 and query.stride(1) == query.size(2)
    return (query.size(0), query.size(1), query.size(2), block_tables.size(1), query.stride(0),
            key_cache.stride(0), key_cache.stride(1))


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                       seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
                       v_scale, tp_rank=0, blocksparse_local_blocks=0, blocksparse_vert_stride=0,
                       blocksparse_block_size=64, blocksparse_head_sliding_step=0):
    B, H, D, maxb, qs, kbs, khs = _attn_common(query, key_cache, block_tables)
    assert out.is_contiguous() and block_tables.dtype == torch.int32 and seq_lens.dtype == torch.int32
    rc = lib().oracle_paged_attention_v1(
        _p(out), _p(query), _p(key_cache), _p(value_cache), _p(block_tables), _p(seq_lens),
        _p(alibi_slopes), B, H, int(num_kv_heads), D, int(block_size), int(max_seq_len), maxb,
        _i64(qs), _i64(kbs), _i64(khs), _f(scale), _f(k_scale), _f(v_scale), _DT[query.dtype],
        _KV[kv_cache_dtype],
        _bs5(tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
             blocksparse_head_sliding_step))
    assert rc == 0


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
                       alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank=0,
                       blocksparse_local_blocks=0, blocksparse_vert_stride=0,
                       blocksparse_block_size=64, blocksparse_head_sliding_step=0):
    B, H, D, maxb, qs, kbs, khs = _attn_common(query, key_cache, block_tables)
    assert exp_sums.is_contiguous() and max_logits.is_contiguous() and tmp_out.is_contiguous()
    rc = lib().oracle_paged_attention_v2(
        _p(out), _p(exp_sums), _p(max_logits), _p(tmp_out), _p(query), _p(key_cache),
        _p(value_cache), _p(block_tables), _p(seq_lens), _p(alibi_slopes), B, H, int(num_kv_heads),
        D, int(block_size), int(max_seq_len), maxb, int(exp_sums.size(2)), _i64(qs), _i64(kbs),
        _i64(khs), _f(scale), _f(k_scale), _f(v_scale), _DT[query.dtype], _KV[kv_cache_dtype],
        _bs5(tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
             blocksparse_head_sliding_step))
    assert rc == 0


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale,
                      v_scale):
    assert key_cache.is_contiguous() and value_cache.is_contiguous() and slot_mapping.dtype == torch.int64
    rc = lib().oracle_reshape_and_cache(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping), key.size(0),
        key.size(1), key.size(2), key_cache.size(3), key_cache.size(4), _i64(key.stride(0)),
        _i64(value.stride(0)), _f(k_scale), _f(v_scale), _DT[key.dtype], _KV[kv_cache_dtype])
    assert rc == 0


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype,
                            k_scale, v_scale):
    assert key_cache.stride(0) == value_cache.stride(0)
    rc = lib().oracle_reshape_and_cache_flash(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping), key.size(0),
        key.size(1), key.size(2), key_cache.size(1), _i64(key_cache.stride(0)),
        _i64(key.stride(0)), _i64(value.stride(0)), _f(k_scale), _f(v_scale), _DT[key.dtype],
        _KV[kv_cache_dtype])
    assert rc == 0


def copy_blocks(key_caches, value_caches, block_mapping):
    n = len(key_caches)
    assert n == len(value_caches)
    if n == 0 or block_mapping.numel() == 0:
        return
    kp = (ctypes.c_void_p * n)(*[t.data_ptr() for t in key_caches])
    vp = (ctypes.c_void_p * n)(*[t.data_ptr() for t in value_caches])
    bm = block_mapping.contiguous()
    block_bytes = key_caches[0].element_size() * key_caches[0][0].numel()
    rc = lib().oracle_copy_blocks(kp, vp, n, _p(bm), bm.size(0), _i64(block_bytes))
    assert rc == 0


def swap_blocks(src, dst, block_mapping):
    bm = block_mapping.contiguous()
    if bm.numel() == 0:
        return
    block_bytes = src.element_size() * src[0].numel()
    rc = lib().oracle_swap_blocks(_p(src), _p(dst), _p(bm), bm.size(0), _i64(block_bytes))
    assert rc == 0


def convert_fp8(output, input, scale=1.0, kv_dtype="fp8"):
    to_fp8 = input.element_size() > 1
    wide = input.dtype if to_fp8 else output.dtype
    rc = lib().oracle_convert_fp8(_p(output), _p(input), _i64(input.numel()), _f(scale), _DT[wide],
                                  _KV[kv_dtype], 1 if to_fp8 else 0)
    assert rc == 0


# ---------------------------------------------------------------- per-step ops (8f) ----
def advance_step(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids,
                 input_positions, seq_lens, slot_mapping, block_tables):
    assert input_tokens.dtype == torch.int64 and seq_lens.dtype == torch.int32
    assert block_tables.dtype == torch.int32 and block_tables.is_contiguous()
    rc = lib().oracle_advance_step(
        int(num_seqs), int(num_queries), int(block_size), _p(input_tokens), _p(sampled_token_ids),
        _p(input_positions), _p(seq_lens), _p(slot_mapping), _p(block_tables),
        _i64(block_tables.stride(0)))
    assert rc == 0


def rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim=None,
                     cos_sin_cache_offsets=None):
    """rotary_embedding / batched_rotary_embedding (with offsets) of the reference, in place."""
    num_tokens = query.numel() // query.shape[-1]
    assert positions.dtype == torch.int64 and query.stride(-1) == 1 and key.stride(-1) == 1
    rc = lib().oracle_rotary_embedding(
        _p(positions), _p(query), _p(key), _p(cos_sin_cache), _p(cos_sin_cache_offsets),
        int(num_tokens), query.shape[-1] // head_size, key.shape[-1] // head_size, int(head_size),
        int(rot_dim if rot_dim is not None else cos_sin_cache.shape[1]), _i64(query.stride(-2)),
        _i64(key.stride(-2)), 1 if is_neox else 0, _DT[query.dtype])
    assert rc == 0


def batched_rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim,
                             cos_sin_cache_offsets):
    rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim,
                     cos_sin_cache_offsets)


def context_attention_fwd(q, k, v, o, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len,
                          max_input_len, alibi_slopes=None, sliding_window=None):
    """Argument order of the reference's ``context_attention_fwd``
    (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-686)."""
    for t in (q, k, v, o):
        assert t.stride(2) == 1
    assert b_loc.dtype == torch.int32 and b_seq_len.dtype == torch.int32 and b_ctx_len.dtype == torch.int32
    start = b_start_loc.to(torch.int32).contiguous()
    rc = lib().oracle_context_attention_fwd(
        _p(q), _p(k), _p(v), _p(o), _p(k_cache), _p(v_cache), _p(b_loc), _p(start), _p(b_seq_len),
        _p(b_ctx_len), b_seq_len.numel(), q.size(1), k.size(1), q.size(2), v_cache.size(3),
        b_loc.size(1), _i64(q.stride(0)), _i64(q.stride(1)), _i64(k.stride(0)), _i64(k.stride(1)),
        _i64(v.stride(0)), _i64(v.stride(1)), _i64(o.stride(0)), _i64(o.stride(1)),
        _i64(k_cache.stride(0)), _i64(k_cache.stride(1)), _i64(v_cache.stride(0)),
        _i64(v_cache.stride(1)), _p(alibi_slopes),
        int(sliding_window) if sliding_window and sliding_window > 0 else 0, _DT[q.dtype])
    assert rc == 0
