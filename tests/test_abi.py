"""CPU: the C-ABI library loads and exports every symbol include/b200_paged.h declares; the torch
binding registers the reference's op schemas; without a GPU the ops fail loudly (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200_paged.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from light_vllm_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 16
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in include/b200_paged.h but not exported"
    assert set(names) == set(_lib.ABI_SYMBOLS)
    assert _lib.lib.b200_abi_version() == 2


def test_parse_kv_cache_dtype():
    from light_vllm_b200 import _lib
    f = _lib.lib.b200_parse_kv_cache_dtype
    assert [f(s) for s in (b"auto", b"fp8", b"fp8_e4m3", b"fp8_e5m2", b"int8", b"")] == [0, 1, 1, 2, -1, -1]


def _sig(op):
    """'name(type arg, ...) -> ret' with alias annotations reduced to a bare '!' marker."""
    s = str(op.default._schema)
    return re.sub(r"\((?:\$\d+|a\d*)!(?: -> )?\)", "!", s)


def test_op_schemas_match_reference():
    """Schemas of csrc/torch_bindings.cpp:24-48,227-265, argument for argument."""
    from light_vllm_b200 import _custom_ops  # noqa: F401
    tail = ("Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale, "
            "Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len, "
            "Tensor? alibi_slopes, str kv_cache_dtype, float k_scale, float v_scale, int tp_rank, "
            "int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size, "
            "int blocksparse_head_sliding_step) -> ()")
    assert _sig(torch.ops._C.paged_attention_v1) == "_C::paged_attention_v1(Tensor! out, " + tail
    assert _sig(torch.ops._C.paged_attention_v2) == (
        "_C::paged_attention_v2(Tensor! out, Tensor exp_sums, Tensor max_logits, Tensor tmp_out, "
        + tail)
    c = torch.ops._C_cache_ops
    assert _sig(c.swap_blocks) == "_C_cache_ops::swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()"
    assert _sig(c.copy_blocks) == ("_C_cache_ops::copy_blocks(Tensor[]! key_caches, "
                                   "Tensor[]! value_caches, Tensor block_mapping) -> ()")
    for name in ("reshape_and_cache", "reshape_and_cache_flash"):
        assert _sig(getattr(c, name)) == (
            f"_C_cache_ops::{name}(Tensor key, Tensor value, Tensor! key_cache, "
            "Tensor! value_cache, Tensor slot_mapping, str kv_cache_dtype, float k_scale, "
            "float v_scale) -> ()")
    assert _sig(c.convert_fp8) == ("_C_cache_ops::convert_fp8(Tensor! dst_cache, Tensor src_cache, "
                                   "float scale, str kv_cache_dtype) -> ()")
    assert _custom_ops.is_custom_op_supported("_C::paged_attention_v1")


def test_no_cpu_fallback():
    """CPU tensors must not be served by anything: the ops exist for the CUDA key only."""
    from light_vllm_b200 import _custom_ops as ops
    k = torch.zeros(2, 1, 64)
    kc = torch.zeros(1, 1, 8, 16, 8)
    vc = torch.zeros(1, 1, 64, 16)
    with pytest.raises((NotImplementedError, RuntimeError)):
        ops.reshape_and_cache(k, k, kc, vc, torch.zeros(2, dtype=torch.int64), "auto", 1.0, 1.0)
    with pytest.raises((NotImplementedError, RuntimeError)):
        ops.paged_attention_v1(torch.zeros(1, 1, 64), torch.zeros(1, 1, 64), kc, vc, 1, 1.0,
                               torch.zeros(1, 1, dtype=torch.int32), torch.ones(1, dtype=torch.int32),
                               16, 1, None, "auto", 1.0, 1.0)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_c_abi_fails_loudly_without_gpu():
    from light_vllm_b200 import _lib
    buf = (ctypes.c_char * 1024)()
    rc = _lib.lib.b200_convert_fp8(buf, buf, ctypes.c_int64(16), ctypes.c_float(1.0), 1, 1, 1, 0, None)
    assert rc == 3  # B200_ERR_CUDA
    assert b"failed" in _lib.lib.b200_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "light_vllm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".inc", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.replace("the oracle", "").lower() or f == "Makefile", \
                    f"{f} mentions the oracle"
