"""Loader for the native libraries of the B200 paged-KV hot path.

Two shared objects live next to this file (built in-tree by ``csrc/Makefile`` or
``__graft_entry__.build()``):

* ``libb200_paged.so`` -- the C-ABI library declared in ``include/b200_paged.h`` (hand-written
  sm_100a kernels; no torch types cross that boundary);
* ``_C.so`` -- the thin torch binding that registers ``torch.ops._C``, ``torch.ops._C_cache_ops``
  and ``torch.ops._C_cuda_utils`` with the reference's schemas
  (reference: ``csrc/torch_bindings.cpp:24-48,227-280``) and forwards to the C ABI.

There is no CPU or pure-PyTorch fallback: if either library is missing, importing this module
raises, and every op raises without a CUDA device.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200_paged.so")
BINDING_PATH = os.path.join(_HERE, "_C.so")

# Every symbol include/b200_paged.h declares (checked by tests/test_abi.py).
ABI_SYMBOLS = (
    "b200_last_error",
    "b200_abi_version",
    "b200_parse_kv_cache_dtype",
    "b200_paged_attention_v1",
    "b200_paged_attention_v2",
    "b200_paged_attention_v2_static",
    "b200_reshape_and_cache",
    "b200_reshape_and_cache_flash",
    "b200_copy_blocks",
    "b200_swap_blocks",
    "b200_convert_fp8",
    "b200_advance_step",
    "b200_rotary_embedding",
    "b200_rotary_reshape_and_cache",
    "b200_get_device_attribute",
    "b200_get_max_shared_memory_per_block_device_attribute",
    "b200_paged_attention_plan",
    "b200_paged_attention_tuned",
    "b200_swap_blocks_mode",
    "b200_launch_count",
    "b200_ctx_exchange_bytes",
    "b200_paged_attention_ctx_shard",
    "b200_ctx_exchange_status",
    "b200_ctx_exchange_status_async",
    "b200_ctx_exchange_set_timeout",
    "b200_ctx_exchange_set_host_mirror",
    "b200_ctx_exchange_trace",
    "b200_ctx_exchange_clear_status",
    "b200_peer_alloc",
    "b200_peer_free",
    "b200_peer_export",
    "b200_peer_open",
    "b200_peer_close",
    "b200_context_attention_fwd",
    "b200_context_attention_fwd_path",
    "b200_context_attention_scratch_bytes",
)


class AttnPlan(ctypes.Structure):
    """Mirror of ``b200_attn_plan_t``."""
    _fields_ = [(n, ctypes.c_int32) for n in ("cluster_size", "num_warps", "num_stages",
                                                "smem_bytes", "ctas_per_sm", "grid_x", "grid_y",
                                                "grid_z")]


class AttnTuning(ctypes.Structure):
    """Mirror of ``b200_attn_tuning_t`` (per-call launch overrides; 0 = library default)."""
    _fields_ = [(n, ctypes.c_int32) for n in ("num_warps", "num_stages", "split", "head_rows")]


def _load():
    for p in (LIB_PATH, BINDING_PATH):
        if not os.path.exists(p):
            raise ImportError(
                f"{p} is missing: build the native extension first "
                "(`make -C light_vllm_b200/csrc` or `python -c 'import __graft_entry__ as g; "
                "g.build()'`). light_vllm_b200 has no CPU / PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_launch_count.restype = ctypes.c_int64
    lib.b200_get_device_attribute.restype = ctypes.c_int64
    lib.b200_get_device_attribute.argtypes = [ctypes.c_int64, ctypes.c_int64]
    lib.b200_get_max_shared_memory_per_block_device_attribute.restype = ctypes.c_int64
    lib.b200_get_max_shared_memory_per_block_device_attribute.argtypes = [ctypes.c_int64]
    lib.b200_parse_kv_cache_dtype.argtypes = [ctypes.c_char_p]
    lib.b200_paged_attention_plan.argtypes = [ctypes.c_int] * 10 + [ctypes.POINTER(AttnTuning),
                                                                    ctypes.POINTER(AttnPlan)]
    lib.b200_ctx_exchange_bytes.restype = ctypes.c_int64
    lib.b200_ctx_exchange_bytes.argtypes = [ctypes.c_int] * 3
    vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
    lib.b200_paged_attention_ctx_shard.argtypes = (
        [vp] * 9 + [i32] * 8 + [i64] * 3 + [f32] * 3 + [i32] * 2 +
        [ctypes.POINTER(vp), i32, i32, i32, ctypes.c_uint32, i32, i32, vp])
    lib.b200_paged_attention_tuned.argtypes = (
        [i32] + [vp] * 10 + [i32] * 8 + [i64] * 3 + [f32] * 3 + [i32] * 2 +
        [vp, ctypes.POINTER(AttnTuning), i32, vp])
    # out, query, key_cache, value_cache, block_tables, seq_lens, alibi | 7 ints | 3 strides | 3 floats |
    # q_dtype, kv_dtype | blocksparse | device | stream
    lib.b200_paged_attention_v2_static.argtypes = (
        [vp] * 7 + [i32] * 7 + [i64] * 3 + [ctypes.c_float] * 3 + [i32, i32, vp, i32, vp])
    lib.b200_swap_blocks_mode.argtypes = [vp, vp, vp, i32, i64, i32, i32, i32, vp]
    lib.b200_context_attention_fwd.argtypes = (
        [vp] * 10 + [i32] * 7 + [i64] * 12 + [vp, i32, f32, vp, i64, i64, i32, i32, vp])
    lib.b200_context_attention_fwd_path.argtypes = lib.b200_context_attention_fwd.argtypes + [i32]
    lib.b200_context_attention_scratch_bytes.restype = ctypes.c_int64
    lib.b200_context_attention_scratch_bytes.argtypes = [i64, i32, i32, i32, i32]
    lib.b200_ctx_exchange_status.argtypes = [vp, i32, ctypes.POINTER(ctypes.c_int32), i32]
    lib.b200_ctx_exchange_status_async.argtypes = [vp, i32, vp, i32, vp]
    lib.b200_ctx_exchange_set_timeout.argtypes = [vp, i32, i32, i32]
    lib.b200_ctx_exchange_set_host_mirror.argtypes = [vp, i32, vp, i32]
    lib.b200_ctx_exchange_trace.argtypes = [vp, i32, vp, i32]
    lib.b200_ctx_exchange_clear_status.argtypes = [vp, i32, i32]
    lib.b200_peer_alloc.argtypes = [ctypes.POINTER(vp), i64, i32]
    lib.b200_peer_free.argtypes = [vp, i32]
    lib.b200_peer_export.argtypes = [vp, ctypes.c_char_p, i32]
    lib.b200_peer_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(vp), i32]
    lib.b200_peer_close.argtypes = [vp, i32]
    torch.ops.load_library(BINDING_PATH)
    return lib


lib = _load()


def launch_count() -> int:
    """Kernels launched by libb200_paged.so from this process (bench.py's ``gpu_launches``)."""
    return int(lib.b200_launch_count())


def last_error() -> str:
    return lib.b200_last_error().decode()


PREFIX_PATHS = {"auto": 0, "mma_sync": 1, "tcgen05": 2, "tcgen05_gather": 3}  # B200_PREFIX_PATH_*
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def _tuning(tuning):
    if tuning is None:
        return None
    w, s, c, *rest = tuning
    return ctypes.byref(AttnTuning(int(w), int(s), int(c), int(rest[0]) if rest else 0))


def attention_plan(version, num_seqs, num_heads, num_kv_heads, head_size, block_size, max_seq_len,
                   dtype, kv_cache_dtype="auto", device=0, tuning=None) -> dict:
    """Launch geometry the library would pick (needs a CUDA device).  ``tuning`` = optional
    (warps, stages, split[, head_rows]) per-call override, as ``paged_attention_tuned`` takes."""
    plan = AttnPlan()
    kv = lib.b200_parse_kv_cache_dtype(kv_cache_dtype.encode())
    rc = lib.b200_paged_attention_plan(version, num_seqs, num_heads, num_kv_heads, head_size,
                                       block_size, max_seq_len, _DT[dtype], kv, device,
                                       _tuning(tuning), ctypes.byref(plan))
    if rc != 0:
        raise RuntimeError(last_error())
    return {n: getattr(plan, n) for n, _ in AttnPlan._fields_}


def paged_attention_tuned(version, out, scratch, query, key_cache, value_cache, num_kv_heads, scale,
                          block_tables, seq_lens, block_size, max_seq_len, kv_cache_dtype="auto",
                          k_scale=1.0, v_scale=1.0, tuning=None) -> None:
    """``paged_attention_v1`` (version 1; ``scratch`` may be None) / ``_v2`` (``scratch`` =
    (exp_sums, max_logits, tmp_out)) with an explicit launch plan ``tuning`` = (warps, stages, split),
    straight through the C ABI on the current stream.  Benchmarking / autotuning only
    (tools/sweep_attn.py); the override is an argument of this one call -- nothing is remembered."""
    es, ml, tmp = scratch if scratch is not None else (None, None, None)
    ptr = lambda t: None if t is None else t.data_ptr()
    B, H, D = query.shape
    kv = lib.b200_parse_kv_cache_dtype(kv_cache_dtype.encode())
    f = ctypes.c_float
    dev = query.device.index if query.device.index is not None else torch.cuda.current_device()
    rc = lib.b200_paged_attention_tuned(
        int(version), out.data_ptr(), ptr(es), ptr(ml), ptr(tmp), query.data_ptr(),
        key_cache.data_ptr(), value_cache.data_ptr(), block_tables.data_ptr(), seq_lens.data_ptr(),
        None, B, H, int(num_kv_heads), D, int(block_size), int(max_seq_len), block_tables.shape[1],
        0 if es is None else es.shape[2], query.stride(0), key_cache.stride(0), key_cache.stride(1),
        f(scale), f(k_scale), f(v_scale), _DT[query.dtype], kv, None, _tuning(tuning), dev,
        torch.cuda.current_stream(query.device).cuda_stream)
    if rc != 0:
        raise RuntimeError(last_error())


def swap_blocks_mode(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor,
                     mode: int) -> None:
    """``swap_blocks`` with the copy mechanism forced (0 auto, 1 memcpy runs, 2 kernel) -- a per-call
    argument of the C ABI; tests and benchmarks only."""
    assert block_mapping.device.type == "cpu" and block_mapping.dtype == torch.int64
    block_mapping = block_mapping.contiguous()
    block_bytes = src[0].numel() * src.element_size()
    sd = src.device.index if src.is_cuda else -1
    dd = dst.device.index if dst.is_cuda else -1
    dev = src.device if src.is_cuda else dst.device
    rc = lib.b200_swap_blocks_mode(src.data_ptr(), dst.data_ptr(), block_mapping.data_ptr(),
                                   block_mapping.shape[0], block_bytes, sd, dd, int(mode),
                                   torch.cuda.current_stream(dev).cuda_stream)
    if rc != 0:
        raise RuntimeError(last_error())
