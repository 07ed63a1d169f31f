"""Prefill over a cached prefix: the reference's ``context_attention_fwd`` entry point
(light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-686), same name, argument order
and meaning, served by the sm_100a kernel behind ``b200_context_attention_fwd``
(``csrc/prefix_prefill.cu``) instead of Triton.  CUDA tensors only; there is no fallback."""
from typing import Optional

import torch

from .. import _lib


def context_attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor,
                          k_cache: torch.Tensor, v_cache: torch.Tensor, b_loc: torch.Tensor,
                          b_start_loc: torch.Tensor, b_seq_len: torch.Tensor,
                          b_ctx_len: torch.Tensor, max_input_len: int,
                          alibi_slopes: Optional[torch.Tensor] = None,
                          sliding_window: Optional[int] = None, use_scratch: bool = True,
                          path: str = "auto") -> None:
    """``path``: ``"auto"`` (default) lets the library pick the tile loop -- the tcgen05 / TMEM kernel
    (``csrc/prefix_prefill_umma.cuh``) for head size 64 / 128 fp16 / bf16 without ALiBi / window once a
    sequence brings >= 128 (token, head) rows, the ``mma.sync`` kernel otherwise; ``"tcgen05"`` /
    ``"mma_sync"`` force one (tests, benchmarks; unsupported combinations raise).

    ``use_scratch`` (default): the new k/v are first rewritten into page layout in a per-call
    ``torch.empty`` scratch (graph-pool safe, like the reference's per-call allocations) and streamed
    through the same bulk-copy ring as the cached context.  That staged mode assumes what
    ``PagedAttention.forward_prefix`` passes: ``b_start_loc`` = packed exclusive prefix sums of the
    query lengths (``query_start_loc[:-1]``).  A non-packed ``b_start_loc`` is memory-safe (indices
    are guarded against the scratch capacity; affected rows come back NaN) but must use
    ``use_scratch=False``, which accepts arbitrary start locations like the reference kernel."""
    if not q.is_cuda:
        raise RuntimeError("context_attention_fwd: CUDA tensors required (no CPU implementation)")
    if q.dtype not in (torch.float16, torch.bfloat16, torch.float32):
        raise RuntimeError(f"Unsupported data type: context_attention_fwd runs fp32 / fp16 / bf16, got {q.dtype}")
    if k_cache.dtype != q.dtype or v_cache.dtype != q.dtype:
        raise RuntimeError("context_attention_fwd: the cache must hold the query dtype "
                           "(the reference kernel has no fp8 path)")
    for t in (q, k, v, o):
        assert t.dim() == 3 and t.stride(2) == 1
    Lq, Lk, Lv = q.shape[-1], k.shape[-1], v.shape[-1]
    assert Lq == Lk and Lk == Lv  # prefix_prefill.py:697-698
    i32 = lambda t: t if t.dtype == torch.int32 else t.to(torch.int32)  # noqa: E731
    b_loc, b_start_loc, b_seq_len, b_ctx_len = (i32(b_loc), i32(b_start_loc).contiguous(),
                                                i32(b_seq_len).contiguous(), i32(b_ctx_len).contiguous())
    assert b_loc.stride(1) == 1
    if alibi_slopes is not None:
        alibi_slopes = alibi_slopes.to(torch.float32).contiguous()
    import ctypes
    # scratch for the staged mode (new k/v rewritten into page layout; see csrc/prefix_prefill.cu)
    need = _lib.lib.b200_context_attention_scratch_bytes(k.shape[0], b_seq_len.shape[0], k.shape[1], Lq,
                                                         v_cache.shape[3])
    scratch = torch.empty(max(need, 16) if use_scratch else 16, dtype=torch.uint8, device=q.device)
    rc = _lib.lib.b200_context_attention_fwd_path(
        q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
        b_loc.data_ptr(), b_start_loc.data_ptr(), b_seq_len.data_ptr(), b_ctx_len.data_ptr(),
        b_seq_len.shape[0], q.shape[1], k.shape[1], Lq, v_cache.shape[3], b_loc.stride(0),
        int(max_input_len), q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0), v.stride(1),
        o.stride(0), o.stride(1), k_cache.stride(0), k_cache.stride(1), v_cache.stride(0),
        v_cache.stride(1), alibi_slopes.data_ptr() if alibi_slopes is not None else None,
        int(sliding_window) if sliding_window and sliding_window > 0 else 0,
        ctypes.c_float(1.0 / (Lq ** 0.5)), scratch.data_ptr() if use_scratch else None, scratch.numel(),
        k.shape[0], _lib._DT[q.dtype], q.device.index,
        torch.cuda.current_stream(q.device).cuda_stream, _lib.PREFIX_PATHS[path])
    if rc != 0:
        raise RuntimeError(_lib.last_error())
