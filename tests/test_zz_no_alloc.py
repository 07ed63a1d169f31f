"""b200_paged_attention_v2_static: paged_attention_v2 with the partition scratch in a library-owned static
device array (no allocation by the callee; no counterpart in the reference).
Called straight through the C ABI and compared with b200_paged_attention_v2 (caller scratch) on
the same inputs, two calls in a row on the shared array.  Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest
import torch

from kv_util import make_case, to_device

STATIC_SCRATCH_BYTES = 16 * 1024 * 1024  # B200_STATIC_SCRATCH_BYTES (include/b200_paged.h)
KV_AUTO = 0  # B200_KV_AUTO
TOL = {torch.float16: 1e-3, torch.bfloat16: 8e-3}  # the paged-attention tests' tolerances; same kernels: expect 0


def _call_both(lib, _lib, d, case, H, KH, D, BS, max_len, dtype):
    q, kc, vc = d["query"], d["key_cache"], d["value_cache"]
    bt, sl = d["block_tables"], d["seq_lens"]
    B = q.shape[0]
    P = max(1, (max_len + 511) // 512)
    f = ctypes.c_float
    vp, i64 = ctypes.c_void_p, ctypes.c_int64  # b200_paged_attention_v2 has no argtypes: wrap explicitly
    stream = vp(torch.cuda.current_stream().cuda_stream)
    common_tail = (vp(bt.data_ptr()), vp(sl.data_ptr()), None, B, H, KH, D, BS, max_len, bt.shape[1])
    strides = (i64(q.stride(0)), i64(kc.stride(0)), i64(kc.stride(1)), f(case["scale"]), f(1.0), f(1.0),
               _lib._DT[dtype], KV_AUTO, None, 0, stream)
    # reference arm: the existing entry point with caller-owned scratch
    es = torch.empty(B, H, P, dtype=torch.float32, device="cuda")
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=dtype, device="cuda")
    want = torch.full_like(q, float("nan"))
    rc = lib.b200_paged_attention_v2(
        ctypes.c_void_p(want.data_ptr()), ctypes.c_void_p(es.data_ptr()), ctypes.c_void_p(ml.data_ptr()),
        ctypes.c_void_p(tmp.data_ptr()), ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(kc.data_ptr()),
        ctypes.c_void_p(vc.data_ptr()), *common_tail, P, *strides)
    assert rc == 0, _lib.last_error()
    got = torch.full_like(q, float("nan"))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    rc = lib.b200_paged_attention_v2_static(
        vp(got.data_ptr()), vp(q.data_ptr()), vp(kc.data_ptr()), vp(vc.data_ptr()), *common_tail, *strides)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    after = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    return want, got, before, after


@pytest.mark.gpu
def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    from light_vllm_b200 import _lib
    lib = _lib.lib
    H, KH, D, BS = 32, 8, 128, 16
    # call 1 fills more of the shared array than call 2 uses, with other lengths: call 2 runs on stale contents
    shapes = [([3000, 17, 5000, 700, 1, 2049], torch.float16, 5),
              ([900, 4100, 0, 513], torch.bfloat16, 6),
              ([3000, 17, 5000, 700, 1, 2049], torch.float16, 5)]
    for n, (lens, dtype, seed) in enumerate(shapes):
        case = make_case(len(lens), H, KH, D, BS, lens, dtype, seed=seed)
        d = to_device(case, "cuda:0")
        want, got, before, after = _call_both(lib, _lib, d, case, H, KH, D, BS, max(lens), dtype)
        live = torch.tensor([x > 0 for x in lens], device="cuda")
        assert not torch.isnan(want[live].float()).any() and not torch.isnan(got[live].float()).any()
        err = (got[live].float() - want[live].float()).abs().max().item()
        assert err <= TOL[dtype], (n, err)
        if n > 0:  # the first call per device resolves the array's address (context set-up may move free memory)
            assert after[0] == before[0] and after[1] >= before[1], (before, after)


@pytest.mark.gpu
def test_static_scratch_rejects_sizes_beyond_the_array():
    from light_vllm_b200 import _lib
    lib = _lib.lib
    H, KH, D, BS = 32, 8, 128, 16
    lens = [64, 64]
    case = make_case(2, H, KH, D, BS, lens, torch.float16, seed=1)
    d = to_device(case, "cuda:0")
    q, kc, vc = d["query"], d["key_cache"], d["value_cache"]
    out = torch.zeros_like(q)
    f = ctypes.c_float
    # claim a context whose partitions cannot fit: 2 * 32 * ceil(max_seq_len / 512) * (8 + 256) bytes > 16 MiB
    too_long = 512 * (STATIC_SCRATCH_BYTES // (2 * H * (8 + 2 * D)) + 2)
    rc = lib.b200_paged_attention_v2_static(
        out.data_ptr(), q.data_ptr(), kc.data_ptr(), vc.data_ptr(), d["block_tables"].data_ptr(),
        d["seq_lens"].data_ptr(), None, 2, H, KH, D, BS, too_long, d["block_tables"].shape[1], q.stride(0),
        kc.stride(0), kc.stride(1), f(case["scale"]), f(1.0), f(1.0), _lib._DT[torch.float16], KV_AUTO, None, 0,
        torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = _lib.last_error()
    assert "scratch" in msg and "built-in array holds" in msg, msg
    torch.cuda.synchronize()
    assert (out == 0).all()  # nothing was launched
