"""The tcgen05 / TMEM tile loop of prefix prefill (csrc/prefix_prefill_umma.cuh) forced through the C ABI
(``b200_context_attention_fwd_path``, path = TCGEN05) against an fp64 ground truth, the CPU oracle and the
mma.sync tile loop, at the same bars as tests/test_prefix_prefill.py (2e-3 fp16 / 1e-2 bf16).
Reference: light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811."""
import functools
import os
import random
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
sys.path.insert(0, HERE)
from make_golden_prefix import make_prefix_case  # noqa: E402
from test_prefix_prefill import TOL, _call, _truth_gpu  # noqa: E402


def _same_as_gather(c, got, tol):
    """BS = 16 calls load K / V tiles with TMA tensor copies; the per-thread 16-byte gather is the other way to fill
    them (and the only one for other block sizes): same tile loop, same answer."""
    gat = _run(c, "tcgen05_gather")
    assert torch.isfinite(gat.float()).all()
    assert (got.float() - gat.float()).abs().max().item() <= tol


def _run(c, path, out=None):
    from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    if out is None:
        out = torch.full_like(d["query"], float("nan"))
    _call(functools.partial(context_attention_fwd, path=path), d, out)
    torch.cuda.synchronize()
    return out.cpu()


@pytest.mark.gpu
@pytest.mark.parametrize("D", [128, 64])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("H,KH,BS", [(32, 8, 16), (28, 4, 16), (8, 8, 32), (6, 2, 8), (16, 1, 16), (4, 2, 32),
                                     (3, 1, 16), (10, 2, 16)])
def test_tcgen05_matches_truth_and_oracle(oracle, dtype, H, KH, BS, D):
    """Ragged batch: no context, contexts that end inside a page / inside a 128-key tile / on a tile edge,
    query lengths below, at and above one 256-row tile, an empty query."""
    ctx = [0, 5, 16, 129, 700, 48, 256]
    ql = [70, 1, 33, 64, 130, 0, 257]
    c = make_prefix_case(ctx, ql, H, KH, D, BS, dtype, seed=H + BS, q_pad=8)
    got = _run(c, "tcgen05")
    assert torch.isfinite(got.float()).all()
    want = _truth_gpu(c)
    assert (got.double() - want).abs().max().item() <= TOL[dtype]
    ora = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]))
    assert (got.float() - ora.float()).abs().max().item() <= TOL[dtype]
    legacy = _run(c, "mma_sync")
    assert (got.float() - legacy.float()).abs().max().item() <= TOL[dtype]
    _same_as_gather(c, got, TOL[dtype])


@pytest.mark.gpu
@pytest.mark.parametrize("D", [128, 64])
def test_tcgen05_nan_poisoned_cache_tails(D):
    """Slots past ctx_len in the last context page (and every unused block) hold NaN: they must not
    reach the output through 0 * NaN in P.V."""
    c = make_prefix_case([3, 77, 130, 1000], [40, 200, 64, 300], 8, 2, D, 16, torch.float16, seed=9)
    BS = 16
    used = set()
    for b, n in enumerate(c["context_lens"].tolist()):
        for i in range((n + BS - 1) // BS):
            used.add(int(c["block_tables"][b, i]))
    for blk in range(c["key_cache"].shape[0]):
        if blk not in used:
            c["key_cache"][blk] = float("nan")
            c["value_cache"][blk] = float("nan")
    for b, n in enumerate(c["context_lens"].tolist()):
        if n % BS:
            blk = int(c["block_tables"][b, n // BS])
            c["key_cache"][blk, :, :, n % BS:, :] = float("nan")
            c["value_cache"][blk, :, :, n % BS:] = float("nan")
    got = _run(c, "tcgen05")
    assert torch.isfinite(got.float()).all()
    assert (got.double() - _truth_gpu(c)).abs().max().item() <= 2e-3
    _same_as_gather(c, got, 2e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(24))
def test_tcgen05_fuzz(seed):
    r = random.Random(7000 + seed)
    BS = r.choice([8, 16, 16, 32])
    G = r.choice([1, 2, 3, 4, 4, 7, 8, 16, 32])
    KH = r.choice([1, 2, 3])
    dtype = r.choice([torch.float16, torch.bfloat16])
    B = r.randint(1, 5)
    ctx = [r.choice([0, r.randint(1, 40), r.randint(100, 1500), 128 * r.randint(1, 6)]) for _ in range(B)]
    ql = [r.choice([1, r.randint(2, 40), r.randint(60, 600)]) for _ in range(B)]
    D = r.choice([64, 128])
    c = make_prefix_case(ctx, ql, G * KH, KH, D, BS, dtype, seed=seed, q_pad=r.choice([0, 8]))
    got = _run(c, "tcgen05")
    assert torch.isfinite(got.float()).all()
    err = (got.double() - _truth_gpu(c)).abs().max().item()
    assert err <= TOL[dtype], (seed, D, BS, G, KH, dtype, ctx, ql, err)
    _same_as_gather(c, got, TOL[dtype])


@pytest.mark.gpu
def test_tcgen05_large_dominant_logits():
    """Row maxima that grow by far more than 2^8 between tiles (the lazy rescale must fire) and rows
    dominated by one key."""
    c = make_prefix_case([512, 300], [256, 129], 8, 2, 128, 16, torch.float16, seed=3)
    c["key"][::37] *= 6.0
    c["query"][::5] *= 4.0
    blk = int(c["block_tables"][0, 20])
    c["key_cache"][blk] *= 8.0
    got = _run(c, "tcgen05")
    assert torch.isfinite(got.float()).all()
    assert (got.double() - _truth_gpu(c)).abs().max().item() <= 2e-3


@pytest.mark.gpu
def test_tcgen05_writes_only_its_rows():
    c = make_prefix_case([0, 37, 300], [45, 1, 270], 8, 2, 128, 16, torch.float16, seed=41, q_pad=8)
    T, H, D = c["query"].shape
    big = torch.full((T + 4, H * D + 24), 777.0, dtype=torch.float16, device="cuda")
    out = big[2:-2, 8:8 + H * D].view(T, H, D)
    _run(c, "tcgen05", out=out)
    assert (big[:2] == 777.0).all() and (big[-2:] == 777.0).all()
    assert (big[:, :8] == 777.0).all() and (big[:, 8 + H * D:] == 777.0).all()
    assert (out.double().cpu() - _truth_gpu(c)).abs().max().item() <= 2e-3


@pytest.mark.gpu
def test_tcgen05_rejects_what_it_cannot_run():
    from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
    c = make_prefix_case([16], [8], 4, 2, 96, 16, torch.float16, seed=1)  # head sizes other than 64 / 128
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    with pytest.raises(RuntimeError, match="tcgen05"):
        _call(functools.partial(context_attention_fwd, path="tcgen05"), d, torch.empty_like(d["query"]))
    c = make_prefix_case([16], [8], 4, 2, 128, 16, torch.float16, seed=1)
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    with pytest.raises(RuntimeError, match="tcgen05"):
        _call(functools.partial(context_attention_fwd, path="tcgen05"), d, torch.empty_like(d["query"]),
              None, 5)  # sliding window
    with pytest.raises(RuntimeError, match="tcgen05"):
        _call(functools.partial(context_attention_fwd, path="tcgen05", use_scratch=False), d,
              torch.empty_like(d["query"]))


@pytest.mark.gpu
def test_tcgen05_prefill_heavy_shape():
    """The shape the tensor roofline is quoted on, shrunk 4x in batch: 4 x (4096 cached + 512 new)."""
    c = make_prefix_case([4096] * 4, [512] * 4, 32, 8, 128, 16, torch.bfloat16, seed=11)
    got = _run(c, "tcgen05")
    assert (got.double() - _truth_gpu(c)).abs().max().item() <= 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["tcgen05", "tcgen05_gather"])
def test_tcgen05_graph_capture(path):
    """Staging pre-pass + tile loop (its programmatic dependent; the TMA variant carries four tensor maps as kernel
    parameters) captured in a CUDA graph and replayed on other queries / new tokens: equal to the eager call."""
    from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
    c = make_prefix_case([700, 48, 256, 0], [130, 64, 257, 40], 8, 2, 128, 16, torch.bfloat16, seed=4)
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    fn = functools.partial(context_attention_fwd, path=path)
    out = torch.zeros_like(d["query"])
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        _call(fn, d, out)
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            _call(fn, d, out)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.Generator(device="cuda").manual_seed(9)
    for name in ("query", "key", "value"):
        d[name].copy_((torch.rand(d[name].shape, generator=g, device="cuda") * 2 - 1).to(d[name].dtype) * 0.09)
    out.fill_(float("nan"))
    graph.replay()
    torch.cuda.synchronize()
    got = out.clone()
    want = _call(fn, d, torch.zeros_like(out))
    torch.cuda.synchronize()
    assert torch.isfinite(got.float()).all()
    assert torch.equal(got, want)
    assert (got.double().cpu() - _truth_gpu({k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in d.items()})).abs().max().item() <= TOL[torch.bfloat16]
