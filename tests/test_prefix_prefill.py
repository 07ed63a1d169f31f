"""Prefix / chunked-prefill attention over the paged cache:
``context_attention_fwd`` / ``PagedAttention.forward_prefix``
(light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811, ops/paged_attn.py:193-225).

CPU: the C oracle against tests/golden/prefix.pt -- outputs of the reference's own Triton kernel run
through Triton's interpreter (tests/golden/make_golden_prefix.py).
GPU: the sm_100a kernel (C ABI ``b200_context_attention_fwd``) against the golden vectors, the
oracle and an fp64 ground truth."""
import os
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_prefix import make_prefix_case  # noqa: E402

GOLD = os.path.join(HERE, "golden", "prefix.pt")
TOL = {torch.float32: 2e-6, torch.float16: 2e-3, torch.bfloat16: 1e-2}


def _call(fn, c, out, slopes=None, window=0):
    fn(c["query"], c["key"], c["value"], out, c["key_cache"], c["value_cache"], c["block_tables"],
       c["query_start_loc"][:-1], c["seq_lens"], c["context_lens"], c["max_query_len"], slopes,
       window if window else None)
    return out


def ground_truth(c, slopes=None, window=0):
    """fp64 softmax over [cached context | causal new tokens] per (sequence, head)."""
    q, k, v = c["query"].double(), c["key"].double(), c["value"].double()
    T, H, D = q.shape
    KH, BS = c["num_kv_heads"], c["block_size"]
    out = torch.zeros(T, H, D, dtype=torch.float64)
    for b in range(len(c["context_lens"])):
        ctx = int(c["context_lens"][b])
        ql = int(c["seq_lens"][b]) - ctx
        s0 = int(c["query_start_loc"][b])
        nblk = (ctx + BS - 1) // BS
        blks = c["block_tables"][b, :nblk].long()
        kc = c["key_cache"][blks].double().permute(1, 0, 3, 2, 4).reshape(KH, nblk * BS, D)[:, :ctx]
        vc = c["value_cache"][blks].double().permute(1, 0, 3, 2).reshape(KH, nblk * BS, D)[:, :ctx]
        for h in range(H):
            kvh = h // (H // KH)
            keys = torch.cat([kc[kvh], k[s0:s0 + ql, kvh]])
            vals = torch.cat([vc[kvh], v[s0:s0 + ql, kvh]])
            pos_k = torch.arange(ctx + ql)
            for i in range(ql):
                pos_q = ctx + i
                logits = (keys @ q[s0 + i, h]) * D ** -0.5
                if slopes is not None:
                    logits = logits + float(slopes[h]) * (pos_k - pos_q)
                mask = pos_k > pos_q
                if window:
                    mask = mask | (pos_q - pos_k >= window)
                logits = logits.masked_fill(mask, float("-inf"))
                out[s0 + i, h] = torch.softmax(logits, 0) @ vals
    return out


def test_oracle_matches_reference_golden(oracle):
    gold = torch.load(GOLD)
    assert len(gold) >= 6
    for g in gold:
        c = make_prefix_case(g["ctx_lens"], g["query_lens"], g["H"], g["KH"], g["D"], g["BS"],
                             g["dtype"], g["seed"])
        o = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]), g["alibi_slopes"],
                  g["sliding_window"])
        err = (o.float() - g["out"].float()).abs().max().item()
        assert err <= (2e-6 if g["dtype"] == torch.float32 else 2.5e-4), (g["dtype"], err)
        gt = ground_truth(c, g["alibi_slopes"], g["sliding_window"])
        assert (o.double() - gt).abs().max().item() <= TOL[g["dtype"]]


def test_oracle_edge_cases(oracle):
    # no context at all, one-token queries, bf16, strided q, empty (ctx only) sequence
    c = make_prefix_case([0, 15, 16, 40], [7, 1, 1, 0], 6, 2, 64, 16, torch.bfloat16, seed=5, q_pad=8)
    o = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]))
    assert (o.double() - ground_truth(c)).abs().max().item() <= 1e-2


# ---------------------------------------------------------------- product path (GPU) ----
def _gpu(c, slopes=None, window=0, use_scratch=True):
    import functools
    from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    out = torch.full_like(d["query"], float("nan"))
    _call(functools.partial(context_attention_fwd, use_scratch=use_scratch), d, out,
          slopes.cuda() if slopes is not None else None, window)
    torch.cuda.synchronize()
    return out.cpu()


@pytest.mark.gpu
def test_kernel_matches_reference_golden():
    """Every golden case (fp16 on the tensor-core kernel, fp32 on the CUDA-core kernel) directly."""
    n = 0
    for g in torch.load(GOLD):
        c = make_prefix_case(g["ctx_lens"], g["query_lens"], g["H"], g["KH"], g["D"], g["BS"],
                             g["dtype"], g["seed"])
        o = _gpu(c, g["alibi_slopes"], g["sliding_window"])
        err = (o.float() - g["out"].float()).abs().max().item()
        assert err <= (2e-3 if g["dtype"] == torch.float16 else 2e-5), (g["dtype"], g["D"], g["BS"], err)
        n += 1
    assert n >= 6


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("H,KH,D,BS", [(32, 8, 128, 16), (28, 4, 128, 16), (14, 2, 64, 16),
                                       (8, 8, 64, 32), (4, 4, 256, 16), (6, 2, 120, 8),
                                       (16, 1, 96, 16), (4, 2, 192, 32)])
def test_kernel_matches_oracle(oracle, dtype, H, KH, D, BS):
    ctx = [0, 5, 16, 129, 700, 48]
    ql = [70, 1, 33, 64, 130, 0]
    c = make_prefix_case(ctx, ql, H, KH, D, BS, dtype, seed=H + D, q_pad=8)
    want = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]))
    got = _gpu(c)
    assert torch.isfinite(got.float()).all()
    assert (got.float() - want.float()).abs().max().item() <= TOL[dtype]
    assert (got.double() - ground_truth(c)).abs().max().item() <= TOL[dtype]


@pytest.mark.gpu
@pytest.mark.parametrize("window", [1, 7, 16, 100, 4096])
def test_kernel_sliding_window(oracle, window):
    c = make_prefix_case([300, 20, 0], [90, 40, 50], 8, 2, 128, 16, torch.float16, seed=window)
    want = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]), None, window)
    got = _gpu(c, None, window)
    assert (got.float() - want.float()).abs().max().item() <= 2e-3
    assert (got.double() - ground_truth(c, None, window)).abs().max().item() <= 2e-3


@pytest.mark.gpu
def test_kernel_alibi(oracle):
    H = 8
    slopes = torch.tensor([2.0 ** (-(j + 1) * 8.0 / H) for j in range(H)], dtype=torch.float32)
    c = make_prefix_case([200, 0, 33], [60, 100, 17], H, 2, 128, 16, torch.bfloat16, seed=9)
    want = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]), slopes)
    got = _gpu(c, slopes)
    assert (got.float() - want.float()).abs().max().item() <= 1e-2
    assert (got.double() - ground_truth(c, slopes)).abs().max().item() <= 1e-2


@pytest.mark.gpu
def test_kernel_alibi_ignores_sliding_window(oracle):
    """As the reference: _fwd_kernel_alibi has no SLIDING_WINDOW (prefix_prefill.py:712-760)."""
    H = 4
    slopes = torch.tensor([2.0 ** (-(j + 1) * 8.0 / H) for j in range(H)], dtype=torch.float32)
    c = make_prefix_case([90, 3], [20, 40], H, 2, 64, 16, torch.float16, seed=31)
    want = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]), slopes, 11)
    assert (want.double() - ground_truth(c, slopes, 0)).abs().max().item() <= 2e-3
    got = _gpu(c, slopes, 11)
    assert (got.float() - want.float()).abs().max().item() <= 2e-3


@pytest.mark.gpu
def test_kernel_poisoned_cache_tail_and_spare_rows():
    """Cache slots past ctx_len (where the new tokens will be written) may hold anything."""
    c = make_prefix_case([21, 40], [30, 9], 8, 2, 128, 16, torch.float16, seed=2)
    BS = 16
    for b, ctx in enumerate([21, 40]):
        if ctx % BS:
            blk = int(c["block_tables"][b, ctx // BS])
            c["key_cache"][blk, :, :, ctx % BS:, :] = float("nan")
            c["value_cache"][blk, :, :, ctx % BS:] = float("nan")
    got = _gpu(c)
    assert torch.isfinite(got.float()).all()
    assert (got.double() - ground_truth(c)).abs().max().item() <= 2e-3


@pytest.mark.gpu
def test_plugin_prefix_prefill_then_decode(oracle):
    """PagedB200AttentionImpl.forward with a prefix-cache hit: the new tokens are written to the cache
    and attend the cached prefix (forward_prefix); a following decode step sees all of them."""
    from light_vllm_b200.attention.backend import PagedB200AttentionImpl, PagedB200AttentionMetadata
    H, KH, D, BS = 8, 2, 128, 16
    ctx, ql = [32, 5], [20, 11]
    c = make_prefix_case(ctx, ql, H, KH, D, BS, torch.float16, seed=12)
    dev = "cuda:0"
    NB = c["key_cache"].shape[0]
    kv_cache = torch.stack([c["key_cache"].reshape(NB, -1), c["value_cache"].reshape(NB, -1)]).to(dev)
    T = sum(ql)
    slots = []
    for b in range(2):
        for i in range(ql[b]):
            pos = ctx[b] + i
            slots.append(int(c["block_tables"][b, pos // BS]) * BS + pos % BS)
    meta = PagedB200AttentionMetadata(
        num_prefills=2, num_prefill_tokens=T, num_decode_tokens=0,
        slot_mapping=torch.tensor(slots, dtype=torch.int64, device=dev),
        seq_lens=[a + b for a, b in zip(ctx, ql)], seq_lens_tensor=c["seq_lens"].to(dev),
        max_query_len=max(ql), max_prefill_seq_len=max(a + b for a, b in zip(ctx, ql)),
        max_decode_seq_len=0, query_start_loc=c["query_start_loc"].to(dev),
        seq_start_loc=torch.tensor([0, ctx[0] + ql[0], sum(ctx) + sum(ql)], dtype=torch.int32, device=dev),
        context_lens_tensor=c["context_lens"].to(dev),
        block_tables=c["block_tables"].to(dev), use_cuda_graph=False)
    impl = PagedB200AttentionImpl(H, D, D ** -0.5, KH, None, None, "auto")
    q = c["query"].reshape(T, H * D).to(dev)
    k = c["key"].reshape(T, KH * D).to(dev)
    v = c["value"].reshape(T, KH * D).to(dev)
    out = impl.forward(q, k, v, kv_cache, meta).view(T, H, D).cpu()
    assert (out.double() - ground_truth(c)).abs().max().item() <= 2e-3


@pytest.mark.gpu
@pytest.mark.parametrize("window", [0, 9])
@pytest.mark.parametrize("H,KH,D,BS", [(32, 8, 128, 16), (6, 2, 120, 8), (4, 2, 192, 32)])
def test_kernel_without_scratch_stages_in_cta(oracle, window, H, KH, D, BS):
    """kv_scratch = NULL: every CTA stages the new k/v itself (the other mode is the default)."""
    c = make_prefix_case([0, 129, 40], [70, 64, 5], H, KH, D, BS, torch.float16, seed=77)
    want = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]), None, window)
    a = _gpu(c, None, window, use_scratch=False)
    b = _gpu(c, None, window, use_scratch=True)
    assert (a.float() - want.float()).abs().max().item() <= 2e-3
    assert (b.float() - want.float()).abs().max().item() <= 2e-3


@pytest.mark.gpu
@pytest.mark.parametrize("H,KH,D,BS,window", [(8, 2, 128, 16, 0), (6, 6, 80, 8, 0), (4, 1, 256, 32, 0),
                                              (8, 2, 64, 16, 40), (4, 2, 120, 16, 3)])
def test_kernel_fp32(oracle, H, KH, D, BS, window):
    c = make_prefix_case([0, 5, 129, 300, 48], [70, 1, 33, 64, 0], H, KH, D, BS, torch.float32, seed=D + BS, q_pad=4)
    want = _call(oracle.context_attention_fwd, c, torch.zeros_like(c["query"]), None, window)
    got = _gpu(c, None, window)
    assert (got - want).abs().max().item() <= 2e-5
    assert (got.double() - ground_truth(c, None, window)).abs().max().item() <= 2e-5


def _truth_gpu(c, slopes=None, window=0, dev="cuda:0"):
    """ground_truth() vectorised per sequence in fp64 on the GPU (same definition; for larger cases)."""
    q, k, v = (c[n].to(dev).double() for n in ("query", "key", "value"))
    T, H, D = q.shape
    KH, BS = c["num_kv_heads"], c["block_size"]
    G = H // KH
    kcache, vcache = c["key_cache"].to(dev), c["value_cache"].to(dev)
    out = torch.zeros(T, H, D, dtype=torch.float64, device=dev)
    for b in range(len(c["context_lens"])):
        ctx = int(c["context_lens"][b])
        ql = int(c["seq_lens"][b]) - ctx
        if ql == 0:
            continue
        s0 = int(c["query_start_loc"][b])
        nblk = (ctx + BS - 1) // BS
        blks = c["block_tables"][b, :nblk].long().to(dev)
        kc = kcache[blks].double().permute(1, 0, 3, 2, 4).reshape(KH, nblk * BS, D)[:, :ctx]
        vc = vcache[blks].double().permute(1, 0, 3, 2).reshape(KH, nblk * BS, D)[:, :ctx]
        keys = torch.cat([kc, k[s0:s0 + ql].transpose(0, 1)], dim=1)          # [KH, ctx+ql, D]
        vals = torch.cat([vc, v[s0:s0 + ql].transpose(0, 1)], dim=1)
        qq = q[s0:s0 + ql].view(ql, KH, G, D)
        logits = torch.einsum("tkgd,ksd->kgts", qq, keys) * D ** -0.5           # [KH, G, ql, ctx+ql]
        pos_k = torch.arange(ctx + ql, device=dev)
        pos_q = ctx + torch.arange(ql, device=dev)
        rel = (pos_k[None, :] - pos_q[:, None])                                 # [ql, S]
        if slopes is not None:
            logits = logits + slopes.to(dev).double().view(KH, G, 1, 1) * rel
        mask = rel > 0
        if window:
            mask = mask | (-rel >= window)
        logits = logits.masked_fill(mask, float("-inf"))
        o = torch.einsum("kgts,ksd->tkgd", torch.softmax(logits, -1), vals)
        out[s0:s0 + ql] = o.reshape(ql, H, D)
    return out.cpu()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(40))
def test_kernel_fuzz_against_fp64_truth(seed):
    """Random shapes: group sizes that do not divide the row tile, odd page counts through the tight
    interior loops, random windows / ALiBi, all three block sizes, with and without the staging pass."""
    import random
    r = random.Random(1000 + seed)
    D = r.choice([64, 80, 96, 112, 120, 128, 192, 256])
    BS = r.choice([8, 16, 16, 32])
    G = r.choice([1, 2, 3, 4, 7, 8, 16])
    KH = r.choice([1, 2, 3])
    H = G * KH
    dtype = r.choice([torch.float16, torch.bfloat16])
    B = r.randint(1, 4)
    ctx = [r.choice([0, r.randint(1, 40), r.randint(100, 1500)]) for _ in range(B)]
    ql = [r.choice([1, r.randint(2, 40), r.randint(60, 260)]) for _ in range(B)]
    window = r.choice([0, 0, r.randint(1, 50), r.randint(100, 900)])
    slopes = None
    if r.random() < 0.3:
        slopes = torch.tensor([2.0 ** (-(j + 1) * 8.0 / H) for j in range(H)], dtype=torch.float32)
    c = make_prefix_case(ctx, ql, H, KH, D, BS, dtype, seed=seed, q_pad=r.choice([0, 8]))
    # the reference's ALiBi kernel takes no sliding window (prefix_prefill.py:712-760): window ignored
    want = _truth_gpu(c, slopes, 0 if slopes is not None else window)
    for use_scratch in (True, False):
        got = _gpu(c, slopes, window, use_scratch=use_scratch)
        assert torch.isfinite(got.float()).all()
        err = (got.double() - want).abs().max().item()
        assert err <= TOL[dtype], (seed, D, BS, G, KH, dtype, ctx, ql, window, slopes is not None,
                                   use_scratch, err)


def test_vectorised_truth_matches_loop_truth():
    H = 6
    slopes = torch.tensor([2.0 ** (-(j + 1) * 8.0 / H) for j in range(H)], dtype=torch.float32)
    c = make_prefix_case([0, 21, 40], [9, 1, 17], H, 2, 64, 8, torch.float16, seed=77, q_pad=8)
    for sl, w in [(None, 0), (slopes, 0), (None, 5), (slopes, 12)]:
        a, b = ground_truth(c, sl, w), _truth_gpu(c, sl, w, dev="cpu")
        assert (a - b).abs().max().item() <= 1e-7  # the loop version rounds the ALiBi bias to fp32


@pytest.mark.gpu
@pytest.mark.parametrize("H,KH,D,BS", [(32, 8, 128, 16), (7, 1, 80, 8), (4, 2, 256, 32)])
def test_kernel_writes_only_its_output_rows(H, KH, D, BS):
    """Guard rows and guard columns around a strided output keep their sentinel; q/k/v, the caches
    and the block tables are bit-identical after the call (the staging pass writes only scratch)."""
    from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
    c = make_prefix_case([0, 37, 300], [45, 1, 70], H, KH, D, BS, torch.float16, seed=41, q_pad=8)
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    T = d["query"].shape[0]
    big = torch.full((T + 4, H * D + 24), 777.0, dtype=torch.float16, device="cuda")
    out = big[2:-2, 8:8 + H * D].view(T, H, D)
    before = {n: d[n].clone() for n in ("query", "key", "value", "key_cache", "value_cache", "block_tables")}
    _call(context_attention_fwd, d, out)
    torch.cuda.synchronize()
    assert (big[:2] == 777.0).all() and (big[-2:] == 777.0).all()
    assert (big[:, :8] == 777.0).all() and (big[:, 8 + H * D:] == 777.0).all()
    for n, t in before.items():
        assert torch.equal(t.view(torch.uint8) if t.dtype != torch.int32 else t,
                           d[n].view(torch.uint8) if t.dtype != torch.int32 else d[n]), n
    assert (out.double().cpu() - ground_truth(c)).abs().max().item() <= 2e-3
