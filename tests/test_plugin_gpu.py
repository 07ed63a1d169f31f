"""GPU: the attention plugin and cache engine end to end (write-then-read on one stream,
prefill + decode batches, fp8 KV, swap-out / copy-on-write / swap-in round trip), checked against
the CPU oracle and an fp64 ground truth."""
import pytest
import torch

from light_vllm_b200.attention import (PagedAttention, PagedB200AttentionBackend,
                                       PagedB200AttentionMetadata)
from light_vllm_b200.cache_engine import CacheEngine

pytestmark = pytest.mark.gpu


def decode_meta(slots, seq_lens, block_tables, dev):
    return PagedB200AttentionMetadata(
        num_prefills=0, num_prefill_tokens=0, num_decode_tokens=len(seq_lens),
        slot_mapping=torch.tensor(slots, dtype=torch.long), seq_lens=None,
        seq_lens_tensor=torch.tensor(seq_lens, dtype=torch.int), max_query_len=1,
        max_prefill_seq_len=0, max_decode_seq_len=max(seq_lens), query_start_loc=None,
        seq_start_loc=None, context_lens_tensor=None,
        block_tables=torch.tensor(block_tables, dtype=torch.int), use_cuda_graph=False).to(dev)


def ref_attention(q, ks, vs, scale):
    """fp64 causal-free attention of one query over its own sequence. q [H,D], ks/vs [n,KH,D]."""
    H, KH = q.shape[0], ks.shape[1]
    out = torch.empty(q.shape, dtype=torch.float64)
    for h in range(H):
        k, v = ks[:, h // (H // KH)].double(), vs[:, h // (H // KH)].double()
        p = torch.softmax((k @ q[h].double()) * scale, 0)
        out[h] = p @ v
    return out


@pytest.mark.parametrize("dtype,kv,tol", [(torch.float16, "auto", 2e-3), (torch.bfloat16, "auto", 1e-2),
                                          (torch.bfloat16, "fp8", 3e-2)])
def test_prefill_then_decode_steps(dtype, kv, tol):
    """Run the plugin like the model runner does: one prefill batch that fills the cache, then
    several pure-decode steps; every step's output must equal attention over all K/V so far."""
    torch.manual_seed(0)
    dev = "cuda"
    H, KH, D, BS, NB = 8, 2, 128, 16, 64
    scale = D ** -0.5
    ksc, vsc = (1.0, 1.0) if kv == "auto" else (0.02, 0.02)
    impl = PagedB200AttentionBackend.get_impl_cls()(H, D, scale, KH, None, None, kv)
    cache_dtype = dtype if kv == "auto" else torch.uint8
    kv_cache = torch.zeros(PagedB200AttentionBackend.get_kv_cache_shape(NB, BS, KH, D),
                           dtype=cache_dtype, device=dev)
    prompt_lens = [5, 33, 16]
    tables = [[7, 3, 50], [1, 9, 22, 40], [12, 13, 14]]
    T = sum(prompt_lens)
    qkv = (torch.rand(T, (H + 2 * KH) * D, device=dev) * 2 - 1).to(dtype) * scale
    q, k, v = qkv[:, :H * D], qkv[:, H * D:(H + KH) * D], qkv[:, (H + KH) * D:]
    slots, hist_k, hist_v, pos = [], [], [], 0
    for n, tb in zip(prompt_lens, tables):
        slots += [tb[i // BS] * BS + i % BS for i in range(n)]
        hist_k.append(k[pos:pos + n].view(n, KH, D).float().cpu())
        hist_v.append(v[pos:pos + n].view(n, KH, D).float().cpu())
        pos += n
    starts = torch.tensor([0] + list(torch.tensor(prompt_lens).cumsum(0)), dtype=torch.int32)
    meta = PagedB200AttentionMetadata(
        num_prefills=3, num_prefill_tokens=T, num_decode_tokens=0,
        slot_mapping=torch.tensor(slots, dtype=torch.long), seq_lens=prompt_lens,
        seq_lens_tensor=torch.tensor(prompt_lens, dtype=torch.int), max_query_len=max(prompt_lens),
        max_prefill_seq_len=max(prompt_lens), max_decode_seq_len=0, query_start_loc=starts,
        seq_start_loc=starts, context_lens_tensor=torch.zeros(3, dtype=torch.int),
        block_tables=torch.zeros(3, 0, dtype=torch.int), use_cuda_graph=False).to(dev)
    out = impl.forward(q, k, v, kv_cache, meta, ksc, vsc)
    assert out.shape == (T, H * D)
    # prefill = causal attention over the prompt itself: check last token of each prompt
    pos = 0
    for i, n in enumerate(prompt_lens):
        want = ref_attention(q[pos + n - 1].view(H, D).float().cpu(), hist_k[i], hist_v[i], scale)
        got = out[pos + n - 1].view(H, D).double().cpu()
        assert (got - want).abs().max() <= 1e-2
        pos += n
    lens = list(prompt_lens)
    for step in range(3):
        B = len(lens)
        qkv = (torch.rand(B, (H + 2 * KH) * D, device=dev) * 2 - 1).to(dtype) * scale
        q, k, v = qkv[:, :H * D], qkv[:, H * D:(H + KH) * D], qkv[:, (H + KH) * D:]
        slots = [tables[i][lens[i] // BS] * BS + lens[i] % BS for i in range(B)]
        lens = [n + 1 for n in lens]
        width = max(len(t) for t in tables)
        bt = [t + [0] * (width - len(t)) for t in tables]
        out = impl.forward(q, k, v, kv_cache, decode_meta(slots, lens, bt, dev), ksc, vsc)
        for i in range(B):
            hist_k[i] = torch.cat([hist_k[i], k[i].view(1, KH, D).float().cpu()])
            hist_v[i] = torch.cat([hist_v[i], v[i].view(1, KH, D).float().cpu()])
            want = ref_attention(q[i].view(H, D).float().cpu(), hist_k[i], hist_v[i], scale)
            err = (out[i].view(H, D).double().cpu() - want).abs().max().item()
            assert err <= tol, f"step {step} seq {i}: {err}"


def test_profile_run_without_cache():
    """kv_cache=None (memory-profiling run, model_runner.py:139-143): prefill only, no cache write."""
    impl = PagedB200AttentionBackend.get_impl_cls()(4, 64, 0.125, 4, None, None, "auto")
    T = 6
    q = torch.randn(T, 256, device="cuda", dtype=torch.float16)
    starts = torch.tensor([0, 6], dtype=torch.int32)
    meta = PagedB200AttentionMetadata(
        num_prefills=1, num_prefill_tokens=T, num_decode_tokens=0,
        slot_mapping=torch.full((T,), -1, dtype=torch.long), seq_lens=[T],
        seq_lens_tensor=torch.tensor([T], dtype=torch.int), max_query_len=T, max_prefill_seq_len=T,
        max_decode_seq_len=0, query_start_loc=starts, seq_start_loc=starts,
        context_lens_tensor=torch.zeros(1, dtype=torch.int),
        block_tables=torch.zeros(1, 0, dtype=torch.int), use_cuda_graph=False).to("cuda")
    out = impl.forward(q, q, q, None, meta)
    assert out.shape == (T, 256) and torch.isfinite(out).all()


@pytest.mark.parametrize("cache_dtype", ["auto", "fp8"])
def test_cache_engine_preemption_round_trip(cache_dtype):
    """swap a sequence out to the pinned pool, fork a block (copy-on-write), swap it back in at
    different block ids: pool bytes must survive bit-exactly (config-5 exercise)."""
    L, KH, D, BS = 3, 2, 64, 16
    eng = CacheEngine(L, KH, D, BS, num_gpu_blocks=24, num_cpu_blocks=10, dtype=torch.float16,
                      cache_dtype=cache_dtype)
    g = torch.Generator().manual_seed(3)
    for t in eng.gpu_cache:
        t.view(torch.uint8).copy_(torch.randint(0, 255, t.view(torch.uint8).shape, dtype=torch.uint8, generator=g))
    before = [t.clone().view(torch.uint8) for t in eng.gpu_cache]
    victim = [5, 17, 2, 9]
    out_map = torch.tensor([[b, i] for i, b in enumerate(victim)], dtype=torch.int64)
    eng.apply(blocks_to_swap_out=out_map.pin_memory())
    eng.apply(blocks_to_copy=torch.tensor([[9, 20], [9, 21]], dtype=torch.int64).pin_memory())
    new_home = [11, 3, 19, 23]
    in_map = torch.tensor([[i, b] for i, b in enumerate(new_home)], dtype=torch.int64)
    for t in eng.gpu_cache:  # the freed blocks get reused by someone else meanwhile
        t[:, new_home] = 0
    eng.apply(blocks_to_swap_in=in_map)
    torch.cuda.synchronize()
    gpu = [t.view(torch.uint8) for t in eng.gpu_cache]   # compare bytes (random bits include NaNs)
    cpu = [t.view(torch.uint8) for t in eng.cpu_cache]
    for layer in range(L):
        for kvi in range(2):
            for old, new in zip(victim, new_home):
                assert torch.equal(gpu[layer][kvi, new], before[layer][kvi, old])
            assert torch.equal(gpu[layer][kvi, 20], before[layer][kvi, 9])
            assert torch.equal(gpu[layer][kvi, 21], before[layer][kvi, 9])
            assert torch.equal(cpu[layer][kvi, 1], before[layer][kvi, 17].cpu())
            untouched = [b for b in range(24) if b not in new_home + [20, 21]]
            assert torch.equal(gpu[layer][kvi, untouched], before[layer][kvi, untouched])


def test_forward_decode_v2_path_and_scratch_reuse():
    """B*H <= 512 and ctx > 512 takes the v2 branch of the reference heuristic."""
    from kv_util import make_case, naive_attention, to_device
    lens = [900, 1300]
    c = make_case(2, 8, 2, 128, 16, lens, torch.float16, seed=9)
    g = to_device(c, "cuda")
    assert not PagedAttention.use_v1(max(lens), 2, 8)
    for _ in range(2):
        out = PagedAttention.forward_decode(g["query"], g["key_cache"], g["value_cache"],
                                            g["block_tables"], g["seq_lens"], max(lens), "auto", 2,
                                            c["scale"], None, 1.0, 1.0)
    assert (out.double().cpu() - naive_attention(c)).abs().max() <= 2e-3


@pytest.mark.parametrize("dtype,tol", [(torch.float16, 2e-3), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("window,alibi", [(None, False), (24, False), (100, False), (None, True), (24, True)])
def test_prefill_sliding_window_and_alibi(dtype, tol, window, alibi):
    """Prompt-only prefill (no cached prefix) with prompt_len > window: the reference passes
    window_size=(W, W) to flash_attn_varlen_func (flash_attn.py:419-420,522-535), i.e. query i sees
    keys i-W <= j <= i; the slot mapping drops older tokens, so decode agrees afterwards."""
    H, KH, D = 8, 2, 64
    lens = [70, 1, 33, 130]
    T = sum(lens)
    g = torch.Generator().manual_seed(5)
    qkv = ((torch.rand(T, (H + 2 * KH) * D, generator=g) * 2 - 1) * 0.5).to(dtype).cuda()
    q, k, v = qkv.split([H * D, KH * D, KH * D], dim=-1)
    slopes = [2.0 ** (-(i + 1)) for i in range(H)] if alibi else None
    impl = PagedB200AttentionBackend.get_impl_cls()(H, D, D ** -0.5, KH, slopes, window, "auto")
    starts = torch.tensor([0] + lens, dtype=torch.int32).cumsum(0).int()
    meta = PagedB200AttentionMetadata(
        num_prefills=len(lens), num_prefill_tokens=T, num_decode_tokens=0,
        slot_mapping=torch.full((T,), -1, dtype=torch.long), seq_lens=lens,
        seq_lens_tensor=torch.tensor(lens, dtype=torch.int), max_query_len=max(lens),
        max_prefill_seq_len=max(lens), max_decode_seq_len=0, query_start_loc=starts,
        seq_start_loc=starts, context_lens_tensor=torch.zeros(len(lens), dtype=torch.int),
        block_tables=torch.zeros(len(lens), 0, dtype=torch.int), use_cuda_graph=False).to("cuda")
    out = impl.forward(q, k, v, None, meta).view(T, H, D).double().cpu()
    qd, kd, vd = (t.double().cpu() for t in (q.reshape(T, H, D), k.reshape(T, KH, D), v.reshape(T, KH, D)))
    pos = 0
    for n in lens:
        i = torch.arange(n)
        rel = (i[None, :] - i[:, None]).double()                      # j - i
        mask = rel > 0
        if window:
            mask = mask | (rel < -window)
        for h in range(H):
            logits = (qd[pos:pos + n, h] @ kd[pos:pos + n, h // (H // KH)].T) * D ** -0.5
            if alibi:
                logits = logits + slopes[h] * rel
            p_ = torch.softmax(logits.masked_fill(mask, float("-inf")), dim=-1)
            want = p_ @ vd[pos:pos + n, h // (H // KH)]
            err = (out[pos:pos + n, h] - want).abs().max().item()
            assert err <= tol, f"len {n} head {h}: {err}"
        pos += n
