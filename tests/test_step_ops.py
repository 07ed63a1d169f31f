"""Per-step ops either side of paged attention: advance_step, rotary
embedding (plain / batched) and the fused rotary + reshape_and_cache.

CPU: the C oracle against golden vectors produced by the reference's own Python
``RotaryEmbedding.forward_native`` (tests/golden/make_golden_rope.py).
GPU: product == oracle (bit-exact), == the reference CUDA kernels recompiled for sm_100a
(bit-exact), fused op == the two-op sequence (bit-exact, whole pools compared)."""
import os

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = torch.load(os.path.join(HERE, "golden", "rope.pt"))


def bits(t):
    return t.contiguous().view(torch.uint8)


def cos_sin_cache(max_pos, rot_dim, dtype, base=10000.0):
    inv = 1.0 / (base ** (torch.arange(0, rot_dim, 2, dtype=torch.float) / rot_dim))
    fr = torch.outer(torch.arange(max_pos, dtype=torch.float), inv)
    return torch.cat((fr.cos(), fr.sin()), dim=-1).to(dtype)


# ------------------------------------------------------------------------------ CPU ----
@pytest.mark.parametrize("idx", range(len(GOLDEN)))
def test_oracle_rope_matches_reference_python(oracle, idx):
    c = GOLDEN[idx]
    q, k = c["query"].clone(), c["key"].clone()
    oracle.rotary_embedding(c["positions"], q, k, c["head_size"], c["cos_sin_cache"], c["is_neox"],
                            c["rot_dim"], c["offsets"])
    if c["dtype"] == torch.float32:
        # the CUDA kernel (and the oracle) contract to FMA; torch on CPU does not: <= 1 ulp apart
        assert (q - c["query_out"]).abs().max().item() <= 1e-6
        assert (k - c["key_out"]).abs().max().item() <= 1e-6
    else:
        assert torch.equal(bits(q), bits(c["query_out"]))
        assert torch.equal(bits(k), bits(c["key_out"]))


def test_oracle_advance_step_semantics(oracle):
    g = torch.Generator().manual_seed(0)
    B, BS, MAXB = 9, 16, 12
    bt = torch.randint(0, 1000, (B, MAXB), generator=g, dtype=torch.int32)
    seq_lens = torch.randint(1, BS * MAXB - 1, (B,), generator=g, dtype=torch.int32)
    seq_lens[0], seq_lens[1] = BS - 1, BS        # crossing into / starting a new block
    sampled = torch.randint(0, 50000, (B, 1), generator=g)
    toks, pos, slots = (torch.full((B,), -7, dtype=torch.int64) for _ in range(3))
    before = seq_lens.clone()
    nq = B - 2                                   # trailing (padding) sequences stay untouched
    oracle.advance_step(B, nq, BS, toks, sampled, pos, seq_lens, slots, bt)
    for i in range(B):
        if i >= nq:
            assert (toks[i], pos[i], slots[i], seq_lens[i]) == (-7, -7, -7, before[i])
            continue
        p = int(before[i])                       # position of the token being fed = old seq_len
        assert toks[i] == sampled[i, 0] and seq_lens[i] == p + 1 and pos[i] == p
        assert slots[i] == int(bt[i, p // BS]) * BS + p % BS


# ------------------------------------------------------------------------------ GPU ----
def _ref_C():
    from oracle import ref
    if not ref.have_ref_cuda():
        return None
    rC, _ = ref.ref_cuda()
    return rC if hasattr(rC, "rotary_embedding") else None


@pytest.mark.gpu
@pytest.mark.parametrize("idx", range(len(GOLDEN)))
def test_rope_golden_gpu(ops, oracle, idx):
    """Product on the golden inputs: == oracle bit for bit (all dtypes), == the reference's Python
    output bit for bit (16-bit)."""
    c = GOLDEN[idx]
    q, k = c["query"].cuda(), c["key"].cuda()
    pos, cache = c["positions"].cuda(), c["cos_sin_cache"].cuda()
    if c["offsets"] is None:
        assert c["rot_dim"] == cache.shape[1]
        ops.rotary_embedding(pos, q, k, c["head_size"], cache, c["is_neox"])
    else:
        ops.batched_rotary_embedding(pos, q, k, c["head_size"], cache, c["is_neox"], c["rot_dim"],
                                     c["offsets"].cuda())
    qo, ko = c["query"].clone(), c["key"].clone()
    oracle.rotary_embedding(c["positions"], qo, ko, c["head_size"], c["cos_sin_cache"],
                            c["is_neox"], c["rot_dim"], c["offsets"])
    assert torch.equal(bits(q.cpu()), bits(qo)) and torch.equal(bits(k.cpu()), bits(ko))
    if c["dtype"] != torch.float32:
        assert torch.equal(bits(q.cpu()), bits(c["query_out"]))
        assert torch.equal(bits(k.cpu()), bits(c["key_out"]))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("is_neox", [True, False])
def test_rope_vs_reference_cuda(ops, dtype, is_neox):
    rC = _ref_C()
    if rC is None:
        pytest.skip("oracle/_ref/libref_cuda.so (with rotary_embedding) not built")
    g = torch.Generator().manual_seed(3)
    for (T, H, KH, D, rot) in ((64, 32, 8, 128, 128), (7, 14, 2, 64, 64), (33, 8, 8, 80, 40)):
        cache = cos_sin_cache(4096, rot, dtype).cuda()
        qkv = torch.randn(T, (H + 2 * KH) * D, generator=g).to(dtype).cuda()
        pos = torch.randint(0, 2048, (T,), generator=g).cuda()
        off = torch.randint(0, 2048, (T,), generator=g).cuda()
        for batched in (False, True):
            a, b = qkv.clone(), qkv.clone()   # q / k are strided slices of the fused qkv output
            qa, ka = a[:, :H * D], a[:, H * D:(H + KH) * D]
            qb, kb = b[:, :H * D], b[:, H * D:(H + KH) * D]
            if batched:
                ops.batched_rotary_embedding(pos, qa, ka, D, cache, is_neox, rot, off)
                rC.batched_rotary_embedding(pos, qb, kb, D, cache, is_neox, rot, off)
            else:
                ops.rotary_embedding(pos, qa, ka, D, cache, is_neox)
                rC.rotary_embedding(pos, qb, kb, D, cache, is_neox)
            torch.cuda.synchronize()
            assert torch.equal(bits(a), bits(b)), (T, H, KH, D, rot, batched)


@pytest.mark.gpu
def test_advance_step_gpu(ops, oracle):
    g = torch.Generator().manual_seed(1)
    for (B, nq, BS, MAXB) in ((1, 1, 16, 4), (257, 257, 16, 64), (600, 512, 32, 40), (8, 0, 8, 4)):
        bt = torch.randint(0, 100000, (B, MAXB), generator=g, dtype=torch.int32)
        seq_lens = torch.randint(1, BS * MAXB - 1, (B,), generator=g, dtype=torch.int32)
        sampled = torch.randint(0, 150000, (max(nq, 1), 1), generator=g)[:nq]
        toks = torch.randint(0, 9, (B,), generator=g)
        pos = torch.randint(0, 9, (B,), generator=g)
        slots = torch.randint(0, 9, (B,), generator=g)
        want = [t.clone() for t in (toks, pos, seq_lens, slots)]
        oracle.advance_step(B, nq, BS, want[0], sampled, want[1], want[2], want[3], bt)
        got = [t.cuda() for t in (toks, pos, seq_lens, slots)]
        ops.advance_step(B, nq, BS, got[0], sampled.cuda(), got[1], got[2], got[3], bt.cuda())
        for a, b in zip(got, want):
            assert torch.equal(a.cpu(), b)
        rC = _ref_C()
        if rC is not None and nq > 0:
            ref = [t.cuda() for t in (toks, pos, seq_lens, slots)]
            rC.advance_step(B, nq, BS, ref[0], sampled.cuda(), ref[1], ref[2], ref[3], bt.cuda())
            for a, b in zip(got, ref):
                assert torch.equal(a, b)
    with pytest.raises(RuntimeError, match="not as expected"):
        ops.advance_step(4, 4, 16, torch.zeros(4, dtype=torch.int32).cuda(),
                         torch.zeros(4, 1, dtype=torch.int64).cuda(),
                         torch.zeros(4, dtype=torch.int64).cuda(),
                         torch.zeros(4, dtype=torch.int32).cuda(),
                         torch.zeros(4, dtype=torch.int64).cuda(),
                         torch.zeros(4, 2, dtype=torch.int32).cuda())


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,kv,ks,vs", [
    (torch.float16, "auto", 1.0, 1.0), (torch.bfloat16, "auto", 1.0, 1.0),
    (torch.float32, "auto", 1.0, 1.0), (torch.float16, "fp8", 0.02, 0.05),
    (torch.bfloat16, "fp8_e5m2", 0.1, 0.01), (torch.float32, "fp8", 0.02, 0.02)])
@pytest.mark.parametrize("is_neox", [True, False])
def test_fused_rope_cache_write_equals_two_ops(ops, dtype, kv, ks, vs, is_neox):
    """rotary_embedding_reshape_and_cache == rotary_embedding ; reshape_and_cache, bit for bit:
    q, k in place and BOTH whole pools (untouched blocks included)."""
    g = torch.Generator().manual_seed(5)
    for (T, H, KH, D, BS, rot) in ((40, 32, 8, 128, 16, 128), (9, 14, 2, 64, 16, 64),
                                   (21, 8, 4, 96, 32, 48), (5, 4, 4, 80, 8, 80)):
        cdt = dtype if kv == "auto" else torch.uint8
        x = 16 // torch.empty((), dtype=cdt).element_size()
        NB = 24
        cache = cos_sin_cache(1024, rot, dtype).cuda()
        qkv = (torch.randn(T, (H + 2 * KH) * D, generator=g) * 0.5).to(dtype).cuda()
        pos = torch.randint(0, 1024, (T,), generator=g).cuda()
        slots = torch.randperm(NB * BS, generator=g)[:T]
        slots[::6] = -1
        slots = slots.cuda()
        kc0 = torch.randint(0, 255, (NB, KH, D // x, BS, x), generator=g, dtype=torch.uint8)
        vc0 = torch.randint(0, 255, (NB, KH, D, BS), generator=g, dtype=torch.uint8)
        if cdt != torch.uint8:   # random *valid* values so equality of bits is meaningful
            kc0 = torch.randn(NB, KH, D // x, BS, x, generator=g).to(cdt)
            vc0 = torch.randn(NB, KH, D, BS, generator=g).to(cdt)
        res = []
        for fused in (False, True):
            a = qkv.clone()
            q, k, v = a[:, :H * D], a[:, H * D:(H + KH) * D], a[:, (H + KH) * D:]
            kc, vc = kc0.cuda(), vc0.cuda()
            if fused:
                ops.rotary_embedding_reshape_and_cache(pos, q, k, v, D, cache, is_neox, kc, vc,
                                                       slots, kv, ks, vs)
            else:
                ops.rotary_embedding(pos, q, k, D, cache, is_neox)
                ops.reshape_and_cache(k.view(T, KH, D), v.view(T, KH, D), kc, vc, slots, kv, ks, vs)
            torch.cuda.synchronize()
            res.append((a, kc, vc))
        for u, w in zip(*res):
            assert torch.equal(bits(u), bits(w)), (T, H, KH, D, BS, rot)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(16))
def test_rope_fuzz_bit_exact(ops, oracle, seed):
    """Random token counts, head counts (GQA), head sizes, partial rotary dims, both styles, all
    dtypes, with and without per-token cache offsets: output bits == the oracle's."""
    import random
    r = random.Random(11000 + seed)
    dtype = r.choice([torch.float16, torch.bfloat16, torch.float32])
    D = r.choice([64, 80, 96, 128, 256])
    rot = r.choice([D, D // 2])
    H = r.choice([1, 4, 14, 32])
    KH = r.choice([k for k in (1, 2, 4, 14, 32) if k <= H and H % k == 0])
    T = r.choice([1, 3, r.randint(4, 200), r.randint(1000, 4000)])
    neox = r.random() < 0.5
    max_pos = 4096
    g = torch.Generator().manual_seed(seed)
    q = (torch.rand(T, H * D, generator=g) * 2 - 1).to(dtype)
    k = (torch.rand(T, KH * D, generator=g) * 2 - 1).to(dtype)
    batched = r.random() < 0.4
    n_caches = 3 if batched else 1
    cache = torch.cat([cos_sin_cache(max_pos, rot, dtype, base=10000.0 * (i + 1)) for i in range(n_caches)])
    pos = torch.randint(0, max_pos, (T,), generator=g, dtype=torch.int64)
    offs = (torch.randint(0, n_caches, (T,), generator=g, dtype=torch.int64) * max_pos) if batched else None
    qo, ko = q.clone(), k.clone()
    oracle.rotary_embedding(pos, qo, ko, D, cache, neox, rot, offs)
    qg, kg = q.cuda(), k.cuda()
    if batched:
        ops.batched_rotary_embedding(pos.cuda(), qg, kg, D, cache.cuda(), neox, rot, offs.cuda())
    else:
        ops.rotary_embedding(pos.cuda(), qg, kg, D, cache.cuda(), neox)
    info = (seed, dtype, D, rot, H, KH, T, neox, batched)
    assert torch.equal(bits(qg.cpu()), bits(qo)), info
    assert torch.equal(bits(kg.cpu()), bits(ko)), info
