"""Synthetic inputs for the paged-KV parity tests.

Follows the reference's own recipe for random KV caches (light_vllm/utils.py:559-613
``create_kv_caches_with_random``: K, V, Q ~ U(-s, s) with s = head_size ** -0.5; fp8 caches are
produced by quantising that tensor) and SURVEY.md 8(d): scattered (randperm) block tables, int32
``seq_lens``, spare blocks that must stay untouched.
"""
import math

import torch

PARTITION_SIZE = 512


def elem_x(cache_dtype: torch.dtype) -> int:
    """x of the K layout [nb, kh, d/x, bs, x]: 16 bytes worth of cache elements."""
    return 16 // torch.empty((), dtype=cache_dtype).element_size()


def make_case(num_seqs, num_heads, num_kv_heads, head_size, block_size, seq_lens, dtype,
              seed=0, spare_blocks=3, cache_dtype=None, nan_padding=False, q_pad=0):
    """CPU tensors for one paged-attention call.

    Returns a dict with query [B,H,D] (row stride H*D + q_pad), key_cache [NB,KH,D/x,BS,x],
    value_cache [NB,KH,D,BS], block_tables [B,maxb] int32 (randperm, padded with 0), seq_lens [B]
    int32, scale.  With ``nan_padding`` the slots past each sequence's end are NaN-poisoned
    (attention_kernels.cu:420-430 requires them not to contribute).
    """
    g = torch.Generator().manual_seed(seed)
    s = head_size ** -0.5
    cache_dtype = cache_dtype or dtype
    seq_lens = [int(v) for v in seq_lens]
    assert len(seq_lens) == num_seqs
    blocks_per_seq = [(n + block_size - 1) // block_size for n in seq_lens]
    maxb = max(1, max(blocks_per_seq))
    nb = sum(blocks_per_seq) + spare_blocks
    x = elem_x(cache_dtype)

    def rnd(*shape):
        return (torch.rand(*shape, generator=g, dtype=torch.float32) * 2 - 1) * s

    qfull = rnd(num_seqs, num_heads * head_size + q_pad).to(dtype)
    query = qfull[:, :num_heads * head_size].view(num_seqs, num_heads, head_size)
    if q_pad == 0:
        query = query.contiguous()
    key_cache = rnd(nb, num_kv_heads, head_size // x, block_size, x).to(cache_dtype)
    value_cache = rnd(nb, num_kv_heads, head_size, block_size).to(cache_dtype)
    perm = torch.randperm(nb, generator=g)
    block_tables = torch.zeros(num_seqs, maxb, dtype=torch.int32)
    pos = 0
    for i, nblk in enumerate(blocks_per_seq):
        block_tables[i, :nblk] = perm[pos:pos + nblk].to(torch.int32)
        pos += nblk
    if nan_padding and cache_dtype in (torch.float16, torch.bfloat16, torch.float32):
        for i, n in enumerate(seq_lens):
            if n % block_size:
                blk = int(block_tables[i, n // block_size])
                off = n % block_size
                key_cache[blk, :, :, off:, :] = float("nan")
                value_cache[blk, :, :, off:] = float("nan")
    return dict(query=query, key_cache=key_cache, value_cache=value_cache,
                block_tables=block_tables, seq_lens=torch.tensor(seq_lens, dtype=torch.int32),
                scale=float(s), num_kv_heads=num_kv_heads, block_size=block_size,
                max_seq_len=max(seq_lens) if seq_lens else 0, num_blocks=nb)


def naive_attention(case, k_scale=1.0, v_scale=1.0, alibi_slopes=None, decode=None):
    """fp64 ground truth softmax(q K^T * scale [+ alibi]) V, gathered through the block table.
    ``decode(uint8 tensor) -> float tensor`` turns fp8 cache bytes into values."""
    q = case["query"].double()
    B, H, D = q.shape
    KH = case["num_kv_heads"]
    BS = case["block_size"]
    kc, vc = case["key_cache"], case["value_cache"]
    if decode is not None:
        kc, vc = decode(kc) * k_scale, decode(vc) * v_scale
    kc, vc = kc.double(), vc.double()
    out = torch.zeros(B, H, D, dtype=torch.float64)
    for b in range(B):
        n = int(case["seq_lens"][b])
        if n == 0:
            continue
        nblk = (n + BS - 1) // BS
        blks = case["block_tables"][b, :nblk].long()
        k = kc[blks].permute(1, 0, 3, 2, 4).reshape(KH, nblk * BS, D)[:, :n]   # [KH, n, D]
        v = vc[blks].permute(1, 0, 3, 2).reshape(KH, nblk * BS, D)[:, :n]
        for h in range(H):
            kvh = h // (H // KH)
            logits = (k[kvh] @ q[b, h]) * case["scale"]
            if alibi_slopes is not None:
                logits = logits + float(alibi_slopes[h]) * (torch.arange(n, dtype=torch.float64) - n + 1)
            p = torch.softmax(logits, dim=0)
            out[b, h] = p @ v[kvh]
    return out


def num_partitions(max_seq_len: int) -> int:
    return max(1, math.ceil(max_seq_len / PARTITION_SIZE))


def v2_scratch(case, device="cpu"):
    q = case["query"]
    B, H, D = q.shape
    P = num_partitions(case["max_seq_len"])
    tmp_out = torch.empty(B, H, P, D, dtype=q.dtype, device=device)
    exp_sums = torch.empty(B, H, P, dtype=torch.float32, device=device)
    max_logits = torch.empty_like(exp_sums)
    return exp_sums, max_logits, tmp_out


def to_device(case, device):
    return {k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in case.items()}
