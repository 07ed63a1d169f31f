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
