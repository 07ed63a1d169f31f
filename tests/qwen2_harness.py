"""Harness for BASELINE.json configs[0]: a Qwen2-0.5B-shaped decoder (random init) doing greedy
decode over the PAGED KV cache, with the attention + cache-write ops pluggable:

  * CPU engine  -- the reference's own CPU kernels (oracle/_ref/libref_cpu.so: reshape_and_cache +
                   paged_attention_v1 from csrc/cpu/{cache,attention}.cpp) or the C oracle port;
  * B200 engine -- the product plugin (PagedB200AttentionImpl.forward -> torch.ops._C*).

light-vllm itself has no CPU worker (decoding/worker/gpu_worker.py:84-86 raises), so this is a
harness around the ops, not the reference engine ("C1 concrete").  Everything that
is not attention (embedding, RMSNorm, QKV/O/MLP projections, neox RoPE as
decode_only/modelzoo/qwen2.py:100-170 wires them) is plain torch on both sides.  Test
infrastructure: nothing under light_vllm_b200/ imports this file.
"""
import math
import time
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F


@dataclass
class Qwen2Shape:
    hidden: int = 896
    heads: int = 14
    kv_heads: int = 2
    head_dim: int = 64
    layers: int = 24
    inter: int = 4864
    vocab: int = 151936
    rope_theta: float = 1e6
    eps: float = 1e-6

    @staticmethod
    def tiny():  # same head geometry, fewer / thinner layers: for unit tests
        return Qwen2Shape(hidden=256, heads=14, kv_heads=2, head_dim=64, layers=3, inter=512,
                          vocab=2048)


def init_weights(shape: Qwen2Shape, seed: int = 0) -> Dict[str, torch.Tensor]:
    """fp32 CPU weights ~ N(0, 0.02) (norm gains 1).  The lm_head is NOT tied to the embedding: with
    random weights a tied head just echoes the input token (its own embedding row wins the
    argmax), which would make the greedy continuation independent of attention."""
    g = torch.Generator().manual_seed(seed)

    def n(*s):
        return torch.randn(*s, generator=g, dtype=torch.float32) * 0.02

    qkv_out = (shape.heads + 2 * shape.kv_heads) * shape.head_dim
    w = {"emb": n(shape.vocab, shape.hidden), "head": n(shape.vocab, shape.hidden),
         "norm": torch.ones(shape.hidden)}
    for l in range(shape.layers):
        w[f"l{l}.ln1"] = torch.ones(shape.hidden)
        w[f"l{l}.ln2"] = torch.ones(shape.hidden)
        w[f"l{l}.qkv"] = n(qkv_out, shape.hidden)
        w[f"l{l}.qkv_b"] = n(qkv_out)
        w[f"l{l}.o"] = n(shape.hidden, shape.heads * shape.head_dim)
        w[f"l{l}.gate_up"] = n(2 * shape.inter, shape.hidden)
        w[f"l{l}.down"] = n(shape.hidden, shape.inter)
    return w


def rmsnorm(x, g, eps):
    xf = x.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * g


def rope_neox(x, positions, theta):
    """x [T, heads, D]; rotate halves (neox style, is_neox_style=True in the reference's Qwen2)."""
    D = x.shape[-1]
    inv = 1.0 / (theta ** (torch.arange(0, D, 2, dtype=torch.float32, device=x.device) / D))
    ang = positions.to(torch.float32)[:, None] * inv[None, :]            # [T, D/2]
    cos, sin = ang.cos()[:, None, :].to(x.dtype), ang.sin()[:, None, :].to(x.dtype)
    x1, x2 = x[..., :D // 2], x[..., D // 2:]
    return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1)


class CpuEngine:
    """Paged KV + attention through a CPU op module with the reference's op signatures
    (``oracle.ref.ref_cpu()`` = the reference's compiled CPU kernels, or ``oracle.oracle``)."""

    def __init__(self, mod, shape: Qwen2Shape, num_blocks: int, block_size: int, dtype):
        self.mod, self.shape, self.bs, self.dtype = mod, shape, block_size, dtype
        x = 16 // torch.empty((), dtype=dtype).element_size()
        KH, D = shape.kv_heads, shape.head_dim
        self.k = [torch.zeros(num_blocks, KH, D // x, block_size, x, dtype=dtype)
                  for _ in range(shape.layers)]
        self.v = [torch.zeros(num_blocks, KH, D, block_size, dtype=dtype)
                  for _ in range(shape.layers)]
        self.scale = D ** -0.5

    def prefill(self, layer, q, k, v, seq_lens, slots):
        self.mod.reshape_and_cache(k, v, self.k[layer], self.v[layer], slots, "auto", 1.0, 1.0)
        out = torch.empty_like(q)
        pos = 0
        for n in seq_lens:
            o = F.scaled_dot_product_attention(
                q[pos:pos + n].transpose(0, 1)[None], k[pos:pos + n].transpose(0, 1)[None],
                v[pos:pos + n].transpose(0, 1)[None], is_causal=True, scale=self.scale,
                enable_gqa=True)
            out[pos:pos + n] = o[0].transpose(0, 1)
            pos += n
        return out

    def decode(self, layer, q, k, v, slots, block_tables, seq_lens, max_len):
        self.mod.reshape_and_cache(k, v, self.k[layer], self.v[layer], slots, "auto", 1.0, 1.0)
        out = torch.empty_like(q)
        self.mod.paged_attention_v1(out, q, self.k[layer], self.v[layer], self.shape.kv_heads,
                                    self.scale, block_tables, seq_lens, self.bs, max_len, None,
                                    "auto", 1.0, 1.0, 0, 0, 0, 64, 0)
        return out


class B200Engine:
    """The product path: PagedB200AttentionImpl.forward over per-layer pools on the GPU."""

    def __init__(self, shape: Qwen2Shape, num_blocks: int, block_size: int, dtype, device):
        from light_vllm_b200.attention import PagedB200AttentionBackend as Backend
        self.shape, self.bs, self.device = shape, block_size, device
        kv_shape = Backend.get_kv_cache_shape(num_blocks, block_size, shape.kv_heads, shape.head_dim)
        self.pools = [torch.zeros(kv_shape, dtype=dtype, device=device) for _ in range(shape.layers)]
        self.impl = Backend.get_impl_cls()(shape.heads, shape.head_dim, shape.head_dim ** -0.5,
                                           shape.kv_heads, None, None, "auto")
        self.meta_cls = Backend.get_metadata_cls()
        self._meta = None

    def begin_step(self, *, prefill_lens=None, slots=None, block_tables=None, seq_lens=None,
                   max_len=0):
        """One metadata object per model step, shared by all layers (as the model runner does)."""
        dev = self.device
        if prefill_lens is not None:
            starts = torch.tensor([0] + list(prefill_lens), dtype=torch.int32).cumsum(0).int()
            self._meta = self.meta_cls(
                num_prefills=len(prefill_lens), num_prefill_tokens=sum(prefill_lens),
                num_decode_tokens=0, slot_mapping=slots.to(dev), seq_lens=list(prefill_lens),
                seq_lens_tensor=torch.tensor(prefill_lens, dtype=torch.int32, device=dev),
                max_query_len=max(prefill_lens), max_prefill_seq_len=max(prefill_lens),
                max_decode_seq_len=0, query_start_loc=starts.to(dev), seq_start_loc=starts.to(dev),
                context_lens_tensor=torch.zeros(len(prefill_lens), dtype=torch.int32, device=dev),
                block_tables=torch.empty(len(prefill_lens), 0, dtype=torch.int32, device=dev),
                use_cuda_graph=False)
        else:
            B = seq_lens.numel()
            self._meta = self.meta_cls(
                num_prefills=0, num_prefill_tokens=0, num_decode_tokens=B,
                slot_mapping=slots.to(dev, non_blocking=True), seq_lens=None,
                seq_lens_tensor=seq_lens.to(dev, non_blocking=True), max_query_len=1,
                max_prefill_seq_len=0, max_decode_seq_len=max_len, query_start_loc=None,
                seq_start_loc=None, context_lens_tensor=None,
                block_tables=block_tables.to(dev, non_blocking=True), use_cuda_graph=False)

    def _fwd(self, layer, q, k, v):
        T = q.shape[0]
        return self.impl.forward(q.reshape(T, -1), k.reshape(T, -1), v.reshape(T, -1),
                                 self.pools[layer], self._meta).view_as(q)

    def prefill(self, layer, q, k, v, seq_lens, slots):
        return self._fwd(layer, q, k, v)

    def decode(self, layer, q, k, v, slots, block_tables, seq_lens, max_len):
        return self._fwd(layer, q, k, v)


def _layers_forward(shape, w, engine, x, positions, attn_call, record):
    """x [T, hidden] -> [T, hidden]; ``attn_call(layer, q, k, v)`` is the pluggable attention."""
    H, KH, D = shape.heads, shape.kv_heads, shape.head_dim
    for l in range(shape.layers):
        h = rmsnorm(x, w[f"l{l}.ln1"], shape.eps)
        qkv = F.linear(h, w[f"l{l}.qkv"], w[f"l{l}.qkv_b"])
        q, k, v = qkv.split([H * D, KH * D, KH * D], dim=-1)
        q = rope_neox(q.reshape(-1, H, D), positions, shape.rope_theta)
        k = rope_neox(k.reshape(-1, KH, D), positions, shape.rope_theta)
        v = v.reshape(-1, KH, D).contiguous()
        o = attn_call(l, q, k, v)
        if record is not None and l == shape.layers - 1:
            record.append(o.float().cpu())
        x = x + F.linear(o.reshape(-1, H * D), w[f"l{l}.o"])
        h = rmsnorm(x, w[f"l{l}.ln2"], shape.eps)
        gate, up = F.linear(h, w[f"l{l}.gate_up"]).chunk(2, dim=-1)
        x = x + F.linear(F.silu(gate) * up, w[f"l{l}.down"])
    return x


@torch.inference_mode()
def greedy_decode(shape: Qwen2Shape, weights, engine, prompts: List[List[int]], new_tokens: int,
                  block_size: int, block_tables: torch.Tensor, device="cpu",
                  forced: Optional[torch.Tensor] = None, record: bool = False):
    """Prefill the prompts, then ``new_tokens`` greedy decode steps over the paged cache.

    block_tables [B, max_blocks] int32 (CPU): physical blocks of every sequence.  ``forced``
    [B, new_tokens]: teacher-forced continuation (for dtype comparisons); the argmax is still
    returned.  Returns dict(tokens [B, new_tokens], attn (last layer's attention output per decode
    step, if record), logits (per step, if record), prefill_s, decode_s)."""
    B = len(prompts)
    lens = [len(p) for p in prompts]
    w = weights
    sync = (lambda: torch.cuda.synchronize()) if str(device).startswith("cuda") else (lambda: None)
    bt = block_tables

    def slot(i, pos):
        return int(bt[i, pos // block_size]) * block_size + pos % block_size

    # ---- prefill ----
    toks = torch.tensor([t for p in prompts for t in p], dtype=torch.long, device=device)
    positions = torch.tensor([j for n in lens for j in range(n)], dtype=torch.long, device=device)
    slots = torch.tensor([slot(i, j) for i, n in enumerate(lens) for j in range(n)], dtype=torch.long)
    sync()
    t0 = time.perf_counter()
    if hasattr(engine, "begin_step"):
        engine.begin_step(prefill_lens=lens, slots=slots)
    slots_dev = slots.to(device)
    x = w["emb"][toks]
    x = _layers_forward(shape, w, engine, x, positions,
                        lambda l, q, k, v: engine.prefill(l, q, k, v, lens, slots_dev), None)
    last = torch.tensor(lens, device=device).cumsum(0) - 1
    logits = F.linear(rmsnorm(x[last], w["norm"], shape.eps), w["head"])
    nxt = logits.float().argmax(-1)
    sync()
    prefill_s = time.perf_counter() - t0

    out_tokens, attn_rec, logit_rec = [], ([] if record else None), []
    cur = list(lens)
    t0 = time.perf_counter()
    for step in range(new_tokens):
        out_tokens.append(nxt.cpu())
        if record:
            logit_rec.append(logits.float().cpu())
        inp = forced[:, step].to(device) if forced is not None else nxt
        positions = torch.tensor(cur, dtype=torch.long, device=device)
        slots = torch.tensor([slot(i, cur[i]) for i in range(B)], dtype=torch.long)
        cur = [c + 1 for c in cur]
        seq_lens = torch.tensor(cur, dtype=torch.int32)
        max_len = max(cur)
        nblk = (max_len + block_size - 1) // block_size
        bt_step = bt[:, :nblk].contiguous()
        if hasattr(engine, "begin_step"):
            engine.begin_step(slots=slots, block_tables=bt_step, seq_lens=seq_lens, max_len=max_len)
            args = (None, None, None, max_len)
        else:
            args = (slots, bt_step, seq_lens, max_len)
        x = w["emb"][inp]
        x = _layers_forward(shape, w, engine, x, positions,
                            lambda l, q, k, v: engine.decode(l, q, k, v, *args), attn_rec)
        logits = F.linear(rmsnorm(x, w["norm"], shape.eps), w["head"])
        nxt = logits.float().argmax(-1)
    sync()
    decode_s = time.perf_counter() - t0
    return dict(tokens=torch.stack(out_tokens, dim=1), attn=attn_rec, logits=logit_rec,
                prefill_s=prefill_s, decode_s=decode_s)


def make_prompts(batch: int, prompt_len: int, vocab: int, seed: int = 0) -> List[List[int]]:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab, (batch, prompt_len), generator=g).tolist()


def make_block_tables(batch: int, total_len: int, block_size: int, seed: int = 0):
    """Scattered (randperm) physical blocks + a few spares; returns (tables int32, num_blocks)."""
    per = math.ceil(total_len / block_size)
    nb = batch * per + 4
    g = torch.Generator().manual_seed(seed)
    return torch.randperm(nb, generator=g)[:batch * per].view(batch, per).int(), nb


def dense_reference(shape, weights, prompts, new_tokens):
    """No paging at all: recompute full causal attention over the whole sequence every step
    (fp32, CPU).  Validates the harness itself."""
    B = len(prompts)
    seqs = [list(p) for p in prompts]
    out = []
    for _ in range(new_tokens):
        step = []
        for s in seqs:
            toks = torch.tensor(s)
            pos = torch.arange(len(s))

            def attn(l, q, k, v):
                o = F.scaled_dot_product_attention(q.transpose(0, 1)[None], k.transpose(0, 1)[None],
                                                   v.transpose(0, 1)[None], is_causal=True,
                                                   scale=shape.head_dim ** -0.5, enable_gqa=True)
                return o[0].transpose(0, 1)

            x = _layers_forward(shape, weights, None, weights["emb"][toks], pos, attn, None)
            logits = F.linear(rmsnorm(x[-1:], weights["norm"], shape.eps), weights["head"])
            step.append(int(logits.argmax(-1)))
        for s, t in zip(seqs, step):
            s.append(t)
        out.append(step)
    return torch.tensor(out).t()


class GraphDecoder:
    """One whole decode step as ONE CUDA graph: no host work per token.

    Graph body, all on static device buffers: embedding -> per layer [RMSNorm, QKV, fused
    rotary_embedding_reshape_and_cache, paged_attention (plan fixed by ``max_len``), O, MLP] ->
    logits -> argmax -> ``advance_step`` (next token, position, seq_len and slot computed on the
    GPU from the block table).  Each replay consumes the previous replay's outputs."""

    def __init__(self, shape: Qwen2Shape, weights, engine: "B200Engine", block_size: int,
                 block_tables: torch.Tensor, max_len: int, dtype, device="cuda"):
        from light_vllm_b200 import _custom_ops as ops
        from light_vllm_b200.attention import PagedAttention
        self.shape, self.w, self.eng, self.bs, self.max_len = shape, weights, engine, block_size, max_len
        self.ops, self.PA = ops, PagedAttention
        B = block_tables.shape[0]
        self.B = B
        self.tok = torch.zeros(B, dtype=torch.long, device=device)
        self.pos = torch.zeros(B, dtype=torch.long, device=device)
        self.seq_lens = torch.zeros(B, dtype=torch.int32, device=device)
        self.slots = torch.zeros(B, dtype=torch.long, device=device)
        self.sampled = torch.zeros(B, 1, dtype=torch.long, device=device)
        self.bt = block_tables.to(device).contiguous()
        D = shape.head_dim
        inv = 1.0 / (shape.rope_theta ** (torch.arange(0, D, 2, dtype=torch.float) / D))
        fr = torch.outer(torch.arange(max_len + 1, dtype=torch.float), inv)
        self.cos_sin = torch.cat((fr.cos(), fr.sin()), dim=-1).to(device=device, dtype=dtype)
        self.caches = [PagedAttention.split_kv_cache(p, shape.kv_heads, D) for p in engine.pools]
        self.graph = None

    def _step(self):
        sh, w, ops = self.shape, self.w, self.ops
        H, KH, D = sh.heads, sh.kv_heads, sh.head_dim
        x = w["emb"][self.tok]
        for l in range(sh.layers):
            h = rmsnorm(x, w[f"l{l}.ln1"], sh.eps)
            qkv = F.linear(h, w[f"l{l}.qkv"], w[f"l{l}.qkv_b"])
            q, k, v = qkv.split([H * D, KH * D, KH * D], dim=-1)
            kc, vc = self.caches[l]
            ops.rotary_embedding_reshape_and_cache(self.pos, q, k, v, D, self.cos_sin, True, kc, vc,
                                                   self.slots, "auto", 1.0, 1.0)
            o = self.PA.forward_decode(q.view(-1, H, D), kc, vc, self.bt, self.seq_lens,
                                       self.max_len, "auto", KH, D ** -0.5, None, 1.0, 1.0)
            x = x + F.linear(o.reshape(-1, H * D), w[f"l{l}.o"])
            h = rmsnorm(x, w[f"l{l}.ln2"], sh.eps)
            gate, up = F.linear(h, w[f"l{l}.gate_up"]).chunk(2, dim=-1)
            x = x + F.linear(F.silu(gate) * up, w[f"l{l}.down"])
        logits = F.linear(rmsnorm(x, w["norm"], sh.eps), w["head"])
        self.sampled.copy_(logits.float().argmax(-1, keepdim=True))
        # next step's inputs, on the GPU: tok <- sampled, seq_len += 1, pos, slot from block table
        ops.advance_step(self.B, self.B, self.bs, self.tok, self.sampled, self.pos, self.seq_lens,
                         self.slots, self.bt)

    def start(self, first_tokens: torch.Tensor, lens: List[int]):
        """State after prefill: feed ``first_tokens`` at positions ``lens`` (seq_len becomes len+1)."""
        dev = self.tok.device
        self.tok.copy_(first_tokens)
        self.pos.copy_(torch.tensor(lens, dtype=torch.long))
        self.seq_lens.copy_(torch.tensor([n + 1 for n in lens], dtype=torch.int32))
        bt = self.bt.cpu()
        self.slots.copy_(torch.tensor([int(bt[i, n // self.bs]) * self.bs + n % self.bs
                                       for i, n in enumerate(lens)], dtype=torch.long))
        self._saved = [t.clone() for t in (self.tok, self.pos, self.seq_lens, self.slots)]
        if self.graph is None:
            s = torch.cuda.Stream(device=dev)
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):       # warm-up outside capture (lazy init, autotune)
                self._step()
            torch.cuda.current_stream().wait_stream(s)
            self._restore()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._step()
            self._restore()                  # capture does not execute, but keep state explicit

    def _restore(self):
        for t, s in zip((self.tok, self.pos, self.seq_lens, self.slots), self._saved):
            t.copy_(s)

    @torch.inference_mode()
    def run(self, new_tokens: int) -> torch.Tensor:
        """``new_tokens`` graph replays; returns the sampled tokens [B, new_tokens] (CPU)."""
        out = torch.empty(self.B, new_tokens, dtype=torch.long, device=self.tok.device)
        for step in range(new_tokens):
            self.graph.replay()
            out[:, step].copy_(self.sampled[:, 0])
        torch.cuda.synchronize()
        return out.cpu()


@torch.inference_mode()
def graph_decode(shape, weights, prompts, new_tokens, block_size, block_tables, num_blocks, dtype,
                 device="cuda"):
    """Eager prefill through the plugin, then CUDA-graph decode.  Returns tokens [B, new_tokens]
    (the first is the prefill's argmax, as in ``greedy_decode``) and the decode wall time."""
    eng = B200Engine(shape, num_blocks, block_size, dtype, device)
    pre = greedy_decode(shape, weights, eng, prompts, 1, block_size, block_tables, device=device)
    # greedy_decode(.., 1) ran the prefill and ONE decode step (its KV is rewritten identically below)
    first = pre["tokens"][:, 0]
    lens = [len(p) for p in prompts]
    dec = GraphDecoder(shape, weights, eng, block_size, block_tables,
                       max(lens) + new_tokens, dtype, device)
    dec.start(first.to(device), lens)
    dec.run(2)                               # warm replays
    dec._restore()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rest = dec.run(new_tokens - 1)
    decode_s = time.perf_counter() - t0
    return dict(tokens=torch.cat([first[:, None], rest], dim=1), decode_s=decode_s,
                steps=new_tokens - 1)
